#!/bin/bash
# Same-box A/B of decode-chain tuning flags (b200_chain_create: bits 4-7 ring depth, bits 8-15 stage KiB).  usage: tools/ab_flags.sh flags...
cd "$(dirname "$0")/.."
for r in 1 2; do
  for f in "$@"; do
    python bench.py --no-llama-bench --steps 5 --warmup 3 --chain-flags $f > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err
    python -c "import json;d=json.load(open('gpurun_out/ab_tmp.json'));print('round $r flags $f', d['value'], d['roofline']['frac'])" 2>/dev/null || echo "round $r flags $f FAILED $(tail -1 gpurun_out/ab_tmp.err | cut -c1-150)"
  done
done
