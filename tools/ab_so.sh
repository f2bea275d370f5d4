#!/bin/bash
# Same-box A/B of kernel-library builds: ab/*.so, alternating, N rounds.  usage: tools/ab_so.sh [rounds]
cd "$(dirname "$0")/.."
R=${1:-2}
for r in $(seq 1 $R); do
  for so in ab/*.so; do
    B200_KERNELS_SO=$PWD/$so python bench.py --no-llama-bench --steps 6 --warmup 3 > gpurun_out/ab_tmp.json 2> gpurun_out/ab_tmp.err
    python -c "import json;d=json.load(open('gpurun_out/ab_tmp.json'));print('round $r $so', d['value'], d['roofline']['frac'])" 2>/dev/null || echo "round $r $so FAILED $(tail -1 gpurun_out/ab_tmp.err)"
  done
done
