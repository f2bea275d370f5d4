#!/bin/bash
cd "$(dirname "$0")/.."
run() { echo "== $*"; timeout 90 python tools/bench_gemv.py --reps 40 "$@" 2>&1 | grep -v "^$"; }
for c in out_q6k wo_q4k; do
  run --case $c --flags 0x1
  run --case $c --flags 0x3
  run --case $c --flags 0x4021      # probe 2 x 64 KiB
  run --case $c --flags 0x4023
  run --case $c --flags 0x3031      # probe 3 x 48 KiB
  run --case $c --flags 0x2041      # probe 4 x 32
  run --case $c --flags 0x2043
  run --case $c --flags 0x1061      # 6 x 16
  run --case $c --flags 0x1063
done
