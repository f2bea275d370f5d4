#!/bin/bash
# tuning sweep of the streaming GEMV: ring depth / stage size / probe mode / CTAs per SM
cd "$(dirname "$0")/.."
run() { echo "== $*"; timeout 90 python tools/bench_gemv.py --reps 40 "$@" 2>&1 | grep -v "^$"; }
for c in gateup_q4k down_q4k out_q6k qkv_q4k; do
  run --case $c
  run --case $c --flags 0x1
  run --case $c --flags 0x1041          # probe, 4 stages x 16 KiB
  run --case $c --flags 0x0c61          # probe, 6 stages x 12 KiB
  run --case $c --flags 0x2031          # probe, 3 stages x 32 KiB
  run --case $c --flags 0x1040          # 4 x 16
  run --case $c --flags 0x0c60          # 6 x 12
  run --case $c --flags 0x2030          # 3 x 32
  run --case $c --n-sm 296
  run --case $c --n-sm 296 --flags 0x1
done
