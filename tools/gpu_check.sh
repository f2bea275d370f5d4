#!/bin/bash
# Runs the reference's own parity harness (test-backend-ops) and llama-bench against the plugin.
# usage: tools/gpu_check.sh [ops...]   (on a GPU box, from the repo root)
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so
REF=$PWD/oracle/_ref
OPS=${*:-MUL_MAT MUL_MAT_ID FLASH_ATTN_EXT RMS_NORM ROPE SET_ROWS GET_ROWS GLU ADD MUL SCALE CPY SOFT_MAX}
mkdir -p gpurun_out
for op in $OPS; do
  timeout 600 $REF/test-backend-ops test -b B200_0 -o $op > gpurun_out/tbo_$op.log 2>&1
  rc=$?
  ok=$(grep -c "OK" gpurun_out/tbo_$op.log); fail=$(grep -c "FAIL" gpurun_out/tbo_$op.log); ns=$(grep -c "not supported" gpurun_out/tbo_$op.log)
  echo "== $op rc=$rc ok=$ok fail=$fail not_supported=$ns : $(grep -E 'tests passed|Backend B200' gpurun_out/tbo_$op.log | tr '\n' ' ')"
  grep "FAIL" gpurun_out/tbo_$op.log | head -8
done
