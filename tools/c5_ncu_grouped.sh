#!/bin/bash
# ncu --set full of one grouped MUL_MAT_ID launch of a gpt-oss-120B pp512 run through the plugin
cd "$(dirname "$0")/.."
python - <<'PY'
import os, __graft_entry__ as g
g.load_package()
from llama_mi50_b200 import gguf_synth as gs
p='/tmp/b200_models/gpt-oss-120b-mxfp4.gguf'
os.makedirs('/tmp/b200_models', exist_ok=True)
if not os.path.exists(p): gs.write_gguf(p, gs.SHAPES['gpt-oss-120b-mxfp4'])
PY
export GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so
timeout 400 ncu --set full --clock-control none --import-source on -k regex:mmvq_grouped -s 6 -c 1 -o gpurun_out/c5_grouped \
  oracle/_ref/llama-bench -m /tmp/b200_models/gpt-oss-120b-mxfp4.gguf -p 512 -n 0 -r 1 -fa 1 -ngl 99 > gpurun_out/ncu_c5g.log 2>&1
echo rc=$?
