#!/bin/bash
# Runs the reference's llama-bench against the plugin on a synthetic GGUF (written on the fly).
# usage: tools/llama_bench_plugin.sh [shape-key] [llama-bench args...]   (on a GPU box, from the repo root)
cd "$(dirname "$0")/.."
KEY=${1:-llama3-8b-q4_k_m}; shift
mkdir -p /tmp/b200_models
GGUF=/tmp/b200_models/$KEY.gguf
[ -f "$GGUF" ] || python - "$KEY" "$GGUF" <<'PY'
import sys
import __graft_entry__ as g
g.load_package()
from llama_mi50_b200 import gguf_synth as gs
gs.write_gguf(sys.argv[2], gs.SHAPES[sys.argv[1]])
PY
export GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so
exec oracle/_ref/llama-bench -m "$GGUF" -fa 1 -ngl 99 "${@:--p 512 -n 128 -r 3}"
