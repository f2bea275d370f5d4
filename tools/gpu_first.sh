#!/bin/bash
# First GPU pass: parity tests, the reference's test-backend-ops against the plugin, llama-bench through the plugin.
# usage (on a GPU box, from the repo root): tools/gpu_first.sh
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
echo "== pytest -m gpu"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu.log
echo "== test-backend-ops"
timeout 1500 tools/gpu_check.sh
echo "== gguf"
mkdir -p /tmp/models
python llama-mi50.cpp_b200/gguf_synth.py tinyllama-1.1b-q4_0 /tmp/models/tinyllama.gguf
python llama-mi50.cpp_b200/gguf_synth.py llama3-8b-q4_k_m /tmp/models/llama3-8b.gguf
export GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so
for m in tinyllama llama3-8b; do
  echo "== llama-bench $m (plugin)"
  timeout 600 oracle/_ref/llama-bench -m /tmp/models/$m.gguf -p 512 -n 128 -r 3 -fa 1 -ngl 99 -o json > gpurun_out/lb_$m.json 2> gpurun_out/lb_$m.err
  echo "rc=$?"; grep -E '"n_prompt"|"n_gen"|"avg_ts"' gpurun_out/lb_$m.json | tr -d '\n'; echo; tail -3 gpurun_out/lb_$m.err
done
echo "== sched debug (splits)"
GGML_SCHED_DEBUG=2 timeout 120 oracle/_ref/llama-bench -m /tmp/models/tinyllama.gguf -p 0 -n 4 -r 1 -fa 1 -ngl 99 -v 2>&1 | grep -E "SPLIT|splits" | sort | uniq -c | head -20
