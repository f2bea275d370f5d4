python - <<'PY'
import __graft_entry__ as g; g.load_package()
from llama_mi50_b200 import gguf_synth as gs
gs.write_gguf('/tmp/tiny-moe.gguf', gs.SHAPES['tiny-moe-mxfp4'])
PY
GGML_SCHED_DEBUG=2 GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so oracle/_ref/llama-bench -m /tmp/tiny-moe.gguf -p 0 -n 4 -r 1 -ngl 99 -fa 1 > gpurun_out/c5_sched.log 2>&1
grep -c "## SPLIT" gpurun_out/c5_sched.log; grep "## SPLIT" gpurun_out/c5_sched.log | sort | uniq -c | head -8
python tools/config_bench.py c5 > gpurun_out/c5.log 2>&1; tail -3 gpurun_out/c5.log
