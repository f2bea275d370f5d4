#!/bin/bash
# C5 prefill + decode through the plugin (gpt-oss-120B shape)
cd "$(dirname "$0")/.."
python tools/config_bench.py c5 > gpurun_out/c5.log 2>&1
grep -o "\"avg_ts\": [0-9.]*\|\"label\": \"[^\"]*\"\|\"error\": \"[^\"]*" gpurun_out/c5.log
