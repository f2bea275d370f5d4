#!/bin/bash
# C5 (gpt-oss-120B shape): llama-bench through the plugin, then an ncu launch list of a short decode run (per-kernel time of the per-op path)
cd "$(dirname "$0")/.."
python tools/config_bench.py c5 --only tg128 > gpurun_out/c5.log 2>&1; tail -1 gpurun_out/c5.log | cut -c1-200
export GGML_BACKEND_PATH=$PWD/llama-mi50.cpp_b200/libggml-b200.so
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 3000 --csv --log-file gpurun_out/launches_c5.csv \
  oracle/_ref/llama-bench -m /tmp/b200_models/gpt-oss-120b-mxfp4.gguf -p 0 -n 8 -r 1 -fa 1 -ngl 99 > gpurun_out/ncu_c5.log 2>&1
echo ncu rc=$?
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/launches_c5.csv')) if len(r)>5]
hdr=None; agg=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    if 'Kernel Name' in r: hdr={h:i for i,h in enumerate(r)}; continue
    if hdr is None or r[hdr['Metric Name']]!='gpu__time_duration.sum': continue
    k=r[hdr['Kernel Name']][:90]; v=float(r[hdr['Metric Value']].replace(',','')); u=r[hdr['Metric Unit']]
    v=v/1e3 if u.startswith('n') else v
    agg[k][0]+=1; agg[k][1]+=v
tot=sum(v[1] for v in agg.values()); n=sum(v[0] for v in agg.values())
print('launches',n,'total us',round(tot))
for k,(c,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:22]:
    print(f'{100*t/tot:5.1f}%  {c:5d} x {t/c:8.1f} us  {k}')
PY
