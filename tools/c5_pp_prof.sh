#!/bin/bash
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
ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 1500 --csv --log-file gpurun_out/launches_c5_pp.csv \
  oracle/_ref/llama-bench -m /tmp/b200_models/gpt-oss-120b-mxfp4.gguf -p 512 -n 0 -r 1 -fa 1 -ngl 99 > gpurun_out/ncu_c5pp.log 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/launches_c5_pp.csv')) if len(r)>5]
hdr=None; agg=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    if 'Kernel Name' in r: hdr={h:i for i,h in enumerate(r)}; continue
    if hdr is None or r[hdr['Metric Name']]!='gpu__time_duration.sum': continue
    k=r[hdr['Kernel Name']][:80]; v=float(r[hdr['Metric Value']].replace(',','')); u=r[hdr['Metric Unit']]
    v=v/1e3 if u.startswith('n') else v
    agg[k][0]+=1; agg[k][1]+=v
tot=sum(v[1] for v in agg.values()); n=sum(v[0] for v in agg.values())
print('launches',n,'total us',round(tot))
for k,(c,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:14]:
    print(f'{100*t/tot:5.1f}%  {c:5d} x {t/c:9.1f} us  {k}')
PY
