#!/usr/bin/env python3
"""Summarise one kernel of an .ncu-rep (ncu --set full) into the JSON kept under profiles/.
usage: ncu_summary.py report.ncu-rep "kernel description" "command that produced it" [algorithmic_bytes] > profiles/xxx.json"""
import csv, io, json, subprocess, sys

rep, desc, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
alg = int(sys.argv[4]) if len(sys.argv) > 4 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}


def get(name, scaled=True):
    i = col[name]
    v = float(vals[i].replace(",", "")) if vals[i] else None
    if v is None:
        return None
    return v * SCALE.get(units[i].split("/")[0], 1.0) if scaled else v


out = {"kernel": desc, "command": cmd,
       "duration_ms": get("gpu__time_duration.sum"),
       "dram_bytes_read": get("dram__bytes_read.sum"), "dram_bytes_write": get("dram__bytes_write.sum")}
out["dram_bytes_per_launch"] = out["dram_bytes_read"] + out["dram_bytes_write"]
if alg:
    out["algorithmic_bytes_per_launch"] = alg
for k, name in (("dram_read_pct_of_peak", "dram__bytes_read.sum.pct_of_peak_sustained_elapsed"),
                ("issue_active_pct", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                ("alu_pipe_pct", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                ("tensor_pipe_pct", "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active"),
                ("registers_per_thread", "launch__registers_per_thread"), ("grid", "launch__grid_size"), ("block", "launch__block_size"),
                ("dynamic_smem_kb", "launch__shared_mem_per_block_dynamic")):
    if name in col:
        out[k] = get(name, scaled=False)
st = []
for h in hdr:
    if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
        st.append((get(h, scaled=False) or 0.0, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
out["warp_stalls_per_issue_top"] = [[n, round(v, 3)] for v, n in sorted(st, reverse=True)[:8]]
print(json.dumps(out, indent=1))
