#!/usr/bin/env python3
"""Per-config evidence (BASELINE.json configs C1-C5): the reference's llama-bench driving the plugin on a random-init GGUF of
the named shape; tg / pp tok/s, algorithmic bytes per token and the fraction of the measured HBM peak; for the small configs
the ggml CPU backend next to it.  Writes gpurun_out/config_<name>.json (copied to profiles/ by hand when kept).

usage: tools/config_bench.py <c1|c2|c3|c4|c5> [--cpu] [--reps N] [--devices N] [--sm row|layer|none]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

graft.load_package()
from llama_mi50_b200 import BACKEND_SO, gguf_synth as gs  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref")
MODEL_DIR = os.environ.get("B200_MODEL_DIR", "/tmp/b200_models")

# name -> (shape key, [(label, llama-bench args, n_ctx_avg for the bytes, kv type)])
CONFIGS = {
    "c1": ("tinyllama-1.1b-q4_0", [("tg128", ["-p", "0", "-n", "128"], 64, gs.F16), ("pp512", ["-p", "512", "-n", "0"], 0, gs.F16)]),
    "c2": ("llama3-8b-q4_k_m", [("tg128", ["-p", "0", "-n", "128"], 64, gs.F16), ("pp512", ["-p", "512", "-n", "0"], 0, gs.F16)]),
    "c3": ("qwen2.5-32b-q8_0", [("tg32@d32768 q8_0 KV", ["-p", "0", "-n", "32", "-d", "32768", "-ctk", "q8_0", "-ctv", "q8_0"], 32768 + 16, 8),
                                 ("tg128 q8_0 KV", ["-p", "0", "-n", "128", "-ctk", "q8_0", "-ctv", "q8_0"], 64, 8),
                                 ("pp512", ["-p", "512", "-n", "0"], 0, gs.F16)]),
    "c4": ("llama3-70b-q4_k_m", [("tg128", ["-p", "0", "-n", "128"], 64, gs.F16), ("pp512", ["-p", "512", "-n", "0"], 0, gs.F16)]),
    "c5": ("gpt-oss-120b-mxfp4", [("tg128", ["-p", "0", "-n", "128"], 64, gs.F16), ("pp512", ["-p", "512", "-n", "0"], 0, gs.F16)]),
}


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def llama_bench(gguf, args, reps, plugin, devices=None, sm=None, threads=None, timeout=3000):
    env = dict(os.environ)
    env.pop("GGML_BACKEND_PATH", None)
    cmd = [os.path.join(REF, "llama-bench"), "-m", gguf, "-r", str(reps), "-fa", "1", "-o", "json", "-ngl", "99" if plugin else "0"] + list(args)
    if plugin:
        env["GGML_BACKEND_PATH"] = BACKEND_SO
        if devices:
            env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(devices))
        if sm:
            cmd += ["-sm", sm]
    if threads:
        cmd += ["-t", str(threads)]
    t0 = time.time()
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout)
    if r.returncode != 0:
        return {"error": f"rc={r.returncode}: {r.stderr[-1500:]}", "cmd": " ".join(cmd)}
    out = json.loads(r.stdout)[0]
    return {"avg_ts": out["avg_ts"], "stddev_ts": out["stddev_ts"], "n_prompt": out["n_prompt"], "n_gen": out["n_gen"],
            "n_depth": out.get("n_depth", 0), "type_k": out.get("type_k"), "type_v": out.get("type_v"), "backends": out.get("backends"),
            "wall_s": round(time.time() - t0, 1), "cmd": " ".join(cmd[1:])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=sorted(CONFIGS))
    ap.add_argument("--cpu", action="store_true", help="also time the ggml CPU backend (tg16, pp64) on the host cores")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--devices", type=int, default=1)
    ap.add_argument("--sm", default=None)
    ap.add_argument("--only", default=None, help="substring of the run label")
    a = ap.parse_args()
    key, runs = CONFIGS[a.config]
    shape = gs.SHAPES[key]
    os.makedirs(MODEL_DIR, exist_ok=True)
    gguf = os.path.join(MODEL_DIR, key + ".gguf")
    t0 = time.time()
    if not os.path.exists(gguf):
        gs.write_gguf(gguf + ".tmp", shape)
        os.replace(gguf + ".tmp", gguf)
    gen_s = time.time() - t0
    pk, pk_src = peak()
    res = {"config": a.config, "workload": key, "gguf_bytes": os.path.getsize(gguf), "gguf_write_s": round(gen_s, 1), "devices": a.devices,
           "split_mode": a.sm or ("none" if a.devices == 1 else "layer (llama-bench default)"), "hbm_peak_gbs": pk, "hbm_peak_source": pk_src, "runs": []}
    for label, args, nctx, kvt in runs:
        if a.only and a.only not in label:
            continue
        r = llama_bench(gguf, args, a.reps, True, a.devices, a.sm)
        r["label"] = label
        if "avg_ts" in r and r["n_gen"] > 0:
            b = gs.decode_bytes_per_token(shape, nctx, kvt)
            r["bytes_per_token"] = b
            r["achieved_gbs"] = round(b["total"] * r["avg_ts"] / 1e9, 1)
            r["frac_of_hbm_peak"] = round(b["total"] * r["avg_ts"] / 1e9 / (pk * a.devices), 4)
        elif "avg_ts" in r:
            fl = 2.0 * gs.matmul_weight_params(shape) if hasattr(gs, "matmul_weight_params") else None
            if fl:
                r["matmul_flop_per_token"] = fl
                r["achieved_tflops"] = round(fl * r["avg_ts"] / 1e12, 1)
        res["runs"].append(r)
        print(json.dumps(r), flush=True)
    if a.cpu:
        import psutil
        cores = psutil.cpu_count(logical=False) or os.cpu_count()
        for label, args in (("cpu tg16", ["-p", "0", "-n", "16"]), ("cpu pp64", ["-p", "64", "-n", "0"])):
            r = llama_bench(gguf, args, 1, False, threads=cores)
            r["label"], r["cores"] = label, cores
            res["runs"].append(r)
            print(json.dumps(r), flush=True)
    else:
        res["cpu_baseline"] = "not timed on the CPU for this config (BASELINE.md section 4: minutes per token on host cores)"
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    suffix = "" if a.devices == 1 else f"_n{a.devices}_{a.sm or 'layer'}"
    with open(os.path.join(ROOT, "gpurun_out", f"config_{a.config}{suffix}.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
