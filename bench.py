#!/usr/bin/env python3
"""bench.py — llama-bench tg128 on the B200 backend (BASELINE.json metric), one JSON line.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]

A "step" is one llama-bench tg128 run: 128 tokens generated one at a time from an empty KV cache
on a random-init GGUF-format model of the named architecture (reference
tools/llama-bench/llama-bench.cpp:1962-2010).  N=1 workload: Llama-3-8B-shape Q4_K_M (BASELINE.json
configs[1]).

  value     tok/s with everything resident in HBM (tokens, weights, KV): K steps of 128 CUDA-graph
            replays of the decode step, CUDA events on the launching stream, max over ranks.
  e2e       the same metric through the reference-facing boundary with HOST buffers:
            the reference's own llama-bench binary driving the ggml backend plugin
            (GGML_BACKEND_PATH=libggml-b200.so; token ids H2D and logits D2H every token).
            When that binary is not present the C-ABI session is timed with a pinned-host token
            H2D and a logits D2H per token instead ("via": "c-abi session").
  roofline  the dominant kernel (quantized GEMV): all GEMV launches of one token replayed alone,
            CUDA events; achieved = matmul weight bytes per token / that time.
  cpu_baseline  the reference's ggml CPU backend (oracle/_ref llama-bench, -ngl 0) on a bounded
            sample (tg16) of the same GGUF on this box's host cores.

--impl reference times the reference's own CPU implementation (oracle/_ref/llama-bench, CPU backend,
all host threads) on the same GGUF and the same step (one tg128 run, about 6 s on 16 host cores for the 8B model).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

N_GEN = 128
WORKLOADS = {"llama3-8b": "llama3-8b-q4_k_m", "tinyllama": "tinyllama-1.1b-q4_0", "qwen2.5-32b": "qwen2.5-32b-q8_0",
             "llama3-70b": "llama3-70b-q4_k_m", "tiny": "tiny-q4_k_m"}
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
LLAMA_BENCH = os.path.join(REF_DIR, "llama-bench")
MODEL_DIR = os.environ.get("B200_MODEL_DIR", "/tmp/b200_models")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        if d.get("bf16_tflops_sustained"):
            return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    return 1400.0, "fallback (B200_PROFILING.md sustained bf16)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_ev = index, [], threading.Event()

    def run(self):
        while not self.stop_ev.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self.stop_ev.wait(0.1)

    def summary(self):
        self.stop_ev.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


def ensure_gguf(shape_key):
    from llama_mi50_b200 import gguf_synth
    os.makedirs(MODEL_DIR, exist_ok=True)
    path = os.path.join(MODEL_DIR, shape_key + ".gguf")
    if not os.path.exists(path):
        gguf_synth.write_gguf(path + ".tmp", gguf_synth.SHAPES[shape_key])
        os.replace(path + ".tmp", path)
    return path


def host_cores():
    try:
        import psutil
        return psutil.cpu_count(logical=False) or os.cpu_count() or 1
    except Exception:
        return os.cpu_count() or 1


def run_llama_bench(gguf, n_gen, n_prompt, reps, plugin: bool, threads=None, extra_env=None, timeout=1500, n_devices=None, extra_args=()):
    """Run the reference's llama-bench; returns the parsed JSON list (or raises)."""
    env = dict(os.environ)
    env.pop("GGML_BACKEND_PATH", None)
    if plugin:
        from llama_mi50_b200 import BACKEND_SO
        env["GGML_BACKEND_PATH"] = BACKEND_SO
    env.update(extra_env or {})
    if plugin and n_devices is not None:
        # llama-bench offloads to EVERY device the plugin registers (layer split): on a multi-GPU box an N-GPU measurement
        # must see exactly N devices
        vis = [x for x in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if x.strip()]
        env["CUDA_VISIBLE_DEVICES"] = ",".join(vis[:n_devices]) if vis else ",".join(str(i) for i in range(n_devices))
    cmd = [LLAMA_BENCH, "-m", gguf, "-p", str(n_prompt), "-n", str(n_gen), "-r", str(reps), "-fa", "1", "-o", "json",
           "-ngl", "99" if plugin else "0"]
    if threads:
        cmd += ["-t", str(threads)]
    cmd += list(extra_args)
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"llama-bench failed rc={r.returncode}: {r.stderr[-2000:]}")
    return json.loads(r.stdout)


def workload_config(key, world):
    """`config` of the JSON line: the workload, identical for the b200 arm and the reference arm."""
    return {"workload": key, "n_prompt": 0, "n_gen": N_GEN, "n_kv": 256, "kv_type": "f16", "flash_attn": 1,
            "parallelism": f"row-split tp{world}" if world > 1 else "single device",
            "l2": "inputs larger than L2 (every weight byte is read once per token; the model is >> 126 MB)"}


def cpu_baseline(gguf, n_gen=16):
    cores = host_cores()
    if os.path.exists(LLAMA_BENCH):
        res = run_llama_bench(gguf, n_gen, 0, 1, plugin=False, threads=cores)
        tg = [x for x in res if x["n_gen"] == n_gen][0]
        return {"value": round(tg["avg_ts"], 3), "unit": "tok/s", "cores": cores, "kind": "reference",
                "sample": f"llama-bench tg{n_gen} -r 1 -ngl 0 -t {cores} on the same GGUF (ggml CPU backend, {tg.get('cpu_info', '?')})"}
    return {"value": None, "unit": "tok/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref/llama-bench not built"}


# ------------------------------------------------------------------------------------------------ reference arm
def main_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    graft.load_package()
    from llama_mi50_b200 import gguf_synth
    key = WORKLOADS[a.workload]
    gguf = ensure_gguf(key)
    cores = host_cores()
    n_gen = N_GEN          # the b200 arm's step: one llama-bench tg128 run (about 6 s per step on 16 host cores for the 8B model)
    t0 = time.time()
    res = run_llama_bench(gguf, n_gen, 0, a.steps + a.warmup, plugin=False, threads=cores)
    tg = [x for x in res if x["n_gen"] == n_gen][0]
    ns = tg["samples_ns"][a.warmup:] if len(tg["samples_ns"]) > a.warmup else tg["samples_ns"]
    ms = sum(ns) / len(ns) / 1e6
    val = n_gen / (ms / 1e3)
    line = {"impl": "reference", "metric": "llama-bench tg128 tok/s", "value": round(val, 3), "unit": "tok/s", "n_gpus": a.gpus,
            "steps": len(ns), "warmup": a.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 x int4/6 (ggml CPU vec_dot, Q8_K activations)", "data": "synthetic",
            "config": workload_config(key, a.gpus),
            "cpu_baseline": {"value": round(val, 3), "unit": "tok/s", "cores": cores, "kind": "reference",
                             "sample": f"llama-bench tg{n_gen} x {len(ns)} reps, ggml CPU backend, -t {cores}, {tg.get('cpu_info', '?')}"},
            "e2e": {"value": round(val, 3), "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": round(time.time() - t0, 1)}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------ own arm
def main_b200(a):
    import torch
    import torch.distributed as dist
    graft.load_package()
    from llama_mi50_b200 import capi, gguf_synth
    from llama_mi50_b200.session import DecodeSession

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its version banner on stdout when the first communicator comes up: keep stdout for the JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    key = WORKLOADS[a.workload]
    shape = gguf_synth.SHAPES[key]
    n_ctx = 256

    gather = None
    if world > 1:
        from llama_mi50_b200.rowsplit import make_gather
        gather = make_gather(dist, world, rank)
    # N > 1: the chain runs row-split over NVLink peer memory (outputs stored into every rank's buffers by the GEMV
    # epilogue, barriers spanning the ranks); --rowsplit nccl keeps one launch + one all-gather per matmul group
    mode = {"chain": "chain", "fused": True, "basic": False}[a.mode if world == 1 or a.mode != "chain" or a.rowsplit == "peer" else "fused"]
    sess = DecodeSession(shape, n_ctx=n_ctx, rank=rank, world=world, gather=gather, fused=mode, peer_dist=dist if world > 1 else None)
    sess.chain_flags = a.chain_flags
    sess.capture(sess.n_kv_for(N_GEN - 1))
    lib = capi.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def tg_step():
        sess.reset()
        for _ in range(N_GEN):
            sess.graph.replay()

    # ---- value: device-resident
    for _ in range(a.warmup):
        tg_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = lib.b200_launch_count()
    ev0.record()
    for _ in range(a.steps):
        tg_step()
    ev1.record()
    barrier()
    clocks = sampler.summary() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / a.steps
    value = N_GEN / (ms_per_step / 1e3)
    gpu_launches = sess.launches_per_token * N_GEN * a.steps

    # ---- e2e through the C ABI with host buffers (token H2D from pinned memory, logits D2H, per token)
    sess.resident_tokens = False
    sess.capture(sess.n_kv_for(N_GEN - 1))
    host_tok = torch.randint(0, shape.n_vocab, (N_GEN,), dtype=torch.int32).pin_memory()
    host_logits = torch.empty(shape.n_vocab, dtype=torch.float32).pin_memory()

    def tg_step_host():
        sess.reset()
        for i in range(N_GEN):
            sess.tok.copy_(host_tok[i:i + 1], non_blocking=True)
            sess.graph.replay()
            host_logits.copy_(sess.logits.view(-1), non_blocking=True)
            torch.cuda.current_stream().synchronize()

    for _ in range(min(a.warmup, 3)):
        tg_step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        tg_step_host()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_cabi = N_GEN * a.steps / e2e_s
    h2d, d2h = 4 * N_GEN, shape.n_vocab * 4 * N_GEN

    # ---- roofline of the dominant kernel
    hbm_peak, peak_src = peaks()
    bpt = gguf_synth.decode_bytes_per_token(shape, n_ctx_avg=(N_GEN - 1) / 2.0)
    wbytes = sess.weight_bytes_local
    chain_mode = mode == "chain"
    gemv_ms, gemv_n = sess.time_gemv_only(reps=max(20, a.steps * 4))       # one streaming-GEMV launch per matmul group, alone
    gemv_alone = {"achieved": round(wbytes / (gemv_ms / 1e3) / 1e9, 1), "launches_per_token": gemv_n,
                  "bytes_per_launch_avg": int(wbytes / max(gemv_n, 1)), "us_per_launch_avg": round(gemv_ms * 1e3 / max(gemv_n, 1), 3)}
    traffic = None
    try:   # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch from the committed ncu --set full capture (profiles/)
        # latest committed `ncu --set full` capture of the 1-GPU chain kernel (tools/ncu_summary.py)
        caps = sorted(x for x in os.listdir(os.path.join(ROOT, "profiles")) if x.endswith("_chain_ncu_summary.json"))
        with open(os.path.join(ROOT, "profiles", caps[-1])) as f:
            traffic = json.load(f).get("dram_bytes_per_launch") if chain_mode and world == 1 else None   # the capture is the 1-GPU chain
    except Exception:
        traffic = None
    if chain_mode:
        # the whole token is ONE launch of the persistent chain kernel (plus a get_rows and a 644 B memset): its average
        # duration over the timed region is ms_per_step / N_GEN (CUDA events on the launching stream, bracketing K steps)
        us_launch = ms_per_step * 1e3 / N_GEN
        # N > 1: this rank's row slices of the weights; the KV cache, embedding row and logits are per rank (replicated attention)
        alg = bpt["total"] if world == 1 else wbytes + bpt["total"] - bpt["weights"]
        roofline = {"bound": "hbm", "kernel": "gemv_stream_kernel<CHAIN> (decode chain: every matmul, RoPE, KV write and attention of one token in one cooperative launch)",
                    "achieved": round(alg / us_launch / 1e3, 1), "peak": hbm_peak, "unit": "GB/s", "frac": round(alg / us_launch / 1e3 / hbm_peak, 4),
                    "traffic": traffic, "peak_source": peak_src, "launches_per_token": 1, "algorithmic_bytes_per_launch": alg,
                    "us_per_launch_avg": round(us_launch, 2), "frac_of_nominal_8000": round(alg / us_launch / 1e3 / 8000.0, 4),
                    "matmul_launches_alone": gemv_alone}
    else:
        roofline = {"bound": "hbm", "kernel": "quantized GEMV (gemv_stream_kernel / b200_mul_mat_vec_q family), every GEMV launch of one token replayed alone",
                    "achieved": gemv_alone["achieved"], "peak": hbm_peak, "unit": "GB/s", "frac": round(gemv_alone["achieved"] / hbm_peak, 4),
                    "traffic": traffic, "peak_source": peak_src, "launches_per_token": gemv_n,
                    "bytes_per_launch_avg": gemv_alone["bytes_per_launch_avg"], "us_per_launch_avg": gemv_alone["us_per_launch_avg"],
                    "step": {"bytes_per_token": bpt["total"] // world if world > 1 else bpt["total"],
                             "achieved": round(bpt["total"] / world * value / 1e9, 1),
                             "frac": round(bpt["total"] / world * value / 1e9 / hbm_peak, 4),
                             "frac_of_nominal_8000": round(bpt["total"] / world * value / 1e9 / 8000.0, 4)}}

    line = {"metric": "llama-bench tg128 tok/s", "value": round(value, 2), "unit": "tok/s", "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int8 x int4/6 dp4a, f32 accumulate", "data": "synthetic",
            "config": workload_config(key, world), "issue": a.mode,
            "exchange": ("decode chain over NVLink peer memory (no collective)" if mode == "chain" else "NCCL all-gather per matmul") if world > 1 else None,
            "roofline": roofline, "clocks": clocks, "gpu_launches": gpu_launches,
            "e2e": {"value": round(e2e_cabi, 2), "unit": "tok/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "via": "c-abi session (pinned-host token H2D + logits D2H + sync per token)"}}

    # ---- rank 0, N=1: the reference's own harness through the plugin, and the CPU baseline
    if rank == 0 and world == 1 and not a.no_llama_bench:
        del sess
        torch.cuda.empty_cache()
        try:
            gguf = ensure_gguf(key)
            if os.path.exists(LLAMA_BENCH):
                res = run_llama_bench(gguf, N_GEN, 512, a.steps + min(a.warmup, 2), plugin=True, extra_env={"GGML_B200_TIMING": "0"}, n_devices=1)
                tg = [x for x in res if x["n_gen"] == N_GEN][0]
                pp = [x for x in res if x["n_prompt"] == 512][0]
                w = min(a.warmup, 2)
                tg_ns = tg["samples_ns"][w:] or tg["samples_ns"]
                tg_ts = N_GEN / (sum(tg_ns) / len(tg_ns) / 1e9)
                line["e2e_cabi"] = line["e2e"]
                line["e2e"] = {"value": round(tg_ts, 2), "unit": "tok/s", "h2d_bytes_per_step": h2d + shape.n_embd * 4 * N_GEN,
                               "d2h_bytes_per_step": d2h,
                               "via": "reference llama-bench (-p 0 -n 128 -fa 1 -ngl 99) driving libggml-b200.so via GGML_BACKEND_PATH"}
                line["llama_bench"] = {"tg128": round(tg_ts, 2), "pp512": round(pp["avg_ts"], 2), "backends": tg.get("backends"),
                                       "gpu_info": tg.get("gpu_info")}
                # the prefill half of the metric: matmul flops per prompt token (2 per weight of every MUL_MAT; attention and glue
                # are not counted) against the measured dense bf16 peak of this pool's B200s (the tensor-core GEMM multiplies
                # 16-bit operands, csrc/kernels/mmq_tc.cu); the sustained figure, since the kernel runs inside a long step
                flop_tok = 2.0 * gguf_synth.matmul_weight_params(shape)
                tf_peak, tf_src = tensor_peak()
                tflops = flop_tok * pp["avg_ts"] / 1e12
                line["roofline_prefill"] = {"bound": "tensor", "kernel": "mmq_tc_kernel (tcgen05.mma kind::f16 over in-smem expanded GGUF blocks)",
                                            "flops": flop_tok * 512, "flop_per_token": flop_tok, "achieved": round(tflops, 1), "peak": tf_peak,
                                            "unit": "TFLOP/s", "frac": round(tflops / tf_peak, 4), "peak_source": tf_src,
                                            "tok_s": round(pp["avg_ts"], 2), "via": "reference llama-bench pp512 driving the plugin"}
            line["cpu_baseline"] = cpu_baseline(gguf)
        except Exception as e:  # the C-ABI numbers above stand on their own
            line["llama_bench_error"] = str(e)[-400:]
            line.setdefault("cpu_baseline", {"value": None, "unit": "tok/s", "cores": 0, "kind": "reference", "sample": "failed: " + str(e)[-200:]})
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        # tearing the NCCL communicator down while captured graphs still hold its kernels was observed to hang
        # (2 x B200, torch 2.11): leave together and skip the teardown
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush(); sys.stderr.flush()
        os._exit(0)
    return 0


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="llama3-8b", choices=sorted(WORKLOADS))
    ap.add_argument("--no-llama-bench", action="store_true", help="skip the llama-bench/plugin and CPU-baseline legs")
    ap.add_argument("--mode", default="chain", choices=["chain", "fused", "basic"],
                    help="chain: one cooperative launch per token; fused: one streaming-GEMV launch per matmul group; basic: unfused ops")
    ap.add_argument("--chain-flags", type=lambda x: int(x, 0), default=0)
    ap.add_argument("--rowsplit", default="peer", choices=["peer", "nccl"],
                    help="N > 1: peer = decode chain with peer-memory stores; nccl = per-matmul launches + all-gather")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    sys.exit(main_reference(args) if args.impl == "reference" else main_b200(args))
