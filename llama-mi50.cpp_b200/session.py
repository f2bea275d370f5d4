"""Decode session: the token-generation graph of a llama-architecture model, issued as C-ABI calls.

This is the measurement harness for `bench.py` (and the N>1 row-split path): it issues, for one
token, the same op sequence libllama builds for a llama/qwen2 decode ubatch (reference
src/models/llama.cpp:3-165) — but every op is a call into
libb200_kernels.so through include/b200_kernels.h.  torch supplies device memory, the stream and
CUDA-graph capture/replay only; no torch math runs on the token path except integer bookkeeping
(position / KV-slot counters).

Weights are random-init tensors in GGUF wire format (gguf_synth.tensor_plan / make_tensor_bytes:
the same bytes gguf_synth.write_gguf streams to a file for llama-bench).

Row split (world > 1): every matmul weight is partitioned by output rows
across ranks exactly like the reference's split buffer type (ggml-cuda.cu:791-804 get_row_split,
boundaries rounded to 128 rows); activations are replicated, each rank computes its row slice and
the slices are concatenated with an all-gather (never summed), so the N-GPU logits are
bit-identical to the 1-GPU logits.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import capi, gguf_synth
from .gguf_synth import F16, F32, Q8_0

ROW_ROUND = 128  # reference: split boundaries rounded down to mmq_y rows (ggml-cuda.cu:778-789)


class _Addr:
    """A bare device address where a tensor is expected (capi only calls data_ptr())."""

    def __init__(self, ptr):
        self._p = ptr

    def data_ptr(self):
        return self._p


def row_split(n_rows: int, world: int, rank: int) -> tuple[int, int]:
    """[lo, hi) of the rows rank owns; equal tensor_split fractions, boundaries rounded down to 128."""
    def bound(r):
        if r >= world:
            return n_rows
        b = n_rows * r // world
        return b - b % ROW_ROUND
    return bound(rank), bound(rank + 1)


class DecodeSession:
    def __init__(self, shape: gguf_synth.ModelShape, n_ctx: int = 256, kv_type: int = F16, seed: int = 1234,
                 rank: int = 0, world: int = 1, device=None, fused: bool = True, gather=None, peer_dist=None):
        import torch
        self.torch = torch
        self.s = shape
        self.rank, self.world = rank, world
        self.dev = device or torch.device("cuda", torch.cuda.current_device())
        self.kv_type = kv_type
        self.n_ctx = n_ctx
        self.fused = fused
        self.gather = gather            # callable(local [n_loc] f32, full [n] f32, lo, hi) for world > 1
        # row split over peer memory (fused == "chain", world > 1): activations live in a symmetric arena that the
        # other ranks' GEMV epilogues write into; peer_dist = torch.distributed (handle exchange + barrier only)
        self.arena = None
        if fused == "chain" and (world > 1 or peer_dist is not None):
            assert peer_dist is not None, "row-split decode chain needs peer_dist (torch.distributed) for the arena exchange"
            self.arena = capi.PeerArena(16 << 20, peer_dist, rank, world, self.dev)
        self.L = capi.lib()
        self._bind()
        self.keep = []                  # descriptors / arrays that must outlive captured launches
        self.weight_bytes_local = 0
        self._load_weights(seed)
        self._alloc_state()
        self.graph = None
        self.launches_per_token = 0
        self.chain, self.chain_n_kv, self.chain_keep, self.chain_flags = None, -1, [], 0

    # ------------------------------------------------------------------ ctypes plumbing
    def _bind(self):
        L, vp = self.L, C.c_void_p
        L.b200_get_rows.argtypes = [vp] * 4
        L.b200_rms_norm.argtypes = [vp] * 4 + [C.c_float, vp]
        L.b200_bin_bcast.argtypes = [C.c_int] + [vp] * 4
        L.b200_glu.argtypes = [C.c_int, vp, vp, vp, C.c_int, C.c_float, C.c_float, vp]
        L.b200_rope.argtypes = [vp] * 8
        L.b200_set_rows.argtypes = [vp] * 4
        L.b200_fattn_scratch_size.restype = C.c_size_t
        L.b200_fattn_scratch_size.argtypes = [vp]
        L.b200_flash_attn_ext.argtypes = [vp, vp]
        L.b200_gemv_fused.argtypes = [vp, vp]

    def _t(self, x, ttype=None, ne=None, nb=None):
        """descriptor of a torch CUDA tensor (kept alive)."""
        if ttype is None:
            d = capi.from_torch(x)
        else:
            d = capi.make_tensor(x.data_ptr(), ttype, ne, nb)
        self.keep.append((d, x))
        return d

    # ------------------------------------------------------------------ weights
    def _load_weights(self, seed):
        torch, s = self.torch, self.s
        rng = np.random.default_rng(seed)
        self.w = {}
        for name, t, ne in gguf_synth.tensor_plan(s):
            val = 1.0 if name.endswith("_norm.weight") else None
            data = gguf_synth.make_tensor_bytes(rng, t, ne, 0.02, val)
            is_mat = len(ne) == 2 and name.endswith(".weight")
            if is_mat and name != "token_embd.weight" and self.world > 1:
                lo, hi = row_split(ne[1], self.world, self.rank)       # row slice of this rank
                rb = gguf_synth.row_bytes(t, ne[0])
                data = data[lo * rb:hi * rb]
                ne = (ne[0], hi - lo)
            if t == F32:
                dev = torch.from_numpy(np.ascontiguousarray(data)).to(self.dev).view(torch.float32)
            else:   # >= 16 B of slack after the last row: bulk copies round their end up to 16 B
                dev = torch.zeros(data.nbytes + 64, dtype=torch.uint8, device=self.dev)
                dev[:data.nbytes] = torch.from_numpy(np.ascontiguousarray(data)).to(self.dev)
            self.w[name] = (dev, t, ne)
            if is_mat and name != "token_embd.weight":
                self.weight_bytes_local += data.nbytes

    # ------------------------------------------------------------------ activations / KV cache
    def _alloc_state(self):
        torch, s = self.torch, self.s
        f32 = dict(dtype=torch.float32, device=self.dev)
        E, Fd, QD, KVD = s.n_embd, s.n_ff, s.n_head * s.head_dim, s.n_head_kv * s.head_dim
        self.tok = torch.zeros(1, dtype=torch.int32, device=self.dev)
        # device-resident token stream (llama-bench feeds rand() % n_vocab tokens, llama-bench.cpp:1975-2007)
        g = torch.Generator(device="cpu"); g.manual_seed(4321)
        self.tokens = torch.randint(0, s.n_vocab, (self.n_ctx,), generator=g, dtype=torch.int32).to(self.dev)
        self.resident_tokens = True      # False: the caller writes self.tok before every replay (host-fed tokens)
        self.pos = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.slot = torch.zeros(1, dtype=torch.int64, device=self.dev)

        def act(n):      # matmul outputs: in the symmetric arena when the chain runs row-split over peer memory
            if self.arena is not None:
                return self.arena.carve(n, torch.float32).view(1, n)
            return torch.zeros(1, n, **f32)
        self.x = act(E)
        self.x2 = act(E)
        self.h = torch.zeros(1, E, **f32)
        self.q = act(QD)
        self.k = act(KVD)
        self.v = act(KVD)
        self.qr = torch.zeros(1, QD, **f32)
        self.kr = torch.zeros(1, KVD, **f32)
        self.att = act(QD)
        self.ff = act(Fd)
        self.logits = act(s.n_vocab)
        # per-rank slices of matmul outputs before the gather (world > 1)
        self.loc = torch.zeros(max(s.n_vocab, Fd, QD), **f32) if self.world > 1 else None
        # KV cache: per layer [n_ctx rows of n_head_kv*head_dim], f16 or q8_0 (reference src/llama-kv-cache.cpp)
        rb = gguf_synth.row_bytes(self.kv_type, KVD)
        self.kv_row_bytes = rb
        self.kc = [torch.zeros(self.n_ctx * rb, dtype=torch.uint8, device=self.dev) for _ in range(s.n_layer)]
        self.vc = [torch.zeros(self.n_ctx * rb, dtype=torch.uint8, device=self.dev) for _ in range(s.n_layer)]
        # f16 mask [n_kv] (0 for filled slots, -inf beyond): reference src/llama-graph.cpp:1698-1738
        self.mask = torch.full((1, self.n_ctx), float("-inf"), dtype=torch.float16, device=self.dev)
        self.zero16 = torch.zeros(1, 1, dtype=torch.float16, device=self.dev)
        self.fa_scratch = torch.zeros(8 << 20, dtype=torch.uint8, device=self.dev)
        self.q8 = torch.zeros(4 << 20, dtype=torch.uint8, device=self.dev)

    def reset(self):
        self.pos.zero_(); self.slot.zero_(); self.mask.fill_(float("-inf"))
        for c in self.kc + self.vc:
            c.zero_()

    # ------------------------------------------------------------------ op wrappers (all on the current torch stream)
    def _st(self):
        return self.torch.cuda.current_stream().cuda_stream

    def _chk(self, rc, what):
        capi.check(rc, what)
        self._n += 1

    def _mm(self, wname, x, dst, bias=None, gate=None, glu=capi.GLU_NONE, requant=True):
        """dst = W . x (+bias) or glu(gate.x) * (W.x); under row split the local slice is gathered into dst."""
        wd, t, ne = self.w[wname]
        K, M = ne
        a = capi.MmvqArgs()
        a.w = capi.make_tensor(wd.data_ptr(), t, [K, M, 1, 1])
        Kp = -(-K // 512) * 512
        if requant:
            tx = self._t(x)
            self._chk(self.L.b200_quantize_q8_1(C.byref(tx), self.q8.data_ptr(), Kp, self._st()), "quantize")
        a.xq, a.xq_ne0_padded = self.q8.data_ptr(), Kp
        for i, v in enumerate([K, 1, 1, 1]):
            a.x_ne[i] = v
        split = self.world > 1
        out = self.loc[:M].view(1, M) if split else dst
        a.dst = capi.from_torch(out)
        if bias is not None:
            lo, hi = row_split(dst.shape[-1], self.world, self.rank) if split else (0, M)
            b = bias.view(-1)[lo:hi]
            tb = self._t(b.view(1, -1)); a.bias = C.pointer(tb)
        if gate is not None:
            gd, gt, gne = self.w[gate]
            tg = capi.make_tensor(gd.data_ptr(), gt, [K, M, 1, 1]); self.keep.append(tg); a.gate_w = C.pointer(tg)
        a.glu_op, a.glu_alpha, a.glu_limit = glu, 1.702, 7.0
        self.keep.append(a)
        if M > 0:        # a rank may own no rows of a small matrix (boundaries round down to 128 rows)
            self._chk(self.L.b200_mul_mat_vec_q(C.byref(a), self._st()), "mmvq " + wname)
        if split:
            lo, hi = row_split(dst.shape[-1], self.world, self.rank)
            self.gather(out.view(-1), dst.view(-1), lo, hi)

    def _gemv(self, x, segs, norm=None, glu=capi.GLU_NONE):
        """streaming fused GEMV: segs = [(weight name, bias tensor|None, dst tensor|None)]; all share activation x.
        Under row split every segment's local slice is gathered into its dst."""
        split = self.world > 1
        csegs, post, o = [], [], 0
        for wname, bias, dst in segs:
            wd, t, (K, M) = self.w[wname]
            out = dst
            if dst is not None and split:
                lo, hi = row_split(dst.numel(), self.world, self.rank)
                out = self.loc[o:o + M]; o += M
                post.append((out, dst, lo, hi))
                bias = bias.view(-1)[lo:hi] if bias is not None else None
            if M > 0:    # a rank may own no rows of a small matrix (boundaries round down to 128 rows)
                csegs.append((wd, t, M, bias, out))
        nw = self.w[norm][0] if norm else None
        if csegs:
            a = capi.gemv_fused_args(x.view(-1), csegs, norm_w=nw, eps=self.s.rms_eps, glu_op=glu, keep=self.keep)
            self._chk(self.L.b200_gemv_fused(C.byref(a), self._st()), "gemv_fused")
        for out, dst, lo, hi in post:
            self.gather(out, dst.view(-1), lo, hi)

    def _norm(self, x, wname, dst):
        tx, tw, td = self._t(x), self._t(self.w[wname][0].view(1, -1)), self._t(dst)
        self._chk(self.L.b200_rms_norm(C.byref(tx), C.byref(tw), None, C.byref(td), self.s.rms_eps, self._st()), "rms_norm")

    def _rope(self, x, dst, n_head, cache=None):
        """RoPE; with `cache` the rotated row goes straight into the KV cache slot (fused SET_ROWS, f16 cache)."""
        s = self.s
        mode = 2 if s.arch == "qwen2" else 0
        rp = capi.RopeParams(s.head_dim, mode, s.n_ctx_train, s.rope_base, 1.0, 0.0, 1.0, 32.0, 1.0)
        tx = self._t(x.view(1, n_head, s.head_dim)); td = self._t(dst.view(1, n_head, s.head_dim)); tp = self._t(self.pos)
        self.keep.append(rp)
        ti = tc = None
        if cache is not None:
            ti = self._t(self.slot)
            tc = self._t(cache, self.kv_type, [n_head * s.head_dim, self.n_ctx, 1, 1])
        self._chk(self.L.b200_rope(C.byref(tx), C.byref(tp), None, C.byref(td), C.byref(rp),
                                   C.byref(ti) if ti else None, C.byref(tc) if tc else None, self._st()), "rope")

    def _set_row(self, src, cache):
        KVD = self.s.n_head_kv * self.s.head_dim
        ts, ti = self._t(src), self._t(self.slot)
        td = self._t(cache, self.kv_type, [KVD, self.n_ctx, 1, 1])
        self._chk(self.L.b200_set_rows(C.byref(ts), C.byref(ti), C.byref(td), self._st()), "set_rows")

    def _attn(self, il, n_kv):
        s = self.s
        D, H, Hkv = s.head_dim, s.n_head, s.n_head_kv
        a = capi.FattnArgs()
        a.q = capi.make_tensor(self.qr.data_ptr(), F32, [D, 1, H, 1], [4, 4 * D * H, 4 * D, 4 * D * H])
        hb = gguf_synth.row_bytes(self.kv_type, D)
        ts = 2 if self.kv_type == F16 else 34
        for name, cache in (("k", self.kc[il]), ("v", self.vc[il])):
            setattr(a, name, capi.make_tensor(cache.data_ptr(), self.kv_type, [D, n_kv, Hkv, 1],
                                              [ts, self.kv_row_bytes, hb, self.kv_row_bytes * self.n_ctx]))
        a.dst = capi.make_tensor(self.att.data_ptr(), F32, [D, H, 1, 1])
        tm = self._t(self.mask, F16, [n_kv, 1, 1, 1], [2, 2 * self.n_ctx, 2 * self.n_ctx, 2 * self.n_ctx])
        a.mask = C.pointer(tm)
        a.scale, a.max_bias, a.logit_softcap = 1.0 / math.sqrt(D), 0.0, 0.0
        need = self.L.b200_fattn_scratch_size(C.byref(a))
        assert 0 < need <= self.fa_scratch.numel(), need
        a.scratch, a.scratch_size = self.fa_scratch.data_ptr(), self.fa_scratch.numel()
        self.keep.append(a)
        self._chk(self.L.b200_flash_attn_ext(C.byref(a), self._st()), "flash_attn_ext")

    # ------------------------------------------------------------------ one token
    def _build_chain(self, n_kv):
        """The whole token as ONE cooperative launch: per layer [QKV | attention | wo | gate/up | down], then the logits."""
        s = self.s
        D, H, Hkv = s.head_dim, s.n_head, s.n_head_kv
        assert self.kv_type == F16, "decode chain: f16 KV"
        keep, phases = self.chain_keep, []
        x, x2 = self.x, self.x2
        world, rank = self.world, self.rank

        def seg(name, bias, dst, n_full=None):
            """(weights, type, rows, bias, dst) of this rank's row slice (the whole matrix on one GPU)."""
            wd, t, (K, M) = self.w[name]
            if world > 1:
                lo, hi = row_split(dst.numel() if dst is not None else n_full, world, rank)
                assert hi - lo == M
                bias = bias.view(-1)[lo:hi] if bias is not None and M > 0 else None
                if dst is not None:      # an empty torch slice has no address: a 0-row segment still names where its slice starts
                    dst = dst.view(-1)[lo:hi] if M > 0 else _Addr(dst.data_ptr() + 4 * lo)
            return (wd, t, M, bias, dst)

        def gemv(xin, segs, **kw):
            # under row split a rank may own no rows of a small matrix: the 0-row segment still tells the library which
            # slice of dst this rank does NOT get from itself
            return capi.gemv_phase(xin.view(-1), segs, keep=keep, **kw)

        hb = gguf_synth.row_bytes(self.kv_type, D)
        mode = 2 if s.arch == "qwen2" else 0
        for il in range(s.n_layer):
            p = f"blk.{il}."
            phases.append(gemv(x, [seg(p + "attn_q.weight", self._full_bias(p + "attn_q.bias"), self.q.view(-1)),
                                   seg(p + "attn_k.weight", self._full_bias(p + "attn_k.bias"), self.k.view(-1)),
                                   seg(p + "attn_v.weight", self._full_bias(p + "attn_v.bias"), self.v.view(-1))],
                               norm_w=self.w[p + "attn_norm.weight"][0], eps=s.rms_eps))
            ph = capi.Phase(); ph.kind = capi.PHASE_ATTN
            a = ph.attn
            a.q = capi.make_tensor(self.q.data_ptr(), F32, [D * H, 1, 1, 1])
            a.k_new = capi.make_tensor(self.k.data_ptr(), F32, [D * Hkv, 1, 1, 1])
            a.v_new = capi.make_tensor(self.v.data_ptr(), F32, [D * Hkv, 1, 1, 1])
            nb = [2, self.kv_row_bytes, hb, self.kv_row_bytes * self.n_ctx]
            a.k_cache = capi.make_tensor(self.kc[il].data_ptr(), F16, [D, n_kv, Hkv, 1], nb)
            a.v_cache = capi.make_tensor(self.vc[il].data_ptr(), F16, [D, n_kv, Hkv, 1], nb)
            tp, ts = capi.from_torch(self.pos), capi.from_torch(self.slot)
            tm = capi.make_tensor(self.mask.data_ptr(), F16, [n_kv, 1, 1, 1], [2, 2 * self.n_ctx, 2 * self.n_ctx, 2 * self.n_ctx])
            keep += [tp, ts, tm]
            a.pos, a.slot, a.mask = C.pointer(tp), C.pointer(ts), C.pointer(tm)
            a.rope = capi.RopeParams(D, mode, s.n_ctx_train, s.rope_base, 1.0, 0.0, 1.0, 32.0, 1.0)
            a.scale = 1.0 / math.sqrt(D)
            a.dst = capi.make_tensor(self.att.data_ptr(), F32, [D * H, 1, 1, 1])
            phases.append(ph)
            phases.append(gemv(self.att, [seg(p + "attn_output.weight", x.view(-1), x2.view(-1))]))
            phases.append(gemv(x2, [seg(p + "ffn_gate.weight", None, None, n_full=s.n_ff), seg(p + "ffn_up.weight", None, self.ff.view(-1))],
                               norm_w=self.w[p + "ffn_norm.weight"][0], eps=s.rms_eps, glu_op=capi.GLU_SWIGLU))
            phases.append(gemv(self.ff, [seg(p + "ffn_down.weight", x2.view(-1), x.view(-1))]))
        phases.append(gemv(x, [seg("output.weight", None, self.logits.view(-1))],
                           norm_w=self.w["output_norm.weight"][0], eps=s.rms_eps))
        peer = None
        if self.arena is not None:
            # barrier counters of this chain: fresh (zero) words of the arena, at the same offset on every rank
            sync = self.arena.carve(len(phases), self.torch.int32)
            keep.append(sync)
            peer = self.arena.group(sync)
        return capi.Chain(phases, flags=self.chain_flags, keep=keep, peer=peer)

    def _issue_token_chain(self, n_kv):
        torch = self.torch
        self._n = 0
        self.mask.view(-1).index_copy_(0, self.slot, self.zero16.view(-1))
        if self.resident_tokens:
            torch.index_select(self.tokens, 0, self.slot, out=self.tok)
        emb, et, ene = self.w["token_embd.weight"]
        te, tt, tx = self._t(emb, et, [ene[0], ene[1], 1, 1]), self._t(self.tok), self._t(self.x)
        self._chk(self.L.b200_get_rows(C.byref(te), C.byref(tt), C.byref(tx), self._st()), "get_rows")
        if self.chain is None or self.chain_n_kv != n_kv:
            self.chain, self.chain_n_kv = self._build_chain(n_kv), n_kv
        self.chain.launch()
        self._n += 1
        self.pos.add_(1); self.slot.add_(1)
        return self._n

    def _issue_token(self, n_kv):
        """Issue every kernel of one decode step on the current stream. Returns the launch count."""
        if self.fused == "chain":
            return self._issue_token_chain(n_kv)
        s, torch = self.s, self.torch
        self._n = 0
        # bookkeeping on device: open the mask at the slot being written (integer/f16 plumbing)
        self.mask.view(-1).index_copy_(0, self.slot, self.zero16.view(-1))
        if self.resident_tokens:
            torch.index_select(self.tokens, 0, self.slot, out=self.tok)
        emb, et, ene = self.w["token_embd.weight"]
        te, tt, tx = self._t(emb, et, [ene[0], ene[1], 1, 1]), self._t(self.tok), self._t(self.x)
        self._chk(self.L.b200_get_rows(C.byref(te), C.byref(tt), C.byref(tx), self._st()), "get_rows")
        x, x2 = self.x, self.x2
        for il in range(s.n_layer):
            p = f"blk.{il}."
            if self.fused:
                self._gemv(x, [(p + "attn_q.weight", self._full_bias(p + "attn_q.bias"), self.q),
                               (p + "attn_k.weight", self._full_bias(p + "attn_k.bias"), self.k),
                               (p + "attn_v.weight", self._full_bias(p + "attn_v.bias"), self.v)], norm=p + "attn_norm.weight")
                self._rope(self.q, self.qr, s.n_head)
                if self.kv_type == F16:
                    self._rope(self.k, self.kr, s.n_head_kv, cache=self.kc[il])
                else:
                    self._rope(self.k, self.kr, s.n_head_kv)
                    self._set_row(self.kr, self.kc[il])
                self._set_row(self.v, self.vc[il])
                self._attn(il, n_kv)
                self._gemv(self.att, [(p + "attn_output.weight", x, x2)])
                self._gemv(x2, [(p + "ffn_gate.weight", None, None), (p + "ffn_up.weight", None, self.ff)],
                           norm=p + "ffn_norm.weight", glu=capi.GLU_SWIGLU)
                self._gemv(self.ff, [(p + "ffn_down.weight", x2, x)])
                continue
            self._norm(x, p + "attn_norm.weight", self.h)
            self._mm(p + "attn_q.weight", self.h, self.q, bias=self._full_bias(p + "attn_q.bias"))
            self._mm(p + "attn_k.weight", self.h, self.k, bias=self._full_bias(p + "attn_k.bias"), requant=False)
            self._mm(p + "attn_v.weight", self.h, self.v, bias=self._full_bias(p + "attn_v.bias"), requant=False)
            self._rope(self.q, self.qr, s.n_head)
            self._rope(self.k, self.kr, s.n_head_kv)
            self._set_row(self.kr, self.kc[il])
            self._set_row(self.v, self.vc[il])
            self._attn(il, n_kv)
            self._mm(p + "attn_output.weight", self.att, x2, bias=x)        # residual rides in the GEMV epilogue
            self._norm(x2, p + "ffn_norm.weight", self.h)
            self._mm(p + "ffn_up.weight", self.h, self.ff, gate=p + "ffn_gate.weight", glu=capi.GLU_SWIGLU)
            self._mm(p + "ffn_down.weight", self.ff, x, bias=x2)
        if self.fused:
            self._gemv(x, [("output.weight", None, self.logits)], norm="output_norm.weight")
        else:
            self._norm(x, "output_norm.weight", self.h)
            self._mm("output.weight", self.h, self.logits)
        self.pos.add_(1); self.slot.add_(1)
        return self._n

    def _full_bias(self, name):
        """Full-length bias vector (or None); _mm slices it by the rank's row range under row split."""
        return self.w[name][0] if name in self.w else None

    # ------------------------------------------------------------------ public API
    def n_kv_for(self, n_past: int) -> int:
        """KV length the attention kernel sees: padded to 256 like libllama (src/llama-kv-cache.cpp get_padding)."""
        return min(self.n_ctx, -(-(n_past + 1) // 256) * 256)

    def capture(self, n_kv: int):
        """Capture one decode step into a CUDA graph (replayed per token)."""
        torch = self.torch
        self.keep.clear()
        st = torch.cuda.Stream(device=self.dev)
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            self._issue_token(n_kv)          # warm-up outside capture (lazy module load, scratch sizing)
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        self.reset()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.launches_per_token = self._issue_token(n_kv)
        self.graph, self.graph_n_kv = g, n_kv
        self.reset()

    def decode(self, token: int | None = None):
        """One token through the captured graph. `token` None = the device-resident self.tok is used as is."""
        if token is not None:
            self.tok.fill_(token)
        self.graph.replay()
        return self.logits

    def decode_eager(self, token: int, n_past: int):
        self.tok.fill_(token)
        self._issue_token(self.n_kv_for(n_past))
        return self.logits

    # ------------------------------------------------------------------ dominant-kernel timing (bench.py roofline)
    def time_gemv_only(self, reps: int = 20):
        """Replay every GEMV launch of one token (same weights, same arguments, no glue, no gather)
        `reps` times; returns (ms per token-worth of GEMVs, launches per token).  The weights of one
        pass (GBs) exceed L2, so every launch streams from HBM."""
        torch, s = self.torch, self.s
        gather, self.gather = self.gather, (lambda *a: None)
        mode, self.fused = self.fused, (True if self.fused == "chain" else self.fused)

        def issue():
            self._n = 0
            for il in range(s.n_layer if self.fused else 0):
                p = f"blk.{il}."
                self._gemv(self.x, [(p + "attn_q.weight", None, self.q), (p + "attn_k.weight", None, self.k),
                                    (p + "attn_v.weight", None, self.v)], norm=p + "attn_norm.weight")
                self._gemv(self.att, [(p + "attn_output.weight", self.x, self.x2)])
                self._gemv(self.x2, [(p + "ffn_gate.weight", None, None), (p + "ffn_up.weight", None, self.ff)],
                           norm=p + "ffn_norm.weight", glu=capi.GLU_SWIGLU)
                self._gemv(self.ff, [(p + "ffn_down.weight", self.x2, self.x)])
            if self.fused:
                self._gemv(self.x, [("output.weight", None, self.logits)], norm="output_norm.weight")
                return self._n
            for il in range(s.n_layer):
                p = f"blk.{il}."
                self._mm(p + "attn_q.weight", self.h, self.q, requant=False)
                self._mm(p + "attn_k.weight", self.h, self.k, requant=False)
                self._mm(p + "attn_v.weight", self.h, self.v, requant=False)
                self._mm(p + "attn_output.weight", self.att, self.x2, bias=self.x, requant=False)
                self._mm(p + "ffn_up.weight", self.h, self.ff, gate=p + "ffn_gate.weight", glu=capi.GLU_SWIGLU, requant=False)
                self._mm(p + "ffn_down.weight", self.ff, self.x, bias=self.x2, requant=False)
            self._mm("output.weight", self.h, self.logits, requant=False)
            return self._n

        self.keep.clear()
        st = torch.cuda.Stream(device=self.dev)
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            issue()
        torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            n = issue()
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        self.gather, self.fused = gather, mode
        return e0.elapsed_time(e1) / reps, n
