"""ctypes bindings of include/b200_kernels.h (the C ABI of libb200_kernels.so).

Used by the parity tests and smoke(): device memory comes from torch (plumbing only), every
compute call goes through the C entry points.  No fallback: a missing library raises.
"""
from __future__ import annotations

import ctypes as C
import os

from . import KERNELS_SO

F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, I32, I64, BF16, MXFP4 = 0, 1, 2, 8, 9, 12, 14, 26, 27, 30, 39
GLU_NONE, GLU_REGLU, GLU_GEGLU, GLU_SWIGLU, GLU_SWIGLU_OAI, GLU_GEGLU_ERF, GLU_GEGLU_QUICK = -1, 0, 1, 2, 3, 4, 5


class Tensor(C.Structure):
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("_pad", C.c_int32),
                ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4)]


class MmvqArgs(C.Structure):
    _fields_ = [("w", Tensor), ("xq", C.c_void_p), ("xq_ne0_padded", C.c_int64), ("x_ne", C.c_int64 * 4),
                ("dst", Tensor), ("ids", C.POINTER(Tensor)), ("bias", C.POINTER(Tensor)),
                ("gate_w", C.POINTER(Tensor)), ("gate_bias", C.POINTER(Tensor)),
                ("glu_op", C.c_int32), ("glu_alpha", C.c_float), ("glu_limit", C.c_float),
                ("col_begin", C.c_int32), ("col_count", C.c_int32)]


class MmqArgs(C.Structure):
    _fields_ = [("w", Tensor), ("x", Tensor), ("dst", Tensor), ("xb", C.c_void_p), ("act_fmt", C.c_int32), ("_pad", C.c_int32),
                ("addend", C.POINTER(Tensor))]


_lib = None


def lib() -> C.CDLL:
    """Load libb200_kernels.so (built in-tree by csrc/Makefile). Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(KERNELS_SO):
            raise RuntimeError(f"{KERNELS_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        L = C.CDLL(KERNELS_SO)
        L.b200_last_error.restype = C.c_char_p
        L.b200_launch_count.restype = C.c_uint64
        L.b200_type_size.restype = C.c_size_t
        L.b200_type_size.argtypes = [C.c_int]
        L.b200_blck_size.restype = C.c_int64
        L.b200_blck_size.argtypes = [C.c_int]
        L.b200_row_size.restype = C.c_size_t
        L.b200_row_size.argtypes = [C.c_int, C.c_int64]
        L.b200_q8_1_buffer_size.restype = C.c_size_t
        L.b200_q8_1_buffer_size.argtypes = [C.POINTER(Tensor), C.c_int64]
        L.b200_quantize_q8_1.argtypes = [C.POINTER(Tensor), C.c_void_p, C.c_int64, C.c_void_p]
        L.b200_mmvq_supported.argtypes = [C.POINTER(MmvqArgs)]
        L.b200_mul_mat_vec_q.argtypes = [C.POINTER(MmvqArgs), C.c_void_p]
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise RuntimeError(f"{what}: b200 error {rc}: {lib().b200_last_error().decode()}")


def row_size(ttype: int, ne0: int) -> int:
    return lib().b200_row_size(ttype, ne0)


def make_tensor(ptr: int, ttype: int, ne, nb=None) -> Tensor:
    """Descriptor for `ptr` with ggml shape `ne` (ne[0] fastest); contiguous strides by default."""
    ne = list(ne) + [1] * (4 - len(ne))
    t = Tensor()
    t.data = ptr
    t.type = ttype
    if nb is None:
        ts = lib().b200_type_size(ttype)
        nb = [ts, lib().b200_row_size(ttype, ne[0])]
        nb.append(nb[1] * ne[1])
        nb.append(nb[2] * ne[2])
    for i in range(4):
        t.ne[i] = ne[i]
        t.nb[i] = nb[i]
    return t


_TORCH_TYPE = None


def from_torch(x, ttype: int | None = None, ne=None) -> Tensor:
    """Descriptor of a contiguous CUDA torch tensor. Quantized data is a uint8 tensor + explicit (ttype, ne)."""
    import torch
    assert x.is_cuda and x.is_contiguous()
    if ttype is None:
        ttype = {torch.float32: F32, torch.float16: F16, torch.int32: I32, torch.int64: I64, torch.bfloat16: BF16}[x.dtype]
        ne = list(reversed(x.shape))
    return make_tensor(x.data_ptr(), ttype, ne)


def stream_ptr() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


def quantize_q8_1(x, ne0_padded: int | None = None):
    """x: CUDA f32 torch tensor [..., K] -> uint8 tensor of q8_1 blocks [rows, K_pad/32*36]."""
    import torch
    K = x.shape[-1]
    if ne0_padded is None:
        ne0_padded = -(-K // 32) * 32
    tx = from_torch(x)
    rows = x.numel() // K
    y = torch.empty(rows * (ne0_padded // 32) * 36, dtype=torch.uint8, device=x.device)
    check(lib().b200_quantize_q8_1(C.byref(tx), y.data_ptr(), ne0_padded, stream_ptr()), "quantize_q8_1")
    return y.view(rows, -1)


def mul_mat_vec_q(w_bytes, wtype: int, w_ne, x, ids=None, bias=None, gate_bytes=None, gate_bias=None,
                  glu_op: int = GLU_NONE, alpha: float = 1.702, limit: float = 7.0, ne0_padded: int | None = None,
                  w_nb=None):
    """dst = W . x through b200_quantize_q8_1 + b200_mul_mat_vec_q.

    w_bytes: CUDA uint8 tensor with packed weights of ggml shape w_ne = [K, M, ne02, ne03];
    x: CUDA f32 tensor of torch shape [ne13, ne12, ne11, K] (or fewer leading dims);
    ids: CUDA int32 tensor [n_tokens, n_used] for MUL_MAT_ID.  Returns f32 dst (torch layout).
    """
    import torch
    K = x.shape[-1]
    xs = [1] * (4 - x.dim()) + list(x.shape)       # [ne13, ne12, ne11, K]
    if ne0_padded is None:
        ne0_padded = -(-K // 512) * 512
    xq = quantize_q8_1(x, ne0_padded)
    a = MmvqArgs()
    a.w = make_tensor(w_bytes.data_ptr(), wtype, w_ne, w_nb)
    a.xq = xq.data_ptr()
    a.xq_ne0_padded = ne0_padded
    for i, v in enumerate([K, xs[2], xs[1], xs[0]]):
        a.x_ne[i] = v
    M = w_ne[1]
    if ids is not None:
        n_tok, n_used = ids.shape
        dst = torch.empty(n_tok, n_used, M, dtype=torch.float32, device=x.device)
        tids = from_torch(ids)
        a.ids = C.pointer(tids)
    else:
        dst = torch.empty(xs[0], xs[1], xs[2], M, dtype=torch.float32, device=x.device)
    a.dst = from_torch(dst)
    keep = []
    for name, t in (("bias", bias), ("gate_bias", gate_bias)):
        if t is not None:
            d = from_torch(t)
            keep.append(d)
            setattr(a, name, C.pointer(d))
    if gate_bytes is not None:
        g = make_tensor(gate_bytes.data_ptr(), wtype, w_ne, w_nb)
        keep.append(g)
        a.gate_w = C.pointer(g)
    a.glu_op, a.glu_alpha, a.glu_limit = glu_op, alpha, limit
    check(lib().b200_mul_mat_vec_q(C.byref(a), stream_ptr()), "mul_mat_vec_q")
    return dst


# ----------------------------------------------------------------------------- glue ops + flash attention
class RopeParams(C.Structure):
    _fields_ = [("n_dims", C.c_int32), ("mode", C.c_int32), ("n_ctx_orig", C.c_int32), ("freq_base", C.c_float),
                ("freq_scale", C.c_float), ("ext_factor", C.c_float), ("attn_factor", C.c_float),
                ("beta_fast", C.c_float), ("beta_slow", C.c_float)]


class FattnArgs(C.Structure):
    _fields_ = [("q", Tensor), ("k", Tensor), ("v", Tensor), ("dst", Tensor), ("mask", C.POINTER(Tensor)),
                ("sinks", C.POINTER(Tensor)), ("scale", C.c_float), ("max_bias", C.c_float), ("logit_softcap", C.c_float),
                ("scratch", C.c_void_p), ("scratch_size", C.c_size_t)]


def _opt(t):
    return C.byref(t) if t is not None else None


def rms_norm(x, mul=None, add=None, eps=1e-5):
    import torch
    d = torch.empty_like(x)
    tx, td = from_torch(x), from_torch(d)
    tm = from_torch(mul) if mul is not None else None
    ta = from_torch(add) if add is not None else None
    L = lib()
    L.b200_rms_norm.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_void_p]
    check(L.b200_rms_norm(C.byref(tx), _opt(tm), _opt(ta), C.byref(td), eps, stream_ptr()), "rms_norm")
    return d


def glu(op, gate, up, alpha=1.702, limit=7.0):
    import torch
    d = torch.empty_like(up)
    tg, tu, td = from_torch(gate), from_torch(up), from_torch(d)
    L = lib()
    L.b200_glu.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p]
    check(L.b200_glu(op, C.byref(tg), C.byref(tu), C.byref(td), 0, alpha, limit, stream_ptr()), "glu")
    return d


def rope(x, pos, n_dims, mode, freq_base=10000.0, freq_scale=1.0, freq_factors=None):
    """x: [n_tok, n_head, head_dim] f32 CUDA; pos: [n_tok] int32."""
    import torch
    d = torch.empty_like(x)
    rp = RopeParams(n_dims, mode, 0, freq_base, freq_scale, 0.0, 1.0, 32.0, 1.0)
    tx, tp, td = from_torch(x), from_torch(pos), from_torch(d)
    tf = from_torch(freq_factors) if freq_factors is not None else None
    L = lib()
    L.b200_rope.argtypes = [C.c_void_p] * 8
    check(L.b200_rope(C.byref(tx), C.byref(tp), _opt(tf), C.byref(td), C.byref(rp), None, None, stream_ptr()), "rope")
    return d


def set_rows(src, idx, dst_bytes, dst_type, dst_ne):
    """src f32 [n, ne0]; idx int64 [n]; dst: uint8/typed CUDA tensor of ggml shape dst_ne, updated in place."""
    ts, ti = from_torch(src), from_torch(idx)
    td = make_tensor(dst_bytes.data_ptr(), dst_type, dst_ne)
    L = lib()
    L.b200_set_rows.argtypes = [C.c_void_p] * 4
    check(L.b200_set_rows(C.byref(ts), C.byref(ti), C.byref(td), stream_ptr()), "set_rows")


def get_rows(src_bytes, src_type, src_ne, ids):
    import torch
    d = torch.empty(ids.numel(), src_ne[0], dtype=torch.float32, device=ids.device)
    ts = make_tensor(src_bytes.data_ptr(), src_type, src_ne)
    ti, td = from_torch(ids), from_torch(d)
    L = lib()
    L.b200_get_rows.argtypes = [C.c_void_p] * 4
    check(L.b200_get_rows(C.byref(ts), C.byref(ti), C.byref(td), stream_ptr()), "get_rows")
    return d


def flash_attn(q, k_bytes, v_bytes, kv_type, n_kv, H_kv, mask=None, sinks=None, scale=1.0, max_bias=0.0, softcap=0.0,
               static_scratch=False, dst=None):
    """q: f32 [H, n_q, D] CUDA (ggml [D, n_q, H]); k/v: packed rows, ggml [D, n_kv, H_kv] contiguous;
    mask: f16 [n_q, n_kv]; returns f32 [n_q, H, D].
    static_scratch: call b200_flash_attn_ext_static (the library's own static device scratch; nothing is allocated by the
    call — pass `dst` to keep this wrapper allocation-free too)."""
    import torch
    H, n_q, D = q.shape
    if dst is None:
        dst = torch.empty(n_q, H, D, dtype=torch.float32, device=q.device)
    a = FattnArgs()
    a.q = from_torch(q)
    a.k = make_tensor(k_bytes.data_ptr(), kv_type, [D, n_kv, H_kv, 1])
    a.v = make_tensor(v_bytes.data_ptr(), kv_type, [D, n_kv, H_kv, 1])
    a.dst = from_torch(dst)
    keep = []
    if mask is not None:
        tm = from_torch(mask); keep.append(tm); a.mask = C.pointer(tm)
    if sinks is not None:
        tk = from_torch(sinks); keep.append(tk); a.sinks = C.pointer(tk)
    a.scale, a.max_bias, a.logit_softcap = scale, max_bias, softcap
    L = lib()
    L.b200_fattn_scratch_size.restype = C.c_size_t
    L.b200_fattn_scratch_size.argtypes = [C.c_void_p]
    L.b200_flash_attn_ext.argtypes = [C.c_void_p, C.c_void_p]
    if static_scratch:
        L.b200_flash_attn_ext_static.argtypes = [C.c_void_p, C.c_void_p]
        check(L.b200_flash_attn_ext_static(C.byref(a), stream_ptr()), "flash_attn_ext_static")
        return dst
    n = L.b200_fattn_scratch_size(C.byref(a))
    assert n > 0, "flash_attn: unsupported shape"
    scratch = torch.zeros(n, dtype=torch.uint8, device=q.device)
    a.scratch, a.scratch_size = scratch.data_ptr(), n
    check(L.b200_flash_attn_ext(C.byref(a), stream_ptr()), "flash_attn_ext")
    return dst


# ----------------------------------------------------------------------------- K4s: streaming fused GEMV
class GemvSeg(C.Structure):
    _fields_ = [("w", Tensor), ("bias", C.POINTER(Tensor)), ("dst", Tensor)]


class GemvFusedArgs(C.Structure):
    _fields_ = [("x", Tensor), ("norm_w", C.POINTER(Tensor)), ("eps", C.c_float), ("n_seg", C.c_int32),
                ("seg", GemvSeg * 4), ("glu_op", C.c_int32), ("glu_alpha", C.c_float), ("glu_limit", C.c_float),
                ("xq_out", C.c_void_p), ("n_sm", C.c_int32), ("flags", C.c_int32)]


def gemv_fused_args(x, segs, norm_w=None, eps=0.0, glu_op=GLU_NONE, alpha=1.702, limit=7.0, xq_out=None, n_sm=0, flags=0, keep=None):
    """Build a b200_gemv_fused_args. x: CUDA f32 [K]; segs: list of (w_bytes, wtype, M, bias|None, dst|None).
    `keep` collects the descriptors that must stay alive while launches referencing them are captured."""
    keep = keep if keep is not None else []
    K = x.numel()
    a = GemvFusedArgs()
    a.x = make_tensor(x.data_ptr(), F32, [K, 1, 1, 1])
    if norm_w is not None:
        tn = make_tensor(norm_w.data_ptr(), F32, [K, 1, 1, 1]); keep.append(tn); a.norm_w = C.pointer(tn)
    a.eps, a.n_seg = eps, len(segs)
    for i, (wb, wt, M, bias, dst) in enumerate(segs):
        a.seg[i].w = make_tensor(wb.data_ptr(), wt, [K, M, 1, 1])
        if bias is not None:
            tb = make_tensor(bias.data_ptr(), F32, [M, 1, 1, 1]); keep.append(tb); a.seg[i].bias = C.pointer(tb)
        if dst is not None:
            a.seg[i].dst = make_tensor(dst.data_ptr(), F32, [M, 1, 1, 1])
    a.glu_op, a.glu_alpha, a.glu_limit = glu_op, alpha, limit
    a.xq_out = xq_out.data_ptr() if xq_out is not None else None
    a.n_sm, a.flags = n_sm, flags
    keep.append(a)
    return a


def gemv_fused(x, segs, **kw):
    L = lib()
    L.b200_gemv_fused.argtypes = [C.c_void_p, C.c_void_p]
    L.b200_gemv_fused_supported.argtypes = [C.c_void_p]
    keep = []
    a = gemv_fused_args(x, segs, keep=keep, **kw)
    check(L.b200_gemv_fused(C.byref(a), stream_ptr()), "gemv_fused")


# ----------------------------------------------------------------------------- decode chain
class AttnPhase(C.Structure):
    _fields_ = [("q", Tensor), ("k_new", Tensor), ("v_new", Tensor), ("k_cache", Tensor), ("v_cache", Tensor),
                ("pos", C.POINTER(Tensor)), ("slot", C.POINTER(Tensor)), ("mask", C.POINTER(Tensor)),
                ("rope", RopeParams), ("scale", C.c_float), ("dst", Tensor), ("slot_v", C.POINTER(Tensor)),
                ("sinks", C.POINTER(Tensor))]


class Phase(C.Structure):
    _fields_ = [("kind", C.c_int32), ("_pad", C.c_int32), ("gemv", GemvFusedArgs), ("attn", AttnPhase)]


PHASE_GEMV, PHASE_ATTN, PHASE_NOP = 0, 1, 2
MAX_RANKS = 8


class PeerGroup(C.Structure):
    """b200_peer_group (include/b200_kernels.h)."""
    _fields_ = [("rank", C.c_int32), ("n_ranks", C.c_int32), ("total_ctas", C.c_int32), ("_pad", C.c_int32),
                ("arena", C.c_void_p), ("arena_bytes", C.c_size_t), ("delta", C.c_int64 * MAX_RANKS), ("sync", C.c_void_p),
                ("act_bytes", C.c_size_t), ("ll_offset", C.c_size_t)]


class _RawCuda:
    """Device memory not owned by torch, presented through __cuda_array_interface__."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerArena:
    """Symmetric activation arena of the row-split decode chain: the same allocation, with the same layout, on every
    rank; each rank maps the arenas of the others over CUDA IPC (NVLink peer access).  `dist` is an initialised
    torch.distributed (plumbing: it only carries the 64-byte handles and a barrier).
    Layout: [0, nbytes/4) activations and barrier counters (carve), [nbytes/2, nbytes) the shadow of tagged words."""

    def __init__(self, nbytes: int, dist, rank: int, world: int, device):
        assert nbytes % 1024 == 0
        self.act_bytes, self.ll_offset = nbytes // 4, nbytes // 2
        import torch
        L = lib()
        vp = C.c_void_p
        L.b200_peer_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.b200_peer_export.argtypes = [vp, C.c_char_p]
        L.b200_peer_open.argtypes = [C.c_char_p, C.POINTER(vp)]
        L.b200_peer_close.argtypes = [vp]
        L.b200_peer_free.argtypes = [vp]
        assert 1 <= world <= MAX_RANKS
        self.rank, self.world, self.nbytes, self.dist = rank, world, nbytes, dist
        base = vp()
        check(L.b200_peer_alloc(nbytes, C.byref(base)), "peer_alloc")
        self.base = base.value
        handle = C.create_string_buffer(64)
        check(L.b200_peer_export(self.base, handle), "peer_export")
        handles, grids = [None] * world, [None] * world
        dist.all_gather_object(handles, bytes(handle.raw))
        dist.all_gather_object(grids, int(L.b200_chain_grid()))
        self.total_ctas = sum(grids)
        self.delta, self.mapped = [0] * MAX_RANKS, []
        for r in range(world):
            if r == rank:
                continue
            p = vp()
            check(L.b200_peer_open(handles[r], C.byref(p)), f"peer_open(rank {r})")
            self.mapped.append(p.value)
            self.delta[r] = p.value - self.base
        self.bytes_t = torch.as_tensor(_RawCuda(self.base, nbytes), device=device)
        assert self.bytes_t.data_ptr() == self.base
        self.off = 0
        dist.barrier()               # every arena is zeroed and mapped before any rank can write to it

    def carve(self, n: int, dtype):
        """n elements of `dtype`, 256-byte aligned, at the same offset on every rank (same call sequence)."""
        import torch
        size = n * torch.empty((), dtype=dtype).element_size()
        off = (self.off + 255) & ~255
        assert off + size <= self.act_bytes, "peer arena exhausted"
        self.off = off + size
        return self.bytes_t[off:off + size].view(dtype)

    def group(self, sync) -> PeerGroup:
        g = PeerGroup()
        g.rank, g.n_ranks, g.total_ctas = self.rank, self.world, self.total_ctas
        g.arena, g.arena_bytes = self.base, self.nbytes
        for r in range(MAX_RANKS):
            g.delta[r] = self.delta[r]
        g.sync = sync.data_ptr()
        g.act_bytes, g.ll_offset = self.act_bytes, self.ll_offset
        return g


class Chain:
    """Owns a b200_chain (device-side phase descriptors). launch() is capturable into a CUDA graph.
    `peer`: a PeerGroup — the chain runs row-split over peer memory (b200_chain_create_peer)."""

    def __init__(self, phases, flags=0, keep=None, peer=None):
        L = lib()
        L.b200_chain_create.restype = C.c_void_p
        L.b200_chain_create.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.b200_chain_create_peer.restype = C.c_void_p
        L.b200_chain_create_peer.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.b200_chain_launch.argtypes = [C.c_void_p, C.c_void_p]
        L.b200_chain_destroy.argtypes = [C.c_void_p]
        arr = (Phase * len(phases))(*phases)
        self._keep = (arr, keep, peer)
        self.n = len(phases)
        if peer is not None:
            self.h = L.b200_chain_create_peer(arr, len(phases), flags, C.byref(peer))
        else:
            self.h = L.b200_chain_create(arr, len(phases), flags)
        if not self.h:
            raise RuntimeError("b200_chain_create: " + L.b200_last_error().decode())

    def launch(self):
        check(lib().b200_chain_launch(self.h, stream_ptr()), "chain_launch")

    def __del__(self):
        try:
            if self.h:
                lib().b200_chain_destroy(self.h)
                self.h = None
        except Exception:
            pass


def nop_phase():
    ph = Phase()
    ph.kind = PHASE_NOP
    return ph


def gemv_phase(x, segs, **kw):
    ph = Phase()
    ph.kind = PHASE_GEMV
    ph.gemv = gemv_fused_args(x, segs, **kw)
    return ph


def mul_mat_q(w, wtype: int, K: int, M: int, x, dst, act_fmt: int = 1, xb=None, addend=None):
    """dst[N, M] (torch, ggml [M, N]) = W . x for x [N, K] f32 through the tensor-core prefill path.
    Returns the converted activation buffer so callers can reuse it for matmuls sharing x."""
    import torch
    L = lib()
    L.b200_mmq_act_size.restype = C.c_size_t
    L.b200_mmq_act_size.argtypes = [C.c_int64, C.c_int64]
    N = x.shape[0]
    tx = from_torch(x)
    if xb is None:
        xb = torch.empty(L.b200_mmq_act_size(K, N) + 1024, dtype=torch.uint8, device=x.device)
        off = (-xb.data_ptr()) % 1024
        xb = xb[off:]
        check(L.b200_mmq_convert_act(C.byref(tx), C.c_void_p(xb.data_ptr()), act_fmt, C.c_void_p(stream_ptr())), "mmq_convert_act")
    a = MmqArgs()
    a.w = make_tensor(w.data_ptr(), wtype, [K, M, 1, 1])
    a.x = tx
    a.dst = from_torch(dst)
    a.xb = xb.data_ptr()
    a.act_fmt = act_fmt
    if addend is not None:
        ta = from_torch(addend)
        a.addend = C.pointer(ta)
    check(L.b200_mul_mat_q(C.byref(a), C.c_void_p(stream_ptr())), "mul_mat_q")
    return xb
