// exec.cpp — supports_op and graph_compute of the B200 backend.
// Replaces the op dispatcher ggml_cuda_compute_forward (reference ggml-cuda/ggml-cuda.cu:2422-2762),
// supports_op (:4440-4860) and the graph executor with fusion + CUDA-graph capture/replay
// (:2862-3086, :3394-3922, :3941-3977).  Design differences:
//   * one function translates a node into C-ABI calls; supports_op runs it in validation mode,
//     so an op is claimed iff the very same code path would execute it;
//   * the activation vector of consecutive MUL_MATs that share src1 (wq/wk/wv, gate/up) is
//     quantized to q8_1 ONCE (the reference re-quantizes per matmul, mmvq.cu:700-706);
//   * kernels are chained with programmatic dependent launch inside one captured CUDA graph.
#include "b200_plugin.h"

#include <chrono>
#include <cmath>
#include <cstring>

namespace {

bool env_flag(const char * name) { const char * v = getenv(name); return v && atoi(v) != 0; }

struct exec_state {
    b200_context * ctx = nullptr;
    void * stream = B200_STREAM_VALIDATE;
    bool validate = true;
    // q8_1 activation cache: the quantized form of `q8_src` currently held in ctx->q8_scratch
    const ggml_tensor * q8_src = nullptr;
    const void * q8_src_data = nullptr;
    const ggml_tensor * xb_src = nullptr;     // same for the 16-bit activation tiles of the tensor-core GEMM
    const void * xb_src_data = nullptr;
    int xb_fmt = -1;
    bool no_tc = env_flag("GGML_B200_DISABLE_TC");
    size_t need_q8 = 0, need_fa = 0, need_xb = 0;   // scratch high-water marks found during the sizing pass
    size_t need_split_x = 0;           // row split: bytes of the largest activation copied to the other devices
    bool sizing = false;               // sizing pass: compute need_*, launch nothing
    uint64_t launches = 0, chain_phases = 0;
};

int64_t pad_to(int64_t x, int64_t n) { return (x + n - 1) / n * n; }

bool is_quant_weight(ggml_type t) {
    return t == GGML_TYPE_Q4_0 || t == GGML_TYPE_Q8_0 || t == GGML_TYPE_Q4_K || t == GGML_TYPE_Q6_K || t == GGML_TYPE_MXFP4;
}

int map_unary(ggml_unary_op op) {
    switch (op) {
        case GGML_UNARY_OP_SILU: return B200_UNARY_SILU;
        case GGML_UNARY_OP_GELU: return B200_UNARY_GELU;
        case GGML_UNARY_OP_RELU: return B200_UNARY_RELU;
        case GGML_UNARY_OP_SIGMOID: return B200_UNARY_SIGMOID;
        case GGML_UNARY_OP_TANH: return B200_UNARY_TANH;
        case GGML_UNARY_OP_GELU_QUICK: return B200_UNARY_GELU_QUICK;
        case GGML_UNARY_OP_NEG: return B200_UNARY_NEG;
        case GGML_UNARY_OP_EXP: return B200_UNARY_EXP;
        default: return -1;
    }
}

bool is_split_tensor(const ggml_tensor * t) {
    ggml_backend_buffer_t buf = t ? (t->view_src ? t->view_src->buffer : t->buffer) : nullptr;
    return buf && b200_buft_is_split(buf->buft);
}

// ---- MUL_MAT with a row-split weight -----------------------------------------------------------
// Replaces ggml_cuda_op_mul_mat (reference ggml-cuda.cu:1452-1770): every device multiplies ITS rows of the weight with
// the whole activation and the row slices of dst are concatenated on the device that owns dst (never summed).  Like the
// reference this is driven from the host, once per matmul: the activation is copied to each other device behind an
// event, the slice runs there on a side stream, and the main stream waits for it.  Unlike the reference there is no
// copy back: the kernel of a slice stores straight into dst over NVLink peer memory (dst rows [lo, hi), stride of the
// whole dst).  The decode fast path for row split is the peer-memory decode chain; this path is
// what makes `--split-mode row` work for every graph shape.
int run_mul_mat_split(exec_state & st, const ggml_tensor * dst) {
    const ggml_tensor * w = dst->src[0]; const ggml_tensor * x = dst->src[1];
    if (dst->op != GGML_OP_MUL_MAT || !is_quant_weight(w->type) || x->type != GGML_TYPE_F32 || dst->type != GGML_TYPE_F32) return B200_ERR_UNSUPPORTED;
    if (w->ne[2] != 1 || w->ne[3] != 1 || x->ne[2] != 1 || x->ne[3] != 1 || x->ne[0] != w->ne[0]) return B200_ERR_UNSUPPORTED;
    if (x->nb[0] != 4 || x->nb[1] != (size_t) x->ne[0] * 4 || dst->nb[0] != 4 || dst->nb[1] != (size_t) dst->ne[0] * 4) return B200_ERR_UNSUPPORTED;
    if (x->ne[1] > 65535) return B200_ERR_UNSUPPORTED;
    const int64_t K = x->ne[0], ncols = x->ne[1], M = w->ne[1];
    const int n_dev = b200_n_devices();
    ggml_backend_buffer_type_t sbt = w->buffer->buft;

    // one slice: weight rows [lo, hi) on `id` (stored at wdata) times the activation at xdata -> dst rows [lo, hi)
    auto slice = [&](void * wdata, int64_t lo, int64_t hi, void * xdata, char * q8, size_t q8_size, char * xbuf, size_t xb_size, void * stream, bool probe) -> int {
        ggml_tensor wt = *w; wt.data = wdata; wt.ne[1] = hi - lo;
        ggml_tensor xt = *x; xt.data = xdata;
        ggml_tensor dt = *dst; dt.data = (char *) dst->data + (size_t) lo * 4; dt.ne[0] = hi - lo;
        if (ncols > 8 && !st.no_tc) {
            b200_mmq_args m{};
            m.w = b200_desc(&wt); m.x = b200_desc(&xt); m.dst = b200_desc(&dt); m.act_fmt = b200_mmq_act_fmt((int) w->type);
            if (b200_mmq_supported(&m)) {
                const size_t need = b200_mmq_act_size(K, ncols) + 1024;
                if (probe) { st.need_xb = std::max(st.need_xb, need); return B200_OK; }
                GGML_ASSERT(need <= xb_size);
                char * xb = (char *) (((uintptr_t) xbuf + 1023) & ~(uintptr_t) 1023);
                B200_KCHECK(b200_mmq_convert_act(&m.x, xb, m.act_fmt, stream));
                m.xb = xb;
                B200_KCHECK(b200_mul_mat_q(&m, stream));
                st.launches += 2;
                return B200_OK;
            }
        }
        b200_mmvq_args a{};
        a.w = b200_desc(&wt);
        a.xq_ne0_padded = pad_to(K, B200_MATRIX_ROW_PADDING);
        for (int i = 0; i < 4; ++i) a.x_ne[i] = x->ne[i];
        a.dst = b200_desc(&dt);
        a.glu_op = B200_GLU_NONE;
        const int nwin = (int) ((ncols + 7) / 8);
        for (int wi = 0; wi < nwin; ++wi) {
            if (nwin > 1) { a.col_begin = wi * 8; a.col_count = (int) std::min<int64_t>(8, ncols - wi * 8); }
            if (!b200_mmvq_supported(&a)) return B200_ERR_UNSUPPORTED;
            if (wi > 0 && wi < nwin - 1) wi = nwin - 2;
        }
        const b200_tensor tx = b200_desc(&xt);
        const size_t q8_bytes = b200_q8_1_buffer_size(&tx, a.xq_ne0_padded);
        if (probe) { st.need_q8 = std::max(st.need_q8, q8_bytes); return B200_OK; }
        GGML_ASSERT(q8_bytes <= q8_size);
        B200_KCHECK(b200_quantize_q8_1(&tx, q8, a.xq_ne0_padded, stream));
        a.xq = q8;
        for (int wi = 0; wi < nwin; ++wi) {
            if (nwin > 1) { a.col_begin = wi * 8; a.col_count = (int) std::min<int64_t>(8, ncols - wi * 8); }
            B200_KCHECK(b200_mul_mat_vec_q(&a, stream));
            st.launches++;
        }
        st.launches++;
        return B200_OK;
    };

    if (st.validate || st.sizing) {
        // shapes only: a slice of 128 rows (or the whole matrix) at a dummy 256 B-aligned address
        int64_t lo = 0, hi = 0;
        if (st.validate) {
            // like the reference (ggml-cuda.cu:4568-4581): a device without rows of a small matrix declines the op
            const int dev = st.ctx ? st.ctx->device : -1;
            if (dev >= 0) { b200_split_rows(sbt, M, dev, lo, hi); if (lo == hi) return B200_ERR_UNSUPPORTED; }
        }
        const int rc = slice((void *) 0x1000, 0, M, (void *) 0x1000, nullptr, 0, nullptr, 0, B200_STREAM_VALIDATE, true);
        if (rc == B200_OK && st.sizing) st.need_split_x = std::max(st.need_split_x, (size_t) K * (size_t) ncols * 4);
        return rc;
    }

    const b200_split_extra * e = (const b200_split_extra *) w->extra;
    GGML_ASSERT(e && "row-split weight without per-device slices");
    b200_context * ctx = st.ctx;
    const int main_dev = ctx->device;
    st.q8_src = nullptr; st.xb_src = nullptr;                 // the main device's scratch arenas are reused below
    bool remote = false;
    for (int id = 0; id < n_dev; ++id) remote = remote || (id != main_dev && e->row_hi[id] > e->row_lo[id]);
    if (remote) B200_CUDA_CHECK(cudaEventRecord(ctx->split_fork, (cudaStream_t) st.stream));
    for (int id = 0; id < n_dev; ++id) {
        const int64_t lo = e->row_lo[id], hi = e->row_hi[id];
        if (hi == lo) continue;
        if (id == main_dev) {
            B200_KCHECK(slice(e->data[id], lo, hi, x->data, ctx->q8_scratch, ctx->q8_size, ctx->xb_scratch, ctx->xb_size, st.stream, false));
            continue;
        }
        b200_split_dev & sd = ctx->split[id];
        b200_set_device(id);
        B200_CUDA_CHECK(cudaStreamWaitEvent(sd.stream, ctx->split_fork, 0));      // x is ready and the old contents of dst are dead
        const size_t xbytes = (size_t) K * (size_t) ncols * 4;
        GGML_ASSERT(xbytes <= sd.x_size);
        B200_CUDA_CHECK(cudaMemcpyPeerAsync(sd.x, b200_device(id).id, x->data, b200_device(main_dev).id, xbytes, sd.stream));
        B200_KCHECK(slice(e->data[id], lo, hi, sd.x, sd.q8, sd.q8_size, sd.xb, sd.xb_size, sd.stream, false));
        B200_CUDA_CHECK(cudaEventRecord(sd.done, sd.stream));
    }
    b200_set_device(main_dev);
    for (int id = 0; id < n_dev; ++id)
        if (id != main_dev && e->row_hi[id] > e->row_lo[id]) B200_CUDA_CHECK(cudaStreamWaitEvent((cudaStream_t) st.stream, ctx->split[id].done, 0));
    return B200_OK;
}

// ---- MUL_MAT / MUL_MAT_ID -------------------------------------------------------------------
// quantize src1 into the q8_1 arena (unless it is already there) and run the GEMV in <= 8-column windows
int run_mul_mat(exec_state & st, const ggml_tensor * dst, const ggml_tensor * bias, const ggml_tensor * gate,
                const ggml_tensor * gate_bias, int glu_op, float alpha, float limit) {
    const ggml_tensor * w = dst->src[0]; const ggml_tensor * x = dst->src[1];
    const ggml_tensor * ids = dst->op == GGML_OP_MUL_MAT_ID ? dst->src[2] : nullptr;
    if (is_split_tensor(w)) return (bias || gate || ids) ? B200_ERR_UNSUPPORTED : run_mul_mat_split(st, dst);
    if (!is_quant_weight(w->type) || x->type != GGML_TYPE_F32 || dst->type != GGML_TYPE_F32) return B200_ERR_UNSUPPORTED;
    if (x->nb[0] != sizeof(float) || x->ne[0] != w->ne[0]) return B200_ERR_UNSUPPORTED;
    if (x->ne[1] > 65535 || x->ne[2] * x->ne[3] > 65535) return B200_ERR_UNSUPPORTED;

    b200_mmvq_args a{};
    a.w = b200_desc(w);
    a.xq_ne0_padded = pad_to(x->ne[0], B200_MATRIX_ROW_PADDING);
    for (int i = 0; i < 4; ++i) a.x_ne[i] = x->ne[i];
    a.dst = b200_desc(dst);
    b200_tensor t_ids, t_bias, t_gate, t_gbias;
    if (ids)       { t_ids = b200_desc(ids); a.ids = &t_ids; }
    if (bias)      { t_bias = b200_desc(bias); a.bias = &t_bias; }
    if (gate)      { t_gate = b200_desc(gate); a.gate_w = &t_gate; }
    if (gate_bias) { t_gbias = b200_desc(gate_bias); a.gate_bias = &t_gbias; }
    a.glu_op = gate ? glu_op : B200_GLU_NONE; a.glu_alpha = alpha; a.glu_limit = limit;

    const int64_t ncols = ids ? 1 : x->ne[1];
    // ---- prefill: more than 8 columns go to the tensor cores (reference router: ggml_cuda_mul_mat, ggml-cuda.cu:2183-2266)
    // a `bias` of dst's full shape (ncols > 1) is the residual stream of a {MUL_MAT, ADD} pair: only the tensor-core path takes it
    const bool full_addend = bias && ncols > 1 && ggml_are_same_shape(bias, dst);
    if (!ids && (!bias || full_addend) && !gate && ncols > 8 && !st.no_tc) {
        b200_mmq_args m{};
        m.w = b200_desc(w); m.x = b200_desc(x); m.dst = b200_desc(dst); m.act_fmt = b200_mmq_act_fmt((int) w->type);
        if (full_addend) m.addend = &t_bias;
        if (b200_mmq_supported(&m)) {
            if (st.validate) return B200_OK;
            const size_t need = b200_mmq_act_size(x->ne[0], x->ne[1]) + 1024;
            if (st.sizing) { st.need_xb = std::max(st.need_xb, need); return B200_OK; }
            GGML_ASSERT(need <= st.ctx->xb_size);
            char * xb = (char *) (((uintptr_t) st.ctx->xb_scratch + 1023) & ~(uintptr_t) 1023);
            if (st.xb_src != x || st.xb_src_data != x->data || st.xb_fmt != m.act_fmt) {      // q/k/v and gate/up share one conversion
                B200_KCHECK(b200_mmq_convert_act(&m.x, xb, m.act_fmt, st.stream));
                st.xb_src = x; st.xb_src_data = x->data; st.xb_fmt = m.act_fmt; st.launches++;
            }
            m.xb = xb;
            B200_KCHECK(b200_mul_mat_q(&m, st.stream));
            st.launches++;
            return B200_OK;
        }
    }
    if (full_addend) return B200_ERR_UNSUPPORTED;
    const int nwin = (int) ((ncols + 7) / 8);
    for (int wi = 0; wi < nwin; ++wi) {   // validate every window shape
        if (nwin > 1) { a.col_begin = wi * 8; a.col_count = (int) std::min<int64_t>(8, ncols - wi * 8); }
        if (!b200_mmvq_supported(&a)) return B200_ERR_UNSUPPORTED;
        if (wi > 0 && wi < nwin - 1) wi = nwin - 2;   // interior windows are identical
    }
    if (st.validate) return B200_OK;

    const b200_tensor tx = b200_desc(x);
    const size_t q8_bytes = b200_q8_1_buffer_size(&tx, a.xq_ne0_padded);
    const bool grouped = ids && !bias && !gate && !gate_bias && dst->ne[1] * dst->ne[2] >= 16 &&
                         b200_mul_mat_id_grouped(&a, nullptr, 0, B200_STREAM_VALIDATE) == B200_OK;
    if (st.sizing) {
        st.need_q8 = std::max(st.need_q8, q8_bytes);
        if (grouped) st.need_fa = std::max(st.need_fa, b200_mul_mat_id_grouped_scratch_size(&a));
        return B200_OK;
    }
    GGML_ASSERT(q8_bytes <= st.ctx->q8_size);
    if (st.q8_src != x || st.q8_src_data != x->data) {
        B200_KCHECK(b200_quantize_q8_1(&tx, st.ctx->q8_scratch, a.xq_ne0_padded, st.stream));
        st.q8_src = x; st.q8_src_data = x->data; st.launches++;
    }
    a.xq = st.ctx->q8_scratch;
    if (grouped) {      // MUL_MAT_ID of a batch: pairs bucketed by expert, 8 columns per weight read (scratch shared with flash attention:
                        // both are consumed inside their own op, in stream order)
        B200_KCHECK(b200_mul_mat_id_grouped(&a, st.ctx->fa_scratch, st.ctx->fa_size, st.stream));
        st.launches += 2;
        return B200_OK;
    }
    if (nwin > 1 && !ids && !bias && !gate && !gate_bias && x->ne[2] == 1 && x->ne[3] == 1) {      // all windows in one launch
        a.col_begin = a.col_count = 0;
        if (b200_mul_mat_vec_q_wide(&a, B200_STREAM_VALIDATE) == B200_OK) {
            B200_KCHECK(b200_mul_mat_vec_q_wide(&a, st.stream));
            st.launches++;
            return B200_OK;
        }
    }
    for (int wi = 0; wi < nwin; ++wi) {
        if (nwin > 1) { a.col_begin = wi * 8; a.col_count = (int) std::min<int64_t>(8, ncols - wi * 8); }
        B200_KCHECK(b200_mul_mat_vec_q(&a, st.stream));
        st.launches++;
    }
    return B200_OK;
}

int run_flash_attn(exec_state & st, const ggml_tensor * dst) {
    const ggml_tensor * q = dst->src[0]; const ggml_tensor * k = dst->src[1]; const ggml_tensor * v = dst->src[2];
    const ggml_tensor * mask = dst->src[3]; const ggml_tensor * sinks = dst->src[4];
    b200_fattn_args a{};
    a.q = b200_desc(q); a.k = b200_desc(k); a.v = b200_desc(v); a.dst = b200_desc(dst);
    b200_tensor tm, ts;
    if (mask)  { tm = b200_desc(mask); a.mask = &tm; }
    if (sinks) { ts = b200_desc(sinks); a.sinks = &ts; }
    memcpy(&a.scale, (const float *) dst->op_params + 0, sizeof(float));
    memcpy(&a.max_bias, (const float *) dst->op_params + 1, sizeof(float));
    memcpy(&a.logit_softcap, (const float *) dst->op_params + 2, sizeof(float));
    if (st.validate) return b200_flash_attn_ext(&a, B200_STREAM_VALIDATE);
    const size_t need = b200_fattn_scratch_size(&a);
    if (st.sizing) { st.need_fa = std::max(st.need_fa, need); return B200_OK; }
    GGML_ASSERT(need <= st.ctx->fa_size);
    a.scratch = st.ctx->fa_scratch; a.scratch_size = st.ctx->fa_size;
    B200_KCHECK(b200_flash_attn_ext(&a, st.stream));
    st.launches++;
    return B200_OK;
}

#define RUN(expr)                                                     \
    do {                                                              \
        const int rc__ = (expr);                                      \
        if (st.validate || st.sizing) return rc__;                    \
        if (rc__ != B200_OK) b200_abort(__FILE__, __LINE__, #expr, b200_last_error()); \
        st.launches++;                                                \
        return B200_OK;                                               \
    } while (0)

// Single (unfused) node. In validation/sizing mode nothing is launched.
int run_node(exec_state & st, const ggml_tensor * n) {
    void * s = (st.validate || st.sizing) ? B200_STREAM_VALIDATE : st.stream;
    const ggml_tensor * s0 = n->src[0]; const ggml_tensor * s1 = n->src[1];
    switch (n->op) {
        case GGML_OP_NONE: case GGML_OP_VIEW: case GGML_OP_RESHAPE: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE:
            return B200_OK;
        case GGML_OP_GET_ROWS: { b200_tensor a = b200_desc(s0), b = b200_desc(s1), d = b200_desc(n); RUN(b200_get_rows(&a, &b, &d, s)); }
        case GGML_OP_RMS_NORM: {
            b200_tensor a = b200_desc(s0), d = b200_desc(n); float eps; memcpy(&eps, n->op_params, sizeof(float));
            RUN(b200_rms_norm(&a, nullptr, nullptr, &d, eps, s));
        }
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV: {
            const int op = n->op == GGML_OP_ADD ? B200_BIN_ADD : n->op == GGML_OP_SUB ? B200_BIN_SUB : n->op == GGML_OP_MUL ? B200_BIN_MUL : B200_BIN_DIV;
            b200_tensor a = b200_desc(s0), b = b200_desc(s1), d = b200_desc(n);
            RUN(b200_bin_bcast(op, &a, &b, &d, s));
        }
        case GGML_OP_SCALE: {
            b200_tensor a = b200_desc(s0), d = b200_desc(n); float sc, bi; memcpy(&sc, n->op_params, 4); memcpy(&bi, (const float *) n->op_params + 1, 4);
            RUN(b200_scale(&a, &d, sc, bi, s));
        }
        case GGML_OP_UNARY: {
            const int op = map_unary(ggml_get_unary_op(n));
            if (op < 0) return B200_ERR_UNSUPPORTED;
            b200_tensor a = b200_desc(s0), d = b200_desc(n);
            RUN(b200_unary(op, &a, &d, s));
        }
        case GGML_OP_GLU: {
            b200_tensor a = b200_desc(s0), d = b200_desc(n), b; if (s1) b = b200_desc(s1);
            float alpha, limit; memcpy(&alpha, (const float *) n->op_params + 2, 4); memcpy(&limit, (const float *) n->op_params + 3, 4);
            RUN(b200_glu((int) ggml_get_glu_op(n), &a, s1 ? &b : nullptr, &d, n->op_params[1], alpha, limit, s));
        }
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP: { b200_tensor a = b200_desc(s0), d = b200_desc(n); RUN(b200_cpy(&a, &d, s)); }
        case GGML_OP_SET_ROWS: { b200_tensor a = b200_desc(s0), b = b200_desc(s1), d = b200_desc(n); RUN(b200_set_rows(&a, &b, &d, s)); }
        case GGML_OP_ROPE: {
            b200_rope_params rp{};
            const int32_t * p = n->op_params;
            rp.n_dims = p[1]; rp.mode = p[2]; rp.n_ctx_orig = p[4];
            memcpy(&rp.freq_base, p + 5, 4); memcpy(&rp.freq_scale, p + 6, 4); memcpy(&rp.ext_factor, p + 7, 4);
            memcpy(&rp.attn_factor, p + 8, 4); memcpy(&rp.beta_fast, p + 9, 4); memcpy(&rp.beta_slow, p + 10, 4);
            b200_tensor a = b200_desc(s0), b = b200_desc(s1), d = b200_desc(n), f; if (n->src[2]) f = b200_desc(n->src[2]);
            RUN(b200_rope(&a, &b, n->src[2] ? &f : nullptr, &d, &rp, nullptr, nullptr, s));
        }
        case GGML_OP_SOFT_MAX: {
            b200_tensor a = b200_desc(s0), d = b200_desc(n), m, k; if (s1) m = b200_desc(s1); if (n->src[2]) k = b200_desc(n->src[2]);
            float sc, mb; memcpy(&sc, n->op_params, 4); memcpy(&mb, (const float *) n->op_params + 1, 4);
            RUN(b200_soft_max(&a, s1 ? &m : nullptr, n->src[2] ? &k : nullptr, &d, sc, mb, s));
        }
        case GGML_OP_ARGSORT: { b200_tensor a = b200_desc(s0), d = b200_desc(n); RUN(b200_argsort(&a, &d, n->op_params[0], s)); }
        case GGML_OP_TOP_K:   { b200_tensor a = b200_desc(s0), d = b200_desc(n); RUN(b200_top_k(&a, &d, s)); }
        case GGML_OP_ADD_ID:  { b200_tensor a = b200_desc(s0), b = b200_desc(s1), i = b200_desc(n->src[2]), d = b200_desc(n); RUN(b200_add_id(&a, &b, &i, &d, s)); }
        case GGML_OP_SUM_ROWS: { b200_tensor a = b200_desc(s0), d = b200_desc(n); RUN(b200_sum_rows(&a, &d, s)); }
        case GGML_OP_CLAMP: {
            b200_tensor a = b200_desc(s0), d = b200_desc(n); float lo, hi; memcpy(&lo, n->op_params, 4); memcpy(&hi, (const float *) n->op_params + 1, 4);
            RUN(b200_clamp(&a, &d, lo, hi, s));
        }
        case GGML_OP_FLASH_ATTN_EXT: return run_flash_attn(st, n);
        case GGML_OP_MUL_MAT:
            if (s0->type == GGML_TYPE_F32 || s0->type == GGML_TYPE_F16 || s0->type == GGML_TYPE_BF16) {     // the MoE router matrix
                if (is_split_tensor(s0)) return B200_ERR_UNSUPPORTED;
                b200_tensor a = b200_desc(s0), b = b200_desc(s1), d = b200_desc(n); RUN(b200_mul_mat_f(&a, &b, nullptr, &d, s));
            }
            return run_mul_mat(st, n, nullptr, nullptr, nullptr, B200_GLU_NONE, 0.0f, 0.0f);
        case GGML_OP_MUL_MAT_ID: return run_mul_mat(st, n, nullptr, nullptr, nullptr, B200_GLU_NONE, 0.0f, 0.0f);
        default: return B200_ERR_UNSUPPORTED;
    }
}
#undef RUN

}  // namespace

bool b200_supports_op(int device, const ggml_tensor * op) {
    // every source must live in (or be allocatable in) this device's buffers
    for (int i = 0; i < GGML_MAX_SRC; ++i) {
        const ggml_tensor * s = op->src[i];
        if (!s) continue;
        ggml_backend_buffer_t buf = s->view_src ? s->view_src->buffer : s->buffer;
        if (buf && b200_buft_is_b200(buf->buft) && !b200_buft_is_split(buf->buft) && b200_buft_device(buf->buft) != device) return false;
        if (buf && b200_buft_is_split(buf->buft) && op->op != GGML_OP_MUL_MAT) return false;
    }
    exec_state st;   // validation mode
    b200_context probe_ctx; probe_ctx.device = device;      // only the device id is read in validation mode
    st.ctx = &probe_ctx;
    return run_node(st, op) == B200_OK;
}

// =====================================================================================
// graph executor: fusion matcher + CUDA-graph capture/replay
// =====================================================================================
namespace {

struct node_key {
    const void * node; const void * data; int op; int32_t flags;
    int64_t ne[4]; size_t nb[4];
    const void * src[GGML_MAX_SRC]; const void * src_data[GGML_MAX_SRC];
    int32_t op_params[GGML_MAX_OP_PARAMS / sizeof(int32_t)];
};

void make_key(const ggml_tensor * n, node_key & k) {
    memset(&k, 0, sizeof(k));
    k.node = n; k.data = n->data; k.op = (int) n->op; k.flags = n->flags;
    for (int i = 0; i < 4; ++i) { k.ne[i] = n->ne[i]; k.nb[i] = n->nb[i]; }
    for (int i = 0; i < GGML_MAX_SRC; ++i) { k.src[i] = n->src[i]; k.src_data[i] = n->src[i] ? n->src[i]->data : nullptr; }
    memcpy(k.op_params, n->op_params, sizeof(k.op_params));
}

// the per-token check (a graph is resubmitted unchanged ~1000 nodes x every token): field compares, no copies
bool key_matches(const ggml_tensor * n, const node_key & k) {
    if (k.node != n || k.data != n->data || k.op != (int) n->op || k.flags != n->flags) return false;
    for (int i = 0; i < 4; ++i) if (k.ne[i] != n->ne[i] || k.nb[i] != n->nb[i]) return false;
    for (int i = 0; i < GGML_MAX_SRC; ++i) {
        const ggml_tensor * s = n->src[i];
        if (k.src[i] != s) return false;
        if (s && k.src_data[i] != s->data) return false;
    }
    return memcmp(k.op_params, n->op_params, sizeof(k.op_params)) == 0;
}

}  // namespace

struct b200_chain_entry { b200_chain * chain = nullptr; float * scratch = nullptr; int n_phases = 0; };

struct b200_graph_cache {
    b200_plan plan;                    // the graph `keys` describes, compiled into chain phases + plain nodes
    bool plan_built = false;
    std::vector<b200_chain_entry> chains;
    char * redirect_buf = nullptr;     // plan-private tensors (b200_plan::redirect_bytes)
    void drop_chains() {
        for (b200_chain_entry & e : chains) { b200_chain_destroy(e.chain); if (e.scratch) cudaFree(e.scratch); }
        chains.clear(); plan_built = false;
        if (redirect_buf) { cudaFree(redirect_buf); redirect_buf = nullptr; }
    }
    bool has_split = false;            // the graph multiplies row-split weights: no chain, no CUDA-graph capture (like the
                                       // reference, ggml-cuda.cu:2883) — the slices run on other devices' streams
    std::vector<node_key> keys;        // properties of the graph last seen
    cudaGraphExec_t exec = nullptr;    // instantiated capture of `keys` (or null)
    uint64_t kernels = 0;              // kernel nodes in `exec`
    int seen = 0;                      // consecutive identical submissions
};

namespace {

bool same_shape(const ggml_tensor * a, const ggml_tensor * b) { return ggml_are_same_shape(a, b); }

// returns the number of graph nodes consumed starting at i (>= 1)
int run_fused(exec_state & st, ggml_cgraph * g, int i, bool fusion, int node_limit) {
    ggml_tensor * n = g->nodes[i];
    if (fusion && !st.validate) {
        // {RMS_NORM, MUL[, ADD]}  (reference fusion: ggml-cuda.cu:3844-3854)
        if (n->op == GGML_OP_RMS_NORM) {
            const ggml_op ops3[] = {GGML_OP_RMS_NORM, GGML_OP_MUL, GGML_OP_ADD};
            int cnt = (i + 3 <= node_limit && ggml_can_fuse(g, i, ops3, 3)) ? 3 : (i + 2 <= node_limit && ggml_can_fuse(g, i, ops3, 2)) ? 2 : 1;
            while (cnt > 1) {
                const ggml_tensor * mul = g->nodes[i + 1];
                const ggml_tensor * w = mul->src[0] == n ? mul->src[1] : mul->src[0];
                const ggml_tensor * add = cnt == 3 ? g->nodes[i + 2] : nullptr;
                const ggml_tensor * b = add ? (add->src[0] == mul ? add->src[1] : add->src[0]) : nullptr;
                const ggml_tensor * out = add ? add : mul;
                b200_tensor tx = b200_desc(n->src[0]), tw = b200_desc(w), td = b200_desc(out), tb; if (b) tb = b200_desc(b);
                float eps; memcpy(&eps, n->op_params, 4);
                if (b200_rms_norm(&tx, &tw, b ? &tb : nullptr, &td, eps, B200_STREAM_VALIDATE) == B200_OK) {
                    if (!st.sizing) { B200_KCHECK(b200_rms_norm(&tx, &tw, b ? &tb : nullptr, &td, eps, st.stream)); st.launches++; }
                    return cnt;
                }
                --cnt;
            }
        }
        // {MUL_MAT(gate), MUL_MAT(up), GLU} on one activation column (reference: ggml-cuda.cu:3737-3771)
        if (n->op == GGML_OP_MUL_MAT && !is_split_tensor(n->src[0]) && i + 2 < node_limit && n->ne[1] == 1 && n->ne[2] == 1 && n->ne[3] == 1) {
            ggml_tensor * m2 = g->nodes[i + 1]; ggml_tensor * glu = g->nodes[i + 2];
            if (m2->op == GGML_OP_MUL_MAT && glu->op == GGML_OP_GLU && m2->src[1] == n->src[1] && glu->src[0] && glu->src[1] &&
                ((glu->src[0] == n && glu->src[1] == m2) || (glu->src[0] == m2 && glu->src[1] == n)) &&
                ggml_node_has_n_uses(g, i, 1) && ggml_node_has_n_uses(g, i + 1, 1) && glu->op_params[1] == 0 &&
                same_shape(n, m2) && same_shape(n, glu) && n->src[0]->type == m2->src[0]->type) {
                const ggml_tensor * gate_mm = glu->src[0]; const ggml_tensor * up_mm = glu->src[1];
                float alpha, limit; memcpy(&alpha, (const float *) glu->op_params + 2, 4); memcpy(&limit, (const float *) glu->op_params + 3, 4);
                // run as: dst(glu) = act(gate_w . x) * (up_w . x)
                ggml_tensor tmp = *glu;            // shallow copy describing the fused output
                tmp.op = GGML_OP_MUL_MAT; tmp.src[0] = up_mm->src[0]; tmp.src[1] = up_mm->src[1]; tmp.src[2] = nullptr;
                exec_state probe = st; probe.validate = true;
                if (run_mul_mat(probe, &tmp, nullptr, gate_mm->src[0], nullptr, (int) ggml_get_glu_op(glu), alpha, limit) == B200_OK) {
                    run_mul_mat(st, &tmp, nullptr, gate_mm->src[0], nullptr, (int) ggml_get_glu_op(glu), alpha, limit);
                    if (st.q8_src == &tmp) st.q8_src = nullptr;
                    return 3;
                }
            }
        }
        // {MUL_MAT(float weights), ADD(per-row bias)}: the MoE router (ffn_gate_inp + bias)
        if (n->op == GGML_OP_MUL_MAT && i + 1 < node_limit && !is_split_tensor(n->src[0]) && n->ne[2] == 1 && n->ne[3] == 1 &&
            (n->src[0]->type == GGML_TYPE_F32 || n->src[0]->type == GGML_TYPE_F16 || n->src[0]->type == GGML_TYPE_BF16)) {
            ggml_tensor * add = g->nodes[i + 1];
            if (add->op == GGML_OP_ADD && (add->src[0] == n || add->src[1] == n) && ggml_node_has_n_uses(g, i, 1) && same_shape(add, n)) {
                const ggml_tensor * other = add->src[0] == n ? add->src[1] : add->src[0];
                if (other != n && other->type == GGML_TYPE_F32 && other->ne[0] == n->ne[0] && ggml_nelements(other) == n->ne[0] && other->nb[0] == 4) {
                    b200_tensor tw = b200_desc(n->src[0]), tx = b200_desc(n->src[1]), tb = b200_desc(other), td = b200_desc(add);
                    if (b200_mul_mat_f(&tw, &tx, &tb, &td, B200_STREAM_VALIDATE) == B200_OK) {
                        if (!st.sizing) { B200_KCHECK(b200_mul_mat_f(&tw, &tx, &tb, &td, st.stream)); st.launches++; }
                        return 2;
                    }
                }
            }
        }
        // MoE routing with softmax-weight gating (gpt-oss): {ARGSORT desc, VIEW(first n_used), GET_ROWS(probs, selected), RESHAPE,
        // SOFT_MAX} -> one warp-per-token kernel (reference fusion: ggml_cuda_topk_moe_fusion, ggml-cuda.cu:3122, topk-moe.cu:314)
        if (n->op == GGML_OP_ARGSORT && i + 4 < node_limit && n->op_params[0] == GGML_SORT_ORDER_DESC) {
            const ggml_tensor * probs = n->src[0];
            ggml_tensor * view = g->nodes[i + 1]; ggml_tensor * gr = g->nodes[i + 2]; ggml_tensor * rs = g->nodes[i + 3]; ggml_tensor * sm = g->nodes[i + 4];
            float sm_scale = 0.0f, sm_bias = 1.0f;
            if (sm->op == GGML_OP_SOFT_MAX) { memcpy(&sm_scale, (const float *) sm->op_params + 0, 4); memcpy(&sm_bias, (const float *) sm->op_params + 1, 4); }
            if (view->op == GGML_OP_VIEW && view->src[0] == n && view->view_offs == 0 && view->ne[1] == n->ne[1] && view->ne[2] == 1 && view->ne[3] == 1 &&
                view->nb[1] == n->nb[1] && gr->op == GGML_OP_GET_ROWS && gr->src[1] == view && gr->src[0]->data == probs->data &&
                gr->src[0]->type == GGML_TYPE_F32 && gr->src[0]->ne[0] == 1 && gr->src[0]->ne[1] == probs->ne[0] && gr->src[0]->ne[2] == probs->ne[1] &&
                gr->src[0]->nb[1] == 4 && gr->src[0]->nb[2] == probs->nb[1] &&
                rs->op == GGML_OP_RESHAPE && rs->src[0] == gr && rs->ne[0] == view->ne[0] && rs->ne[1] == view->ne[1] &&
                sm->op == GGML_OP_SOFT_MAX && sm->src[0] == rs && !sm->src[1] && !sm->src[2] && sm_scale == 1.0f && sm_bias == 0.0f &&
                probs->ne[2] == 1 && probs->ne[3] == 1 && ggml_is_contiguous(sm) &&
                ggml_node_has_n_uses(g, i, 1) && ggml_node_has_n_uses(g, i + 2, 1) && ggml_node_get_use_count(g, i + 3) == 1 &&
                !(rs->flags & GGML_TENSOR_FLAG_OUTPUT)) {       // (the RESHAPE is a view: ggml_node_has_n_uses declines views)
                b200_tensor tl = b200_desc(probs), ti = b200_desc(view), tw = b200_desc(sm);
                if (b200_topk_moe(&tl, &ti, &tw, B200_STREAM_VALIDATE) == B200_OK) {
                    if (!st.sizing) { B200_KCHECK(b200_topk_moe(&tl, &ti, &tw, st.stream)); st.launches++; }
                    return 5;
                }
            }
        }
        // weighted sum over the selected experts: {MUL(experts, weights), VIEW x n_used, ADD x (n_used - 1)} (build_moe_ffn)
        if (n->op == GGML_OP_MUL && n->type == GGML_TYPE_F32 && n->src[0] && n->src[1] && n->ne[1] >= 2 && n->ne[1] <= 64 && n->ne[3] == 1 &&
            n->src[1]->ne[0] == 1 && n->src[1]->ne[1] == n->ne[1] && n->src[1]->ne[2] == n->ne[2] && ggml_are_same_shape(n, n->src[0])) {
            const int nu = (int) n->ne[1];
            if (i + 2 * nu - 1 < node_limit) {
                // the graph lists them in build order of the left fold: v0, v1, ADD, v2, ADD, ...
                bool ok = ggml_node_has_n_uses(g, i, nu);
                const ggml_tensor * views[64] = {nullptr};
                const ggml_tensor * acc = nullptr;
                int next = 1, nviews = 0;
                for (int j = i + 1; ok && j < i + 2 * nu; ++j) {
                    const ggml_tensor * t = g->nodes[j];
                    if (t->op == GGML_OP_VIEW) {
                        const size_t e = n->nb[1] ? t->view_offs / n->nb[1] : 0;
                        ok = t->src[0] == n && e < (size_t) nu && !views[e] && t->view_offs == e * n->nb[1] && t->ne[0] == n->ne[0] && t->ne[1] == n->ne[2] &&
                             t->ne[2] == 1 && t->ne[3] == 1 && t->nb[1] == n->nb[2] && ggml_node_get_use_count(g, j) == 1 && !(t->flags & GGML_TENSOR_FLAG_OUTPUT);
                        if (ok) { views[e] = t; ++nviews; }
                    } else if (t->op == GGML_OP_ADD) {
                        const ggml_tensor * lhs = acc ? acc : views[0];
                        ok = lhs && next < nu && views[next] && t->src[0] == lhs && t->src[1] == views[next] && (next == nu - 1 || ggml_node_has_n_uses(g, j, 1));
                        acc = t; ++next;
                    } else ok = false;
                }
                ok = ok && nviews == nu && next == nu && acc;
                if (ok) {
                    // [+ ADD(sum, residual stream)] right behind the fold
                    const ggml_tensor * out = acc; const ggml_tensor * res = nullptr;
                    int consumed = 2 * nu;
                    if (i + consumed < node_limit && ggml_node_has_n_uses(g, i + consumed - 1, 1)) {
                        const ggml_tensor * add = g->nodes[i + consumed];
                        const ggml_tensor * other = add->op == GGML_OP_ADD ? (add->src[0] == acc ? add->src[1] : add->src[1] == acc ? add->src[0] : nullptr) : nullptr;
                        if (other && other != acc && other->type == GGML_TYPE_F32 && ggml_are_same_shape(other, acc) && ggml_are_same_shape(add, acc) && other->nb[0] == 4) { out = add; res = other; ++consumed; }
                    }
                    b200_tensor tx = b200_desc(n->src[0]), tw = b200_desc(n->src[1]), td = b200_desc(out), tr; if (res) tr = b200_desc(res);
                    if (b200_moe_weighted_sum(&tx, &tw, res ? &tr : nullptr, &td, B200_STREAM_VALIDATE) == B200_OK) {
                        if (!st.sizing) { B200_KCHECK(b200_moe_weighted_sum(&tx, &tw, res ? &tr : nullptr, &td, st.stream)); st.launches++; }
                        return consumed;
                    }
                }
            }
        }
        // MoE expert matmuls (reference fusions: ggml-cuda.cu:3783-3837; graph: src/llama-graph.cpp build_moe_ffn)
        if (n->op == GGML_OP_MUL_MAT_ID && !is_split_tensor(n->src[0])) {
            const ggml_tensor * ids = n->src[2];
            // {MUL_MAT_ID(up), ADD_ID(up_b), MUL_MAT_ID(gate), ADD_ID(gate_b), GLU}: one GEMV launch reads the up row and the gate
            // row of an expert, adds both expert biases and applies the GLU (gpt-oss: SWIGLU_OAI)
            if (i + 4 < node_limit) {
                ggml_tensor * a1 = g->nodes[i + 1]; ggml_tensor * m2 = g->nodes[i + 2]; ggml_tensor * a2 = g->nodes[i + 3]; ggml_tensor * glu = g->nodes[i + 4];
                if (a1->op == GGML_OP_ADD_ID && a1->src[0] == n && a1->src[2] == ids &&
                    m2->op == GGML_OP_MUL_MAT_ID && m2->src[1] == n->src[1] && m2->src[2] == ids && m2->src[0]->type == n->src[0]->type &&
                    !is_split_tensor(m2->src[0]) && a2->op == GGML_OP_ADD_ID && a2->src[0] == m2 && a2->src[2] == ids &&
                    glu->op == GGML_OP_GLU && glu->src[0] && glu->src[1] && glu->op_params[1] == 0 &&
                    ((glu->src[0] == a2 && glu->src[1] == a1) || (glu->src[0] == a1 && glu->src[1] == a2)) &&
                    ggml_node_has_n_uses(g, i, 1) && ggml_node_has_n_uses(g, i + 1, 1) && ggml_node_has_n_uses(g, i + 2, 1) && ggml_node_has_n_uses(g, i + 3, 1) &&
                    same_shape(n, m2) && same_shape(n, a1) && same_shape(n, a2) && same_shape(n, glu) && same_shape(n->src[0], m2->src[0])) {
                    const bool gate_second = glu->src[0] == a2;                 // glu(gate, up): src[0] is the activated operand
                    const ggml_tensor * gate_mm = gate_second ? m2 : n;  const ggml_tensor * up_mm = gate_second ? n : m2;
                    const ggml_tensor * gate_b = gate_second ? a2->src[1] : a1->src[1];
                    const ggml_tensor * up_b = gate_second ? a1->src[1] : a2->src[1];
                    float alpha, limit; memcpy(&alpha, (const float *) glu->op_params + 2, 4); memcpy(&limit, (const float *) glu->op_params + 3, 4);
                    ggml_tensor tmp = *glu;
                    tmp.op = GGML_OP_MUL_MAT_ID; tmp.src[0] = up_mm->src[0]; tmp.src[1] = n->src[1]; tmp.src[2] = const_cast<ggml_tensor *>(ids);
                    exec_state probe = st; probe.validate = true;
                    if (run_mul_mat(probe, &tmp, up_b, gate_mm->src[0], gate_b, (int) ggml_get_glu_op(glu), alpha, limit) == B200_OK) {
                        run_mul_mat(st, &tmp, up_b, gate_mm->src[0], gate_b, (int) ggml_get_glu_op(glu), alpha, limit);
                        return 5;
                    }
                }
            }
            // {MUL_MAT_ID, ADD_ID}: the expert bias rides in the GEMV epilogue
            if (i + 1 < node_limit) {
                ggml_tensor * a1 = g->nodes[i + 1];
                if (a1->op == GGML_OP_ADD_ID && a1->src[0] == n && a1->src[2] == ids && ggml_node_has_n_uses(g, i, 1) && same_shape(a1, n)) {
                    ggml_tensor tmp = *a1;
                    tmp.op = GGML_OP_MUL_MAT_ID; tmp.src[0] = n->src[0]; tmp.src[1] = n->src[1]; tmp.src[2] = const_cast<ggml_tensor *>(ids);
                    exec_state probe = st; probe.validate = true;
                    if (run_mul_mat(probe, &tmp, a1->src[1], nullptr, nullptr, B200_GLU_NONE, 0, 0) == B200_OK) {
                        run_mul_mat(st, &tmp, a1->src[1], nullptr, nullptr, B200_GLU_NONE, 0, 0);
                        return 2;
                    }
                }
            }
        }
        // {MUL_MAT, ADD} with many columns: the residual stream rides in the epilogue of the tensor-core GEMM
        if (n->op == GGML_OP_MUL_MAT && !is_split_tensor(n->src[0]) && i + 1 < node_limit && n->ne[1] > 8 && n->ne[2] == 1 && n->ne[3] == 1) {
            ggml_tensor * add = g->nodes[i + 1];
            if (add->op == GGML_OP_ADD && (add->src[0] == n || add->src[1] == n) && ggml_node_has_n_uses(g, i, 1) && same_shape(add, n)) {
                const ggml_tensor * other = add->src[0] == n ? add->src[1] : add->src[0];
                if (other != n && other->type == GGML_TYPE_F32 && same_shape(other, n) && other->nb[0] == 4) {
                    ggml_tensor tmp = *add;
                    tmp.op = GGML_OP_MUL_MAT; tmp.src[0] = n->src[0]; tmp.src[1] = n->src[1]; tmp.src[2] = nullptr;
                    exec_state probe = st; probe.validate = true;
                    if (run_mul_mat(probe, &tmp, other, nullptr, nullptr, B200_GLU_NONE, 0, 0) == B200_OK) {
                        run_mul_mat(st, &tmp, other, nullptr, nullptr, B200_GLU_NONE, 0, 0);
                        return 2;
                    }
                }
            }
        }
        // {MUL_MAT, ADD} with a single column: the addend (bias or residual stream) rides in the GEMV epilogue
        if (n->op == GGML_OP_MUL_MAT && !is_split_tensor(n->src[0]) && i + 1 < node_limit && n->ne[1] == 1 && n->ne[2] == 1 && n->ne[3] == 1) {
            ggml_tensor * add = g->nodes[i + 1];
            if (add->op == GGML_OP_ADD && (add->src[0] == n || add->src[1] == n) && ggml_node_has_n_uses(g, i, 1) && same_shape(add, n)) {
                const ggml_tensor * other = add->src[0] == n ? add->src[1] : add->src[0];
                if (other->type == GGML_TYPE_F32 && other->ne[0] == n->ne[0] && ggml_nelements(other) == n->ne[0] && other->nb[0] == 4) {
                    ggml_tensor tmp = *add;
                    tmp.op = GGML_OP_MUL_MAT; tmp.src[0] = n->src[0]; tmp.src[1] = n->src[1]; tmp.src[2] = nullptr;
                    exec_state probe = st; probe.validate = true;
                    if (run_mul_mat(probe, &tmp, other, nullptr, nullptr, B200_GLU_NONE, 0, 0) == B200_OK) {
                        run_mul_mat(st, &tmp, other, nullptr, nullptr, B200_GLU_NONE, 0, 0);
                        return 2;
                    }
                }
            }
        }
    }
    const int rc = run_node(st, n);
    if (rc != B200_OK) {
        GGML_LOG_ERROR("ggml-b200: node %d (%s, %s) is not supported but reached graph_compute\n", i, ggml_op_name(n->op), n->name);
        GGML_ABORT("ggml-b200: unsupported node in graph_compute (supports_op must have declined it)");
    }
    return 1;
}

void walk_graph(exec_state & st, ggml_cgraph * g, bool fusion) {
    for (int i = 0; i < g->n_nodes;) {
        ggml_tensor * n = g->nodes[i];
        if (ggml_is_empty(n) || (n->flags & GGML_TENSOR_FLAG_COMPUTE) == 0) { ++i; continue; }
        i += run_fused(st, g, i, fusion, g->n_nodes);
    }
}

// one chain = a maximal run of consecutive phases of the plan; built on first use, replayed afterwards
void flush_chain(exec_state & st, b200_graph_cache & c, std::vector<b200_plan_item *> & pending, size_t & seq) {
    if (pending.empty()) return;
    if (!st.sizing) {
        if (seq >= c.chains.size()) {
            b200_chain_entry e;
            if (c.plan.scratch_floats > 0) B200_CUDA_CHECK(cudaMalloc((void **) &e.scratch, (size_t) c.plan.scratch_floats * sizeof(float)));
            std::vector<b200_phase> ph;
            std::deque<b200_tensor> fixed;                     // biases whose data pointer is a plan-private address
            auto fix = [&](void *& p) { if (b200_is_redirect(p)) p = c.redirect_buf + ((uintptr_t) p - B200_REDIRECT_VBASE); };
            for (b200_plan_item * it : pending) {
                b200_phase p = it->phase;
                if (p.kind == B200_PHASE_GEMV) {
                    fix(p.gemv.x.data);
                    for (int s = 0; s < p.gemv.n_seg; ++s) {
                        fix(p.gemv.seg[s].dst.data);
                        if (p.gemv.seg[s].bias && b200_is_redirect(p.gemv.seg[s].bias->data)) {
                            fixed.push_back(*p.gemv.seg[s].bias); fix(fixed.back().data); p.gemv.seg[s].bias = &fixed.back();
                        }
                    }
                }
                if (it->scratch_rel) {
                    char * base = (char *) e.scratch;
                    if (p.kind == B200_PHASE_GEMV) { for (int s = 0; s < p.gemv.n_seg; ++s) p.gemv.seg[s].dst.data = base + (intptr_t) p.gemv.seg[s].dst.data; }
                    else { p.attn.q.data = base + (intptr_t) p.attn.q.data; p.attn.k_new.data = base + (intptr_t) p.attn.k_new.data; p.attn.v_new.data = base + (intptr_t) p.attn.v_new.data; }
                }
                ph.push_back(p);
            }
            e.chain = b200_chain_create(ph.data(), (int) ph.size(), 0);
            if (!e.chain) b200_abort(__FILE__, __LINE__, "b200_chain_create", b200_last_error());
            e.n_phases = (int) ph.size();
            c.chains.push_back(e);
        }
        B200_KCHECK(b200_chain_launch(c.chains[seq].chain, st.stream));
        st.launches++; st.chain_phases += (uint64_t) c.chains[seq].n_phases;
    }
    ++seq;
    pending.clear();
}

void walk_plan(exec_state & st, ggml_cgraph * g, b200_graph_cache & c, bool fusion) {
    std::vector<b200_plan_item *> pending;
    size_t seq = 0;
    int next_node = 0;
    std::vector<b200_plan_item> & items = c.plan.items;
    for (size_t k = 0; k < items.size(); ++k) {
        b200_plan_item & it = items[k];
        if (it.kind == B200_ITEM_PHASE) { pending.push_back(&it); continue; }
        if (it.hoisted) {      // runs alone, out of index order (n_nodes > 1: a fusable run such as RMS_NORM + MUL)
            flush_chain(st, c, pending, seq);
            for (int n0 = it.first_node, n1 = it.first_node + std::max(it.n_nodes, 1); n0 < n1;) n0 += run_fused(st, g, n0, it.n_nodes > 1, n1);
            continue;
        }
        if (it.first_node < next_node) continue;                  // consumed by a fusion of the previous plain node
        {   // nodes that launch nothing (views, reshapes, empty or non-compute tensors) must not cut the chain in two
            const ggml_tensor * n = g->nodes[it.first_node];
            if (it.n_nodes <= 1 && (n->op == GGML_OP_NONE || n->op == GGML_OP_VIEW || n->op == GGML_OP_RESHAPE || n->op == GGML_OP_PERMUTE ||
                                    n->op == GGML_OP_TRANSPOSE || ggml_is_empty(n) || (n->flags & GGML_TENSOR_FLAG_COMPUTE) == 0)) continue;
        }
        flush_chain(st, c, pending, seq);
        int limit = g->n_nodes;                                   // plain-node fusions stop where the next phase begins
        for (size_t j = k + 1; j < items.size(); ++j) if (items[j].kind == B200_ITEM_PHASE || items[j].hoisted) { limit = items[j].first_node; break; }
        next_node = it.first_node + run_fused(st, g, it.first_node, fusion, std::max(limit, it.first_node + 1));
    }
    flush_chain(st, c, pending, seq);
}

void grow(b200_context * ctx, char *& p, size_t & cur, size_t need, bool zero) {
    if (need <= cur) return;
    B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    if (p) B200_CUDA_CHECK(cudaFree(p));
    cur = (need + (1u << 20)) & ~((size_t) (1u << 20) - 1);
    B200_CUDA_CHECK(cudaMalloc((void **) &p, cur));
    if (zero) B200_CUDA_CHECK(cudaMemset(p, 0, cur));
    if (ctx->cache) {   // arena addresses are baked into captured graphs
        if (ctx->cache->exec) { cudaGraphExecDestroy(ctx->cache->exec); ctx->cache->exec = nullptr; }
        ctx->cache->seen = 0;
    }
}

}  // namespace

void b200_timing_collect(b200_context * ctx) {
    if (!ctx->timing || !ctx->ev_pending) return;
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end) == cudaSuccess) {
        std::lock_guard<std::mutex> lock(b200_stats().mu);
        b200_stats().s.device_ms_total += ms; b200_stats().s.device_ms_last = ms;
    } else cudaGetLastError();
    ctx->ev_pending = false;
}

ggml_status b200_graph_compute(b200_context * ctx, ggml_cgraph * g) {
    const auto t_host0 = std::chrono::steady_clock::now();
    b200_set_device(ctx->device);
    static const bool no_graphs = env_flag("GGML_B200_DISABLE_GRAPHS"), no_fusion = env_flag("GGML_B200_DISABLE_FUSION");
    static const bool no_chain = env_flag("GGML_B200_DISABLE_CHAIN") || no_fusion;
    if (!ctx->cache) ctx->cache = new b200_graph_cache;
    b200_graph_cache & c = *ctx->cache;

    // 1. has this exact graph been submitted before?
    bool same = (size_t) g->n_nodes == c.keys.size();
    for (int i = 0; same && i < g->n_nodes; ++i) same = key_matches(g->nodes[i], c.keys[(size_t) i]);
    if (!same) {
        c.keys.resize((size_t) g->n_nodes);
        c.has_split = false;
        for (int i = 0; i < g->n_nodes; ++i) {
            make_key(g->nodes[i], c.keys[(size_t) i]);
            c.has_split = c.has_split || (g->nodes[i]->op == GGML_OP_MUL_MAT && is_split_tensor(g->nodes[i]->src[0]));
        }
        c.seen = 0;
        if (c.exec) { cudaGraphExecDestroy(c.exec); c.exec = nullptr; }
        if (!c.chains.empty()) B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));   // a chain of the previous graph may still be running
        c.drop_chains();
    }
    c.seen++;

    if (ctx->timing) {   // device time of the previous graph must be read before its events are re-recorded
        if (ctx->ev_pending) { B200_CUDA_CHECK(cudaEventSynchronize(ctx->ev_end)); b200_timing_collect(ctx); }
        B200_CUDA_CHECK(cudaEventRecord(ctx->ev_begin, ctx->stream));
    }

    uint64_t launched = 0; bool replay = false;
    if (c.exec) {
        B200_CUDA_CHECK(cudaGraphLaunch(c.exec, ctx->stream));
        launched = c.kernels; replay = true;
    } else {
        // 2. size the scratch arenas (no launches)
        const bool use_chain = !no_chain && !c.has_split;
        if (use_chain && !c.plan_built) {
            b200_plan_build(g, c.plan); c.plan_built = true;
            if (env_flag("GGML_B200_PRINT_PLAN")) {
                int phases = 0, chains = 0, plain = 0; bool in_chain = false;
                for (const b200_plan_item & it : c.plan.items) {
                    if (it.kind == B200_ITEM_PHASE) { ++phases; if (!in_chain) { ++chains; in_chain = true; } continue; }
                    const ggml_tensor * n = g->nodes[it.first_node];
                    if (n->op == GGML_OP_NONE || n->op == GGML_OP_VIEW || n->op == GGML_OP_RESHAPE || n->op == GGML_OP_PERMUTE || n->op == GGML_OP_TRANSPOSE) continue;
                    in_chain = false; ++plain;
                    fprintf(stderr, "ggml-b200 plan: plain node %d %s %s [%lld,%lld,%lld]\n", it.first_node, ggml_op_name(n->op), n->name, (long long) n->ne[0], (long long) n->ne[1], (long long) n->ne[2]);
                }
                fprintf(stderr, "ggml-b200 plan: %d nodes -> %d phases in %d chains, %d plain compute nodes, redirect %lld B\n", g->n_nodes, phases, chains, plain, (long long) c.plan.redirect_bytes);
            }
            if (c.plan.redirect_bytes > 0) B200_CUDA_CHECK(cudaMalloc((void **) &c.redirect_buf, (size_t) c.plan.redirect_bytes));
        }
        auto walk = [&](exec_state & s) { if (!use_chain) walk_graph(s, g, !no_fusion); else walk_plan(s, g, c, true); };
        exec_state sz; sz.ctx = ctx; sz.validate = false; sz.sizing = true;
        walk(sz);
        grow(ctx, ctx->q8_scratch, ctx->q8_size, sz.need_q8, false);
        grow(ctx, ctx->fa_scratch, ctx->fa_size, sz.need_fa, true);
        grow(ctx, ctx->xb_scratch, ctx->xb_size, sz.need_xb, false);
        if (sz.need_split_x > 0) {      // row split: a stream, an event and the same scratch arenas on every other device
            if (!ctx->split_fork) B200_CUDA_CHECK(cudaEventCreateWithFlags(&ctx->split_fork, cudaEventDisableTiming));
            auto grow_on = [&](char *& p, size_t & cur, size_t need) {
                if (need <= cur) return;
                B200_CUDA_CHECK(cudaDeviceSynchronize());
                if (p) B200_CUDA_CHECK(cudaFree(p));
                cur = (need + (1u << 20)) & ~((size_t) (1u << 20) - 1);
                B200_CUDA_CHECK(cudaMalloc((void **) &p, cur));
            };
            for (int id = 0; id < b200_n_devices(); ++id) {
                if (id == ctx->device) continue;
                b200_split_dev & sd = ctx->split[id];
                b200_set_device(id);
                if (!sd.stream) {
                    B200_CUDA_CHECK(cudaStreamCreateWithFlags(&sd.stream, cudaStreamNonBlocking));
                    B200_CUDA_CHECK(cudaEventCreateWithFlags(&sd.done, cudaEventDisableTiming));
                }
                grow_on(sd.x, sd.x_size, sz.need_split_x);
                grow_on(sd.q8, sd.q8_size, sz.need_q8);
                grow_on(sd.xb, sd.xb_size, sz.need_xb);
            }
            b200_set_device(ctx->device);
        }

        // 3. run, capturing into a CUDA graph from the second identical submission on
        const bool capture = !no_graphs && !c.has_split && c.seen >= 2;
        exec_state st; st.ctx = ctx; st.validate = false; st.stream = ctx->stream;
        if (capture) B200_CUDA_CHECK(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed));
        walk(st);
        launched = st.launches;
        if (capture) {
            cudaGraph_t graph = nullptr;
            B200_CUDA_CHECK(cudaStreamEndCapture(ctx->stream, &graph));
            B200_CUDA_CHECK(cudaGraphInstantiate(&c.exec, graph, 0));
            B200_CUDA_CHECK(cudaGraphDestroy(graph));
            c.kernels = launched;
            B200_CUDA_CHECK(cudaGraphLaunch(c.exec, ctx->stream));
        }
    }
    if (ctx->timing) { B200_CUDA_CHECK(cudaEventRecord(ctx->ev_end, ctx->stream)); ctx->ev_pending = true; }
    {
        std::lock_guard<std::mutex> lock(b200_stats().mu);
        ggml_b200_stats & s = b200_stats().s;
        s.graphs_computed++; s.graphs_replayed += replay; s.kernels_launched += launched; s.nodes_last_graph = launched;
        s.host_ms_total += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host0).count();
    }
    return GGML_STATUS_SUCCESS;
}

void b200_context_free(b200_context * ctx) {
    b200_set_device(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (env_flag("GGML_B200_PRINT_STATS")) {
        b200_timing_collect(ctx);
        std::lock_guard<std::mutex> lock(b200_stats().mu);
        const ggml_b200_stats & s = b200_stats().s;
        fprintf(stderr, "ggml-b200 stats: graphs %llu (replayed %llu), kernel launches %llu, host ms in graph_compute %.2f (%.1f us/graph), device ms %.2f\n",
                s.graphs_computed, s.graphs_replayed, s.kernels_launched, s.host_ms_total, s.graphs_computed ? 1e3 * s.host_ms_total / s.graphs_computed : 0.0, s.device_ms_total);
    }
    if (ctx->cache) { if (ctx->cache->exec) cudaGraphExecDestroy(ctx->cache->exec); ctx->cache->drop_chains(); delete ctx->cache; }
    if (ctx->q8_scratch) cudaFree(ctx->q8_scratch);
    if (ctx->fa_scratch) cudaFree(ctx->fa_scratch);
    if (ctx->xb_scratch) cudaFree(ctx->xb_scratch);
    for (int id = 0; id < B200_MAX_DEVICES; ++id) {
        b200_split_dev & sd = ctx->split[id];
        if (!sd.stream) continue;
        b200_set_device(id);
        cudaStreamSynchronize(sd.stream);
        if (sd.x) cudaFree(sd.x);
        if (sd.q8) cudaFree(sd.q8);
        if (sd.xb) cudaFree(sd.xb);
        cudaEventDestroy(sd.done); cudaStreamDestroy(sd.stream);
    }
    b200_set_device(ctx->device);
    if (ctx->split_fork) cudaEventDestroy(ctx->split_fork);
    if (ctx->ev_begin) cudaEventDestroy(ctx->ev_begin);
    if (ctx->ev_end) cudaEventDestroy(ctx->ev_end);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}
