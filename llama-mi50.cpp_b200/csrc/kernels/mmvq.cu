// mmvq.cu — K4: quantized GEMV for decode MUL_MAT / MUL_MAT_ID (C-ABI b200_mul_mat_vec_q).
// Replaces ggml_cuda_mul_mat_vec_q (reference ggml-cuda/mmvq.cu:638-736).
//
//   MUL_MAT    : dst[m,n,c,s] = sum_k W[k,m,c/r2,s/r3] * x[k,n,c,s]         (n = ne11 <= 8)
//   MUL_MAT_ID : dst[m,e,t]   = sum_k W[k,m,ids[e,t]]  * x[k,e % ne11,t]    (one column per (e,t))
//
// This file holds the GENERIC kernel: any strides, batch/broadcast, ids, fused bias/gate/GLU,
// byte-granular loads, one warp per output row.  The streaming fast path for contiguous decode
// weights is a separate entry point (b200_gemv_fused / the decode chain, gemv_stream.cu).
#include "vecdot.cuh"

namespace b200 {

struct mmvq_params {
    const uint8_t * w; const uint8_t * gate_w;
    const blk_q8_1 * xq;
    float * dst;
    const int32_t * ids;
    const float * bias; const float * gate_bias;
    int64_t K, M;
    int64_t w_nb1, w_nb2, w_nb3;
    int64_t xq_col, xq_chan, xq_samp;      // strides in q8_1 blocks
    int64_t dst_col, dst_chan, dst_samp;   // strides in floats
    int64_t ids_nb1;                       // ids stride between tokens, in i32 elements
    int64_t bias_chan;                     // bias stride per expert (MUL_MAT_ID), floats
    int32_t ne11;                          // activation channels per token (MUL_MAT_ID broadcast)
    int32_t r2, r3;                        // broadcast ratios ne12/ne02, ne13/ne03
    int32_t glu_op; float glu_alpha, glu_limit;
};

__device__ __forceinline__ float apply_glu(int op, float g, float x, float alpha, float limit) {
    switch (op) {
        case B200_GLU_REGLU:  return fmaxf(g, 0.0f) * x;
        case B200_GLU_GEGLU: {
            const float c = 0.79788456080286535587989211986876f, a = 0.044715f;
            return 0.5f * g * (1.0f + tanhf(c * g * (1.0f + a * g * g))) * x;
        }
        case B200_GLU_SWIGLU: return g / (1.0f + expf(-g)) * x;
        case B200_GLU_SWIGLU_OAI: {
            g = fminf(g, limit);
            x = fmaxf(fminf(x, limit), -limit);
            return g / (1.0f + expf(-g * alpha)) * (1.0f + x);
        }
        case B200_GLU_GEGLU_ERF:   return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)) * x;
        case B200_GLU_GEGLU_QUICK: return g / (1.0f + expf(-1.702f * g)) * x;
        default: return x;
    }
}

// grid: (ceil(M/NW), channels, samples|tokens); block: (32, NW). One warp = one row, NCOLS columns.
template <int TYPE, int NCOLS, bool IDS, bool FUSED, int NW>
__global__ void __launch_bounds__(32 * NW) mmvq_generic_kernel(const mmvq_params p) {
    using F = fmt<TYPE>;
    pdl_launch_dependents();
    const int lane = threadIdx.x;
    const int64_t row = (int64_t) blockIdx.x * NW + threadIdx.y;
    const int chan = blockIdx.y, samp = blockIdx.z;
    if (row >= p.M) return;   // whole warp exits together

    pdl_wait();
    int64_t w_off, x_off, bias_off = 0;
    if (IDS) {   // chan = expert slot e, samp = token t
        const int expert = p.ids[chan + (int64_t) samp * p.ids_nb1];
        w_off = (int64_t) expert * p.w_nb2;
        x_off = (int64_t) (chan % p.ne11) * p.xq_chan + (int64_t) samp * p.xq_samp;
        bias_off = (int64_t) expert * p.bias_chan;
    } else {
        w_off = (int64_t) (chan / p.r2) * p.w_nb2 + (int64_t) (samp / p.r3) * p.w_nb3;
        x_off = (int64_t) chan * p.xq_chan + (int64_t) samp * p.xq_samp;
    }
    const uint8_t * wrow = p.w + w_off + row * p.w_nb1;
    const uint8_t * grow = FUSED && p.gate_w ? p.gate_w + w_off + row * p.w_nb1 : nullptr;
    const blk_q8_1 * x = p.xq + x_off;

    float acc[NCOLS], accg[NCOLS];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) { acc[c] = 0.0f; accg[c] = 0.0f; }

    const int64_t nunits = p.K / 32;
#pragma unroll 4                                       // a lane's units are independent: their weight bytes are requested together
    for (int64_t u = lane; u < nunits; u += 32) {
        unit_w w;
        F::unpack(wrow, u, w);
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) acc[c] += unit_dot<F::has_min>(w, x + c * p.xq_col + u);
        if (FUSED && grow) {
            F::unpack(grow, u, w);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) accg[c] += unit_dot<F::has_min>(w, x + c * p.xq_col + u);
        }
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        acc[c] = warp_sum(acc[c]);
        if (FUSED && grow) accg[c] = warp_sum(accg[c]);
    }
    if (lane == 0) {
        float * d = p.dst + (int64_t) chan * p.dst_chan + (int64_t) samp * p.dst_samp + row;
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            float r = acc[c];
            if (FUSED) {
                if (p.bias) r += p.bias[bias_off + row];
                if (grow) {
                    float g = accg[c];
                    if (p.gate_bias) g += p.gate_bias[bias_off + row];
                    r = apply_glu(p.glu_op, g, r, p.glu_alpha, p.glu_limit);   // glu(gate) * up
                }
            }
            d[c * p.dst_col] = r;
        }
    }
}

// ---------------------------------------------------------------- MUL_MAT_ID with many tokens: (expert slot, token) pairs grouped by expert
// The ids path above multiplies one column per pair: a 512-token batch of a 4-of-128 MoE layer read every expert matrix 16
// times.  Here the pairs are bucketed by expert first (reference mm_ids_helper, ggml-cuda/mmid.cu:28-116) and cut into windows
// of up to 8 pairs of ONE expert; a warp then reads a weight row once per window and multiplies it with the window's 8 activation
// columns (reference: mmq over expert_bounds, mmq.cu:160-217; here still the dp4a GEMV, 8 columns wide).  Every pair's value is
// computed independently of its window mates, so the bucket order (atomics) does not change a bit of the result.
struct moe_windows {            // device scratch filled by moe_group_kernel
    int32_t n_win; int32_t _pad[3];
    // followed by: int32 win_expert[w_max], win_start[w_max], win_count[w_max], order[n_pairs]
};
__global__ void __launch_bounds__(1024) moe_group_kernel(const int32_t * ids, int64_t ids_nb1, int n_used, int n_tok, int n_expert, int w_max, int32_t * scratch) {
    extern __shared__ int32_t s_cnt[];       // [n_expert] counts, then [n_expert] offsets, then [n_expert] fill
    int32_t * s_off = s_cnt + n_expert; int32_t * s_fill = s_off + n_expert;
    moe_windows * hdr = (moe_windows *) scratch;
    int32_t * win_expert = scratch + 4; int32_t * win_start = win_expert + w_max; int32_t * win_count = win_start + w_max; int32_t * order = win_count + w_max;
    pdl_launch_dependents();
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) { s_cnt[e] = 0; s_fill[e] = 0; }
    pdl_wait();
    __syncthreads();
    const int n_pairs = n_used * n_tok;
    for (int i = threadIdx.x; i < n_pairs; i += blockDim.x) {
        const int e = ids[(i % n_used) + (int64_t) (i / n_used) * ids_nb1];
        if (e >= 0 && e < n_expert) atomicAdd(&s_cnt[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int off = 0, nw = 0;
        for (int e = 0; e < n_expert; ++e) {
            s_off[e] = off;
            for (int c = 0; c < s_cnt[e]; c += 8) { win_expert[nw] = e; win_start[nw] = off + c; win_count[nw] = min(8, s_cnt[e] - c); ++nw; }
            off += s_cnt[e];
        }
        hdr->n_win = nw;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_pairs; i += blockDim.x) {
        const int e = ids[(i % n_used) + (int64_t) (i / n_used) * ids_nb1];
        if (e >= 0 && e < n_expert) order[s_off[e] + atomicAdd(&s_fill[e], 1)] = i;
    }
}

// A warp owns RT rows of one window.  K is walked in passes of 32 units (lane = unit); in a pass the lane holds the q8_1 unit of
// all 8 columns in registers and runs the RT rows against them, so an activation word is loaded once per RT dot products (the
// first version reloaded it for every row: one load per dp4a, 2.9 ms per MUL_MAT_ID of a 512-token gpt-oss-120B batch).  Units
// are met in the same order as in the one-column-per-pair kernel and summed by the same warp reduction: bit-identical results.
template <int TYPE, int NW, int RT>
__global__ void __launch_bounds__(32 * NW) mmvq_grouped_kernel(const mmvq_params p, const int32_t * scratch, int w_max, int n_used) {
    using F = fmt<TYPE>;
    pdl_launch_dependents();
    const int lane = threadIdx.x;
    const int64_t row0 = ((int64_t) blockIdx.x * NW + threadIdx.y) * RT;
    pdl_wait();
    const int w = blockIdx.y;
    if (row0 >= p.M || w >= scratch[0]) return;
    const int32_t * win_expert = scratch + 4; const int32_t * win_start = win_expert + w_max; const int32_t * win_count = win_start + w_max; const int32_t * order = win_count + w_max;
    const int expert = win_expert[w], start = win_start[w], cnt = win_count[w];
    const uint8_t * wbase = p.w + (int64_t) expert * p.w_nb2 + row0 * p.w_nb1;
    int64_t xo[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int pair = order[start + min(c, cnt - 1)];
        xo[c] = (int64_t) ((pair % n_used) % p.ne11) * p.xq_chan + (int64_t) (pair / n_used) * p.xq_samp;
    }
    float acc[RT][8];
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.0f;
    const int64_t nunits = p.K / 32;
    for (int64_t u = lane; u < nunits; u += 32) {
        int xv[8][8]; float xd[8]; int su[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const blk_q8_1 * xb = p.xq + xo[c] + u;
            const int * q = (const int *) xb->qs;
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[c][i] = q[i];
            xd[c] = __low2float(xb->ds);
            su[c] = 0;
            if (F::has_min) {
#pragma unroll
                for (int i = 0; i < 8; ++i) su[c] = __dp4a(0x01010101, xv[c][i], su[c]);
            }
        }
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            if (row0 + r < p.M) {
                unit_w uw;
                F::unpack(wbase + r * p.w_nb1, u, uw);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) { s0 = __dp4a(uw.v[i], xv[c][i], s0); s1 = __dp4a(uw.v[4 + i], xv[c][4 + i], s1); }
                    float t = uw.wd[0] * (float) s0 + uw.wd[1] * (float) s1;          // same expression as unit_dot
                    if (F::has_min) t -= uw.wm * (float) su[c];
                    acc[r][c] += xd[c] * t;
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = warp_sum(acc[r][c]);
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (c < cnt) {
                const int pair = order[start + c];
                float * d = p.dst + (int64_t) (pair % n_used) * p.dst_chan + (int64_t) (pair / n_used) * p.dst_samp + row0;
#pragma unroll
                for (int r = 0; r < RT; ++r) if (row0 + r < p.M) d[r] = acc[r][c];
            }
        }
    }
}

// MUL_MAT with many columns that the tensor-core GEMM cannot take (rows that are not 16-byte multiples: K = 2880 Q8_0 / MXFP4):
// all 8-column windows in ONE launch, a warp owning RT rows of a window with the activation units in registers (as above).
template <int TYPE, int NW, int RT>
__global__ void __launch_bounds__(32 * NW) mmvq_wide_kernel(const mmvq_params p, int n_cols) {
    using F = fmt<TYPE>;
    pdl_launch_dependents();
    const int lane = threadIdx.x;
    const int64_t row0 = ((int64_t) blockIdx.x * NW + threadIdx.y) * RT;
    pdl_wait();
    const int c0 = blockIdx.y * 8, cnt = min(8, n_cols - c0);
    if (row0 >= p.M) return;
    const uint8_t * wbase = p.w + row0 * p.w_nb1;
    float acc[RT][8];
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.0f;
    const int64_t nunits = p.K / 32;
    for (int64_t u = lane; u < nunits; u += 32) {
        int xv[8][8]; float xd[8]; int su[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const blk_q8_1 * xb = p.xq + (int64_t) (c0 + min(c, cnt - 1)) * p.xq_col + u;
            const int * q = (const int *) xb->qs;
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[c][i] = q[i];
            xd[c] = __low2float(xb->ds);
            su[c] = 0;
            if (F::has_min) {
#pragma unroll
                for (int i = 0; i < 8; ++i) su[c] = __dp4a(0x01010101, xv[c][i], su[c]);
            }
        }
#pragma unroll
        for (int r = 0; r < RT; ++r) {
            if (row0 + r < p.M) {
                unit_w uw;
                F::unpack(wbase + r * p.w_nb1, u, uw);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) { s0 = __dp4a(uw.v[i], xv[c][i], s0); s1 = __dp4a(uw.v[4 + i], xv[c][4 + i], s1); }
                    float t = uw.wd[0] * (float) s0 + uw.wd[1] * (float) s1;
                    if (F::has_min) t -= uw.wm * (float) su[c];
                    acc[r][c] += xd[c] * t;
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < RT; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = warp_sum(acc[r][c]);
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
            if (c < cnt) {
                float * d = p.dst + (int64_t) (c0 + c) * p.dst_col + row0;
#pragma unroll
                for (int r = 0; r < RT; ++r) if (row0 + r < p.M) d[r] = acc[r][c];
            }
    }
}

template <int TYPE, int NCOLS>
static int launch_generic(const mmvq_params & p, bool ids, bool fused, dim3 grid_cs, cudaStream_t st) {
    constexpr int NW = 4;
    dim3 grid((unsigned) ((p.M + NW - 1) / NW), grid_cs.y, grid_cs.z), block(32, NW);
    if (ids) {
        if (NCOLS != 1) return B200_ERR_UNSUPPORTED;
        if (fused) mmvq_generic_kernel<TYPE, 1, true, true, NW><<<grid, block, 0, st>>>(p);
        else       mmvq_generic_kernel<TYPE, 1, true, false, NW><<<grid, block, 0, st>>>(p);
    } else {
        if (fused) mmvq_generic_kernel<TYPE, NCOLS, false, true, NW><<<grid, block, 0, st>>>(p);
        else       mmvq_generic_kernel<TYPE, NCOLS, false, false, NW><<<grid, block, 0, st>>>(p);
    }
    count_launch();
    return check_launch("mmvq_generic");
}

template <int TYPE>
static int launch_type(const mmvq_params & p, int ncols, bool ids, bool fused, dim3 g, cudaStream_t st) {
    switch (ncols) {
        case 1: return launch_generic<TYPE, 1>(p, ids, fused, g, st);
        case 2: return launch_generic<TYPE, 2>(p, ids, fused, g, st);
        case 3: return launch_generic<TYPE, 3>(p, ids, fused, g, st);
        case 4: return launch_generic<TYPE, 4>(p, ids, fused, g, st);
        case 5: return launch_generic<TYPE, 5>(p, ids, fused, g, st);
        case 6: return launch_generic<TYPE, 6>(p, ids, fused, g, st);
        case 7: return launch_generic<TYPE, 7>(p, ids, fused, g, st);
        case 8: return launch_generic<TYPE, 8>(p, ids, fused, g, st);
        default: return B200_ERR_UNSUPPORTED;
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_mmvq_supported(const b200_mmvq_args * a) {
    if (!a) return 0;
    const b200_tensor & w = a->w; const b200_tensor & d = a->dst;
    switch (w.type) { case B200_Q4_0: case B200_Q8_0: case B200_MXFP4: case B200_Q4_K: case B200_Q6_K: break; default: return 0; }
    const int64_t K = w.ne[0], M = w.ne[1];
    if (K <= 0 || M <= 0 || K % b200_blck_size(w.type) != 0 || K != a->x_ne[0]) return 0;
    if (w.nb[0] != (int64_t) b200_type_size(w.type) || d.type != B200_F32 || d.nb[0] != 4 || d.ne[0] != M) return 0;
    if (a->xq_ne0_padded % 32 != 0 || a->xq_ne0_padded < K) return 0;
    if (a->ids) {
        const b200_tensor & ids = *a->ids;   // [n_used, n_tokens]
        if (ids.type != B200_I32 || ids.nb[0] != 4 || ids.nb[1] % 4 != 0) return 0;
        if (d.ne[1] != ids.ne[0] || d.ne[2] != ids.ne[1] || d.ne[3] != 1 || w.ne[3] != 1) return 0;
        if (a->x_ne[2] != d.ne[2] || (a->x_ne[1] != 1 && a->x_ne[1] != d.ne[1]) || a->x_ne[3] != 1) return 0;
        if (d.ne[1] > 65535 || d.ne[2] > 65535 || a->col_begin || a->col_count) return 0;
    } else {
        if (a->x_ne[1] < 1 || d.ne[1] != a->x_ne[1]) return 0;
        if (a->col_begin < 0 || a->col_count < 0 || a->col_begin + a->col_count > a->x_ne[1]) return 0;
        if ((a->col_count ? a->col_count : a->x_ne[1]) > 8) return 0;
        if (d.ne[2] != a->x_ne[2] || d.ne[3] != a->x_ne[3]) return 0;
        if (w.ne[2] < 1 || w.ne[3] < 1 || d.ne[2] % w.ne[2] != 0 || d.ne[3] % w.ne[3] != 0) return 0;
        if (d.ne[2] > 65535 || d.ne[3] > 65535) return 0;
    }
    if (a->gate_w) {
        const b200_tensor & g = *a->gate_w;
        if (g.type != w.type) return 0;
        for (int i = 0; i < 4; ++i) if (g.ne[i] != w.ne[i] || g.nb[i] != w.nb[i]) return 0;
        if (a->glu_op < B200_GLU_REGLU || a->glu_op > B200_GLU_GEGLU_QUICK) return 0;
    }
    if (a->bias || a->gate_w) {   // fused epilogue: single column per (channel)
        if (a->ids ? d.ne[2] != 1 : d.ne[1] != 1) return 0;
        if (!a->ids && (d.ne[2] != 1 || d.ne[3] != 1)) return 0;
    }
    if (a->bias && (a->bias->type != B200_F32 || a->bias->ne[0] != M || a->bias->nb[0] != 4)) return 0;
    if (a->gate_bias && (!a->gate_w || a->gate_bias->type != B200_F32 || a->gate_bias->ne[0] != M || a->gate_bias->nb[0] != 4)) return 0;
    return 1;
}

extern "C" int b200_mul_mat_vec_q(const b200_mmvq_args * a, void * stream) {
    if (!b200_mmvq_supported(a)) { set_error("b200_mul_mat_vec_q: unsupported arguments"); return B200_ERR_UNSUPPORTED; }
    const b200_tensor & w = a->w; const b200_tensor & d = a->dst;
    if (nelements(d) == 0) return B200_OK;

    mmvq_params p{};
    p.w = (const uint8_t *) w.data; p.gate_w = a->gate_w ? (const uint8_t *) a->gate_w->data : nullptr;
    p.xq = (const blk_q8_1 *) a->xq; p.dst = (float *) d.data;
    p.ids = a->ids ? (const int32_t *) a->ids->data : nullptr;
    p.bias = a->bias ? (const float *) a->bias->data : nullptr;
    p.gate_bias = a->gate_bias ? (const float *) a->gate_bias->data : nullptr;
    p.K = w.ne[0]; p.M = w.ne[1];
    p.w_nb1 = w.nb[1]; p.w_nb2 = w.nb[2]; p.w_nb3 = w.nb[3];
    const int64_t bpr = a->xq_ne0_padded / 32;   // q8_1 layout: [bpr, ne11, ne12, ne13]
    const int64_t x1 = bpr, x2 = bpr * a->x_ne[1], x3 = x2 * a->x_ne[2];
    p.glu_op = a->glu_op; p.glu_alpha = a->glu_alpha; p.glu_limit = a->glu_limit;
    const bool fused = a->bias || a->gate_w;
    int ncols; dim3 g(1, 1, 1);
    if (a->ids) {
        ncols = 1;
        p.xq_col = 0; p.xq_chan = x1; p.xq_samp = x2;
        p.dst_col = 0; p.dst_chan = d.nb[1] / 4; p.dst_samp = d.nb[2] / 4;
        p.ids_nb1 = a->ids->nb[1] / 4; p.ne11 = (int32_t) a->x_ne[1];
        p.bias_chan = a->bias ? a->bias->nb[1] / 4 : (a->gate_bias ? a->gate_bias->nb[1] / 4 : 0);
        p.r2 = p.r3 = 1;
        g.y = (unsigned) d.ne[1]; g.z = (unsigned) d.ne[2];
    } else {
        ncols = a->col_count ? a->col_count : (int) a->x_ne[1];
        p.xq += (int64_t) a->col_begin * x1;
        p.dst += (int64_t) a->col_begin * (d.nb[1] / 4);
        p.xq_col = x1; p.xq_chan = x2; p.xq_samp = x3;
        p.dst_col = d.nb[1] / 4; p.dst_chan = d.nb[2] / 4; p.dst_samp = d.nb[3] / 4;
        p.r2 = (int32_t) (d.ne[2] / w.ne[2]); p.r3 = (int32_t) (d.ne[3] / w.ne[3]); p.ne11 = 1;
        g.y = (unsigned) d.ne[2]; g.z = (unsigned) d.ne[3];
    }
    cudaStream_t st = (cudaStream_t) stream;
    switch (w.type) {
        case B200_Q4_0:  return launch_type<B200_Q4_0>(p, ncols, a->ids != nullptr, fused, g, st);
        case B200_Q8_0:  return launch_type<B200_Q8_0>(p, ncols, a->ids != nullptr, fused, g, st);
        case B200_MXFP4: return launch_type<B200_MXFP4>(p, ncols, a->ids != nullptr, fused, g, st);
        case B200_Q4_K:  return launch_type<B200_Q4_K>(p, ncols, a->ids != nullptr, fused, g, st);
        case B200_Q6_K:  return launch_type<B200_Q6_K>(p, ncols, a->ids != nullptr, fused, g, st);
        default: return B200_ERR_UNSUPPORTED;
    }
}

// ---- MUL_MAT_ID, pairs grouped by expert (prefill).  Same arguments as b200_mul_mat_vec_q with ids set and no fused epilogue;
// scratch: b200_mul_mat_id_grouped_scratch_size() bytes of device memory, 16-byte aligned.
static int64_t moe_w_max(const b200_mmvq_args * a) { return (a->dst.ne[1] * a->dst.ne[2] + 7) / 8 + a->w.ne[2]; }
extern "C" size_t b200_mul_mat_id_grouped_scratch_size(const b200_mmvq_args * a) {
    if (!a || !a->ids) return 0;
    return (size_t) (16 + 4 * (3 * moe_w_max(a) + a->dst.ne[1] * a->dst.ne[2]));
}
extern "C" int b200_mul_mat_id_grouped(const b200_mmvq_args * a, void * scratch, size_t scratch_size, void * stream) {
    if (!b200_mmvq_supported(a) || !a->ids || a->bias || a->gate_w || a->gate_bias) return B200_ERR_UNSUPPORTED;
    const b200_tensor & w = a->w; const b200_tensor & d = a->dst;
    const int64_t n_used = d.ne[1], n_tok = d.ne[2], n_expert = w.ne[2], w_max = moe_w_max(a);
    if (n_expert < 1 || n_expert > 4096 || n_used * n_tok > (1 << 24) || w_max > 65535) return B200_ERR_UNSUPPORTED;
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    if (!scratch || ((uintptr_t) scratch) % 16 || scratch_size < b200_mul_mat_id_grouped_scratch_size(a)) { set_error("mul_mat_id_grouped: scratch too small"); return B200_ERR_ARG; }
    if (nelements(d) == 0) return B200_OK;
    mmvq_params p{};
    p.w = (const uint8_t *) w.data; p.xq = (const blk_q8_1 *) a->xq; p.dst = (float *) d.data;
    p.K = w.ne[0]; p.M = w.ne[1]; p.w_nb1 = w.nb[1]; p.w_nb2 = w.nb[2];
    const int64_t bpr = a->xq_ne0_padded / 32;
    p.xq_chan = bpr; p.xq_samp = bpr * a->x_ne[1]; p.dst_chan = d.nb[1] / 4; p.dst_samp = d.nb[2] / 4; p.ne11 = (int32_t) a->x_ne[1];
    cudaStream_t st = (cudaStream_t) stream;
    moe_group_kernel<<<1, 1024, (size_t) 3 * n_expert * 4, st>>>((const int32_t *) a->ids->data, a->ids->nb[1] / 4, (int) n_used, (int) n_tok, (int) n_expert, (int) w_max, (int32_t *) scratch);
    count_launch();
    constexpr int NW = 4, RT = 4;
    dim3 grid((unsigned) ((p.M + NW * RT - 1) / (NW * RT)), (unsigned) w_max), block(32, NW);
    switch (w.type) {
        case B200_Q4_0:  mmvq_grouped_kernel<B200_Q4_0, NW, RT><<<grid, block, 0, st>>>(p, (const int32_t *) scratch, (int) w_max, (int) n_used); break;
        case B200_Q8_0:  mmvq_grouped_kernel<B200_Q8_0, NW, RT><<<grid, block, 0, st>>>(p, (const int32_t *) scratch, (int) w_max, (int) n_used); break;
        case B200_MXFP4: mmvq_grouped_kernel<B200_MXFP4, NW, RT><<<grid, block, 0, st>>>(p, (const int32_t *) scratch, (int) w_max, (int) n_used); break;
        case B200_Q4_K:  mmvq_grouped_kernel<B200_Q4_K, NW, RT><<<grid, block, 0, st>>>(p, (const int32_t *) scratch, (int) w_max, (int) n_used); break;
        case B200_Q6_K:  mmvq_grouped_kernel<B200_Q6_K, NW, RT><<<grid, block, 0, st>>>(p, (const int32_t *) scratch, (int) w_max, (int) n_used); break;
        default: return B200_ERR_UNSUPPORTED;
    }
    count_launch();
    return check_launch("mmvq_grouped");
}

// ---- MUL_MAT, 2-D weights, any number of columns, no epilogue: every 8-column window in one launch (the fallback of the
// tensor-core GEMM for rows that are not 16-byte multiples).  `a` as for b200_mul_mat_vec_q without ids / bias / gate and without
// a column window; x_ne[2] == x_ne[3] == 1.
extern "C" int b200_mul_mat_vec_q_wide(const b200_mmvq_args * a, void * stream) {
    if (!a || a->ids || a->bias || a->gate_w || a->gate_bias || a->col_begin || a->col_count) return B200_ERR_UNSUPPORTED;
    const b200_tensor & w = a->w; const b200_tensor & d = a->dst;
    switch (w.type) { case B200_Q4_0: case B200_Q8_0: case B200_MXFP4: case B200_Q4_K: case B200_Q6_K: break; default: return B200_ERR_UNSUPPORTED; }
    const int64_t K = w.ne[0], M = w.ne[1], N = a->x_ne[1];
    if (K <= 0 || M <= 0 || N < 1 || K % b200_blck_size(w.type) != 0 || K != a->x_ne[0] || w.ne[2] != 1 || w.ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (a->x_ne[2] != 1 || a->x_ne[3] != 1 || w.nb[0] != (int64_t) b200_type_size(w.type) || d.type != B200_F32 || d.nb[0] != 4) return B200_ERR_UNSUPPORTED;
    if (d.ne[0] != M || d.ne[1] != N || d.ne[2] != 1 || d.ne[3] != 1 || a->xq_ne0_padded % 32 != 0 || a->xq_ne0_padded < K || (N + 7) / 8 > 65535) return B200_ERR_UNSUPPORTED;
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    mmvq_params p{};
    p.w = (const uint8_t *) w.data; p.xq = (const blk_q8_1 *) a->xq; p.dst = (float *) d.data;
    p.K = K; p.M = M; p.w_nb1 = w.nb[1]; p.xq_col = a->xq_ne0_padded / 32; p.dst_col = d.nb[1] / 4;
    constexpr int NW = 4, RT = 4;
    dim3 grid((unsigned) ((M + NW * RT - 1) / (NW * RT)), (unsigned) ((N + 7) / 8)), block(32, NW);
    cudaStream_t st = (cudaStream_t) stream;
    switch (w.type) {
        case B200_Q4_0:  mmvq_wide_kernel<B200_Q4_0, NW, RT><<<grid, block, 0, st>>>(p, (int) N); break;
        case B200_Q8_0:  mmvq_wide_kernel<B200_Q8_0, NW, RT><<<grid, block, 0, st>>>(p, (int) N); break;
        case B200_MXFP4: mmvq_wide_kernel<B200_MXFP4, NW, RT><<<grid, block, 0, st>>>(p, (int) N); break;
        case B200_Q4_K:  mmvq_wide_kernel<B200_Q4_K, NW, RT><<<grid, block, 0, st>>>(p, (int) N); break;
        default:         mmvq_wide_kernel<B200_Q6_K, NW, RT><<<grid, block, 0, st>>>(p, (int) N); break;
    }
    count_launch();
    return check_launch("mmvq_wide");
}
