// ops_moe.cu — row G glue of the MoE decode graph (gpt-oss, BASELINE.json configs[4]): expert routing and the small float
// matmul in front of it.  Replaces reference ggml-cuda/argsort.cu (bitonic sort), top-k.cu, add-id.cu (:28), sumrows.cu,
// clamp.cu and the float-weight path of ggml_cuda_mul_mat (ggml-cuda.cu:2183-2266; mmvf.cu for few columns) for the router
// matrix `ffn_gate_inp` (f32 [n_embd, n_expert]).  Graph of reference src/llama-graph.cpp build_moe_ffn:
//   logits = gate_inp . x (+ bias) ; selected = view(argsort_desc(logits), n_used) ; weights = softmax(get_rows(logits, selected))
//   up/gate/down = mul_mat_id(..., selected) (+ add_id(bias, selected)) ; out = sum_e weights[e] * down[e]
#include <cuda_bf16.h>

#include <mutex>
#include <type_traits>

#include "common.cuh"

namespace b200 {

#define VALIDATE_ONLY(stream) ((stream) == B200_STREAM_VALIDATE)

// ------------------------------------------------------------------ ARGSORT / TOP_K
// One CTA per row: indices in shared memory, bitonic network over the row padded to a power of two (padding sorts last).
// order: 0 ascending, 1 descending (ggml_sort_order).  Only the first n_out indices of the sorted row are written.
template <int ORDER>
__global__ void __launch_bounds__(1024) argsort_kernel(const float * x, int32_t * dst, int n, int n_pad, int64_t x_row, int64_t d_row, int n_out) {
    extern __shared__ int32_t s_idx[];
    float * s_val = (float *) (s_idx + n_pad);                   // the row itself: 28 compare steps at n = 128 would each go to global memory
    pdl_launch_dependents();
    const float * row = x + (int64_t) blockIdx.x * x_row;
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x) s_idx[i] = i;
    pdl_wait();
    const bool staged = n_pad <= 16384;                          // longer rows: indices alone fill the shared memory, values stay in global
    if (staged) for (int i = threadIdx.x; i < n; i += blockDim.x) s_val[i] = row[i];
    __syncthreads();
    // a < b in the target order; indices >= n (padding) are larger than everything
    auto before = [&](int a, int b) {
        if (a >= n) return false;
        if (b >= n) return true;
        const float va = staged ? s_val[a] : row[a], vb = staged ? s_val[b] : row[b];
        return ORDER == 0 ? va < vb : va > vb;
    };
    for (int k = 2; k <= n_pad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n_pad; i += blockDim.x) {
                const int l = i ^ j;
                if (l > i) {
                    const int a = s_idx[i], b = s_idx[l];
                    const bool up = (i & k) == 0;           // this pair sorts in the target order
                    if (up ? before(b, a) : before(a, b)) { s_idx[i] = b; s_idx[l] = a; }
                }
            }
            __syncthreads();
        }
    }
    int32_t * out = dst + (int64_t) blockIdx.x * d_row;
    for (int i = threadIdx.x; i < n_out; i += blockDim.x) out[i] = s_idx[i];
}

static int argsort_common(const b200_tensor * x, const b200_tensor * d, int order, int n_out, void * stream, const char * what) {
    if (!x || !d || x->type != B200_F32 || d->type != B200_I32 || x->nb[0] != 4 || d->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    if (!is_contiguous(*x) || !is_contiguous(*d) || x->ne[0] < 1 || x->ne[0] > 32768) return B200_ERR_UNSUPPORTED;
    if (d->ne[0] != n_out || d->ne[1] != x->ne[1] || d->ne[2] != x->ne[2] || d->ne[3] != x->ne[3] || n_out > x->ne[0] || n_out < 1) return B200_ERR_UNSUPPORTED;
    if (order != 0 && order != 1) return B200_ERR_UNSUPPORTED;
    const int64_t rows = nrows(*x);
    if (rows > 0x7fffffff) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (rows == 0) return B200_OK;
    int n_pad = 1; while (n_pad < x->ne[0]) n_pad <<= 1;
    const size_t smem = (size_t) n_pad * (n_pad <= 16384 ? 8 : 4);      // indices (+ values)
    static std::once_flag once[64]; int dev = 0; B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_ARG;
    cudaError_t rc = cudaSuccess;
    std::call_once(once[dev], [&] {
        rc = cudaFuncSetAttribute(argsort_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
        if (rc == cudaSuccess) rc = cudaFuncSetAttribute(argsort_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    });
    B200_CUDA_TRY(rc);
    const int threads = n_pad >= 1024 ? 1024 : (n_pad < 32 ? 32 : n_pad);
    if (order == 0) argsort_kernel<0><<<(unsigned) rows, threads, smem, (cudaStream_t) stream>>>((const float *) x->data, (int32_t *) d->data, (int) x->ne[0], n_pad, x->ne[0], n_out, n_out);
    else            argsort_kernel<1><<<(unsigned) rows, threads, smem, (cudaStream_t) stream>>>((const float *) x->data, (int32_t *) d->data, (int) x->ne[0], n_pad, x->ne[0], n_out, n_out);
    count_launch();
    return check_launch(what);
}

// ------------------------------------------------------------------ ADD_ID: dst[:, e, t] = a[:, e, t] + b[:, ids[e, t]]
__global__ void __launch_bounds__(256) add_id_kernel(const char * a, const char * b, const char * ids, char * d, int64_t ne0, int64_t ne1,
                                                     int64_t a_nb1, int64_t a_nb2, int64_t b_nb1, int64_t i_nb0, int64_t i_nb1, int64_t d_nb1, int64_t d_nb2, int vec) {
    // one element group per thread (grid.y walks the row): a decode call is 4 rows of 2880 floats, and the bias rows come from
    // HBM: one round trip instead of a dozen dependent ones
    pdl_launch_dependents();
    const int64_t e = blockIdx.x % ne1, t = blockIdx.x / ne1;
    pdl_wait();
    const int32_t id = *(const int32_t *) (ids + e * i_nb0 + t * i_nb1);
    const float * pa = (const float *) (a + e * a_nb1 + t * a_nb2);
    const float * pb = (const float *) (b + (int64_t) id * b_nb1);
    float * pd = (float *) (d + e * d_nb1 + t * d_nb2);
    const int64_t i = ((int64_t) blockIdx.y * blockDim.x + threadIdx.x) * (vec ? 4 : 1);
    if (i >= ne0) return;
    if (vec) {
        const float4 x = *(const float4 *) (pa + i), y = *(const float4 *) (pb + i);
        *(float4 *) (pd + i) = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    } else pd[i] = pa[i] + pb[i];
}

// ------------------------------------------------------------------ SUM_ROWS / CLAMP
__global__ void __launch_bounds__(256) sum_rows_kernel(const char * x, char * d, int64_t ne0, int64_t ne1, int64_t ne2,
                                                       int64_t x_nb1, int64_t x_nb2, int64_t x_nb3, int64_t d_nb1, int64_t d_nb2, int64_t d_nb3) {
    pdl_launch_dependents();
    const int64_t r = blockIdx.x, i1 = r % ne1, i2 = (r / ne1) % ne2, i3 = r / (ne1 * ne2);
    pdl_wait();
    const float * row = (const float *) (x + i1 * x_nb1 + i2 * x_nb2 + i3 * x_nb3);
    float s = 0.0f;
    for (int64_t i = threadIdx.x; i < ne0; i += blockDim.x) s += row[i];
    s = warp_sum(s);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < (int) (blockDim.x >> 5); ++w) t += part[w];
        *(float *) (d + i1 * d_nb1 + i2 * d_nb2 + i3 * d_nb3) = t;
    }
}
__global__ void __launch_bounds__(256) clamp_kernel(const float * x, float * d, float lo, float hi, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    pdl_wait();
    if (i < n) { const float v = x[i]; d[i] = v < lo ? lo : (v > hi ? hi : v); }
}

// ------------------------------------------------------------------ float-weight matmul: dst[m, n] = sum_k W[k, m] * x[k, n]
// one CTA per weight row and window of up to 8 columns (the router matrix is 1.5 MB: it lives in L2 across the windows)
template <typename T> __device__ __forceinline__ float w_as_f32(const T & v);
template <> __device__ __forceinline__ float w_as_f32<float>(const float & v) { return v; }
template <> __device__ __forceinline__ float w_as_f32<__half>(const __half & v) { return __half2float(v); }
template <> __device__ __forceinline__ float w_as_f32<__nv_bfloat16>(const __nv_bfloat16 & v) { return __bfloat162float(v); }
template <typename T>
__global__ void __launch_bounds__(128) mul_mat_f_kernel(const char * w, const float * x, const float * bias, float * d, int64_t K, int64_t N, int64_t w_nb1, int64_t x_ld, int64_t d_ld) {
    pdl_launch_dependents();
    const int64_t m = blockIdx.x, n0 = (int64_t) blockIdx.y * 8;
    const int nc = (int) min((int64_t) 8, N - n0);
    const T * wr = (const T *) (w + m * w_nb1);
    pdl_wait();
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (std::is_same<T, float>::value && nc == 1 && K % 4 == 0 && (((uintptr_t) wr | (uintptr_t) (x + n0 * x_ld)) % 16) == 0) {
        // decode: the router matrix of a MoE layer (f32 [n_embd, n_expert]) times one column: 128-bit loads, four in flight
        const float4 * w4 = (const float4 *) wr; const float4 * x4 = (const float4 *) (x + n0 * x_ld);
        const int64_t K4 = K / 4;
        float a0 = 0.0f;
#pragma unroll 4
        for (int64_t k = threadIdx.x; k < K4; k += blockDim.x) {
            const float4 a = w4[k], b = x4[k];
            a0 += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
        }
        acc[0] = a0;
    } else
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        const float wv = w_as_f32<T>(wr[k]);
#pragma unroll
        for (int c = 0; c < 8; ++c) if (c < nc) acc[c] += wv * x[(n0 + c) * x_ld + k];
    }
    __shared__ float part[4][8];
#pragma unroll
    for (int c = 0; c < 8; ++c) { const float s = warp_sum(acc[c]); if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5][c] = s; }
    __syncthreads();
    if (threadIdx.x < nc) {
        float r = part[0][threadIdx.x] + part[1][threadIdx.x] + part[2][threadIdx.x] + part[3][threadIdx.x];
        if (bias) r = __fadd_rn(r, bias[m]);               // the ADD of a per-row bias that follows (MoE router: ffn_gate_inp.bias)
        d[(n0 + threadIdx.x) * d_ld + m] = r;
    }
}

// ------------------------------------------------------------------ fused routing of a MoE layer (softmax-weight gating)
// Replaces {ARGSORT desc, VIEW(first n_used), GET_ROWS(probs, selected), SOFT_MAX} (reference fusion: ggml-cuda/topk-moe.cu:314,
// "delayed softmax" variant; graph: src/llama-graph.cpp build_moe_ffn, gpt-oss).  One warp per token: the logits of all experts
// live in registers (lane l holds experts l, l + 32, ...), n_used rounds of warp arg-max (the lowest index wins a tie) pick the
// experts in descending order, then a softmax over the picked values.  ids go to the first n_used slots of the ARGSORT row
// (the only part of it the graph reads), weights to the SOFT_MAX output.
template <int EPT>
__global__ void __launch_bounds__(128) topk_moe_kernel(const float * logits, int32_t * ids, float * weights, int n_expert, int n_used, int64_t n_tok,
                                                       int64_t l_row, int64_t i_row, int64_t w_row) {
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31;
    const int64_t t = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (t >= n_tok) return;
    pdl_wait();
    const float * row = logits + t * l_row;
    float v[EPT];
#pragma unroll
    for (int i = 0; i < EPT; ++i) { const int e = lane + 32 * i; v[i] = e < n_expert ? row[e] : -INFINITY; }
    float mine = -INFINITY;                              // lane k keeps the k-th picked value (n_used <= 32)
    for (int k = 0; k < n_used; ++k) {
        float bv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll
        for (int i = 0; i < EPT; ++i) if (v[i] > bv) { bv = v[i]; bi = lane + 32 * i; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
#pragma unroll
        for (int i = 0; i < EPT; ++i) if (lane + 32 * i == bi) v[i] = -INFINITY;      // taken
        if (lane == 0) ids[t * i_row + k] = bi;
        if (lane == k) mine = bv;
    }
    const float top = __shfl_sync(0xffffffffu, mine, 0);                              // the row maximum
    const float e = lane < n_used ? expf(mine - top) : 0.0f;
    const float sum = warp_sum(e);
    if (lane < n_used) weights[t * w_row + lane] = e / sum;
}

// ------------------------------------------------------------------ weighted sum over the selected experts
// Replaces {MUL(experts, weights), VIEW x n_used, ADD x (n_used - 1)} at the end of build_moe_ffn: dst[:, t] = sum_e experts[:, e, t] * w[e, t],
// products rounded first and added left to right, exactly like the separate nodes.
__global__ void __launch_bounds__(256) moe_weighted_sum_kernel(const char * x, const char * w, const char * res, char * d, int64_t ne0, int n_used,
                                                               int64_t x_nb1, int64_t x_nb2, int64_t w_nb1, int64_t w_nb2, int64_t r_nb1, int64_t d_nb1) {
    pdl_launch_dependents();
    const int64_t t = blockIdx.y, i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ne0) return;
    pdl_wait();
    float acc = 0.0f;
    for (int e = 0; e < n_used; ++e) {
        const float p = __fmul_rn(*(const float *) (x + i * 4 + e * x_nb1 + t * x_nb2), *(const float *) (w + e * w_nb1 + t * w_nb2));
        acc = e == 0 ? p : __fadd_rn(acc, p);
    }
    if (res) acc = __fadd_rn(acc, *(const float *) (res + i * 4 + t * r_nb1));      // the residual stream (the ADD that follows the fold)
    *(float *) (d + i * 4 + t * d_nb1) = acc;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_moe_weighted_sum(const b200_tensor * x, const b200_tensor * w, const b200_tensor * res, const b200_tensor * d, void * stream) {
    if (!x || !w || !d) return B200_ERR_ARG;
    if (res && (res->type != B200_F32 || res->nb[0] != 4 || res->ne[0] != d->ne[0] || res->ne[1] != d->ne[1] || res->ne[2] != 1 || res->ne[3] != 1)) return B200_ERR_UNSUPPORTED;
    if (x->type != B200_F32 || w->type != B200_F32 || d->type != B200_F32 || x->nb[0] != 4 || d->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    const int64_t E = x->ne[0], n_used = x->ne[1], n_tok = x->ne[2];
    if (x->ne[3] != 1 || w->ne[0] != 1 || w->ne[1] != n_used || w->ne[2] != n_tok || w->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (d->ne[0] != E || d->ne[1] != n_tok || d->ne[2] != 1 || d->ne[3] != 1 || n_used < 1 || n_used > 1024 || n_tok > 65535) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (E == 0 || n_tok == 0) return B200_OK;
    dim3 grid((unsigned) ((E + 255) / 256), (unsigned) n_tok);
    moe_weighted_sum_kernel<<<grid, 256, 0, (cudaStream_t) stream>>>((const char *) x->data, (const char *) w->data, res ? (const char *) res->data : nullptr, (char *) d->data,
                                                                    E, (int) n_used, x->nb[1], x->nb[2], w->nb[1], w->nb[2], res ? res->nb[1] : 0, d->nb[1]);
    count_launch();
    return check_launch("moe_weighted_sum");
}

extern "C" int b200_topk_moe(const b200_tensor * logits, const b200_tensor * ids, const b200_tensor * weights, void * stream) {
    if (!logits || !ids || !weights) return B200_ERR_ARG;
    if (logits->type != B200_F32 || ids->type != B200_I32 || weights->type != B200_F32 || logits->nb[0] != 4 || ids->nb[0] != 4 || weights->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    const int64_t n_expert = logits->ne[0], n_tok = logits->ne[1], n_used = ids->ne[0];
    if (logits->ne[2] != 1 || logits->ne[3] != 1 || ids->ne[1] != n_tok || ids->ne[2] != 1 || ids->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (weights->ne[0] != n_used || weights->ne[1] != n_tok || weights->ne[2] != 1 || weights->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (n_expert < 1 || n_expert > 512 || n_used < 1 || n_used > 32 || n_used > n_expert || n_tok < 1) return B200_ERR_UNSUPPORTED;
    if (logits->nb[1] % 4 || ids->nb[1] % 4 || weights->nb[1] % 4) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    const unsigned blocks = (unsigned) ((n_tok + 3) / 4);
    cudaStream_t st = (cudaStream_t) stream;
    const float * pl = (const float *) logits->data; int32_t * pi = (int32_t *) ids->data; float * pw = (float *) weights->data;
    const int64_t lr = logits->nb[1] / 4, ir = ids->nb[1] / 4, wr = weights->nb[1] / 4;
    if (n_expert <= 32)       topk_moe_kernel<1><<<blocks, 128, 0, st>>>(pl, pi, pw, (int) n_expert, (int) n_used, n_tok, lr, ir, wr);
    else if (n_expert <= 128) topk_moe_kernel<4><<<blocks, 128, 0, st>>>(pl, pi, pw, (int) n_expert, (int) n_used, n_tok, lr, ir, wr);
    else if (n_expert <= 256) topk_moe_kernel<8><<<blocks, 128, 0, st>>>(pl, pi, pw, (int) n_expert, (int) n_used, n_tok, lr, ir, wr);
    else                      topk_moe_kernel<16><<<blocks, 128, 0, st>>>(pl, pi, pw, (int) n_expert, (int) n_used, n_tok, lr, ir, wr);
    count_launch();
    return check_launch("topk_moe");
}

extern "C" int b200_argsort(const b200_tensor * x, const b200_tensor * dst, int order, void * stream) {
    return argsort_common(x, dst, order, dst ? (int) dst->ne[0] : 0, stream, "argsort");
}
extern "C" int b200_top_k(const b200_tensor * x, const b200_tensor * dst, void * stream) {
    return argsort_common(x, dst, 1, dst ? (int) dst->ne[0] : 0, stream, "top_k");
}

extern "C" int b200_add_id(const b200_tensor * a, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * d, void * stream) {
    if (!a || !b || !ids || !d) return B200_ERR_ARG;
    if (a->type != B200_F32 || b->type != B200_F32 || d->type != B200_F32 || ids->type != B200_I32) return B200_ERR_UNSUPPORTED;
    if (a->nb[0] != 4 || b->nb[0] != 4 || d->nb[0] != 4 || a->ne[3] != 1 || d->ne[3] != 1 || b->ne[2] != 1 || b->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (b->ne[0] != a->ne[0] || ids->ne[0] != a->ne[1] || ids->ne[1] != a->ne[2] || ids->ne[2] != 1 || ids->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    for (int i = 0; i < 4; ++i) if (d->ne[i] != a->ne[i]) return B200_ERR_UNSUPPORTED;
    if (a->ne[1] * a->ne[2] > 0x7fffffff) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*d) == 0) return B200_OK;
    const int vec = a->ne[0] % 4 == 0 && (((uintptr_t) a->data | (uintptr_t) b->data | (uintptr_t) d->data) % 16) == 0 &&
                    (a->nb[1] | a->nb[2] | b->nb[1] | d->nb[1] | d->nb[2]) % 16 == 0;
    const int64_t per_block = 256 * (vec ? 4 : 1);
    if ((a->ne[0] + per_block - 1) / per_block > 65535) return B200_ERR_UNSUPPORTED;
    dim3 grid((unsigned) (a->ne[1] * a->ne[2]), (unsigned) ((a->ne[0] + per_block - 1) / per_block));
    add_id_kernel<<<grid, 256, 0, (cudaStream_t) stream>>>((const char *) a->data, (const char *) b->data, (const char *) ids->data, (char *) d->data,
        a->ne[0], a->ne[1], a->nb[1], a->nb[2], b->nb[1], ids->nb[0], ids->nb[1], d->nb[1], d->nb[2], vec);
    count_launch();
    return check_launch("add_id");
}

extern "C" int b200_sum_rows(const b200_tensor * x, const b200_tensor * d, void * stream) {
    if (!x || !d || x->type != B200_F32 || d->type != B200_F32 || x->nb[0] != 4 || d->ne[0] != 1) return B200_ERR_UNSUPPORTED;
    if (d->ne[1] != x->ne[1] || d->ne[2] != x->ne[2] || d->ne[3] != x->ne[3] || nrows(*x) > 0x7fffffff) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nrows(*x) == 0) return B200_OK;
    sum_rows_kernel<<<(unsigned) nrows(*x), 256, 0, (cudaStream_t) stream>>>((const char *) x->data, (char *) d->data, x->ne[0], x->ne[1], x->ne[2],
        x->nb[1], x->nb[2], x->nb[3], d->nb[1], d->nb[2], d->nb[3]);
    count_launch();
    return check_launch("sum_rows");
}

extern "C" int b200_clamp(const b200_tensor * x, const b200_tensor * d, float lo, float hi, void * stream) {
    if (!x || !d || x->type != B200_F32 || d->type != B200_F32 || !is_contiguous(*x) || !is_contiguous(*d) || nelements(*x) != nelements(*d)) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    const int64_t n = nelements(*x);
    if (n == 0) return B200_OK;
    clamp_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, (cudaStream_t) stream>>>((const float *) x->data, (float *) d->data, lo, hi, n);
    count_launch();
    return check_launch("clamp");
}

extern "C" int b200_mul_mat_f(const b200_tensor * w, const b200_tensor * x, const b200_tensor * bias, const b200_tensor * d, void * stream) {
    if (!w || !x || !d) return B200_ERR_ARG;
    if (bias && (bias->type != B200_F32 || bias->nb[0] != 4 || bias->ne[0] != w->ne[1] || nelements(*bias) != w->ne[1])) return B200_ERR_UNSUPPORTED;
    if (w->type != B200_F32 && w->type != B200_F16 && w->type != B200_BF16) return B200_ERR_UNSUPPORTED;
    if (x->type != B200_F32 || d->type != B200_F32 || w->nb[0] != (int64_t) b200_type_size(w->type) || x->nb[0] != 4 || d->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    if (w->ne[2] != 1 || w->ne[3] != 1 || x->ne[2] != 1 || x->ne[3] != 1 || d->ne[2] != 1 || d->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (x->ne[0] != w->ne[0] || d->ne[0] != w->ne[1] || d->ne[1] != x->ne[1] || x->nb[1] % 4 || d->nb[1] % 4 || w->nb[1] % (int64_t) b200_type_size(w->type)) return B200_ERR_UNSUPPORTED;
    if (w->ne[1] > 0x7fffffff || (x->ne[1] + 7) / 8 > 65535) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*d) == 0) return B200_OK;
    dim3 grid((unsigned) w->ne[1], (unsigned) ((x->ne[1] + 7) / 8));
    cudaStream_t st = (cudaStream_t) stream;
    const char * pw = (const char *) w->data; const float * px = (const float *) x->data; float * pd = (float *) d->data;
    const float * pb = bias ? (const float *) bias->data : nullptr;
    switch (w->type) {
        case B200_F32: mul_mat_f_kernel<float><<<grid, 128, 0, st>>>(pw, px, pb, pd, w->ne[0], x->ne[1], w->nb[1], x->nb[1] / 4, d->nb[1] / 4); break;
        case B200_F16: mul_mat_f_kernel<__half><<<grid, 128, 0, st>>>(pw, px, pb, pd, w->ne[0], x->ne[1], w->nb[1], x->nb[1] / 4, d->nb[1] / 4); break;
        default:       mul_mat_f_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(pw, px, pb, pd, w->ne[0], x->ne[1], w->nb[1], x->nb[1] / 4, d->nb[1] / 4); break;
    }
    count_launch();
    return check_launch("mul_mat_f");
}
