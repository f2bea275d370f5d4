// fattn.cu — K7: FLASH_ATTN_EXT for decode (few query columns), f16 or q8_0 KV cache, split-KV
// across SMs with an in-kernel "last block combines" reduction (no second launch).
// Replaces reference ggml-cuda/fattn-vec.cuh:19-507 + flash_attn_combine_results
// (fattn-common.cuh:728-779) + launch_fattn (:781-1036).
//
// Layout: a block owns one KV split of one (kv-head, column chunk, sequence).  The NCOLS columns of
// a chunk are the GQA query heads (x query positions) that share the kv-head, so K/V bytes are read
// once per group.  Inside a warp, LPR = D/8 lanes own one K/V row (each lane 8 consecutive
// elements: one 128-bit load for f16, 8 B + scale for q8_0) and 32/LPR "row slots" run in parallel,
// each with a private online-softmax state; states are merged once per block in shared memory.
#include <algorithm>

#include "common.cuh"

namespace b200 {

struct fattn_params {
    const char * q; const char * k; const char * v; const char * mask; const float * sinks; char * dst;
    int64_t q_nb1, q_nb2, q_nb3, k_nb1, k_nb2, k_nb3, v_nb1, v_nb2, v_nb3;
    int64_t m_nb1, m_nb2, m_nb3, m_ne2, m_ne3, d_nb1, d_nb2, d_nb3;
    int32_t n_q, H, H_kv, gqa, n_kv, n_split, kv_per_split, n_chunks, ncols_total;
    float scale, softcap, max_bias, m0, m1; uint32_t n_head_log2;
    float * part; int * counters;
};

// raw K/V bytes of one lane's 8 elements, kept unconverted while the loads are in flight
template <int KVT> struct kvraw;
template <> struct kvraw<B200_F16> {
    uint4 r;
    __device__ __forceinline__ void load(const char * row, int sub) { r = *(const uint4 *) (row + 16 * sub); }
    __device__ __forceinline__ void zero() { r = make_uint4(0u, 0u, 0u, 0u); }
    __device__ __forceinline__ void expand(float f[8]) const {
        const __half2 * h = (const __half2 *) &r;
#pragma unroll
        for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
};
template <> struct kvraw<B200_Q8_0> {
    uint32_t q0, q1; float d;          // 8 int8 + the block scale
    __device__ __forceinline__ void load(const char * row, int sub) {
        const char * b = row + (sub >> 2) * 34;
        d = __half2float(*(const __half *) b);
        const uint16_t * qs = (const uint16_t *) (b + 2 + 8 * (sub & 3));      // 2-byte aligned
        q0 = (uint32_t) qs[0] | ((uint32_t) qs[1] << 16); q1 = (uint32_t) qs[2] | ((uint32_t) qs[3] << 16);
    }
    __device__ __forceinline__ void zero() { q0 = q1 = 0u; d = 0.0f; }
    __device__ __forceinline__ void expand(float f[8]) const {
        // int8 -> f16 without I2F: (q ^ 0x80) | 0x6400 is exactly 1024 + 128 + q; minus 1152
        const __half2 off = __float2half2_rn(1152.0f);
        const uint32_t w[2] = {q0 ^ 0x80808080u, q1 ^ 0x80808080u};
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const uint32_t lo = __byte_perm(w[i], 0x64646464u, 0x4140), hi = __byte_perm(w[i], 0x64646464u, 0x4342);
            const float2 a = __half22float2(__hsub2(*(const __half2 *) &lo, off)), b = __half22float2(__hsub2(*(const __half2 *) &hi, off));
            f[4 * i] = d * a.x; f[4 * i + 1] = d * a.y; f[4 * i + 2] = d * b.x; f[4 * i + 3] = d * b.y;
        }
    }
};

// U = row steps of a warp whose K / V / mask requests are issued together before any arithmetic (bytes in flight per
// warp = U * 32 lanes * 32 B): the loop is bound by HBM latency otherwise.  One online-softmax update per U rows.
template <int D, int KVT, int NCOLS>
__global__ void __launch_bounds__(128) fattn_vec_kernel(const fattn_params p) {
    constexpr int LPR = D / 8, RPI = 32 / LPR, NW = 4, NSTATE = NW * RPI, ST = D + 2;
    constexpr int U = NCOLS >= 3 ? 2 : 4;
    extern __shared__ float sm[];   // [NSTATE][NCOLS][ST]: acc[D], m, l
    __shared__ int s_last;
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane % LPR, rsel = lane / LPR;
    const int split = blockIdx.x, grp = blockIdx.y, seq = blockIdx.z;
    const int hkv = grp / p.n_chunks, chunk = grp % p.n_chunks;
    const int d0 = 8 * sub;

    // column c of this chunk -> (query position j, head)
    int cj[NCOLS], ch[NCOLS]; bool cact[NCOLS]; float slope[NCOLS];
    float qreg[NCOLS][8];
    const char * mrow[NCOLS];        // mask row of the column (null: no mask)
    pdl_wait();
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        const int ci = chunk * NCOLS + c;
        cact[c] = ci < p.ncols_total;
        cj[c] = cact[c] ? ci / p.gqa : 0;
        ch[c] = hkv * p.gqa + (cact[c] ? ci % p.gqa : 0);
        const float * qp = (const float *) (p.q + (int64_t) cj[c] * p.q_nb1 + (int64_t) ch[c] * p.q_nb2 + (int64_t) seq * p.q_nb3);
#pragma unroll
        for (int i = 0; i < 8; ++i) qreg[c][i] = cact[c] ? qp[d0 + i] * p.scale : 0.0f;
        slope[c] = 1.0f;
        if (p.max_bias > 0.0f) {
            const uint32_t h = (uint32_t) ch[c];
            slope[c] = h < p.n_head_log2 ? powf(p.m0, (float) (h + 1)) : powf(p.m1, (float) (2 * (h - p.n_head_log2) + 1));
        }
        mrow[c] = p.mask && cact[c] ? p.mask + (int64_t) cj[c] * p.m_nb1 + (int64_t) (ch[c] % p.m_ne2) * p.m_nb2 + (int64_t) (seq % p.m_ne3) * p.m_nb3 : nullptr;
    }
    // decode (one query position, mask not per head): every column reads the same mask row -> one load per row step
    bool mask_uniform = true;
#pragma unroll
    for (int c = 1; c < NCOLS; ++c) mask_uniform = mask_uniform && (!cact[c] || mrow[c] == mrow[0]);
    // "nothing seen yet" is a large finite value, not -inf: exp(m_old - m_new) and exp(s - m) then need no special cases
    // (a masked score is -inf and gives exp(-inf) = 0); the state written out below turns it back into -inf
    constexpr float M_NONE = -1.0e30f;
    float m[NCOLS], l[NCOLS], acc[NCOLS][8];
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) { m[c] = M_NONE; l[c] = 0.0f;
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[c][i] = 0.0f; }

    const int kv0 = split * p.kv_per_split, kv1 = min(p.n_kv, kv0 + p.kv_per_split);
    const char * kb = p.k + (int64_t) hkv * p.k_nb2 + (int64_t) seq * p.k_nb3;
    const char * vb = p.v + (int64_t) hkv * p.v_nb2 + (int64_t) seq * p.v_nb3;

    // software pipeline: the requests of the NEXT U row steps are issued before the arithmetic of the current ones, so a
    // warp always has 2 x U row steps of K/V bytes in flight (3 warps per scheduler do not hide HBM latency otherwise:
    // ncu showed long-scoreboard as the top stall at 41 % issue utilisation)
    kvraw<KVT> kn[U], vn[U];
    float mn[U];                     // uniform mask value of the row (+ -inf past the end), fetched with the row
    auto request = [&](int base) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int r0 = base + u * RPI + rsel;
            const int r = max(min(r0, kv1 - 1), kv0);
            kn[u].load(kb + (int64_t) r * p.k_nb1, sub); vn[u].load(vb + (int64_t) r * p.v_nb1, sub);
            mn[u] = r0 < kv1 ? 0.0f : -INFINITY;
            if (p.mask && mask_uniform && mrow[0]) mn[u] += __half2float(*(const __half *) (mrow[0] + (int64_t) r * 2));
        }
    };
    if (kv0 < kv1) request(kv0 + warp * (U * RPI));
    for (int base = kv0 + warp * (U * RPI); base < kv1; base += NW * U * RPI) {
        kvraw<KVT> kr[U], vr[U];
        float mk[U][NCOLS];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            kr[u] = kn[u]; vr[u] = vn[u];
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) mk[u][c] = slope[c] * mn[u];
        }
        if (p.mask && !mask_uniform) {       // prefill-style calls (several query positions / per-head masks): fetched here
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = max(min(base + u * RPI + rsel, kv1 - 1), kv0);
#pragma unroll
                for (int c = 0; c < NCOLS; ++c) if (mrow[c]) mk[u][c] += slope[c] * __half2float(*(const __half *) (mrow[c] + (int64_t) r * 2));
            }
        }
        if (base + NW * U * RPI < kv1) request(base + NW * U * RPI);
        // ---- scores of the U x NCOLS (row, column) pairs
        float s[U][NCOLS];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float kf[8];
            kr[u].expand(kf);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) {
                float t = 0.0f;
#pragma unroll
                for (int i = 0; i < 8; ++i) t += qreg[c][i] * kf[i];
                s[u][c] = t;
            }
        }
#pragma unroll
        for (int o = LPR / 2; o > 0; o >>= 1) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
                for (int c = 0; c < NCOLS; ++c) s[u][c] += __shfl_xor_sync(0xffffffffu, s[u][c], o);
            }
        }
        // ---- one (branch-free) online-softmax update per column for the U rows; columns past ncols_total carry q = 0
        // and are never written out
#pragma unroll
        for (int c = 0; c < NCOLS; ++c) {
            float mx = m[c];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float t = s[u][c];
                if (p.softcap != 0.0f) t = p.softcap * tanhf(t);
                t += mk[u][c];
                s[u][c] = t;
                mx = fmaxf(mx, t);
            }
            const float corr = __expf(m[c] - mx);
            m[c] = mx;
            float ls = 0.0f;
#pragma unroll
            for (int u = 0; u < U; ++u) { s[u][c] = __expf(s[u][c] - mx); ls += s[u][c]; }
            l[c] = l[c] * corr + ls;
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[c][i] *= corr;
        }
        // ---- V accumulation
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float vf[8];
            vr[u].expand(vf);
#pragma unroll
            for (int c = 0; c < NCOLS; ++c) {
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[c][i] += s[u][c] * vf[i];
            }
        }
    }
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) if (m[c] <= 0.5f * M_NONE) m[c] = -INFINITY;      // no unmasked row: the merges below test for -inf

    // ---- merge the NSTATE per-slot states of this block
    const int state = warp * RPI + rsel;
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        float * o = sm + ((int64_t) state * NCOLS + c) * ST;
#pragma unroll
        for (int i = 0; i < 8; ++i) o[d0 + i] = acc[c][i];
        if (sub == 0) { o[D] = m[c]; o[D + 1] = l[c]; }
    }
    __syncthreads();
    const int64_t grp_lin = (int64_t) seq * gridDim.y + grp;
    float * part = p.part + (grp_lin * p.n_split + split) * NCOLS * ST;
    for (int e = threadIdx.x; e < NCOLS * D; e += 128) {
        const int c = e / D, d = e % D;
        float M = -INFINITY;
        for (int s = 0; s < NSTATE; ++s) M = fmaxf(M, sm[((int64_t) s * NCOLS + c) * ST + D]);
        float A = 0.0f, L = 0.0f;
        for (int s = 0; s < NSTATE; ++s) {
            const float * o = sm + ((int64_t) s * NCOLS + c) * ST;
            const float w = o[D] == -INFINITY ? 0.0f : expf(o[D] - M);
            A += w * o[d]; L += w * o[D + 1];
        }
        if (p.n_split == 1) {
            const int ci = chunk * NCOLS + c;
            if (ci < p.ncols_total) {
                const int j = ci / p.gqa, h = hkv * p.gqa + ci % p.gqa;
                if (p.sinks) {
                    const float sk = p.sinks[h], Mn = fmaxf(M, sk);
                    const float w = M == -INFINITY ? 0.0f : expf(M - Mn);
                    A *= w; L = L * w + expf(sk - Mn);
                }
                ((float *) (p.dst + (int64_t) h * p.d_nb1 + (int64_t) j * p.d_nb2 + (int64_t) seq * p.d_nb3))[d] = A / L;
            }
        } else {
            part[c * ST + d] = A;
            if (d == 0) { part[c * ST + D] = M; part[c * ST + D + 1] = L; }
        }
    }
    if (p.n_split == 1) return;

    // ---- split-KV: the last block of this group to arrive combines all partials
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const int prev = atomicAdd(p.counters + grp_lin, 1);
        s_last = prev == p.n_split - 1;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // The partials of the n_split blocks are independent loads: a first pass (threads over splits) turns the per-split
    // maxima into weights in shared memory, a second pass (threads over output elements) sums with 8 loads in flight.
    // (The first version walked the splits one dependent L2 round trip at a time: ~40 us of an 87 us launch at 64 splits.)
    const float * all = p.part + grp_lin * p.n_split * NCOLS * ST;
    float * s_w = sm;                                   // [NCOLS][n_split] weights; reuses the state area (all reads of it are done)
    __shared__ float s_red[NCOLS][4], s_M[NCOLS], s_L[NCOLS];
    __syncthreads();
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        float M = -INFINITY;
        for (int sp = threadIdx.x; sp < p.n_split; sp += 128) M = fmaxf(M, all[((int64_t) sp * NCOLS + c) * ST + D]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
        if (lane == 0) s_red[c][warp] = M;
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < NCOLS; ++c) {
        float M = fmaxf(fmaxf(s_red[c][0], s_red[c][1]), fmaxf(s_red[c][2], s_red[c][3]));
        const int ci = chunk * NCOLS + c;
        const int h = hkv * p.gqa + (ci < p.ncols_total ? ci % p.gqa : 0);
        float Lsink = 0.0f;
        if (p.sinks && ci < p.ncols_total) { const float sk = p.sinks[h]; const float Mn = fmaxf(M, sk); Lsink = expf(sk - Mn); M = Mn; }
        float L = 0.0f;
        for (int sp = threadIdx.x; sp < p.n_split; sp += 128) {
            const float * o = all + ((int64_t) sp * NCOLS + c) * ST;
            const float w = o[D] == -INFINITY ? 0.0f : expf(o[D] - M);
            s_w[c * p.n_split + sp] = w;
            L += w * o[D + 1];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) L += __shfl_xor_sync(0xffffffffu, L, o);
        __syncthreads();                                // s_red of this column has been read by everyone
        if (lane == 0) s_red[c][warp] = L;
        __syncthreads();
        if (threadIdx.x == 0) { s_M[c] = M; s_L[c] = s_red[c][0] + s_red[c][1] + s_red[c][2] + s_red[c][3] + Lsink; }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < NCOLS * D; e += 128) {
        const int c = e / D, d = e % D;
        const int ci = chunk * NCOLS + c;
        if (ci >= p.ncols_total) continue;
        const int j = ci / p.gqa, h = hkv * p.gqa + ci % p.gqa;
        const float * col = all + (int64_t) c * ST + d;
        const float * w = s_w + c * p.n_split;
        float A = 0.0f;
        int sp = 0;
        for (; sp + 8 <= p.n_split; sp += 8) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = col[(int64_t) (sp + u) * NCOLS * ST];
#pragma unroll
            for (int u = 0; u < 8; ++u) A += w[sp + u] * v[u];
        }
        for (; sp < p.n_split; ++sp) A += w[sp] * col[(int64_t) sp * NCOLS * ST];
        ((float *) (p.dst + (int64_t) h * p.d_nb1 + (int64_t) j * p.d_nb2 + (int64_t) seq * p.d_nb3))[d] = A / s_L[c];
    }
}

struct fattn_plan { int ncols, n_chunks, n_split, kv_per_split, ncols_total; int64_t n_groups; size_t part_bytes, counter_bytes; };

static bool fattn_make_plan(const b200_fattn_args * a, fattn_plan & pl) {
    const b200_tensor & q = a->q; const b200_tensor & k = a->k; const b200_tensor & v = a->v; const b200_tensor & d = a->dst;
    const int64_t D = q.ne[0], n_q = q.ne[1], H = q.ne[2], n_seq = q.ne[3], n_kv = k.ne[1], H_kv = k.ne[2];
    if (q.type != B200_F32 || d.type != B200_F32 || q.nb[0] != 4 || d.nb[0] != 4) return false;
    if (D != 64 && D != 128 && D != 256) return false;
    if (k.type != v.type || (k.type != B200_F16 && k.type != B200_Q8_0)) return false;
    if (k.ne[0] != D || v.ne[0] != D || v.ne[1] != n_kv || v.ne[2] != H_kv || k.ne[3] != n_seq || v.ne[3] != n_seq) return false;
    if (H_kv < 1 || H % H_kv != 0 || n_kv < 1 || n_q < 1) return false;
    if (d.ne[0] != D || d.ne[1] != H || d.ne[2] != n_q || d.ne[3] != n_seq) return false;
    if (k.nb[0] != (int64_t) b200_type_size(k.type) || v.nb[0] != k.nb[0]) return false;
    if (k.type == B200_F16) {
        if (((uintptr_t) k.data | (uintptr_t) v.data) % 16) return false;
        for (int i = 1; i < 4; ++i) if (k.nb[i] % 16 || v.nb[i] % 16) return false;
    }
    const int64_t gqa = H / H_kv;
    // any column count is accepted: prefill-sized n_q runs as many 8-column chunks (correct, not fast)
    if (a->mask) {
        const b200_tensor & m = *a->mask;
        if (m.type != B200_F16 || m.nb[0] != 2 || m.ne[0] < n_kv || m.ne[1] < n_q || m.ne[2] < 1 || m.ne[3] < 1) return false;
        if (H % m.ne[2] != 0 || n_seq % m.ne[3] != 0) return false;
    }
    if (a->sinks && (a->sinks->type != B200_F32 || a->sinks->ne[0] != H || a->sinks->nb[0] != 4)) return false;
    pl.ncols_total = (int) (gqa * n_q);
    pl.ncols = pl.ncols_total >= 8 ? 8 : pl.ncols_total;   // 1..8: a GQA group of 5 (Qwen2.5-32B) or 6 pays for no idle columns
    pl.n_chunks = (pl.ncols_total + pl.ncols - 1) / pl.ncols;
    pl.n_groups = (int64_t) H_kv * pl.n_chunks * n_seq;
    if (H_kv * pl.n_chunks > 65535 || n_seq > 65535) return false;
    // enough blocks for ~6 per SM (bytes in flight), at least 256 KV positions per split
    const int64_t target = 6 * 148;
    int64_t ns = (target + pl.n_groups - 1) / pl.n_groups;
    const int64_t max_ns = (n_kv + 255) / 256;
    if (ns > max_ns) ns = max_ns;
    if (ns > 512) ns = 512;            // the combine keeps one weight per (column, split) in shared memory
    if (ns < 1) ns = 1;
    pl.kv_per_split = (int) (((n_kv + ns - 1) / ns + 31) / 32 * 32);
    pl.n_split = (int) ((n_kv + pl.kv_per_split - 1) / pl.kv_per_split);
    pl.part_bytes = pl.n_split > 1 ? (size_t) pl.n_groups * pl.n_split * pl.ncols * (D + 2) * 4 : 0;
    pl.counter_bytes = (size_t) ((pl.n_groups * 4 + 255) / 256 * 256);
    return true;
}

template <int D, int KVT>
static void fattn_launch_nc(const fattn_params & p, int ncols, dim3 grid, cudaStream_t st) {
    constexpr int RPI = 32 / (D / 8);
    const size_t smem = std::max((size_t) 4 * RPI * ncols * (D + 2) * 4, (size_t) ncols * p.n_split * 4);
    switch (ncols) {
        case 1: fattn_vec_kernel<D, KVT, 1><<<grid, 128, smem, st>>>(p); break;
        case 2: fattn_vec_kernel<D, KVT, 2><<<grid, 128, smem, st>>>(p); break;
        case 3: fattn_vec_kernel<D, KVT, 3><<<grid, 128, smem, st>>>(p); break;
        case 4: fattn_vec_kernel<D, KVT, 4><<<grid, 128, smem, st>>>(p); break;
        case 5: fattn_vec_kernel<D, KVT, 5><<<grid, 128, smem, st>>>(p); break;
        case 6: fattn_vec_kernel<D, KVT, 6><<<grid, 128, smem, st>>>(p); break;
        case 7: fattn_vec_kernel<D, KVT, 7><<<grid, 128, smem, st>>>(p); break;
        default: fattn_vec_kernel<D, KVT, 8><<<grid, 128, smem, st>>>(p); break;
    }
}

}  // namespace b200

using namespace b200;

// scratch = [arrival counters (zeroed by the launcher when n_split > 1)] [partials]
extern "C" size_t b200_fattn_scratch_size(const b200_fattn_args * a) {
    fattn_plan pl;
    if (!a || !fattn_make_plan(a, pl)) return 0;
    return pl.counter_bytes + pl.part_bytes;
}

extern "C" int b200_flash_attn_ext(const b200_fattn_args * a, void * stream) {
    fattn_plan pl;
    if (!a) return B200_ERR_ARG;
    if (!fattn_make_plan(a, pl)) return B200_ERR_UNSUPPORTED;
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    if (a->scratch_size < pl.counter_bytes + pl.part_bytes || !a->scratch) { set_error("flash_attn_ext: scratch too small"); return B200_ERR_ARG; }
    const b200_tensor & q = a->q; const b200_tensor & k = a->k; const b200_tensor & v = a->v; const b200_tensor & d = a->dst;
    fattn_params p{};
    p.q = (const char *) q.data; p.k = (const char *) k.data; p.v = (const char *) v.data; p.dst = (char *) d.data;
    p.mask = a->mask ? (const char *) a->mask->data : nullptr; p.sinks = a->sinks ? (const float *) a->sinks->data : nullptr;
    p.q_nb1 = q.nb[1]; p.q_nb2 = q.nb[2]; p.q_nb3 = q.nb[3];
    p.k_nb1 = k.nb[1]; p.k_nb2 = k.nb[2]; p.k_nb3 = k.nb[3];
    p.v_nb1 = v.nb[1]; p.v_nb2 = v.nb[2]; p.v_nb3 = v.nb[3];
    if (a->mask) { p.m_nb1 = a->mask->nb[1]; p.m_nb2 = a->mask->nb[2]; p.m_nb3 = a->mask->nb[3]; p.m_ne2 = a->mask->ne[2]; p.m_ne3 = a->mask->ne[3]; }
    p.d_nb1 = d.nb[1]; p.d_nb2 = d.nb[2]; p.d_nb3 = d.nb[3];
    p.n_q = (int) q.ne[1]; p.H = (int) q.ne[2]; p.H_kv = (int) k.ne[2]; p.gqa = p.H / p.H_kv; p.n_kv = (int) k.ne[1];
    p.n_split = pl.n_split; p.kv_per_split = pl.kv_per_split; p.n_chunks = pl.n_chunks; p.ncols_total = pl.ncols_total;
    p.softcap = a->logit_softcap; p.scale = a->logit_softcap != 0.0f ? a->scale / a->logit_softcap : a->scale;
    p.max_bias = a->max_bias;
    p.n_head_log2 = 1u << (uint32_t) floorf(log2f((float) p.H));
    p.m0 = powf(2.0f, -a->max_bias / p.n_head_log2); p.m1 = powf(2.0f, -(a->max_bias / 2.0f) / p.n_head_log2);
    p.counters = (int *) a->scratch; p.part = (float *) ((char *) a->scratch + pl.counter_bytes);
    dim3 grid((unsigned) pl.n_split, (unsigned) (p.H_kv * pl.n_chunks), (unsigned) q.ne[3]);
    cudaStream_t st = (cudaStream_t) stream;
    if (pl.n_split > 1)   // arrival counters of the split-KV combine: zeroed per launch (a memset node under graph capture)
        B200_CUDA_TRY(cudaMemsetAsync(p.counters, 0, pl.counter_bytes, st));
    const int D = (int) q.ne[0];
    if (k.type == B200_F16) {
        if (D == 64) fattn_launch_nc<64, B200_F16>(p, pl.ncols, grid, st);
        else if (D == 128) fattn_launch_nc<128, B200_F16>(p, pl.ncols, grid, st);
        else fattn_launch_nc<256, B200_F16>(p, pl.ncols, grid, st);
    } else {
        if (D == 64) fattn_launch_nc<64, B200_Q8_0>(p, pl.ncols, grid, st);
        else if (D == 128) fattn_launch_nc<128, B200_Q8_0>(p, pl.ncols, grid, st);
        else fattn_launch_nc<256, B200_Q8_0>(p, pl.ncols, grid, st);
    }
    count_launch();
    return check_launch("flash_attn_ext");
}
