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
#include <cstdlib>

#include <mutex>

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

// ---------------------------------------------------------------- shared tail of the decode kernels
// in: sm = NSTATE online-softmax states per column, [state][column][acc[D], m, l] (m = -inf: nothing seen).  Merges them,
// writes the output (one split) or the partial of this split, and lets the last block of the group combine all splits.
template <int D, int NCOLS, int NSTATE>
__device__ __forceinline__ void fattn_tail(const fattn_params & p, float * sm, int split, int grp, int seq, int hkv, int chunk, int lane, int warp) {
    constexpr int ST = D + 2;
    __shared__ int s_last;
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
    fattn_tail<D, NCOLS, NSTATE>(p, sm, split, grp, seq, hkv, chunk, lane, warp);
}

// ---------------------------------------------------------------- prefill attention on the tensor cores
// Replaces, for many query positions and an f16 cache, the reference's mma kernel (fattn-mma-f16.cuh; selector fattn.cu:371-393).
// FlashAttention-2 forward: a CTA of 4 warps owns 64 query positions of one head (16 per warp, Q fragments in registers as
// f16), K and V tiles of 64 positions are staged in shared memory (cp.async, rows padded by 16 B so ldmatrix is
// conflict-free), S = Q.K^T and O += P.V run as mma.sync.m16n8k16 (f16 in, f32 accumulate), the accumulator fragment of S
// is reused as the A operand of P.V, V fragments come from ldmatrix.trans.  Mask f16 (added to the scores), scale; softcap,
// ALiBi and sinks stay on the vector kernel.  (tcgen05 would need M = 128 query rows per instruction and TMEM round trips
// for the softmax; at 4.3 GFLOP per layer this op is not the tensor-pipe-bound one.)
__device__ __forceinline__ void fa_ldsm4(uint32_t addr, uint32_t r[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void fa_ldsm4_t(uint32_t addr, uint32_t r[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void fa_mma(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t fa_pack(float x, float y) { const __half2 h = __floats2half2_rn(x, y); return *(const uint32_t *) &h; }

template <int D, int KVT>
__global__ void __launch_bounds__(128) fattn_mma_kernel(const fattn_params p) {
    constexpr int BN = 64, ROW = D * 2 + 16, KS = D / 16, NT = BN / 8, DT = D / 8, CPR = D / 8;   // CPR: 16 B chunks per row
    extern __shared__ __align__(16) uint8_t fsm[];
    uint8_t * sK = fsm, * sV = fsm + BN * ROW;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int h = blockIdx.y, seq = blockIdx.z, hkv = h / p.gqa;
    const int qa = blockIdx.x * 64 + warp * 16 + g, qb = qa + 8;
    const int qa_c = min(qa, p.n_q - 1), qb_c = min(qb, p.n_q - 1);      // rows past n_q compute on a clamped row and are not stored

    // ---- Q fragments (scaled, f16): A[kk] of m16n8k16, rows g / g+8, dims kk*16 + t*2 (+8)
    uint32_t qf[KS][4];
    {
        const float * qpa = (const float *) (p.q + (int64_t) qa_c * p.q_nb1 + (int64_t) h * p.q_nb2 + (int64_t) seq * p.q_nb3);
        const float * qpb = (const float *) (p.q + (int64_t) qb_c * p.q_nb1 + (int64_t) h * p.q_nb2 + (int64_t) seq * p.q_nb3);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
            const int d = kk * 16 + t * 2;
            const float2 a0 = *(const float2 *) (qpa + d), a1 = *(const float2 *) (qpb + d), a2 = *(const float2 *) (qpa + d + 8), a3 = *(const float2 *) (qpb + d + 8);
            qf[kk][0] = fa_pack(a0.x * p.scale, a0.y * p.scale); qf[kk][1] = fa_pack(a1.x * p.scale, a1.y * p.scale);
            qf[kk][2] = fa_pack(a2.x * p.scale, a2.y * p.scale); qf[kk][3] = fa_pack(a3.x * p.scale, a3.y * p.scale);
        }
    }
    const char * mra = p.mask ? p.mask + (int64_t) qa_c * p.m_nb1 + (int64_t) (h % p.m_ne2) * p.m_nb2 + (int64_t) (seq % p.m_ne3) * p.m_nb3 : nullptr;
    const char * mrb = p.mask ? p.mask + (int64_t) qb_c * p.m_nb1 + (int64_t) (h % p.m_ne2) * p.m_nb2 + (int64_t) (seq % p.m_ne3) * p.m_nb3 : nullptr;
    const char * kb = p.k + (int64_t) hkv * p.k_nb2 + (int64_t) seq * p.k_nb3;
    const char * vb = p.v + (int64_t) hkv * p.v_nb2 + (int64_t) seq * p.v_nb3;

    constexpr float M_NONE = -1.0e30f;
    float ma = M_NONE, mb = M_NONE, la = 0.0f, lb = 0.0f;       // l: this lane's share of the row sum (4 lanes per row)
    float o[DT][4];
#pragma unroll
    for (int i = 0; i < DT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.0f; }
    const uint32_t sK_a = (uint32_t) __cvta_generic_to_shared(sK), sV_a = (uint32_t) __cvta_generic_to_shared(sV);

    for (int kv0 = 0; kv0 < p.n_kv; kv0 += BN) {
        if (mra) {      // a tile that is masked for all 64 query positions of the CTA (causal prompt: about half of them) is skipped
            bool alive = false;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int kc = min(kv0 + nt * 8 + t * 2, p.n_kv - 2);
                const float2 fa = __half22float2(*(const __half2 *) (mra + (int64_t) kc * 2)), fb = __half22float2(*(const __half2 *) (mrb + (int64_t) kc * 2));
                alive = alive || fa.x > -INFINITY || fa.y > -INFINITY || fb.x > -INFINITY || fb.y > -INFINITY;
            }
            if (!__syncthreads_or(alive)) continue;
        }
        __syncthreads();                                         // the previous tiles have been consumed
        if (KVT == B200_Q8_0) {
            // q8_0 cache: blocks (half d, 32 int8; 2-byte aligned) are expanded to f16 on the way into shared memory, so the
            // MMA part is the same as for an f16 cache (the reference dequantises the whole cache to an f16 copy first,
            // fattn-common.cuh:830-890)
            constexpr int BPR = D / 32;
            for (int i = tid; i < 2 * BN * BPR; i += 128) {
                const bool isv = i >= BN * BPR;
                const int ii = isv ? i - BN * BPR : i, r = ii / BPR, j = ii % BPR, kr = kv0 + r;
                uint4 * dstp = (uint4 *) ((isv ? sV : sK) + r * ROW + j * 64);
                if (kr < p.n_kv) {
                    const uintptr_t a = (uintptr_t) ((isv ? vb + (int64_t) kr * p.v_nb1 : kb + (int64_t) kr * p.k_nb1) + 34 * j);
                    const uint32_t sh = (uint32_t) (a & 3) * 8;
                    const uint32_t * q = (const uint32_t *) (a & ~(uintptr_t) 3);
                    uint32_t w[10], x[9];
#pragma unroll
                    for (int u = 0; u < 9; ++u) w[u] = q[u];
                    w[9] = 0u;
#pragma unroll
                    for (int u = 0; u < 9; ++u) x[u] = __funnelshift_r(w[u], w[u + 1], sh);      // x: the block from its first byte
                    const uint16_t dbits = (uint16_t) (x[0] & 0xFFFFu);
                    const __half2 d2 = __half2half2(*(const __half *) &dbits), off = __float2half2_rn(1152.0f);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        uint32_t o[4];
#pragma unroll
                        for (int v = 0; v < 2; ++v) {
                            const uint32_t y = __funnelshift_r(x[2 * u + v], x[2 * u + v + 1], 16) ^ 0x80808080u;    // 4 int8 + 128
                            const uint32_t lo = __byte_perm(y, 0x64646464u, 0x4140), hi = __byte_perm(y, 0x64646464u, 0x4342);
                            const __half2 hl = __hmul2(__hsub2(*(const __half2 *) &lo, off), d2), hh = __hmul2(__hsub2(*(const __half2 *) &hi, off), d2);
                            o[2 * v] = *(const uint32_t *) &hl; o[2 * v + 1] = *(const uint32_t *) &hh;
                        }
                        dstp[u] = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < 4; ++u) dstp[u] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        } else
        for (int i = tid; i < BN * CPR; i += 128) {
            const int r = i / CPR, c = i % CPR, kr = kv0 + r;
            const uint32_t dk = sK_a + r * ROW + c * 16, dv = sV_a + r * ROW + c * 16;
            if (kr < p.n_kv) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dk), "l"(kb + (int64_t) kr * p.k_nb1 + c * 16) : "memory");
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dv), "l"(vb + (int64_t) kr * p.v_nb1 + c * 16) : "memory");
            } else {
                *(uint4 *) (sK + r * ROW + c * 16) = make_uint4(0u, 0u, 0u, 0u);
                *(uint4 *) (sV + r * ROW + c * 16) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();

        // ---- S = Q . K^T  (16 query rows x 64 positions per warp)
        float sc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) { sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.0f; }
#pragma unroll
        for (int kk = 0; kk < KS; kk += 2) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                uint32_t b[4];      // matrices: (positions nt*8.., dims kk*16 + 8j), j = 0..3
                fa_ldsm4(sK_a + (nt * 8 + (lane & 7)) * ROW + (kk * 16 + 8 * (lane >> 3)) * 2, b);
                fa_mma(sc[nt], qf[kk], b[0], b[1]);
                fa_mma(sc[nt], qf[kk + 1], b[2], b[3]);
            }
        }
        // ---- mask, online softmax (rows g and g+8; a row lives in the 4 lanes of a quad)
        float mxa = ma, mxb = mb;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int k = kv0 + nt * 8 + t * 2;
            float2 fa = make_float2(0.0f, 0.0f), fb = fa;
            if (mra) {                                          // positions past n_kv are discarded below: read an in-range pair
                const int kc = min(k, p.n_kv - 2);
                fa = __half22float2(*(const __half2 *) (mra + (int64_t) kc * 2)); fb = __half22float2(*(const __half2 *) (mrb + (int64_t) kc * 2));
            }
            sc[nt][0] = k     < p.n_kv ? sc[nt][0] + fa.x : -INFINITY; sc[nt][1] = k + 1 < p.n_kv ? sc[nt][1] + fa.y : -INFINITY;
            sc[nt][2] = k     < p.n_kv ? sc[nt][2] + fb.x : -INFINITY; sc[nt][3] = k + 1 < p.n_kv ? sc[nt][3] + fb.y : -INFINITY;
            mxa = fmaxf(mxa, fmaxf(sc[nt][0], sc[nt][1])); mxb = fmaxf(mxb, fmaxf(sc[nt][2], sc[nt][3]));
        }
        mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 1)); mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 2));
        mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 1)); mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 2));
        const float ca = __expf(ma - mxa), cb = __expf(mb - mxb);
        ma = mxa; mb = mxb;
        float sa = 0.0f, sb = 0.0f;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            sc[nt][0] = __expf(sc[nt][0] - mxa); sc[nt][1] = __expf(sc[nt][1] - mxa);
            sc[nt][2] = __expf(sc[nt][2] - mxb); sc[nt][3] = __expf(sc[nt][3] - mxb);
            sa += sc[nt][0] + sc[nt][1]; sb += sc[nt][2] + sc[nt][3];
        }
        la = la * ca + sa; lb = lb * cb + sb;
#pragma unroll
        for (int i = 0; i < DT; ++i) { o[i][0] *= ca; o[i][1] *= ca; o[i][2] *= cb; o[i][3] *= cb; }
        // ---- O += P . V  (the accumulator fragments of S are the A fragments of P)
#pragma unroll
        for (int j = 0; j < BN / 16; ++j) {
            uint32_t pa[4];
            pa[0] = fa_pack(sc[2 * j][0], sc[2 * j][1]);         pa[1] = fa_pack(sc[2 * j][2], sc[2 * j][3]);
            pa[2] = fa_pack(sc[2 * j + 1][0], sc[2 * j + 1][1]); pa[3] = fa_pack(sc[2 * j + 1][2], sc[2 * j + 1][3]);
#pragma unroll
            for (int dn = 0; dn < DT; dn += 2) {
                uint32_t b[4];      // matrices: (positions 16j + 8*(i&1).., dims (dn + (i>>1))*8..), transposed on load
                const int mi = lane >> 3;
                fa_ldsm4_t(sV_a + (16 * j + 8 * (mi & 1) + (lane & 7)) * ROW + ((dn + (mi >> 1)) * 8) * 2, b);
                fa_mma(o[dn], pa, b[0], b[1]);
                fa_mma(o[dn + 1], pa, b[2], b[3]);
            }
        }
    }
    // ---- normalise and store: dst f32 [D, H, n_q]
    la += __shfl_xor_sync(0xffffffffu, la, 1); la += __shfl_xor_sync(0xffffffffu, la, 2);
    lb += __shfl_xor_sync(0xffffffffu, lb, 1); lb += __shfl_xor_sync(0xffffffffu, lb, 2);
    const float ia = la > 0.0f ? 1.0f / la : 0.0f, ib = lb > 0.0f ? 1.0f / lb : 0.0f;
    float * da = (float *) (p.dst + (int64_t) h * p.d_nb1 + (int64_t) qa_c * p.d_nb2 + (int64_t) seq * p.d_nb3);
    float * db = (float *) (p.dst + (int64_t) h * p.d_nb1 + (int64_t) qb_c * p.d_nb2 + (int64_t) seq * p.d_nb3);
#pragma unroll
    for (int i = 0; i < DT; ++i) {
        if (qa < p.n_q) *(float2 *) (da + i * 8 + t * 2) = make_float2(o[i][0] * ia, o[i][1] * ia);
        if (qb < p.n_q) *(float2 *) (db + i * 8 + t * 2) = make_float2(o[i][2] * ib, o[i][3] * ib);
    }
}

// ---------------------------------------------------------------- decode attention on the tensor cores (contexts >= 1024 positions)
// Long-context decode with GQA is instruction-bound on the CUDA cores.  Here the <= 8 query columns of a
// chunk (the heads of a GQA group) are the rows of the A operand (rows 8..15 of m16n8k16 stay zero), every warp owns KV
// tiles of 32 positions in its private shared memory (cp.async for f16, block expansion for q8_0 — the loaders of the
// prefill kernel), S = Q.K^T and O += P.V are mma.sync, and each warp ends as one online-softmax state of the shared
// tail (fattn_tail: in-block merge, split-KV partials, last-block combine).  Measured at 32 k positions against the vector
// kernel (profiles/r01c_fattn_decode.txt): Llama-3-70B f16 (gqa 8) 189 -> 99 us, gpt-oss q8_0 130 -> 67 us, Qwen2.5-32B q8_0
// (C3) 127 -> 107 us, Llama-3-8B f16 89 -> 72 us.  Q and P are rounded to f16 (NMSE ~1e-7 against the f32 oracle).
template <int D, int KVT, int NCOLS>
__global__ void __launch_bounds__(128) fattn_dec_mma_kernel(const fattn_params p) {
    constexpr int BN = 32, ROW = D * 2 + 16, KS = D / 16, NT = BN / 8, DT = D / 8, CPR = D / 8, ST = D + 2, NW = 4;
    extern __shared__ __align__(16) uint8_t dsm[];
    pdl_launch_dependents();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int split = blockIdx.x, grp = blockIdx.y, seq = blockIdx.z;
    const int hkv = grp / p.n_chunks, chunk = grp % p.n_chunks;
    uint8_t * sK = dsm + warp * (2 * BN * ROW), * sV = sK + BN * ROW;
    float * sm = (float *) (dsm + NW * 2 * BN * ROW);
    const uint32_t sK_a = (uint32_t) __cvta_generic_to_shared(sK), sV_a = (uint32_t) __cvta_generic_to_shared(sV);

    // row g of the A operand = column g of the chunk (rows of inactive columns carry q = 0 and are never written out)
    const bool cvalid = g < NCOLS && chunk * NCOLS + g < p.ncols_total;
    const int ci = cvalid ? chunk * NCOLS + g : chunk * NCOLS;
    const int cj = ci / p.gqa, ch = hkv * p.gqa + ci % p.gqa;
    pdl_wait();
    uint32_t qf[KS][4];
    {
        const float * qp = (const float *) (p.q + (int64_t) cj * p.q_nb1 + (int64_t) ch * p.q_nb2 + (int64_t) seq * p.q_nb3);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
            const int d = kk * 16 + t * 2;
            const float2 a0 = *(const float2 *) (qp + d), a2 = *(const float2 *) (qp + d + 8);
            qf[kk][0] = cvalid ? fa_pack(a0.x * p.scale, a0.y * p.scale) : 0u; qf[kk][1] = 0u;
            qf[kk][2] = cvalid ? fa_pack(a2.x * p.scale, a2.y * p.scale) : 0u; qf[kk][3] = 0u;
        }
    }
    const char * mr = p.mask ? p.mask + (int64_t) cj * p.m_nb1 + (int64_t) (ch % p.m_ne2) * p.m_nb2 + (int64_t) (seq % p.m_ne3) * p.m_nb3 : nullptr;
    const int kv0 = split * p.kv_per_split, kv1 = min(p.n_kv, kv0 + p.kv_per_split);
    const char * kb = p.k + (int64_t) hkv * p.k_nb2 + (int64_t) seq * p.k_nb3;
    const char * vb = p.v + (int64_t) hkv * p.v_nb2 + (int64_t) seq * p.v_nb3;

    constexpr float M_NONE = -1.0e30f;
    float ma = M_NONE, la = 0.0f;
    float o[DT][4];
#pragma unroll
    for (int i = 0; i < DT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.0f; }

    for (int base = kv0 + warp * BN; base < kv1; base += NW * BN) {
        __syncwarp();                                            // the warp has consumed its previous tile
        if (KVT == B200_Q8_0) {
            constexpr int BPR = D / 32;
            for (int i = lane; i < 2 * BN * BPR; i += 32) {
                const bool isv = i >= BN * BPR;
                const int ii = isv ? i - BN * BPR : i, r = ii / BPR, j = ii % BPR, kr = base + r;
                uint4 * dstp = (uint4 *) ((isv ? sV : sK) + r * ROW + j * 64);
                if (kr < kv1) {
                    const uintptr_t a = (uintptr_t) ((isv ? vb + (int64_t) kr * p.v_nb1 : kb + (int64_t) kr * p.k_nb1) + 34 * j);
                    const uint32_t sh = (uint32_t) (a & 3) * 8;
                    const uint32_t * q = (const uint32_t *) (a & ~(uintptr_t) 3);
                    uint32_t w[10], x[9];
#pragma unroll
                    for (int u = 0; u < 9; ++u) w[u] = q[u];
                    w[9] = 0u;
#pragma unroll
                    for (int u = 0; u < 9; ++u) x[u] = __funnelshift_r(w[u], w[u + 1], sh);
                    const uint16_t dbits = (uint16_t) (x[0] & 0xFFFFu);
                    const __half2 d2 = __half2half2(*(const __half *) &dbits), off = __float2half2_rn(1152.0f);
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        uint32_t ov[4];
#pragma unroll
                        for (int v = 0; v < 2; ++v) {
                            const uint32_t y = __funnelshift_r(x[2 * u + v], x[2 * u + v + 1], 16) ^ 0x80808080u;
                            const uint32_t lo = __byte_perm(y, 0x64646464u, 0x4140), hi = __byte_perm(y, 0x64646464u, 0x4342);
                            const __half2 hl = __hmul2(__hsub2(*(const __half2 *) &lo, off), d2), hh = __hmul2(__hsub2(*(const __half2 *) &hi, off), d2);
                            ov[2 * v] = *(const uint32_t *) &hl; ov[2 * v + 1] = *(const uint32_t *) &hh;
                        }
                        dstp[u] = make_uint4(ov[0], ov[1], ov[2], ov[3]);
                    }
                } else {
#pragma unroll
                    for (int u = 0; u < 4; ++u) dstp[u] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        } else {
            for (int i = lane; i < BN * CPR; i += 32) {
                const int r = i / CPR, c = i % CPR, kr = base + r;
                const uint32_t dk = sK_a + r * ROW + c * 16, dv = sV_a + r * ROW + c * 16;
                if (kr < kv1) {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dk), "l"(kb + (int64_t) kr * p.k_nb1 + c * 16) : "memory");
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dv), "l"(vb + (int64_t) kr * p.v_nb1 + c * 16) : "memory");
                } else {
                    *(uint4 *) (sK + r * ROW + c * 16) = make_uint4(0u, 0u, 0u, 0u);
                    *(uint4 *) (sV + r * ROW + c * 16) = make_uint4(0u, 0u, 0u, 0u);
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();

        float sc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) { sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.0f; }
#pragma unroll
        for (int kk = 0; kk < KS; kk += 2) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                uint32_t b[4];
                fa_ldsm4(sK_a + (nt * 8 + (lane & 7)) * ROW + (kk * 16 + 8 * (lane >> 3)) * 2, b);
                fa_mma(sc[nt], qf[kk], b[0], b[1]);
                fa_mma(sc[nt], qf[kk + 1], b[2], b[3]);
            }
        }
        float mx = ma;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int k = base + nt * 8 + t * 2;
            float2 f = make_float2(0.0f, 0.0f);
            if (mr) f = __half22float2(*(const __half2 *) (mr + (int64_t) min(k, p.n_kv - 2) * 2));
            sc[nt][0] = k     < kv1 ? sc[nt][0] + f.x : -INFINITY;
            sc[nt][1] = k + 1 < kv1 ? sc[nt][1] + f.y : -INFINITY;
            mx = fmaxf(mx, fmaxf(sc[nt][0], sc[nt][1]));
        }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1)); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float ca = __expf(ma - mx);
        ma = mx;
        float sa = 0.0f;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            sc[nt][0] = __expf(sc[nt][0] - mx); sc[nt][1] = __expf(sc[nt][1] - mx);
            sa += sc[nt][0] + sc[nt][1];
        }
        la = la * ca + sa;
#pragma unroll
        for (int i = 0; i < DT; ++i) { o[i][0] *= ca; o[i][1] *= ca; }
#pragma unroll
        for (int j = 0; j < BN / 16; ++j) {
            uint32_t pa[4];
            pa[0] = fa_pack(sc[2 * j][0], sc[2 * j][1]); pa[1] = 0u;
            pa[2] = fa_pack(sc[2 * j + 1][0], sc[2 * j + 1][1]); pa[3] = 0u;
#pragma unroll
            for (int dn = 0; dn < DT; dn += 2) {
                uint32_t b[4];
                const int mi = lane >> 3;
                fa_ldsm4_t(sV_a + (16 * j + 8 * (mi & 1) + (lane & 7)) * ROW + ((dn + (mi >> 1)) * 8) * 2, b);
                fa_mma(o[dn], pa, b[0], b[1]);
                fa_mma(o[dn + 1], pa, b[2], b[3]);
            }
        }
    }
    la += __shfl_xor_sync(0xffffffffu, la, 1); la += __shfl_xor_sync(0xffffffffu, la, 2);
    // ---- every warp is one state of the shared tail
    if (g < NCOLS) {
        float * os = sm + ((int64_t) warp * NCOLS + g) * ST;
#pragma unroll
        for (int i = 0; i < DT; ++i) { os[i * 8 + t * 2] = o[i][0]; os[i * 8 + t * 2 + 1] = o[i][1]; }
        if (t == 0) { os[D] = ma <= 0.5f * M_NONE ? -INFINITY : ma; os[D + 1] = la; }
    }
    __syncthreads();
    fattn_tail<D, NCOLS, NW>(p, sm, split, grp, seq, hkv, chunk, lane, warp);
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

// many query positions, f16 cache, plain scale + mask: the tensor-core kernel
static bool fattn_use_mma(const b200_fattn_args * a) {
    const b200_tensor & q = a->q; const b200_tensor & k = a->k; const b200_tensor & v = a->v; const b200_tensor & d = a->dst;
    const int64_t D = q.ne[0], n_q = q.ne[1], n_kv = k.ne[1];
    if (n_q < 16 || (D != 64 && D != 128) || a->sinks || a->logit_softcap != 0.0f || a->max_bias != 0.0f) return false;
    // this kernel has no split-KV: a CTA walks all n_kv positions of its 64 query rows.  A small batch (16-63 tokens of
    // speculative / parallel decode) at long context would leave most SMs idle: it stays on the split-KV decode path
    if (n_kv > 2048 && ((n_q + 63) / 64) * q.ne[2] * q.ne[3] < 74) return false;
    if (k.type == B200_Q8_0) {          // blocks are read as aligned words around their 2-byte aligned start
        if (((uintptr_t) k.data | (uintptr_t) v.data) % 4) return false;
        for (int i = 1; i < 4; ++i) if (k.nb[i] % 4 || v.nb[i] % 4) return false;
    } else if (k.type != B200_F16) return false;
    if (n_kv < 2 || n_kv % 2 || q.ne[2] > 65535 || q.ne[3] > 65535) return false;
    if (((uintptr_t) q.data | (uintptr_t) d.data) % 8) return false;
    for (int i = 1; i < 4; ++i) if (q.nb[i] % 8 || d.nb[i] % 8) return false;
    if (a->mask && (((uintptr_t) a->mask->data) % 4 || a->mask->nb[1] % 4 || a->mask->nb[2] % 4 || a->mask->nb[3] % 4)) return false;
    return true;
}

// long-context decode (>= 1024 positions) of up to 8 columns on the tensor cores; B200_FATTN_DECODE_MMA=0 keeps the vector kernel
static bool fattn_use_dec_mma(const b200_fattn_args * a) {
    static const bool on = [] { const char * e = getenv("B200_FATTN_DECODE_MMA"); return !(e && atoi(e) == 0); }();   // =0: vector kernel only
    if (!on) return false;
    const b200_tensor & q = a->q; const b200_tensor & k = a->k; const b200_tensor & v = a->v;
    const int64_t D = q.ne[0], n_kv = k.ne[1];
    // from 1024 positions on; with four or more query columns per KV head (GQA) from 256 on: the vector kernel with 4-8 columns is
    // ~8 k instructions of unrolled code that runs cold in every layer (the SM instruction cache holds ~32 KB): gpt-oss decode
    // (gqa 8, 256 positions) spent 31 us per layer in it
    const int64_t cols = q.ne[1] * (q.ne[2] / k.ne[2]);
    if ((D != 64 && D != 128) || a->logit_softcap != 0.0f || a->max_bias != 0.0f || n_kv < (cols >= 4 ? 256 : 1024) || n_kv % 2) return false;
    if (k.type == B200_Q8_0) {
        if (((uintptr_t) k.data | (uintptr_t) v.data) % 4) return false;
        for (int i = 1; i < 4; ++i) if (k.nb[i] % 4 || v.nb[i] % 4) return false;
    } else if (k.type != B200_F16) return false;
    if (((uintptr_t) q.data) % 8) return false;
    for (int i = 1; i < 4; ++i) if (q.nb[i] % 8) return false;
    if (a->mask && (((uintptr_t) a->mask->data) % 4 || a->mask->nb[1] % 4 || a->mask->nb[2] % 4 || a->mask->nb[3] % 4)) return false;
    return true;
}
template <int D, int KVT, int NCOLS>
static int fattn_launch_dec_mma_n(const fattn_params & p, dim3 grid, cudaStream_t st) {
    const size_t smem = (size_t) 4 * 2 * 32 * (D * 2 + 16) + std::max((size_t) 4 * NCOLS * (D + 2) * 4, (size_t) NCOLS * p.n_split * 4);
    static std::once_flag attr_once[64];                       // per device; launches may come from several host threads
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_ARG;
    cudaError_t attr_rc = cudaSuccess;
    std::call_once(attr_once[dev], [&] { attr_rc = cudaFuncSetAttribute(fattn_dec_mma_kernel<D, KVT, NCOLS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024); });
    B200_CUDA_TRY(attr_rc);
    fattn_dec_mma_kernel<D, KVT, NCOLS><<<grid, 128, smem, st>>>(p);
    return B200_OK;
}
template <int D, int KVT>
static int fattn_launch_dec_mma(const fattn_params & p, int ncols, dim3 grid, cudaStream_t st) {
    switch (ncols) {
        case 1: return fattn_launch_dec_mma_n<D, KVT, 1>(p, grid, st);
        case 2: return fattn_launch_dec_mma_n<D, KVT, 2>(p, grid, st);
        case 3: return fattn_launch_dec_mma_n<D, KVT, 3>(p, grid, st);
        case 4: return fattn_launch_dec_mma_n<D, KVT, 4>(p, grid, st);
        case 5: return fattn_launch_dec_mma_n<D, KVT, 5>(p, grid, st);
        case 6: return fattn_launch_dec_mma_n<D, KVT, 6>(p, grid, st);
        case 7: return fattn_launch_dec_mma_n<D, KVT, 7>(p, grid, st);
        default: return fattn_launch_dec_mma_n<D, KVT, 8>(p, grid, st);
    }
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
    cudaStream_t st = (cudaStream_t) stream;
    if (fattn_use_mma(a)) {
        const int D = (int) q.ne[0];
        dim3 grid((unsigned) ((p.n_q + 63) / 64), (unsigned) p.H, (unsigned) q.ne[3]);
        const size_t smem = (size_t) 2 * 64 * (D * 2 + 16);
        if (k.type == B200_F16) { if (D == 64) fattn_mma_kernel<64, B200_F16><<<grid, 128, smem, st>>>(p); else fattn_mma_kernel<128, B200_F16><<<grid, 128, smem, st>>>(p); }
        else                    { if (D == 64) fattn_mma_kernel<64, B200_Q8_0><<<grid, 128, smem, st>>>(p); else fattn_mma_kernel<128, B200_Q8_0><<<grid, 128, smem, st>>>(p); }
        count_launch();
        return check_launch("flash_attn_ext (mma)");
    }
    dim3 grid((unsigned) pl.n_split, (unsigned) (p.H_kv * pl.n_chunks), (unsigned) q.ne[3]);
    if (pl.n_split > 1)   // arrival counters of the split-KV combine: zeroed per launch (a memset node under graph capture)
        B200_CUDA_TRY(cudaMemsetAsync(p.counters, 0, pl.counter_bytes, st));
    const int D = (int) q.ne[0];
    if (fattn_use_dec_mma(a)) {
        int rc;
        if (k.type == B200_F16) rc = D == 64 ? fattn_launch_dec_mma<64, B200_F16>(p, pl.ncols, grid, st) : fattn_launch_dec_mma<128, B200_F16>(p, pl.ncols, grid, st);
        else                    rc = D == 64 ? fattn_launch_dec_mma<64, B200_Q8_0>(p, pl.ncols, grid, st) : fattn_launch_dec_mma<128, B200_Q8_0>(p, pl.ncols, grid, st);
        if (rc != B200_OK) return rc;
        count_launch();
        return check_launch("flash_attn_ext (decode mma)");
    }
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

// ---- FLASH_ATTN_EXT without any allocation on the caller's side -------------------------------------------------------
// Same kernels as b200_flash_attn_ext; the split-KV scratch (arrival counters + partials) is a file-scope device array
// instead of a caller-provided buffer.  Every GPU that loads the library holds B200_FATTN_STATIC_SCRATCH_BYTES until exit.
// All calls on a device share the array: calls that overlap on the GPU (different streams) race.  It keeps the previous
// call's contents; the only part that must start at zero, the arrival counters, is cleared by the launcher on the stream
// (cudaMemsetAsync on existing memory: not an allocation).
static __device__ __align__(256) unsigned char fattn_static_scratch[B200_FATTN_STATIC_SCRATCH_BYTES];

extern "C" int b200_flash_attn_ext_static(const b200_fattn_args * a, void * stream) {
    fattn_plan pl;
    if (!a) return B200_ERR_ARG;
    if (!fattn_make_plan(a, pl)) return B200_ERR_UNSUPPORTED;
    const size_t need = pl.counter_bytes + pl.part_bytes;
    if (need > (size_t) B200_FATTN_STATIC_SCRATCH_BYTES) {
        set_error("flash_attn_ext_static: this shape needs %zu bytes of scratch, the static array holds %zu (use b200_flash_attn_ext with a caller buffer)",
                  need, (size_t) B200_FATTN_STATIC_SCRATCH_BYTES);
        return B200_ERR_ARG;
    }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that
    // device current.
    static std::mutex mu;
    static void * dev_ptr[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_ARG;
    void * ptr;
    {
        std::lock_guard<std::mutex> g(mu);
        if (!dev_ptr[dev]) {
            void * sym = nullptr;
            B200_CUDA_TRY(cudaGetSymbolAddress(&sym, fattn_static_scratch));
            dev_ptr[dev] = sym;
        }
        ptr = dev_ptr[dev];
    }
    b200_fattn_args b = *a;
    b.scratch = ptr; b.scratch_size = (size_t) B200_FATTN_STATIC_SCRATCH_BYTES;
    return b200_flash_attn_ext(&b, stream);
}
