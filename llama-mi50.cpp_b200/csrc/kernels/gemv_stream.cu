// gemv_stream.cu — K4s: streaming fused GEMV for decode (C-ABI b200_gemv_fused).
//
// Replaces, in ONE launch, the reference's per-matmul sequence
//   rms_norm (norm.cu:476) -> mul -> quantize_q8_1 (quantize.cu:4-48) -> mul_mat_vec_q (mmvq.cu:140-356) [-> add | glu]
// for up to 4 weight matrices that share one activation column.
//
// B200 design (see DESIGN.md "K4s"):
//   * persistent CTAs, one (or two co-resident across consecutive launches) per SM; the rows of all
//     segments are split across CTAs by BYTES, so every SM streams the same amount of HBM;
//   * a producer warp moves weight rows HBM -> shared memory with 1-D TMA bulk copies
//     (cp.async.bulk ... mbarrier::complete_tx) into a 3-stage ring; alignment of the odd block formats
//     (18/34/17/210-byte blocks) is paid on shared memory reads, never on HBM transactions;
//   * 8 consumer warps do dp4a dot products out of shared memory against the q8_1 activation held in
//     shared memory (quantised once per CTA in the prologue, bank-conflict-free padded layout);
//   * a finalizer warp reduces the per-lane partial sums and applies the epilogue (bias/residual, GLU)
//     so consumers never wait on a shuffle chain or a CTA-wide barrier in the main loop;
//   * programmatic dependent launch: the producer starts prefetching weights before the previous kernel
//     in the stream has finished; only the activation read waits (griddepcontrol.wait).
#include <algorithm>
#include <cstring>

#include "quantize.cuh"
#include "vecdot.cuh"

namespace b200 {

constexpr int GS_NW          = 8;                      // consumer warps
constexpr int GS_THREADS     = (GS_NW + 2) * 32;       // + producer warp + finalizer warp
constexpr int GS_PRO_THREADS = (GS_NW + 1) * 32;       // warps that run the prologue (all but the producer)
constexpr int GS_STAGES      = 3;                      // default ring depth
constexpr int GS_MAX_STAGES  = 8;
constexpr int GS_STAGE_BYTES = 24 * 1024;              // preferred payload capacity of a stage (runtime: gs_params::stage_stride)
constexpr int GS_STAGE_SLACK = 128;                    // slack for 16 B alignment of both ends / funnel over-read
constexpr int GS_RED_FLOATS  = 1280;                   // per parity: rows * (KS*32 + 4) <= GS_RED_FLOATS
constexpr int GS_SMEM_2CTA   = 112 * 1024;             // budget that still lets two CTAs share an SM (PDL overlap)
constexpr int GS_SMEM_MAX    = 224 * 1024;
constexpr int GS_RB          = 2;                      // rows per warp pass (activation registers reused across them)

struct gs_seg {
    const uint8_t * w[2];     // [1] = paired "up" matrix of a GLU work segment (else null)
    const float *   bias;
    float *         dst;
    int64_t         byte_prefix;   // bytes of the work segments before this one
    int32_t M, rb, type, nsub;     // rows, row bytes, ggml type, 1 (or 2 for a GLU pair)
    int32_t R, KS, spr, ipr;       // rows per chunk, warps along K, steps per row, items per row
    int32_t spw, _pad;             // K-steps per warp (1, 2 or 4): the per-step activation registers are hoisted
};
struct gs_params {
    gs_seg seg[B200_GEMV_MAX_SEG];
    int32_t n_ws;                  // work segments
    int32_t glu_op; float glu_alpha, glu_limit;
    const float * x; const float * norm_w; float eps; int32_t K;
    blk_q8_1 * xq_out;
    int64_t total_bytes;
    int32_t act_qs_bytes;
    int32_t stage_stride;          // bytes between ring stages (payload + GS_STAGE_SLACK, multiple of 128)
    int32_t n_stages;              // ring depth
    int32_t flags;                 // tuning/debug: bit 0 = consumers skip the dot products (bandwidth probe)
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t * b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t * b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t * b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * b, uint32_t parity) {
    const uint32_t a = smem_u32(b);
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier; weights are read once: evict-first in L2
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// N consecutive little-endian u32 words starting at an arbitrarily aligned shared-memory byte address
template <int N>
__device__ __forceinline__ void lds_words(const uint8_t * p, uint32_t * out) {
    const uint32_t a  = smem_u32(p);
    const uint32_t sh = (a & 3u) * 8u;
    const uint32_t * q = (const uint32_t *) (p - (a & 3u));
    uint32_t w[N + 1];
#pragma unroll
    for (int i = 0; i <= N; ++i) w[i] = q[i];
#pragma unroll
    for (int i = 0; i < N; ++i) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}

// activation layout in shared memory: unit u (32 int8) at byte u*32 + (u/4)*16  (16 B pad every 128 B:
// the 8 lanes of a quarter-warp then hit 8 distinct 16-byte bank groups for every item shape used below)
__device__ __forceinline__ int act_off(int u) { return u * 32 + (u >> 2) * 16; }

// ---------------------------------------------------------------- per-type ITEMS
// An ITEM is the work of one lane in one K-step of a row: 64 weights (k-quants) or 32 weights.  The lane
// handles the SAME item index for every row it ever sees in a segment, so everything that depends only on
// the item — the q8_1 activation values, their scales, field offsets, scale-decode shifts — is loaded ONCE
// per segment into registers (item<TYPE>::init); the per-row work is then loads of weight bytes + dp4a.
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {      // unsigned bytes x signed bytes
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

template <int TYPE> struct item;

// ---- Q4_K: item = (super-block sb, 32-byte nibble group p): sub-blocks 2p (low nibbles) and 2p+1 (high nibbles)
template <> struct item<B200_Q4_K> {
    static constexpr int wpi = 64, max_spw = 2;
    int au[16]; float d0, s0, d1, s1;     // d1 is pre-divided by 16: the high nibbles are used unshifted (x16)
    int hoff, qoff; uint32_t shs, shm, m1, m2; bool hi, ok;
    __device__ __forceinline__ void init(int it, int ipr, const uint8_t * aq, const float2 * ads) {
        ok = it < ipr;
        const int itc = ok ? it : 0;
        const int sb = itc >> 2, p = itc & 3, u0 = sb * 8 + 2 * p;
        const int4 a0 = *(const int4 *) (aq + act_off(u0)),      a1 = *(const int4 *) (aq + act_off(u0) + 16);
        const int4 a2 = *(const int4 *) (aq + act_off(u0) + 32), a3 = *(const int4 *) (aq + act_off(u0) + 48);
        au[0] = a0.x; au[1] = a0.y; au[2] = a0.z; au[3] = a0.w; au[4] = a1.x; au[5] = a1.y; au[6] = a1.z; au[7] = a1.w;
        au[8] = a2.x; au[9] = a2.y; au[10] = a2.z; au[11] = a2.w; au[12] = a3.x; au[13] = a3.y; au[14] = a3.z; au[15] = a3.w;
        const float4 ds = *(const float4 *) (ads + u0);
        d0 = ds.x; s0 = ds.y; d1 = ds.z * 0.0625f; s1 = ds.w;
        hoff = sb * 144; qoff = hoff + 16 + 32 * p;      // super-block header {dm, scales[12]}; this item's 32 nibble bytes
        // 6-bit (scale, min) of sub-blocks 2p, 2p+1 (get_scale_min_k4, ggml-quants.c): branch-free per-lane constants
        hi = p >= 2;
        shs = hi ? 16 * (p - 2) : 16 * p;    // byte pair position inside the source word
        shm = hi ? shs + 4 : shs;
        m1 = hi ? 0x0F0Fu : 0x3F3Fu; m2 = hi ? 0x3030u : 0u;
    }
    template <int RB>
    __device__ __forceinline__ void dot(const uint8_t * const * rowp, float * acc) const {
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint4 hdr = *(const uint4 *) (rowp[j] + hoff);
            const uint4 q0 = *(const uint4 *) (rowp[j] + qoff), q1 = *(const uint4 *) (rowp[j] + qoff + 16);
            const uint32_t q[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
            int sl = 0, sh = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                sl = __dp4a((int) (q[i] & 0x0F0F0F0Fu), au[i], sl);
                sh = dp4a_us(q[i] & 0xF0F0F0F0u, au[8 + i], sh);
            }
            const uint32_t xs = hi ? hdr.w : hdr.y, xm = hi ? hdr.w : hdr.z;
            const uint32_t sc = ((xs >> shs) & m1) | ((hdr.y >> (shs + 2)) & m2);
            const uint32_t mn = ((xm >> shm) & m1) | ((hdr.z >> (shs + 2)) & m2);
            const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
            const float t = (float) ((int) (sc & 0xFF) * sl) * d0 + (float) ((int) (sc >> 8) * sh) * d1;
            const float m = (float) (mn & 0xFF) * s0 + (float) (mn >> 8) * s1;
            acc[j] += dm.x * t - dm.y * m;
        }
    }
};

// ---- Q6_K: item = (super-block, half h, 16-column group c): 16 weights of each of the 4 sub-blocks of the half.
// q = lo4 | hi2<<4 (0..63), value q-32:  sum (q-32)u = sum lo4*u + 16*sum hi2*u - 32*sum u; nibbles / bit pairs are used
// unshifted with unsigned dp4a and the power-of-two factors folded into per-item float constants.
template <> struct item<B200_Q6_K> {
    static constexpr int wpi = 64, max_spw = 2;
    int a[4][4]; float dl[4], dh[4], su[4];   // per sub-block r: d8/16^(r>>1), d8*16/4^r, 32*d8*sum(u over the 16 columns)
    int off_l, off_h, off_s, off_d; bool ok;
    __device__ __forceinline__ void init(int it, int ipr, const uint8_t * aq, const float2 * ads) {
        // lane -> quad (sb, h) in the low 4 lane bits, column group c in bit 4 (keeps activation reads conflict-free)
        const int step = it >> 5, l = it & 31;
        int quad = step * 16 + (l & 15);
        const int c = l >> 4;
        ok = quad < (ipr >> 1);
        if (!ok) quad = 0;
        const int sb = quad >> 1, h = quad & 1, u0 = sb * 8 + 4 * h;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int4 t = *(const int4 *) (aq + act_off(u0 + r) + 16 * c);
            a[r][0] = t.x; a[r][1] = t.y; a[r][2] = t.z; a[r][3] = t.w;
            const float d8 = ads[u0 + r].x;
            int s = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) s = __dp4a(0x01010101, a[r][i], s);
            dl[r] = (r >> 1) ? d8 * 0.0625f : d8;
            dh[r] = d8 * 16.0f / (float) (1 << (2 * r));
            su[r] = 32.0f * d8 * (float) s;
        }
        off_l = sb * 210 + 64 * h + 16 * c; off_h = sb * 210 + 128 + 32 * h + 16 * c;
        off_s = sb * 210 + 192 + 8 * h + c; off_d = sb * 210 + 208;
    }
    template <int RB>
    __device__ __forceinline__ void dot(const uint8_t * const * rowp, float * acc) const {
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint8_t * b = rowp[j];
            uint32_t la[4], lb[4], hh[4];
            lds_words<4>(b + off_l, la);            // ql: low nibbles -> r=0, high -> r=2
            lds_words<4>(b + off_l + 32, lb);       //                    r=1,         r=3
            lds_words<4>(b + off_h, hh);            // qh: 2 bits per r
            const int8_t * sc = (const int8_t *) (b + off_s);
            const float d = __half2float(*(const __half *) (b + off_d));
            float part = 0.0f;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                int sl = 0, sh = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t w = (r & 1) ? lb[i] : la[i];
                    sl = dp4a_us(w & ((r >> 1) ? 0xF0F0F0F0u : 0x0F0F0F0Fu), a[r][i], sl);
                    sh = dp4a_us(hh[i] & (0x03030303u << (2 * r)), a[r][i], sh);
                }
                part += (float) sc[2 * r] * ((float) sl * dl[r] + (float) sh * dh[r] - su[r]);
            }
            acc[j] += d * part;
        }
    }
};

// ---- 32-weight block formats: item = block index
template <> struct item<B200_Q4_0> {
    static constexpr int wpi = 32, max_spw = 4;
    int au[8]; float d8, d8h, s8; int off; bool ok;
    __device__ __forceinline__ void init(int it, int ipr, const uint8_t * aq, const float2 * ads) {
        ok = it < ipr; const int u = ok ? it : 0;
        const int4 a0 = *(const int4 *) (aq + act_off(u)), a1 = *(const int4 *) (aq + act_off(u) + 16);
        au[0] = a0.x; au[1] = a0.y; au[2] = a0.z; au[3] = a0.w; au[4] = a1.x; au[5] = a1.y; au[6] = a1.z; au[7] = a1.w;
        const float2 ds = ads[u];
        d8 = ds.x; d8h = ds.x * 0.0625f; s8 = 8.0f * ds.y; off = u * 18;
    }
    template <int RB>
    __device__ __forceinline__ void dot(const uint8_t * const * rowp, float * acc) const {
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint8_t * b = rowp[j] + off;
            uint32_t q[4];
            lds_words<4>(b + 2, q);
            int sl = 0, sh = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                sl = __dp4a((int) (q[i] & 0x0F0F0F0Fu), au[i], sl);
                sh = dp4a_us(q[i] & 0xF0F0F0F0u, au[4 + i], sh);
            }
            const float d = __half2float(*(const __half *) b);
            acc[j] += d * ((float) sl * d8 + (float) sh * d8h - s8);      // sum (q-8)u = sum q*u - 8*sum u
        }
    }
};

template <> struct item<B200_Q8_0> {
    static constexpr int wpi = 32, max_spw = 4;
    int au[8]; float d8; int off; bool ok;
    __device__ __forceinline__ void init(int it, int ipr, const uint8_t * aq, const float2 * ads) {
        ok = it < ipr; const int u = ok ? it : 0;
        const int4 a0 = *(const int4 *) (aq + act_off(u)), a1 = *(const int4 *) (aq + act_off(u) + 16);
        au[0] = a0.x; au[1] = a0.y; au[2] = a0.z; au[3] = a0.w; au[4] = a1.x; au[5] = a1.y; au[6] = a1.z; au[7] = a1.w;
        d8 = ads[u].x; off = u * 34;
    }
    template <int RB>
    __device__ __forceinline__ void dot(const uint8_t * const * rowp, float * acc) const {
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint8_t * b = rowp[j] + off;
            uint32_t q[8];
            lds_words<8>(b + 2, q);
            int s = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) s = __dp4a((int) q[i], au[i], s);
            acc[j] += __half2float(*(const __half *) b) * (float) s * d8;
        }
    }
};

template <> struct item<B200_MXFP4> {
    static constexpr int wpi = 32, max_spw = 4;
    int au[8]; float d8; int off; bool ok;
    __device__ __forceinline__ void init(int it, int ipr, const uint8_t * aq, const float2 * ads) {
        ok = it < ipr; const int u = ok ? it : 0;
        const int4 a0 = *(const int4 *) (aq + act_off(u)), a1 = *(const int4 *) (aq + act_off(u) + 16);
        au[0] = a0.x; au[1] = a0.y; au[2] = a0.z; au[3] = a0.w; au[4] = a1.x; au[5] = a1.y; au[6] = a1.z; au[7] = a1.w;
        d8 = ads[u].x; off = u * 17;
    }
    template <int RB>
    __device__ __forceinline__ void dot(const uint8_t * const * rowp, float * acc) const {
#pragma unroll
        for (int j = 0; j < RB; ++j) {
            const uint8_t * b = rowp[j] + off;
            uint32_t q[4];
            lds_words<4>(b + 1, q);
            int s = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int lo, hi;
                mxfp4_expand(q[i], lo, hi);
                s = __dp4a(lo, au[i], s);
                s = __dp4a(hi, au[4 + i], s);
            }
            acc[j] += e8m0_half_to_f32(b[0]) * (float) s * d8;
        }
    }
};

__device__ __forceinline__ float gs_glu(int op, float g, float x, float alpha, float limit) {
    switch (op) {
        case B200_GLU_REGLU:  return fmaxf(g, 0.0f) * x;
        case B200_GLU_GEGLU: { const float c = 0.79788456080286535587989211986876f, a = 0.044715f;
                               return 0.5f * g * (1.0f + tanhf(c * g * (1.0f + a * g * g))) * x; }
        case B200_GLU_SWIGLU: return g / (1.0f + expf(-g)) * x;
        case B200_GLU_SWIGLU_OAI: g = fminf(g, limit); x = fmaxf(fminf(x, limit), -limit); return g / (1.0f + expf(-g * alpha)) * (1.0f + x);
        case B200_GLU_GEGLU_ERF:   return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)) * x;
        case B200_GLU_GEGLU_QUICK: return g / (1.0f + expf(-1.702f * g)) * x;
        default: return x;
    }
}

// ring / reduction-buffer position carried by every role across the segments of a launch
struct gs_pos {
    int st = 0, full_par = 0;      // ring stage, parity of its "full" phase
    int rp = 0, red_par = 0;       // reduction buffer, parity of its phase
    int ns = GS_STAGES;
    int dbg = 0;
    __device__ __forceinline__ void next() {
        ++dbg;
        if (++st == ns) { st = 0; full_par ^= 1; }
        if (rp) red_par ^= 1;
        rp ^= 1;
    }
};

// timeline probe (flags bit 2): CTA 0 records clock64() stamps per chunk into xq_out viewed as long long[3][4096][4]
#define GS_STAMP(role, n, k) do { if ((p.flags & 4) && blockIdx.x == 0 && (n) < 4096) ((long long *) p.xq_out)[((role) * 4096 + (n)) * 4 + (k)] = clock64(); } while (0)

struct gs_shared {
    uint64_t full[GS_MAX_STAGES], empty[GS_MAX_STAGES], rfull[2], rempty[2];
    float part[GS_NW + 1];
};

// consumer warps: all chunks of one work segment
template <int TYPE, int SPW>
__device__ __forceinline__ void consume_segment(const gs_params & p, const gs_seg & sg, int r0, int r1, gs_pos & pos, gs_shared & sh,
                                                uint8_t * ring, float * red, const uint8_t * aq, const float2 * ads, int warp, int lane) {
    const int KS = sg.KS, ksl = 31 - __clz(KS), ks = warp & (KS - 1), rg = warp >> ksl, RG = GS_NW >> ksl;
    item<TYPE> itm[SPW];
#pragma unroll
    for (int s = 0; s < SPW; ++s) itm[s].init((ks + s * KS) * 32 + lane, sg.ipr, aq, ads);
    const int rstride = KS * 32 + 4, rbase = ks * 32 + lane, vstep = RG * GS_RB, rb = sg.rb, R = sg.R;
    const bool pair = sg.nsub == 2;
    uint32_t ga0 = (uint32_t) ((uintptr_t) (sg.w[0] + (int64_t) r0 * rb) & 15);
    uint32_t ga1 = pair ? (uint32_t) ((uintptr_t) (sg.w[1] + (int64_t) r0 * rb) & 15) : 0u;
    const uint32_t gstep = (uint32_t) R * (uint32_t) rb;
    for (int row = r0; row < r1; row += R) {
        const int nr = min(R, r1 - row), nv = pair ? 2 * nr : nr;
        const uint8_t * sub0 = ring + pos.st * p.stage_stride + ga0;
        const uint8_t * sub1 = sub0 - ga0 + ((ga0 + (uint32_t) nr * rb + 15u) & ~15u) + ga1;
        float * r = red + pos.rp * GS_RED_FLOATS + rbase;
        mbar_wait(&sh.full[pos.st], pos.full_par);
        if (warp == 0 && lane == 0) GS_STAMP(1, pos.dbg, 0);
        bool red_free = false;
        for (int vb = rg * GS_RB; vb < nv; vb += vstep) {
            const uint8_t * rowp[GS_RB];
#pragma unroll
            for (int j = 0; j < GS_RB; ++j) {
                const int v = min(vb + j, nv - 1);                   // clamp: a duplicate row is computed and dropped
                rowp[j] = v < nr ? sub0 + v * rb : sub1 + (v - nr) * rb;
            }
            float acc[GS_RB];
#pragma unroll
            for (int j = 0; j < GS_RB; ++j) acc[j] = 0.0f;
            if (!(p.flags & 1)) {
#pragma unroll
                for (int s = 0; s < SPW; ++s)
                    if (itm[s].ok) itm[s].template dot<GS_RB>(rowp, acc);
            }
            if (!red_free) { mbar_wait(&sh.rempty[pos.rp], pos.red_par ^ 1); red_free = true; }
#pragma unroll
            for (int j = 0; j < GS_RB; ++j)
                if (vb + j < nv) r[(vb + j) * rstride] = acc[j];
        }
        // every warp must observe the reduction buffer being released, even one without rows in this chunk:
        // its arrival below must not land in the buffer's previous phase
        if (!red_free) mbar_wait(&sh.rempty[pos.rp], pos.red_par ^ 1);
        __syncwarp();
        if (warp == 0 && lane == 0) GS_STAMP(1, pos.dbg, 1);
        if (lane == 0) { mbar_arrive(&sh.empty[pos.st]); mbar_arrive(&sh.rfull[pos.rp]); }
        if (warp == 0 && lane == 0) GS_STAMP(1, pos.dbg, 2);
        pos.next();
        ga0 = (ga0 + gstep) & 15u; ga1 = (ga1 + gstep) & 15u;
    }
}

template <int TYPE>
__device__ __forceinline__ void consume_dispatch(const gs_params & p, const gs_seg & sg, int r0, int r1, gs_pos & pos, gs_shared & sh,
                                                 uint8_t * ring, float * red, const uint8_t * aq, const float2 * ads, int warp, int lane) {
    if (sg.spw == 1)      consume_segment<TYPE, 1>(p, sg, r0, r1, pos, sh, ring, red, aq, ads, warp, lane);
    else if (sg.spw == 2) consume_segment<TYPE, 2>(p, sg, r0, r1, pos, sh, ring, red, aq, ads, warp, lane);
    else if (item<TYPE>::max_spw >= 4)
        consume_segment<TYPE, (item<TYPE>::max_spw >= 4 ? 4 : 1)>(p, sg, r0, r1, pos, sh, ring, red, aq, ads, warp, lane);
}

template <int T0, int T1>
__global__ void __launch_bounds__(GS_THREADS, 2) gemv_stream_kernel(const __grid_constant__ gs_params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *  ring = smem;                                                    // GS_STAGES x stage_stride
    float *    red  = (float *) (smem + p.n_stages * p.stage_stride);          // 2 x GS_RED_FLOATS
    uint8_t *  aq   = (uint8_t *) (red + 2 * GS_RED_FLOATS);                   // activation int8, padded layout
    float2 *   ads  = (float2 *) (aq + p.act_qs_bytes);                        // (d, d*sum) per unit
    __shared__ gs_shared sh;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < p.n_stages; ++s) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], GS_NW); }
        mbar_init(&sh.rfull[0], GS_NW); mbar_init(&sh.rfull[1], GS_NW);
        mbar_init(&sh.rempty[0], 1);    mbar_init(&sh.rempty[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    pdl_launch_dependents();          // the next kernel in the stream may begin its own weight prefetch

    // ---- this CTA's share of every work segment, split by bytes
    const int64_t lo = p.total_bytes * (int64_t) blockIdx.x / gridDim.x;
    const int64_t hi = p.total_bytes * (int64_t) (blockIdx.x + 1) / gridDim.x;
    auto row_range = [&](const gs_seg & sg, int & r0, int & r1) {
        const int64_t rbw = (int64_t) sg.rb * sg.nsub, tot = rbw * sg.M;
        const int64_t a = min(max(lo - sg.byte_prefix, (int64_t) 0), tot), b = min(max(hi - sg.byte_prefix, (int64_t) 0), tot);
        r0 = (int) ((a + rbw - 1) / rbw); r1 = (int) ((b + rbw - 1) / rbw);
    };

    if (warp == GS_NW) {
        // ================= producer: TMA bulk copies into the ring
        if (lane == 0) {
            const uint64_t pol = policy_evict_first();
            gs_pos pos; pos.ns = p.n_stages;
            int dbg_n = 0;
            for (int s = 0; s < p.n_ws; ++s) {
                const gs_seg & sg = p.seg[s];
                int r0, r1; row_range(sg, r0, r1);
                for (int row = r0; row < r1; row += sg.R) {
                    const int nr = min(sg.R, r1 - row);
                    mbar_wait(&sh.empty[pos.st], pos.full_par ^ 1);
                    GS_STAMP(0, dbg_n, 0);
                    uint8_t * dstp = ring + pos.st * p.stage_stride;
                    const uintptr_t g0 = (uintptr_t) (sg.w[0] + (int64_t) row * sg.rb);
                    const uint32_t o0 = (uint32_t) (g0 & 15), b0 = (o0 + (uint32_t) nr * sg.rb + 15u) & ~15u;
                    if (sg.nsub == 2) {
                        const uintptr_t g1 = (uintptr_t) (sg.w[1] + (int64_t) row * sg.rb);
                        const uint32_t o1 = (uint32_t) (g1 & 15), b1 = (o1 + (uint32_t) nr * sg.rb + 15u) & ~15u;
                        mbar_expect_tx(&sh.full[pos.st], b0 + b1);
                        bulk_g2s(dstp, (const void *) (g0 - o0), b0, &sh.full[pos.st], pol);
                        bulk_g2s(dstp + b0, (const void *) (g1 - o1), b1, &sh.full[pos.st], pol);
                    } else if (p.flags & 2) {     // experiment: the same bytes as 4 concurrent copies
                        mbar_expect_tx(&sh.full[pos.st], b0);
                        const uint32_t q4 = (b0 / 4 + 15u) & ~15u;
                        for (uint32_t o = 0; o < b0; o += q4)
                            bulk_g2s(dstp + o, (const void *) (g0 - o0 + o), min(q4, b0 - o), &sh.full[pos.st], pol);
                    } else {
                        mbar_expect_tx(&sh.full[pos.st], b0);
                        bulk_g2s(dstp, (const void *) (g0 - o0), b0, &sh.full[pos.st], pol);
                    }
                    GS_STAMP(0, dbg_n, 1);
                    ++dbg_n;
                    pos.next();
                }
            }
        }
        return;
    }

    // ================= prologue (consumers + finalizer): x -> [rms norm * w] -> q8_1 in shared memory
    pdl_wait();                       // x, bias and dst belong to earlier kernels in the stream
    {
        const int K = p.K, nblk = K >> 5;
        const int pw = warp < GS_NW ? warp : GS_NW, ptid = pw * 32 + lane;      // prologue warp / thread index (producer excluded)
        float scale = 1.0f;
        if (p.norm_w) {
            float ss = 0.0f;
            for (int i = ptid * 4; i < K; i += GS_PRO_THREADS * 4) {
                const float4 v = *(const float4 *) (p.x + i);
                ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
            }
            ss = warp_sum(ss);
            if (lane == 0) sh.part[pw] = ss;
            named_bar_sync(1, GS_PRO_THREADS);
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i <= GS_NW; ++i) tot += sh.part[i];
            scale = rsqrtf(tot / (float) K + p.eps);
        }
        const int sub = lane & 7, bw = lane >> 3;                 // 8 lanes per 32-element block, 4 blocks per warp pass
        for (int b0 = pw * 4; b0 < nblk; b0 += (GS_NW + 1) * 4) {
            const int blk = b0 + bw;
            const bool act = blk < nblk;
            float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (act) {
                const float4 t = *(const float4 *) (p.x + blk * 32 + sub * 4);
                v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
                if (p.norm_w) {                                    // same op order as rms_norm then mul: (x*scale)*w
                    const float4 w = *(const float4 *) (p.norm_w + blk * 32 + sub * 4);
                    v[0] = __fmul_rn(__fmul_rn(v[0], scale), w.x); v[1] = __fmul_rn(__fmul_rn(v[1], scale), w.y);
                    v[2] = __fmul_rn(__fmul_rn(v[2], scale), w.z); v[3] = __fmul_rn(__fmul_rn(v[3], scale), w.w);
                }
            }
            float d; int sum;
            const int q = quantize_block32_quad<8>(v, d, sum);
            if (act) {
                *(int *) (aq + act_off(blk) + sub * 4) = q;
                if (sub == 0) {
                    const float dh = __half2float(__float2half_rn(d));          // the value a q8_1 block stores
                    ads[blk] = make_float2(dh, dh * (float) sum);
                }
                if (p.xq_out && blockIdx.x == 0) {
                    int * out = (int *) (p.xq_out + blk);
                    out[1 + sub] = q;
                    if (sub == 0) *(__half2 *) out = q8_1_scales(d, sum);
                }
            }
        }
        named_bar_sync(1, GS_PRO_THREADS);
    }

    if (warp < GS_NW) {
        // ================= consumers
        gs_pos pos; pos.ns = p.n_stages;
        for (int s = 0; s < p.n_ws; ++s) {
            const gs_seg & sg = p.seg[s];
            int r0, r1; row_range(sg, r0, r1);
            if (r0 >= r1) continue;
            if (sg.type == T0) consume_dispatch<T0>(p, sg, r0, r1, pos, sh, ring, red, aq, ads, warp, lane);
            else               consume_dispatch<T1>(p, sg, r0, r1, pos, sh, ring, red, aq, ads, warp, lane);
        }
        return;
    }

    // ================= finalizer: reduce partials, epilogue, store (4 lanes per row, 8 rows per pass)
    {
        gs_pos pos; pos.ns = p.n_stages;
        const int part = lane & 3, rsel = lane >> 2;
        for (int s = 0; s < p.n_ws; ++s) {
            const gs_seg & sg = p.seg[s];
            int r0, r1; row_range(sg, r0, r1);
            const int nvals = sg.KS * 32, rstride = nvals + 4, per = nvals >> 2;      // floats per lane: KS*8
            for (int row = r0; row < r1; row += sg.R) {
                const int nr = min(sg.R, r1 - row);
                float bias0 = 0.0f;                                   // prefetch the epilogue operand of the first pass
                if (sg.bias && rsel < nr && part == 0) bias0 = sg.bias[row + rsel];
                mbar_wait(&sh.rfull[pos.rp], pos.red_par);
                if (lane == 0) GS_STAMP(2, pos.dbg, 0);
                const float * r = red + pos.rp * GS_RED_FLOATS;
                // The buffer is released as soon as its values sit in registers, and by a lane that never stores to
                // global memory: mbarrier.arrive is a release, and a release issued by a storing thread waits for that
                // thread's outstanding global stores (~1 us per chunk when it sat behind the dst store).
                if (nr <= 8) {
                    const int j = rsel;
                    float v0 = 0.0f, v1 = 0.0f;
                    if (j < nr) {
                        const float4 * src = (const float4 *) (r + j * rstride + part * per);
                        for (int i = 0; i < per / 4; ++i) { const float4 t = src[i]; v0 += (t.x + t.y) + (t.z + t.w); }
                        if (sg.nsub == 2) {
                            src = (const float4 *) (r + (nr + j) * rstride + part * per);
                            for (int i = 0; i < per / 4; ++i) { const float4 t = src[i]; v1 += (t.x + t.y) + (t.z + t.w); }
                        }
                    }
                    v0 += __shfl_xor_sync(0xffffffffu, v0, 1); v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
                    v0 += __shfl_xor_sync(0xffffffffu, v0, 2); v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
                    if (lane == 1) mbar_arrive(&sh.rempty[pos.rp]);
                    if (j < nr && part == 0) {
                        float out = sg.nsub == 2 ? gs_glu(p.glu_op, v0, v1, p.glu_alpha, p.glu_limit) : v0;
                        if (sg.bias) out += bias0;
                        sg.dst[row + j] = out;
                    }
                } else {
                    float keep0[4], keep1[4];                       // up to 32 rows: 4 passes held in registers
#pragma unroll
                    for (int ps = 0; ps < 4; ++ps) {
                        const int j = ps * 8 + rsel;
                        float v0 = 0.0f, v1 = 0.0f;
                        if (j < nr) {
                            const float4 * src = (const float4 *) (r + j * rstride + part * per);
                            for (int i = 0; i < per / 4; ++i) { const float4 t = src[i]; v0 += (t.x + t.y) + (t.z + t.w); }
                            if (sg.nsub == 2) {
                                src = (const float4 *) (r + (nr + j) * rstride + part * per);
                                for (int i = 0; i < per / 4; ++i) { const float4 t = src[i]; v1 += (t.x + t.y) + (t.z + t.w); }
                            }
                        }
                        v0 += __shfl_xor_sync(0xffffffffu, v0, 1); v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
                        v0 += __shfl_xor_sync(0xffffffffu, v0, 2); v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
                        keep0[ps] = v0; keep1[ps] = v1;
                    }
                    if (lane == 1) mbar_arrive(&sh.rempty[pos.rp]);
#pragma unroll
                    for (int ps = 0; ps < 4; ++ps) {
                        const int j = ps * 8 + rsel;
                        if (j < nr && part == 0) {
                            float out = sg.nsub == 2 ? gs_glu(p.glu_op, keep0[ps], keep1[ps], p.glu_alpha, p.glu_limit) : keep0[ps];
                            if (sg.bias) out += ps == 0 ? bias0 : sg.bias[row + j];
                            sg.dst[row + j] = out;
                        }
                    }
                }
                if (lane == 0) GS_STAMP(2, pos.dbg, 1);
                pos.next();
            }
        }
    }
}

// ---------------------------------------------------------------- host side
static bool gs_type_ok(int t) { return t == B200_Q4_0 || t == B200_Q8_0 || t == B200_MXFP4 || t == B200_Q4_K || t == B200_Q6_K; }
static int  gs_wpi(int t) { return (t == B200_Q4_K || t == B200_Q6_K) ? 64 : 32; }

static bool gs_build(const b200_gemv_fused_args * a, gs_params & p) {
    if (!a || a->n_seg < 1 || a->n_seg > B200_GEMV_MAX_SEG) return false;
    const b200_tensor & x = a->x;
    const int64_t K = x.ne[0];
    if (x.type != B200_F32 || x.nb[0] != 4 || x.ne[1] * x.ne[2] * x.ne[3] != 1 || K < 32 || K % 32 != 0 || K > 65536) return false;
    if (((uintptr_t) x.data) % 16 != 0) return false;
    if (a->norm_w && (a->norm_w->type != B200_F32 || a->norm_w->nb[0] != 4 || a->norm_w->ne[0] != K ||
                      nelements(*a->norm_w) != K || ((uintptr_t) a->norm_w->data) % 16 != 0)) return false;
    const bool glu = a->glu_op != B200_GLU_NONE;
    if (glu && (a->n_seg != 2 || a->glu_op < B200_GLU_REGLU || a->glu_op > B200_GLU_GEGLU_QUICK)) return false;
    int types[2] = {-1, -1};
    for (int s = 0; s < a->n_seg; ++s) {
        const b200_gemv_seg & g = a->seg[s];
        const b200_tensor & w = g.w;
        if (!gs_type_ok(w.type) || w.ne[0] != K || w.ne[1] < 1 || w.ne[2] != 1 || w.ne[3] != 1) return false;
        if (K % b200_blck_size(w.type) != 0) return false;
        if (w.nb[0] != (int64_t) b200_type_size(w.type) || w.nb[1] != (int64_t) b200_row_size(w.type, K)) return false;
        if (((uintptr_t) w.data) % 16 != 0 || w.ne[1] > (1 << 30)) return false;
        if (!(glu && s == 0)) {
            const b200_tensor & d = g.dst;
            if (d.type != B200_F32 || d.nb[0] != 4 || d.ne[0] != w.ne[1] || nelements(d) != w.ne[1]) return false;
        }
        if (g.bias && (g.bias->type != B200_F32 || g.bias->nb[0] != 4 || g.bias->ne[0] != w.ne[1] || nelements(*g.bias) != w.ne[1])) return false;
        if (types[0] < 0 || types[0] == w.type) types[0] = w.type;
        else if (types[1] < 0 || types[1] == w.type) types[1] = w.type;
        else return false;
    }
    if (glu) {
        const b200_tensor & g = a->seg[0].w; const b200_tensor & u = a->seg[1].w;
        if (g.type != u.type || g.ne[1] != u.ne[1] || a->seg[0].bias) return false;
    }
    // shared-memory budget: ring + reduction buffers + activation
    const int act_qs = (int) ((K + (K / 128 + 1) * 16 + 15) & ~15);
    const int act_ds = (int) (K / 32 * 8);
    int64_t need = 0;                                   // smallest stage payload that holds one row (pair)
    for (int s = 0; s < a->n_seg; ++s) need = std::max<int64_t>(need, a->seg[s].w.nb[1] * (glu ? 2 : 1) + 64);
    const int fixed = 2 * GS_RED_FLOATS * 4 + act_qs + act_ds + 16;
    const int flags = a->flags;
    const int n_stages = ((flags >> 4) & 0xF) ? std::min((flags >> 4) & 0xF, GS_MAX_STAGES) : GS_STAGES;
    const int64_t want = ((flags >> 8) & 0xFF) ? (int64_t) ((flags >> 8) & 0xFF) * 1024 : GS_STAGE_BYTES;
    int64_t payload = std::max<int64_t>(want, need);
    auto total_for = [&](int64_t pl) { return (int64_t) n_stages * (((pl + 127) & ~127) + GS_STAGE_SLACK) + fixed; };
    if (!((flags >> 8) & 0xFF) && total_for(payload) > GS_SMEM_2CTA && need < payload) {          // shrink the stages if that keeps two CTAs per SM
        const int64_t fit = (GS_SMEM_2CTA - fixed) / n_stages - GS_STAGE_SLACK - 128;
        if (fit >= need && fit >= 12 * 1024) payload = fit & ~127;
    }
    if (total_for(payload) > GS_SMEM_MAX) return false;
    const int stage_payload = (int) ((payload + 127) & ~127);

    memset(&p, 0, sizeof(p));
    p.K = (int32_t) K; p.x = (const float *) x.data; p.norm_w = a->norm_w ? (const float *) a->norm_w->data : nullptr; p.eps = a->eps;
    p.glu_op = a->glu_op; p.glu_alpha = a->glu_alpha; p.glu_limit = a->glu_limit;
    p.xq_out = (blk_q8_1 *) a->xq_out; p.act_qs_bytes = act_qs; p.stage_stride = stage_payload + GS_STAGE_SLACK;
    p.n_stages = n_stages; p.flags = flags;
    int64_t prefix = 0;
    p.n_ws = glu ? 1 : a->n_seg;
    for (int s = 0; s < p.n_ws; ++s) {
        gs_seg & g = p.seg[s];
        const b200_gemv_seg & src = glu ? a->seg[1] : a->seg[s];
        g.nsub = glu ? 2 : 1;
        g.w[0] = (const uint8_t *) (glu ? a->seg[0].w.data : src.w.data);
        g.w[1] = glu ? (const uint8_t *) a->seg[1].w.data : nullptr;
        g.bias = src.bias ? (const float *) src.bias->data : nullptr;
        g.dst = (float *) src.dst.data;
        g.M = (int32_t) src.w.ne[1]; g.rb = (int32_t) src.w.nb[1]; g.type = src.w.type;
        g.ipr = (int32_t) (K / gs_wpi(g.type)); g.spr = (g.ipr + 31) / 32;
        int KS = 1; while (KS * 2 <= GS_NW && KS * 2 <= g.spr) KS *= 2;
        g.KS = KS;
        int spw = (g.spr + KS - 1) / KS;
        if (spw == 3) spw = 4;
        if (spw > (gs_wpi(g.type) == 64 ? 2 : 4)) return false;        // row too long for the register-resident activation
        g.spw = spw;
        const int RG = GS_NW / KS;
        // rows per chunk: fit the stage (both sub-blocks + alignment slack), the reduction buffer, one pass granularity
        int R = (stage_payload - 64) / (g.rb * g.nsub);
        const int red_rows = GS_RED_FLOATS / (KS * 32 + 4) / g.nsub;
        if (R > red_rows) R = red_rows;
        if (R > 32) R = 32;
        if (R < 1) return false;                       // a single row does not fit a stage
        const int pass = RG * GS_RB / g.nsub > 0 ? RG * GS_RB / g.nsub : 1;
        if (R > pass) R -= R % pass;
        g.R = R;
        g.byte_prefix = prefix;
        prefix += (int64_t) g.rb * g.nsub * g.M;
    }
    p.total_bytes = prefix;
    return true;
}

static size_t gs_smem(const gs_params & p) {
    return (size_t) p.n_stages * p.stage_stride + 2 * GS_RED_FLOATS * 4 + p.act_qs_bytes + (size_t) p.K / 32 * 8 + 16;
}

template <int T0, int T1>
static int gs_launch(const gs_params & p, int grid, cudaStream_t st) {
    static bool attr_set[64] = {};
    const size_t smem = gs_smem(p);
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_ARG;
    if (!attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(gemv_stream_kernel<T0, T1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GS_SMEM_MAX));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3(GS_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, gemv_stream_kernel<T0, T1>, p));
    count_launch();
    return B200_OK;
}

static int gs_sm_count() {
    static int n[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!n[dev]) { if (cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n[dev] <= 0) n[dev] = 148; }
    return n[dev];
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemv_fused_supported(const b200_gemv_fused_args * a) {
    gs_params p;
    return gs_build(a, p) ? 1 : 0;
}

extern "C" int b200_gemv_fused(const b200_gemv_fused_args * a, void * stream) {
    gs_params p;
    if (!gs_build(a, p)) { set_error("b200_gemv_fused: unsupported arguments"); return B200_ERR_UNSUPPORTED; }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    int t0 = p.seg[0].type, t1 = t0;
    for (int s = 1; s < p.n_ws; ++s) if (p.seg[s].type != t0) t1 = p.seg[s].type;
    // chunks available: never launch more CTAs than there are rows
    int64_t rows = 0;
    for (int s = 0; s < p.n_ws; ++s) rows += p.seg[s].M;
    int grid = a->n_sm > 0 ? a->n_sm : gs_sm_count();
    if (grid > rows) grid = (int) rows;
    cudaStream_t st = (cudaStream_t) stream;
    if (t0 == t1) {
        switch (t0) {
            case B200_Q4_K:  return gs_launch<B200_Q4_K, B200_Q4_K>(p, grid, st);
            case B200_Q6_K:  return gs_launch<B200_Q6_K, B200_Q6_K>(p, grid, st);
            case B200_Q4_0:  return gs_launch<B200_Q4_0, B200_Q4_0>(p, grid, st);
            case B200_Q8_0:  return gs_launch<B200_Q8_0, B200_Q8_0>(p, grid, st);
            case B200_MXFP4: return gs_launch<B200_MXFP4, B200_MXFP4>(p, grid, st);
        }
    }
    const int lo = t0 < t1 ? t0 : t1, hi = t0 < t1 ? t1 : t0;
    if (lo == B200_Q4_K && hi == B200_Q6_K) return gs_launch<B200_Q4_K, B200_Q6_K>(p, grid, st);
    if (lo == B200_Q4_0 && hi == B200_Q6_K) return gs_launch<B200_Q4_0, B200_Q6_K>(p, grid, st);
    if (lo == B200_Q4_0 && hi == B200_Q8_0) return gs_launch<B200_Q4_0, B200_Q8_0>(p, grid, st);
    set_error("b200_gemv_fused: unsupported type pair (%d, %d)", t0, t1);
    return B200_ERR_UNSUPPORTED;
}
