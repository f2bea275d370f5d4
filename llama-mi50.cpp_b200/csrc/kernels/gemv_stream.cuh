// gemv_stream.cuh — the streaming GEMV / decode-chain kernel template (see gemv_stream.cu for the design notes and the host
// side).  The kernel is instantiated per weight-type pair in gemv_stream_i_*.cu so the instantiations compile in parallel.
#pragma once
#include <algorithm>
#include <cstring>
#include <mutex>
#include <type_traits>
#include <vector>

#include "quantize.cuh"
#include "vecdot.cuh"

namespace b200 {

constexpr int GS_NW          = 16;                     // consumer warps (4 per scheduler)
constexpr int GS_THREADS     = (GS_NW + 4) * 32;       // + producer warpgroup: warp GS_NW streams, the other three exit at once.
                                                       // A scheduler partition owns 16 K registers; a fifth warp on it caps every
                                                       // thread at 96.  With a whole warpgroup the producer side can hand its
                                                       // registers over (setmaxnreg.dec) and the consumers take them (setmaxnreg.inc).
                                                       // The pool is what the CTA got at launch (96 per thread), not the register
                                                       // file: an inc that does not fit WAITS FOREVER (measured: a hung launch)
constexpr int GS_REG_LAUNCH = 96, GS_REG_PRODUCER = 32, GS_REG_CONSUMER = 112;
static_assert(GS_NW * 32 * GS_REG_CONSUMER + 4 * 32 * GS_REG_PRODUCER <= GS_THREADS * GS_REG_LAUNCH, "setmaxnreg.inc would never be granted");
constexpr int GS_CON_THREADS = GS_NW * 32;
constexpr int GS_PRO_THREADS = GS_CON_THREADS;         // warps that run the prologue (all but the producer)
constexpr int GS_STAGES      = 3;                      // default ring depth
constexpr int GS_MAX_STAGES  = 8;
constexpr int GS_STAGE_BYTES = 56 * 1024;              // preferred stage payload: HBM efficiency needs >= ~32 KB per bulk copy
constexpr int GS_STAGE_SLACK = 128;                    // slack for 16 B alignment of both ends / funnel over-read
constexpr int GS_RED_FLOATS  = 2304;                   // per parity: rows * (KS*32 + 4) <= GS_RED_FLOATS
constexpr int GS_SMEM_MAX    = 220 * 1024;             // dynamic; ~3 KB of static shared memory on top
constexpr int GS_RB          = 2;                      // rows per warp pass of the K-split segments (KS > 1); KS == 1 segments run passes of four
constexpr int GS_MAX_PHASES  = 1024;

// Inside a chain every value a later phase reads travels as a TAGGED WORD {f32 bits, tag of (launch, producing phase)} in a
// shadow array (`*_ll`): the reader spins on the tag, so a phase starts as soon as ITS inputs exist — no grid barrier, no
// fence (a device-scope release on B200 waits for everything the SM has in flight, including the bulk copies of the next
// phase's weights).  Null `*_ll`: plain memory written before the launch (graph inputs, weights' biases).
struct gs_seg {
    const uint8_t * w[2];     // [1] = paired "up" matrix of a GLU work segment (else null)
    const float *   bias;
    float *         dst;
    unsigned long long * dst_ll;         // tagged shadow of dst (null: nothing in the chain reads it)
    const unsigned long long * bias_ll;  // tagged shadow of bias when an earlier phase of the chain produced it
    int64_t         byte_prefix;   // bytes of the work segments before this one
    int32_t M, rb, type, nsub;     // rows, row bytes, ggml type, 1 (or 2 for a GLU pair)
    int32_t R, KS, spr, ipr;       // rows per chunk, warps along K, steps per row, items per row
    int32_t spw; uint32_t bias_tag;   // K-steps per warp (1, 2 or 4): the per-step activation registers are hoisted; phase that produced bias
};
// attention phase of the decode chain: one query position, f16 KV cache, RoPE + KV-cache write fused in
struct gs_attn {
    const float * q; const float * k_new; const float * v_new;   // raw projections of the current token
    char * k_cache; char * v_cache; const __half * mask; float * dst;
    const int32_t * pos; const int64_t * slot; const int64_t * slot_v;
    int64_t k_nb1, k_nb2, v_nb1, v_nb2;
    int32_t D, H, H_kv, n_kv, rope_mode, n_dims;
    float scale, theta_scale, freq_scale;
    float ext_factor, attn_factor, corr0, corr1;           // YaRN (ggml_rope_ext): ramp between interpolated and extrapolated angles
    const float * sinks;                                    // one extra logit per head (no value), or null
    // tagged shadows of q / k_new / v_new (null: plain) and the index of the phase that produced them; shadow of dst
    const unsigned long long * q_ll; const unsigned long long * k_ll; const unsigned long long * v_ll;
    unsigned long long * dst_ll;
    uint32_t q_tag, k_tag, v_tag, kv_type;                 // kv_type: B200_F16 or B200_Q8_0 (rows of 34-byte blocks)
};
struct alignas(16) gs_phase {
    int32_t kind;                  // 0 = GEMV, 1 = attention
    int32_t n_ws;                  // work segments (GEMV)
    gs_seg seg[B200_GEMV_MAX_SEG];
    int32_t glu_op; float glu_alpha, glu_limit;
    const float * x; const float * norm_w; float eps; int32_t K;
    blk_q8_1 * xq_out;
    int64_t total_bytes;
    // x_ll: tagged shadow of x written in phase x_tag (null: x is plain memory written before the launch).
    // peer_mode (row split): how this phase's outputs reach the other ranks: 0 not at all, 1 tagged words (read by a later
    // phase), 2 plain stores + system-scope arrival (read after the launch: the logits)
    const unsigned long long * x_ll; uint32_t x_tag; int32_t peer_mode;
    uint32_t self_tag; int32_t _pad2[3];      // index of this phase (added to the launch tag for its own tagged stores)
    gs_attn attn;
};
struct gs_params {
    gs_phase inl;                  // the phase of a single-phase launch (b200_gemv_fused)
    const gs_phase * phases;       // device array for a chain (null: use inl)
    unsigned int * exit_count;     // chain: CTAs that have finished the launch (the last one counts the launch in `epoch`)
    const int2 * ranges;           // chain: rows [x, y) of segment s of phase i for CTA c at [(i * grid + c) * B200_GEMV_MAX_SEG + s]
                                   // (computed on the host: 64-bit divisions in every thread of every phase were 5 % of the launch)
    int32_t n_phases;
    int32_t act_qs_bytes;          // activation region sized for the largest K of the chain
    int32_t stage_stride;          // bytes between ring stages (payload + GS_STAGE_SLACK, multiple of 128)
    int32_t n_stages;              // ring depth
    int32_t attn_floats;           // attention scratch (floats) after the activation scales
    int32_t flags;                 // tuning/debug: bit 0 = consumers skip the dot products (bandwidth probe)
    unsigned long long * dbg;      // flags bit 2: globaltimer stamps [phase][cta 0 | last cta][4]
    // ---- row split over peer memory (MODE 2): outputs and barrier arrivals are mirrored to the other ranks
    int32_t n_peers;               // other ranks
    int32_t total_ctas;            // CTAs of all ranks: arrivals per phase and launch
    int64_t peer_delta[B200_MAX_RANKS - 1];   // byte offset from a local arena address to the same location on peer r
    unsigned int * epoch;          // chain: launches completed so far: tags of different launches differ
    unsigned int * sync;           // arena: arrival counters of the peer_mode 2 phases (all ranks' CTAs)
    int32_t n_final; int32_t final_phase[4];   // peer_mode 2 phases the launch waits for before it ends
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

// ---- tagged words (row split over peer memory): a value and the tag of the (launch, phase) that produced it travel in
// ONE 64-bit store, so a reader that sees the tag has the value: one NVLink traversal, no fence round trip, no counter
// (the idea of NCCL's LL protocol).  Spins are bounded: a rank that never delivers becomes a trap, not a hung GPU.
// SYS: system scope (the word crosses NVLink to / from another GPU); else device scope (L2 of this GPU)
template <bool SYS>
__device__ __forceinline__ void ll_store(void * p, float v, unsigned int tag) {
    const unsigned long long w = (unsigned long long) __float_as_uint(v) | ((unsigned long long) tag << 32);
    if (SYS) asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
    else     asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
template <bool SYS>
__device__ __forceinline__ unsigned long long ll_load1(const void * p) {
    unsigned long long w;
    if (SYS) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    else     asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}
template <bool SYS>
__device__ __forceinline__ ulonglong2 ll_load2(const void * p) {      // two elements, 16 B aligned
    ulonglong2 r;
    if (SYS) asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p) : "memory");
    else     asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ bool ll_ok(unsigned long long w, unsigned int tag) { return (unsigned int) (w >> 32) == tag; }
__device__ __forceinline__ void ll_spin_check(unsigned int & spins, unsigned long long & t0) {
    if ((++spins & 0x3FFu) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (!t0) t0 = now; else if (now - t0 > 20000000000ull) __trap();          // 20 s
    }
}
// The spins are out of line (one copy per kernel): the callers' fast path is "load, compare the tag, done".  Code size matters
// here: the chain kernel runs a different stretch of code in every phase, and what does not fit the instruction cache is
// fetched again at every phase boundary, on the critical path.
template <bool SYS>
__device__ __noinline__ unsigned long long ll_spin1(const void * p, unsigned int tag) {
    unsigned int spins = 0; unsigned long long t0 = 0;
    unsigned long long w = ll_load1<SYS>(p);
    while (!ll_ok(w, tag)) { ll_spin_check(spins, t0); w = ll_load1<SYS>(p); }
    return w;
}
template <bool SYS>
__device__ __forceinline__ float ll_settle1(unsigned long long w, const void * p, unsigned int tag) {
    if (!ll_ok(w, tag)) w = ll_spin1<SYS>(p, tag);
    return __uint_as_float((unsigned int) w);
}
template <bool SYS>
__device__ __forceinline__ float ll_read(const void * p, unsigned int tag) { return ll_settle1<SYS>(ll_load1<SYS>(p), p, tag); }
template <bool SYS>
__device__ __noinline__ ulonglong2 ll_spin2(const void * p, unsigned int tag) {
    unsigned int spins = 0; unsigned long long t0 = 0;
    ulonglong2 a = ll_load2<SYS>(p);
    while (!ll_ok(a.x, tag) || !ll_ok(a.y, tag)) { ll_spin_check(spins, t0); a = ll_load2<SYS>(p); }
    return a;
}
template <bool SYS>
__device__ __forceinline__ void ll_settle2(ulonglong2 & a, const void * p, unsigned int tag) {
    if (!ll_ok(a.x, tag) || !ll_ok(a.y, tag)) a = ll_spin2<SYS>(p, tag);
}
// one lane waits (with back-off) until the word at p carries `tag`: the gate in front of a many-thread read, so that CTAs
// waiting for a phase to finish do not flood L2 with retries
template <bool SYS>
__device__ __noinline__ void ll_gate(const void * p, unsigned int tag) {
    unsigned int spins = 0; unsigned long long t0 = 0;
    while (!ll_ok(ll_load1<SYS>(p), tag)) { ll_spin_check(spins, t0); __nanosleep(100); }
}

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

static __device__ __noinline__ float gs_glu(int op, float g, float x, float alpha, float limit) {      // one or two lanes per pass: one copy
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

// debug stamps: the SM's cycle counter.  (%globaltimer costs ~0.3 us per read on B200: five stamps per phase made the stamped
// CTAs the slowest of the grid.)  Stamps of one CTA are comparable with each other, not across CTAs.
__device__ __forceinline__ unsigned long long gs_clk() { return (unsigned long long) clock64(); }
__device__ __forceinline__ unsigned long long gs_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ring position carried by the producer and the consumers across the segments and phases of a launch
struct gs_pos {
    int st = 0, full_par = 0;      // ring stage, parity of its "full" phase
    int ns = GS_STAGES;
    __device__ __forceinline__ void next() { if (++st == ns) { st = 0; full_par ^= 1; } }
};

struct gs_shared {
    uint64_t full[GS_MAX_STAGES], empty[GS_MAX_STAGES];
    float part[GS_NW];
    float slots[2][GS_NW][GS_RB];  // per-warp row sums of a pass, combined across the KS warps of a row group
    unsigned int tag;              // row split: tag of this launch (phase index is added)
};

struct gs_ctx {                    // per-thread view of the launch-wide shared memory carve-up
    uint8_t * ring; uint8_t * aq; float2 * ads; float * attn; gs_shared * sh;
    int stage_stride, flags;
};

__device__ __forceinline__ unsigned int sh_tag_of(const gs_ctx & c) { return c.sh->tag; }

__device__ __forceinline__ void gs_row_range(const gs_seg & sg, int64_t lo, int64_t hi, int & r0, int & r1) {
    const int64_t rbw = (int64_t) sg.rb * sg.nsub, tot = rbw * sg.M;
    const int64_t a = min(max(lo - sg.byte_prefix, (int64_t) 0), tot), b = min(max(hi - sg.byte_prefix, (int64_t) 0), tot);
    r0 = (int) ((a + rbw - 1) / rbw); r1 = (int) ((b + rbw - 1) / rbw);
}

// rows of every segment of a staged phase for this CTA -> rng[0..n_ws): from the host's table (chain) or computed by one lane each
__device__ __forceinline__ void gs_stage_ranges(int2 * rng, const gs_phase * buf, const int2 * table, int ip, int cta, int n_cta, int t) {
    if (t >= B200_GEMV_MAX_SEG) return;
    if (table) { rng[t] = table[((int64_t) ip * n_cta + cta) * B200_GEMV_MAX_SEG + t]; return; }
    int r0 = 0, r1 = 0;
    if (buf->kind == 0 && t < buf->n_ws) {
        const int64_t lo = buf->total_bytes * (int64_t) cta / n_cta, hi = buf->total_bytes * (int64_t) (cta + 1) / n_cta;
        gs_row_range(buf->seg[t], lo, hi, r0, r1);
    }
    rng[t] = make_int2(r0, r1);
}

// warp-uniform walk over the weight chunks of ONE CTA through every GEMV phase of a launch, in consumption order
struct gs_walk { int ip = -1, s = 0, row = 0, r1 = 0; };
// moves w to its next chunk (rows [row, min(row + R, r1)) of segment s of phase ip); the descriptor of a phase and this CTA's
// row ranges are staged into *buf / rng by the whole warp when the walk enters it.  The caller steps `row` by the segment's R.
// false: no chunk left
__device__ __forceinline__ bool gs_walk_next(gs_walk & w, gs_phase * buf, int2 * rng, const gs_phase * phases, const int2 * table, int n_phases, int cta, int n_cta, int lane) {
    for (;;) {
        if (w.row < w.r1) return true;
        if (w.ip >= 0 && buf->kind == 0 && ++w.s < buf->n_ws) { w.row = rng[w.s].x; w.r1 = rng[w.s].y; continue; }
        if (++w.ip >= n_phases) return false;
        __syncwarp();
        for (int i = lane; i < (int) (sizeof(gs_phase) / 16); i += 32) ((uint4 *) buf)[i] = ((const uint4 *) (phases + w.ip))[i];
        __syncwarp();
        gs_stage_ranges(rng, buf, table, w.ip, cta, n_cta, lane);
        __syncwarp();
        w.s = -1; w.row = w.r1 = 0;
    }
}

__device__ __forceinline__ int wt_first_row(const gs_walk & w, const gs_phase &) { return w.row; }      // (debug hook: every chunk qualifies; the stamp keeps the first)

// consumer warps: all chunks of one work segment.  A pass of a warp covers GS_RB rows (for a GLU pair: the gate row and the
// up row of ONE output row); the KS warps of a row group each own a K slice, reduce their lanes by shuffles and meet at a
// named barrier of their own; warp ks == 0 sums the KS values in fixed order, applies the epilogue and stores.
// one output value: the local store and, under row split, the same store into every peer's copy of the buffer
template <bool CHAIN, bool PEER>
__device__ __forceinline__ void gs_put(const gs_params & p, const gs_phase & ph, const gs_shared & sh, float * d, unsigned long long * ll, float v) {
    *d = v;
    if (CHAIN && ll) {                                   // the value a later phase waits for
        const unsigned int tag = sh.tag + ph.self_tag;
        ll_store<PEER>(ll, v, tag);
        if (PEER && ph.peer_mode == 1) {
#pragma unroll
            for (int r = 0; r < B200_MAX_RANKS - 1; ++r)
                if (r < p.n_peers) ll_store<true>((char *) ll + p.peer_delta[r], v, tag);
        }
    }
    if (PEER && ph.peer_mode == 2) {
#pragma unroll
        for (int r = 0; r < B200_MAX_RANKS - 1; ++r)
            if (r < p.n_peers) *(float *) ((char *) d + p.peer_delta[r]) = v;
    }
}

// Rows of a chunk are dealt to the RG row groups (a row group = the KS warps that share a row) in PASSES of RB weight rows,
// round-robin, starting one group further every chunk, so short chunks keep every warp busy over time.
//   RB = 4 (rows whose K fits one warp, KS == 1): four plain rows, or two output rows of a GLU pair (gate0, up0, gate1, up1).
//            The kernel is issue-bound: the per-pass bookkeeping, the shuffle reduction (6 shuffles for 4 rows instead of 12) and
//            the epilogue are paid once per four rows.
//   RB = 2 (KS > 1): two plain rows, or the gate row and the up row of ONE output row; the KS warps of the group meet at a
//            named barrier of their own and warp ks == 0 sums their values in fixed order.
// The row sums of a pass are reduced TOGETHER: every exchange halves the number of values a lane carries, so the totals land
// in lanes 0..RB-1, which then run the epilogue of "their" row.  Results are bit-deterministic (fixed order, no atomics).
template <int TYPE, int SPW, int RB, bool COHERENT, bool PEER>
__device__ __forceinline__ void consume_segment(const gs_params & p, const gs_ctx & c, const gs_phase & ph, const gs_seg & sg, int r0, int r1, gs_pos & pos, int warp, int lane) {
    static_assert(RB == 2 || RB == 4, "passes of two or four rows");
    const int KS = RB == 4 ? 1 : sg.KS, ksl = 31 - __clz(KS), ks = warp & (KS - 1), rg = warp >> ksl, RG = GS_NW >> ksl;
    item<TYPE> itm[SPW];
#pragma unroll
    for (int s = 0; s < SPW; ++s) itm[s].init((ks + s * KS) * 32 + lane, sg.ipr, c.aq, c.ads);
    const int rb = sg.rb, R = sg.R;
    const bool pair = sg.nsub == 2;
    const int opp = pair ? RB / 2 : RB;                    // output rows per pass
    int off = 0;                                           // row group that gets the first pass of the current chunk
    uint32_t ga0 = (uint32_t) ((uintptr_t) (sg.w[0] + (int64_t) r0 * rb) & 15);
    uint32_t ga1 = pair ? (uint32_t) ((uintptr_t) (sg.w[1] + (int64_t) r0 * rb) & 15) : 0u;
    const uint32_t gstep = (uint32_t) R * (uint32_t) rb;
    gs_shared & sh = *c.sh;
    const volatile gs_seg & vs = sg;                       // epilogue fields are re-read where they are used (few lanes per pass):
                                                           // hoisted into registers they cost more in the dot-product loops
    // output row (relative to the pass) whose epilogue this lane runs: the paired reduction leaves row j in lane perm(j)
    //   RB = 2: lane 0 -> row 0, lane 1 -> row 1;   RB = 4: lane 0 -> row 0, lane 1 -> row 2, lane 2 -> row 1, lane 3 -> row 3
    // GLU pair: rows are (gate, up) per output: RB = 2: lane 0 owns the output; RB = 4: lanes 0 / 1 own outputs 0 / 1 (up in lanes 2 / 3)
    const int myj = RB == 2 ? lane : (pair ? lane : ((lane & 1) << 1) | ((lane >> 1) & 1));
    const bool epi = ks == 0 && lane < opp;
    int par = 0;
    for (int row = r0; row < r1; row += R) {
        const int nr = min(R, r1 - row);
        const uint8_t * sub0 = c.ring + pos.st * c.stage_stride + ga0;
        const uint8_t * sub1 = sub0 - ga0 + ((ga0 + (uint32_t) nr * rb + 15u) & ~15u) + ga1;
        mbar_wait(&sh.full[pos.st], pos.full_par);
        const int npass = (nr + opp - 1) / opp;
        int first = rg - off; if (first < 0) first += RG;
        off += npass; while (off >= RG) off -= RG;
        for (int q = first; q < npass; q += RG) {
            const int ob = q * opp;                        // first output row of the pass (relative to the chunk)
            const int cnt = min(opp, nr - ob);             // output rows of this pass (warp-uniform)
            const uint8_t * rowp[RB];
            if (pair) {
#pragma unroll
                for (int j = 0; j < RB / 2; ++j) {
                    const int o = ob + (j < cnt ? j : 0);
                    rowp[2 * j] = sub0 + o * rb; rowp[2 * j + 1] = sub1 + o * rb;
                }
            } else {
#pragma unroll
                for (int j = 0; j < RB; ++j) rowp[j] = sub0 + (ob + (j < cnt ? j : 0)) * rb;
            }
            const int my = ob + myj;                       // the output row whose epilogue this lane runs
            const bool live = epi && myj < cnt;
            float bv = 0.0f; unsigned long long bw = 0ull;      // epilogue operand, requested before the dot products
            if (live && vs.bias) {
                if (COHERENT && vs.bias_ll) bw = ll_load1<PEER>((const unsigned long long *) vs.bias_ll + row + my);
                else bv = COHERENT ? __ldcg((const float *) vs.bias + row + my) : ((const float *) vs.bias)[row + my];
            }
            float acc[RB];
#pragma unroll
            for (int j = 0; j < RB; ++j) acc[j] = 0.0f;
            if (!(c.flags & 1)) {
                // a short last pass computes duplicates of its first row and drops them: one copy of the dot-product code per
                // variant (code that is not executed still spreads the hot code over more instruction-cache lines)
#pragma unroll
                for (int s = 0; s < SPW; ++s)
                    if (itm[s].ok) itm[s].template dot<RB>(rowp, acc);
            }
            float v;
            if (RB == 2) {
                // even lanes end with the total of row 0, odd lanes with the total of row 1
                v = (lane & 1) ? acc[1] : acc[0];
                v += __shfl_xor_sync(0xffffffffu, (lane & 1) ? acc[0] : acc[1], 1);
#pragma unroll
                for (int o = 2; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            } else {
                // lanes with bit 0 clear keep rows 0, 1, the others rows 2, 3; then bit 1 selects within the pair
                const bool b0 = lane & 1, b1 = lane & 2;
                float x0 = b0 ? acc[RB - 2] : acc[0], x1 = b0 ? acc[RB - 1] : acc[1];
                x0 += __shfl_xor_sync(0xffffffffu, b0 ? acc[0] : acc[RB - 2], 1);
                x1 += __shfl_xor_sync(0xffffffffu, b0 ? acc[1] : acc[RB - 1], 1);
                v = b1 ? x1 : x0;
                v += __shfl_xor_sync(0xffffffffu, b1 ? x0 : x1, 2);
#pragma unroll
                for (int o = 4; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                // lane l (0..3) now holds row ((l & 1) << 1) | (l >> 1): 0, 2, 1, 3
            }
            if (RB == 2 && KS > 1) {
                if (lane < RB) sh.slots[par][warp][lane] = v;
                named_bar_sync(4 + rg, KS * 32);
                if (ks == 0 && lane < RB) {
                    float t = 0.0f;
                    for (int k = 0; k < KS; ++k) t += sh.slots[par][warp + k][lane];
                    v = t;
                }
                par ^= 1;
            }
            if (pair) {
                // RB = 2: lane 0 gate, lane 1 up.  RB = 4: rows (gate0, up0, gate1, up1) sit in lanes (0, 2, 1, 3)
                const float up = RB == 2 ? __shfl_sync(0xffffffffu, v, 1) : __shfl_down_sync(0xffffffffu, v, 2);
                if (live) v = gs_glu(ph.glu_op, v, up, ph.glu_alpha, ph.glu_limit);
            }
            if (live) {
                if (COHERENT && vs.bias && vs.bias_ll) bv = ll_settle1<PEER>(bw, (const unsigned long long *) vs.bias_ll + row + my, sh.tag + vs.bias_tag);
                unsigned long long * dll = COHERENT ? (unsigned long long *) vs.dst_ll : nullptr;
                gs_put<COHERENT, PEER>(p, ph, sh, (float *) vs.dst + row + my, dll ? dll + row + my : nullptr, v + bv);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sh.empty[pos.st]);
        pos.next();
        ga0 = (ga0 + gstep) & 15u; ga1 = (ga1 + gstep) & 15u;
    }
}

template <int TYPE, bool COHERENT, bool PEER>
__device__ __forceinline__ void consume_dispatch(const gs_params & p, const gs_ctx & c, const gs_phase & ph, const gs_seg & sg, int r0, int r1, gs_pos & pos, int warp, int lane) {
    if (sg.KS == 1) {                                      // one warp per row: passes of four rows
        if (sg.spw == 1)      consume_segment<TYPE, 1, 4, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
        else if (sg.spw == 2) consume_segment<TYPE, 2, 4, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
        else if (item<TYPE>::max_spw >= 4)
            consume_segment<TYPE, (item<TYPE>::max_spw >= 4 ? 4 : 1), 4, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
        return;
    }
    if (sg.spw == 1)      consume_segment<TYPE, 1, 2, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
    else if (sg.spw == 2) consume_segment<TYPE, 2, 2, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
    else if (item<TYPE>::max_spw >= 4)
        consume_segment<TYPE, (item<TYPE>::max_spw >= 4 ? 4 : 1), 2, COHERENT, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
}

#define GS_STAMP(ip, k) do { if (p.dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) \
    p.dbg[((ip) * 2 + (blockIdx.x ? 1 : 0)) * 8 + (k)] = gs_clk(); } while (0)
#define GS_STAMP_MAX(ip, k) do { if (p.dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) \
    atomicMax(p.dbg + ((ip) * 2 + (blockIdx.x ? 1 : 0)) * 8 + (k), gs_clk()); } while (0)

// ---------------------------------------------------------------- phase pieces
// prologue of a GEMV phase (consumers + finalizer): x -> [rms norm * w] -> q8_1 in shared memory.
// COHERENT selects L2-coherent loads: inside a chain x was written by other SMs during this launch.
template <bool COHERENT>
__device__ __forceinline__ float4 gs_ld4(const float * p) {
    return COHERENT ? __ldcg((const float4 *) p) : *(const float4 *) p;
}
// eight consecutive elements of an activation an earlier phase produced
template <bool SYS>
__device__ __forceinline__ void ll_read8(const unsigned long long * l, unsigned int tag, float * out) {
    ulonglong2 a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = ll_load2<SYS>(l + 2 * i);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        ll_settle2<SYS>(a[i], l + 2 * i, tag);
        out[2 * i] = __uint_as_float((unsigned int) a[i].x); out[2 * i + 1] = __uint_as_float((unsigned int) a[i].y);
    }
}
// Prologue of a GEMV phase: x -> [rms norm * w] -> q8_1 in shared memory, arithmetic of quantize_row_q8_1_ref
// (ggml-quants.c:225-258).  Four lanes own a 32-element block, eight consecutive elements per lane: K = 4096 is ONE pass of
// the 512 prologue threads (every CTA repeats this work for every phase: it was a fifth of the launch's instructions).
template <bool COHERENT, bool PEER>
__device__ __forceinline__ void gs_prologue(const gs_params & p, const gs_ctx & c, const gs_phase & ph, int warp, int lane) {
    const int K = ph.K, nblk = K >> 5;
    // inside a chain x is normally the output of an earlier phase: tagged words, each thread spins on the ones it needs
    const unsigned long long * xl = COHERENT ? ph.x_ll : nullptr;
    const unsigned int xtag = COHERENT ? c.sh->tag + ph.x_tag : 0u;
    const float * __restrict__ x = ph.x;
    const float * __restrict__ nw = ph.norm_w;
    blk_q8_1 * xq_out = blockIdx.x == 0 ? ph.xq_out : nullptr;
    gs_shared & sh = *c.sh;
    const int sub = lane & 3, bw = lane >> 2;                   // 4 lanes per 32-element block, 8 blocks per warp pass
    constexpr int BPP = GS_NW * 8;                              // blocks per pass of the prologue warps

    auto load8 = [&](int e, float * v) {
        if (xl) { ll_read8<PEER>(xl + e, xtag, v); return; }
        const float4 a = gs_ld4<COHERENT>(x + e), b = gs_ld4<COHERENT>(x + e + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    };
    auto quant_store = [&](int blk, bool act, float * v, float scale) {
        if (act && nw) {                                        // same op order as rms_norm then mul: (x*scale)*w
            const float4 w0 = *(const float4 *) (nw + blk * 32 + sub * 8), w1 = *(const float4 *) (nw + blk * 32 + sub * 8 + 4);
            v[0] = __fmul_rn(__fmul_rn(v[0], scale), w0.x); v[1] = __fmul_rn(__fmul_rn(v[1], scale), w0.y);
            v[2] = __fmul_rn(__fmul_rn(v[2], scale), w0.z); v[3] = __fmul_rn(__fmul_rn(v[3], scale), w0.w);
            v[4] = __fmul_rn(__fmul_rn(v[4], scale), w1.x); v[5] = __fmul_rn(__fmul_rn(v[5], scale), w1.y);
            v[6] = __fmul_rn(__fmul_rn(v[6], scale), w1.z); v[7] = __fmul_rn(__fmul_rn(v[7], scale), w1.w);
        }
        float d; int sum;
        const int2 q = quantize_block32_oct(v, d, sum);
        if (act) {
            *(int2 *) (c.aq + act_off(blk) + sub * 8) = q;
            if (sub == 0) {
                const float dh = __half2float(__float2half_rn(d));          // the value a q8_1 block stores
                c.ads[blk] = make_float2(dh, dh * (float) sum);
            }
            if (xq_out) {
                int * out = (int *) (xq_out + blk);
                out[1 + 2 * sub] = q.x; out[2 + 2 * sub] = q.y;
                if (sub == 0) *(__half2 *) out = q8_1_scales(d, sum);
            }
        }
    };

    if (xl) {
        // gate: one lane per warp waits (with back-off) for the first word this warp needs; the other 31 lanes and the
        // CTAs that finished early add no traffic while a phase is still running
        if (lane == 0 && warp * 8 < nblk) ll_gate<PEER>(xl + (warp * 8) * 32, xtag);
        __syncwarp();
    }
    auto reg_path = [&](auto nv_tag) {
        constexpr int NV = decltype(nv_tag)::value;             // passes (of BPP blocks) this row length needs: no predicated-off work
        float xv[NV][8];
#pragma unroll
        for (int it = 0; it < NV; ++it) {
            const int blk = warp * 8 + bw + it * BPP;
#pragma unroll
            for (int i = 0; i < 8; ++i) xv[it][i] = 0.0f;
            if (blk < nblk) load8(blk * 32 + sub * 8, xv[it]);
        }
        if (COHERENT && warp == 0 && lane == 0) GS_STAMP(ph.self_tag, 5);
        float scale = 1.0f;
        if (nw) {
            float ss = 0.0f;
#pragma unroll
            for (int it = 0; it < NV; ++it)
#pragma unroll
                for (int i = 0; i < 8; ++i) ss += xv[it][i] * xv[it][i];
            ss = warp_sum(ss);
            if (lane == 0) sh.part[warp] = ss;
            named_bar_sync(1, GS_PRO_THREADS);
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < GS_NW; ++i) tot += sh.part[i];
            scale = rsqrtf(tot / (float) K + ph.eps);
        }
#pragma unroll
        for (int it = 0; it < NV; ++it) {
            const int b0 = warp * 8 + it * BPP;                 // warp-uniform
            if (b0 < nblk) quant_store(b0 + bw, b0 + bw < nblk, xv[it], scale);
        }
        if (COHERENT && warp == 0 && lane == 0) GS_STAMP(ph.self_tag, 6);
    };
    if (nblk <= BPP)          reg_path(std::integral_constant<int, 1>{});
    else if (nblk <= BPP * 2) reg_path(std::integral_constant<int, 2>{});
    else if (nblk <= BPP * 4) reg_path(std::integral_constant<int, 4>{});
    else {
        // ---- long rows: two passes over x
        float scale = 1.0f;
        if (nw) {
            float ss = 0.0f;
            for (int b0 = warp * 8; b0 < nblk; b0 += BPP) {
                float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                if (b0 + bw < nblk) load8((b0 + bw) * 32 + sub * 8, v);
#pragma unroll
                for (int i = 0; i < 8; ++i) ss += v[i] * v[i];
            }
            ss = warp_sum(ss);
            if (lane == 0) sh.part[warp] = ss;
            named_bar_sync(1, GS_PRO_THREADS);
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < GS_NW; ++i) tot += sh.part[i];
            scale = rsqrtf(tot / (float) K + ph.eps);
        }
        for (int b0 = warp * 8; b0 < nblk; b0 += BPP) {
            float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const bool act = b0 + bw < nblk;
            if (act) load8((b0 + bw) * 32 + sub * 8, v);
            quant_store(b0 + bw, act, v, scale);
        }
    }
    named_bar_sync(1, GS_PRO_THREADS);
}

// ---------------------------------------------------------------- attention phase (decode, one query position)
// Item = one query head.  Replaces, for n_q = 1 and an f16 cache, the reference's ROPE(q), ROPE(k) (+SET_ROWS k, v) and
// FLASH_ATTN_EXT nodes (rope.cu:563, set-rows.cu:318, fattn-vec.cuh:19-507).  The current token's K/V rows are taken
// from the raw projections (RoPE applied, rounded to f16 exactly as the cache stores them) and written to the cache by
// the first query head of each KV group; all other positions are read from the cache.
// one element of q / k_new / v_new: a tagged word when a phase of this chain produced it
template <bool PEER>
__device__ __forceinline__ float gs_attn_ld(const float * base, const unsigned long long * ll, int64_t i, unsigned int tag) {
    return ll ? ll_read<PEER>(ll + i, tag) : __ldcg(base + i);
}
// (cos, sin) of the rotation of pair i at position pos, YaRN included (reference ggml-cpu/ops.cpp rope_yarn); one copy, out of line
static __device__ __noinline__ float2 gs_rope_angle(int pos, float theta_scale, int i, float freq_scale, float ext_factor, float attn_factor, float corr0, float corr1) {
    const float theta_extrap = (float) pos * powf(theta_scale, (float) i);
    float theta = freq_scale * theta_extrap, mscale = attn_factor;
    if (ext_factor != 0.0f) {
        const float y = ((float) i - corr0) / fmaxf(0.001f, corr1 - corr0);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * ext_factor;
        theta = theta * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / freq_scale);
    }
    float sn, cs; sincosf(theta, &sn, &cs);
    return make_float2(cs * mscale, sn * mscale);
}
template <int D, bool PEER>
__device__ __noinline__ void gs_attn_item(unsigned int ltag, unsigned int self_tag, const gs_attn & a, int h, float * sm, int warp, int lane) {
    constexpr int LPP = D / 32, PPB = 32 / LPP, DPL = D / 32;     // lanes per position, positions per batch, dims per lane (V)
    const int tid = warp * 32 + lane;
    const int gqa = a.H / a.H_kv, g = h / gqa;
    const int pos = __ldcg(a.pos), slot = (int) __ldcg(a.slot), slot_v = (int) __ldcg(a.slot_v);
    float * s_q = sm, * s_k = sm + D, * s_v = sm + 2 * D, * s_part = sm + 3 * D;      // s_part: [GS_NW][D + 2]
    const bool q8 = a.kv_type == B200_Q8_0;
    // ---- 1. RoPE(q[h]) * scale, RoPE(k_new[g]) and v_new[g] rounded to f16 (f16 cache) or quantised below (q8_0 cache)
    {
        const int half_d = a.n_dims / 2;
        if (tid < D) {                                            // threads 0..D/2-1: q pairs, D/2..D-1: k pairs
            const bool isk = tid >= D / 2;
            const int i = isk ? tid - D / 2 : tid;
            const float * src = isk ? a.k_new : a.q;
            const int64_t so = isk ? (int64_t) g * D : (int64_t) h * D;
            const unsigned long long * sll = isk ? a.k_ll : a.q_ll;
            const unsigned int stag = ltag + (isk ? a.k_tag : a.q_tag);
            float * dst = isk ? s_k : s_q;
            if (i < half_d) {
                const int i0 = a.rope_mode == 2 ? i : 2 * i, i1 = a.rope_mode == 2 ? i + half_d : 2 * i + 1;
                const float2 csn = gs_rope_angle(pos, a.theta_scale, i, a.freq_scale, a.ext_factor, a.attn_factor, a.corr0, a.corr1);
                const float cs = csn.x, sn = csn.y;
                const float x0 = gs_attn_ld<PEER>(src, sll, so + i0, stag), x1 = gs_attn_ld<PEER>(src, sll, so + i1, stag);
                float y0 = x0 * cs - x1 * sn, y1 = x0 * sn + x1 * cs;
                if (isk) { if (!q8) { y0 = __half2float(__float2half_rn(y0)); y1 = __half2float(__float2half_rn(y1)); } }
                else     { y0 *= a.scale; y1 *= a.scale; }
                dst[i0] = y0; dst[i1] = y1;
            }
            for (int d = a.n_dims + i; d < D; d += D / 2) {       // dims beyond n_dims pass through
                const float y = gs_attn_ld<PEER>(src, sll, so + d, stag);
                dst[d] = isk ? (q8 ? y : __half2float(__float2half_rn(y))) : y * a.scale;
            }
        } else if (tid < 2 * D) {
            const int d = tid - D;
            const float y = gs_attn_ld<PEER>(a.v_new, a.v_ll, (int64_t) g * D + d, ltag + a.v_tag);
            s_v[d] = q8 ? y : __half2float(__float2half_rn(y));
        }
    }
    named_bar_sync(3, GS_CON_THREADS);
    if (q8) {
        // q8_0 cache: a warp quantises one 32-element block of the new K / V row with the arithmetic of quantize_row_q8_0_ref
        // (ggml-quants.c:197-223: d = amax / 127, q = roundf(x / d)); the attention below sees what the cache row decodes to
        if (tid < 2 * D) {
            const bool isv = tid >= D;
            const int d = isv ? tid - D : tid;
            float * sv = isv ? s_v : s_k;
            const float x = sv[d];
            float amax = fabsf(x);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
            const float dd = amax / 127.0f;
            const float id = dd != 0.0f ? 1.0f / dd : 0.0f;
            const int qv = (int) roundf(x * id);
            const __half dh = __float2half_rn(dd);
            sv[d] = __half2float(dh) * (float) qv;
            if (h % gqa == 0) {                                   // one writer per group
                char * blk = (isv ? a.v_cache + (int64_t) slot_v * a.v_nb1 + (int64_t) g * a.v_nb2
                                  : a.k_cache + (int64_t) slot * a.k_nb1 + (int64_t) g * a.k_nb2) + (d >> 5) * 34;
                if (lane == 0) *(__half *) blk = dh;
                ((int8_t *) blk)[2 + lane] = (int8_t) qv;
            }
        }
        named_bar_sync(3, GS_CON_THREADS);
    } else if (h % gqa == 0 && tid < 2 * D) {                     // KV-cache write of the current token (one writer per group)
        const int d = tid < D ? tid : tid - D;
        char * row = tid < D ? a.k_cache + (int64_t) slot * a.k_nb1 + (int64_t) g * a.k_nb2
                             : a.v_cache + (int64_t) slot_v * a.v_nb1 + (int64_t) g * a.v_nb2;
        ((__half *) row)[d] = __float2half_rn(tid < D ? s_k[d] : s_v[d]);
    }
    // ---- 2. scores + online softmax + V accumulation, KV positions strided over the warps
    float qreg[32];
    const int qsub = lane % LPP;
#pragma unroll
    for (int i = 0; i < 32; ++i) qreg[i] = s_q[qsub * 32 + i];
    float m = -INFINITY, l = 0.0f, acc[DPL];
#pragma unroll
    for (int i = 0; i < DPL; ++i) acc[i] = 0.0f;
    const char * kb = a.k_cache + (int64_t) g * a.k_nb2;
    const char * vb = a.v_cache + (int64_t) g * a.v_nb2;
    for (int base = warp * PPB; base < a.n_kv; base += GS_NW * PPB) {
        const int p = base + lane / LPP;
        const float mk = p < a.n_kv ? (a.mask ? __half2float(a.mask[p]) : 0.0f) : -INFINITY;
        float s = 0.0f;
        if (mk != -INFINITY) {
            if (p == slot) {
#pragma unroll
                for (int i = 0; i < 32; ++i) s += qreg[i] * s_k[qsub * 32 + i];
            } else if (q8) {
                // block qsub of row p: 34 bytes from a 2-byte aligned address, read as aligned words + funnel shifts
                const uintptr_t addr = (uintptr_t) (kb + (int64_t) p * a.k_nb1 + qsub * 34);
                const uint32_t sh = (uint32_t) (addr & 3) * 8;
                const uint32_t * wq = (const uint32_t *) (addr & ~(uintptr_t) 3);
                uint32_t ww[10];
#pragma unroll
                for (int u = 0; u < 9; ++u) ww[u] = wq[u];
                ww[9] = 0u;
                uint32_t xx[9];
#pragma unroll
                for (int u = 0; u < 9; ++u) xx[u] = __funnelshift_r(ww[u], ww[u + 1], sh);
                const unsigned short db = (unsigned short) (xx[0] & 0xFFFFu);
#pragma unroll
                for (int v = 0; v < 8; ++v) {
                    const uint32_t y = __funnelshift_r(xx[v], xx[v + 1], 16);
                    s += qreg[4 * v] * (float) (int8_t) (y & 0xFF) + qreg[4 * v + 1] * (float) (int8_t) ((y >> 8) & 0xFF) +
                         qreg[4 * v + 2] * (float) (int8_t) ((y >> 16) & 0xFF) + qreg[4 * v + 3] * (float) (int8_t) (y >> 24);
                }
                s *= __half2float(*(const __half *) &db);
            } else {
                const uint4 * kr = (const uint4 *) (kb + (int64_t) p * a.k_nb1 + qsub * 64);
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4) {
                    const uint4 t = kr[c4];
                    const __half2 * hp = (const __half2 *) &t;
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const float2 f = __half22float2(hp[i]); s += qreg[c4 * 8 + 2 * i] * f.x + qreg[c4 * 8 + 2 * i + 1] * f.y; }
                }
            }
        }
#pragma unroll
        for (int o = 1; o < LPP; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        s = mk == -INFINITY ? -INFINITY : s + mk;
        float bm = s;
#pragma unroll
        for (int o = LPP; o < 32; o <<= 1) bm = fmaxf(bm, __shfl_xor_sync(0xffffffffu, bm, o));
        const float mn = fmaxf(m, bm);
        if (mn == -INFINITY) continue;                            // whole batch masked (warp-uniform)
        const float corr = __expf(m - mn);
        const float pr = s == -INFINITY ? 0.0f : __expf(s - mn);
        float psum = pr;
#pragma unroll
        for (int o = LPP; o < 32; o <<= 1) psum += __shfl_xor_sync(0xffffffffu, psum, o);
        l = l * corr + psum; m = mn;
#pragma unroll
        for (int i = 0; i < DPL; ++i) acc[i] *= corr;
#pragma unroll
        for (int j = 0; j < PPB; ++j) {
            const float pj = __shfl_sync(0xffffffffu, pr, j * LPP);
            const int pp = base + j;
            if (pj != 0.0f) {                                     // warp-uniform
                if (pp == slot_v) {
#pragma unroll
                    for (int i = 0; i < DPL; ++i) acc[i] += pj * s_v[lane * DPL + i];
                } else if (q8) {
                    // dims lane * DPL .. of row pp: DPL int8 inside block (lane * DPL) / 32, and the block's scale
                    const char * blk = vb + (int64_t) pp * a.v_nb1 + ((lane * DPL) >> 5) * 34;
                    const float dv = pj * __half2float(*(const __half *) blk);
                    const unsigned short * qp = (const unsigned short *) (blk + 2 + ((lane * DPL) & 31));
#pragma unroll
                    for (int i = 0; i < DPL / 2; ++i) {
                        const unsigned short t = qp[i];
                        acc[2 * i] += dv * (float) (int8_t) (t & 0xFF); acc[2 * i + 1] += dv * (float) (int8_t) (t >> 8);
                    }
                } else {
                    const __half * vr = (const __half *) (vb + (int64_t) pp * a.v_nb1) + lane * DPL;
                    if (DPL == 4) { const uint2 t = *(const uint2 *) vr; const __half2 * hp = (const __half2 *) &t;
                                    const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]);
                                    acc[0] += pj * f0.x; acc[1] += pj * f0.y; acc[2] += pj * f1.x; acc[3] += pj * f1.y; }
                    else          { const float2 f0 = __half22float2(*(const __half2 *) vr); acc[0] += pj * f0.x; acc[1] += pj * f0.y; }
                }
            }
        }
    }
    // ---- 3. combine the warps' states
    float * mine = s_part + warp * (D + 2);
#pragma unroll
    for (int i = 0; i < DPL; ++i) mine[lane * DPL + i] = acc[i];
    if (lane == 0) { mine[D] = m; mine[D + 1] = l; }
    named_bar_sync(3, GS_CON_THREADS);
    if (tid < D) {
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < GS_NW; ++w) M = fmaxf(M, s_part[w * (D + 2) + D]);
        float A = 0.0f, L = 0.0f;
#pragma unroll
        for (int w = 0; w < GS_NW; ++w) {
            const float mw = s_part[w * (D + 2) + D];
            const float wgt = mw == -INFINITY ? 0.0f : __expf(mw - M);
            A += wgt * s_part[w * (D + 2) + tid]; L += wgt * s_part[w * (D + 2) + D + 1];
        }
        if (a.sinks) {                                            // the sink: one more logit in the softmax, no value
            const float sk = a.sinks[h], M2 = fmaxf(M, sk), c0 = M == -INFINITY ? 0.0f : __expf(M - M2);
            A *= c0; L = L * c0 + __expf(sk - M2);
        }
        const float o = A / L;
        a.dst[(int64_t) h * D + tid] = o;
        if (a.dst_ll) ll_store<PEER>(a.dst_ll + (int64_t) h * D + tid, o, ltag + self_tag);      // every rank computes every head: local only
    }
    named_bar_sync(3, GS_CON_THREADS);                            // scratch is reused by the next item
}

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int * p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int * p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// thread 0 of a CTA: wait until counter `bar` has seen every arrival of this launch.  Bounded by wall time: a rank
// that never arrives (crashed process, mismatched phase lists) becomes a trap, not a hung GPU.
template <bool PEER>
__device__ __forceinline__ void gs_wait_counter(const unsigned int * bar, unsigned int target) {
    if (!PEER) {
        // bounded by WALL time, not by iterations: under compute-sanitizer, ncu replay or GPU time-slicing a healthy wait can
        // outlast any iteration count.  20 s without progress is a deadlock (a trap ends the context instead of hanging the GPU)
        unsigned int spins = 0; unsigned long long t0 = 0;
        while (ld_acquire_gpu(bar) < target) {
            if ((++spins & 0xFFFu) == 0) {
                const unsigned long long now = gs_now();
                if (!t0) t0 = now; else if (now - t0 > 20000000000ull) __trap();
                __nanosleep(64);
            }
        }
    } else {
        unsigned int spins = 0; unsigned long long t0 = 0;
        while ((int) (ld_acquire_sys(bar) - target) < 0) {
            if ((++spins & 0xFFFu) == 0) {
                const unsigned long long now = gs_now();
                if (!t0) t0 = now; else if (now - t0 > 20000000000ull) __trap();      // 20 s
            }
        }
    }
}

// ---------------------------------------------------------------- the kernel
// MODE 0: one phase (b200_gemv_fused, PDL); 1: decode chain on one GPU; 2: decode chain, row split over peer memory
template <int T0, int T1, int MODE>
__global__ void __launch_bounds__(GS_THREADS, 1) gemv_stream_kernel(const __grid_constant__ gs_params p) {
    constexpr bool CHAIN = MODE >= 1, PEER = MODE == 2;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ gs_shared sh;
    // phase descriptors are staged in shared memory: read from global inside the chunk loops they cost an L2 round
    // trip per field per chunk (measured: 1.3 us per chunk in the finalizer)
    __shared__ gs_phase s_ph[2];       // consumers + finalizer (double-buffered: the finalizer may trail by a phase)
    __shared__ gs_phase s_pp;          // producer (runs ahead)
    __shared__ int2 s_rng[2][B200_GEMV_MAX_SEG], s_prng[B200_GEMV_MAX_SEG];      // this CTA's row ranges of the staged phases
    gs_ctx c;
    c.ring = smem;                                                            // n_stages x stage_stride
    c.aq   = smem + p.n_stages * p.stage_stride;                              // activation int8, padded layout
    c.ads  = (float2 *) (c.aq + p.act_qs_bytes);                              // (d, d*sum) per unit; sized with act_qs_bytes/4
    c.attn = (float *) (((uintptr_t) ((uint8_t *) c.ads + (p.act_qs_bytes >> 2) + 64) + 15) & ~(uintptr_t) 15);   // attention scratch (read as float4)
    c.sh = &sh; c.stage_stride = p.stage_stride; c.flags = p.flags;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < p.n_stages; ++s) { mbar_init(&sh.full[s], 1); mbar_init(&sh.empty[s], GS_NW); }
        // row split: tags of this launch = (launches so far + 1) * 1024 + phase index (never 0, the value of fresh memory)
        sh.tag = CHAIN ? (__ldcg(p.epoch) + 1u) << 10 : 0u;
        if ((p.flags & 4) && p.dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1))     // debug accumulators of this launch
            for (int i = 0; i < p.n_phases; ++i) { unsigned long long * dd = p.dbg + (i * 2 + (blockIdx.x ? 1 : 0)) * 8; dd[3] = dd[4] = dd[7] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (!CHAIN) pdl_launch_dependents();      // the next kernel in the stream may begin its own weight prefetch

    const gs_phase * phases = CHAIN ? p.phases : &p.inl;
    const int n_phases = CHAIN ? p.n_phases : 1;
    gs_pos pos; pos.ns = p.n_stages;
    constexpr int PH_VEC = (int) (sizeof(gs_phase) / 16);
    static_assert(sizeof(gs_phase) % 16 == 0, "gs_phase is copied as uint4");

    if (warp >= GS_NW) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(GS_REG_PRODUCER));
        if (warp > GS_NW) return;
        // ================= producer: TMA bulk copies into the ring, straight through all phases.  The whole warp walks
        // the chunk sequence (warp-uniform state); lane 0 issues the copies.
        // (Measured and dropped: a second walker requesting chunks into L2 2-8 chunks ahead of the ring — bulk prefetch
        // or per-line prefetch.global.L2 — ran 1-4 % slower; the consume phases are not short of HBM bytes in flight.)
        const uint64_t pol = policy_evict_first();
        gs_walk wt;
        while (gs_walk_next(wt, &s_pp, s_prng, phases, CHAIN ? p.ranges : nullptr, n_phases, blockIdx.x, gridDim.x, lane)) {
            if (lane == 0) {
                const gs_seg & sg = s_pp.seg[wt.s];
                const int nr = min(sg.R, wt.r1 - wt.row);
                mbar_wait(&sh.empty[pos.st], pos.full_par ^ 1);
                if ((p.flags & 4) && p.dbg && wt.row == wt_first_row(wt, s_pp) && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) {
                    unsigned long long * dd = p.dbg + (wt.ip * 2 + (blockIdx.x ? 1 : 0)) * 8;
                    if (!dd[7]) dd[7] = gs_clk();                  // debug: when the FIRST chunk of this phase was requested
                }
                uint8_t * dstp = c.ring + pos.st * p.stage_stride;
                const uintptr_t g0 = (uintptr_t) (sg.w[0] + (int64_t) wt.row * sg.rb);
                const uint32_t o0 = (uint32_t) (g0 & 15), b0 = (o0 + (uint32_t) nr * sg.rb + 15u) & ~15u;
                if (sg.nsub == 2) {
                    const uintptr_t g1 = (uintptr_t) (sg.w[1] + (int64_t) wt.row * sg.rb);
                    const uint32_t o1 = (uint32_t) (g1 & 15), b1 = (o1 + (uint32_t) nr * sg.rb + 15u) & ~15u;
                    mbar_expect_tx(&sh.full[pos.st], b0 + b1);
                    bulk_g2s(dstp, (const void *) (g0 - o0), b0, &sh.full[pos.st], pol);
                    bulk_g2s(dstp + b0, (const void *) (g1 - o1), b1, &sh.full[pos.st], pol);
                } else {
                    mbar_expect_tx(&sh.full[pos.st], b0);
                    bulk_g2s(dstp, (const void *) (g0 - o0), b0, &sh.full[pos.st], pol);
                }
            }
            __syncwarp();
            wt.row += s_pp.seg[wt.s].R;
            pos.next();
        }
        return;
    }

    // ================= consumers (warps 0..GS_NW-1)
    // There is no grid barrier between the phases of a chain: whatever a phase reads from an earlier one arrives as tagged
    // words (see gs_seg), so a CTA moves on as soon as ITS inputs exist, publishes nothing and fences nothing.  No CTA can be
    // more than one phase ahead of the slowest one: a GEMV prologue needs every row of the phase before it.
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(GS_REG_CONSUMER));
    for (int ip = 0; ip < n_phases; ++ip) {
        const gs_phase & ph = s_ph[ip & 1];
        // the descriptor is double-buffered: slower warps of this CTA may still be reading the previous one
        for (int i = tid; i < PH_VEC; i += GS_CON_THREADS) ((uint4 *) &s_ph[ip & 1])[i] = ((const uint4 *) (phases + ip))[i];
        if (CHAIN) gs_stage_ranges(s_rng[ip & 1], &ph, p.ranges, ip, blockIdx.x, gridDim.x, tid);
        if (ip == 0) pdl_wait();                                   // plain inputs were written by earlier kernels
        named_bar_sync(1, GS_CON_THREADS);                         // descriptor staged; every warp has left phase ip - 1
        if (!CHAIN) { gs_stage_ranges(s_rng[0], &ph, nullptr, 0, blockIdx.x, gridDim.x, tid); named_bar_sync(1, GS_CON_THREADS); }
        if (CHAIN && tid == 0) GS_STAMP(ip, 0);
        if (CHAIN && tid == 0 && p.dbg) p.dbg[16 * (size_t) n_phases + ((size_t) ip * gridDim.x + blockIdx.x) * 2] = gs_now();
        if (ph.kind == 0) {
            gs_prologue<CHAIN, PEER>(p, c, ph, warp, lane);
            if (CHAIN && tid == 0) GS_STAMP(ip, 1);
            for (int s = 0; s < ph.n_ws; ++s) {
                const gs_seg & sg = ph.seg[s];
                const int r0 = s_rng[ip & 1][s].x, r1 = s_rng[ip & 1][s].y;
                if (r0 >= r1) continue;
                if (sg.type == T0) consume_dispatch<T0, CHAIN, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
                else               consume_dispatch<T1, CHAIN, PEER>(p, c, ph, sg, r0, r1, pos, warp, lane);
            }
        } else if (CHAIN && ph.kind == 1) {
            for (int h = blockIdx.x; h < ph.attn.H; h += gridDim.x) {
                unsigned long long * dbg = p.dbg && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1) ? p.dbg + (ip * 2 + (blockIdx.x ? 1 : 0)) * 8 : nullptr;
                if (ph.attn.D == 128) gs_attn_item<128, PEER>(sh.tag, ph.self_tag, ph.attn, h, c.attn, warp, lane);
                else                  gs_attn_item<64, PEER>(sh.tag, ph.self_tag, ph.attn, h, c.attn, warp, lane);
            }
        }
        if (CHAIN && tid == 0) GS_STAMP(ip, 2);
        if (CHAIN && tid == 0 && p.dbg) p.dbg[16 * (size_t) n_phases + ((size_t) ip * gridDim.x + blockIdx.x) * 2 + 1] = gs_now();
        if (PEER && ph.peer_mode == 2) {
            // outputs read after the launch (the logits) went to the other ranks as plain stores: CTA barrier, one
            // system-scope fence, then the arrival is counted on every rank
            named_bar_sync(1, GS_CON_THREADS);
            if (tid == 0) {
                __threadfence_system();
                atomicAdd_system(p.sync + ip, 1u);
#pragma unroll
                for (int r = 0; r < B200_MAX_RANKS - 1; ++r)
                    if (r < p.n_peers) atomicAdd_system((unsigned int *) ((char *) (p.sync + ip) + p.peer_delta[r]), 1u);
            }
        }
    }
    if (CHAIN && tid == 0) {
        if (PEER) {
            // the launch ends only when every rank has delivered its plain-stored rows (the logits) into this rank's buffers
            if (blockIdx.x == 0)
                for (int k = 0; k < p.n_final; ++k) gs_wait_counter<true>(p.sync + p.final_phase[k], (unsigned int) p.total_ctas);
        }
        // the last CTA to leave counts the launch: every CTA read the epoch in its first instructions, and the next launch
        // starts only after this one has exited (same stream)
        if (atomicAdd(p.exit_count, 1u) == gridDim.x - 1) {
            const unsigned int e = sh.tag >> 10;                   // = epoch + 1; tags are (epoch + 1) << 10 | phase: keep them non-zero in 32 bits
            *p.exit_count = 0u; *p.epoch = e >= 0x3FFFFEu ? 0u : e;
        }
    }
}

// launcher of one instantiation (explicitly instantiated in gemv_stream_i_*.cu)
template <int T0, int T1, int MODE>
int gs_launch(const gs_params & p, size_t smem, int grid, cudaStream_t st) {
    constexpr bool CHAIN = MODE >= 1;
    static std::once_flag attr_once[64];                       // per device; launches may come from several host threads
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return B200_ERR_ARG;
    cudaError_t attr_rc = cudaSuccess;
    std::call_once(attr_once[dev], [&] { attr_rc = cudaFuncSetAttribute(gemv_stream_kernel<T0, T1, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, GS_SMEM_MAX); });
    B200_CUDA_TRY(attr_rc);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned) grid); cfg.blockDim = dim3(GS_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    if (CHAIN) { at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1; }     // all CTAs co-resident: the grid barriers need it
    else       { at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1; }
    cfg.attrs = at; cfg.numAttrs = 1;
    B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, gemv_stream_kernel<T0, T1, MODE>, p));
    count_launch();
    return B200_OK;
}

}  // namespace b200
