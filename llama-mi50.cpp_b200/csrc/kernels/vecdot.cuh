// vecdot.cuh — per-format "unit" unpackers for the quantized dot product.
//
// Every supported weight format is consumed in UNITS of 32 consecutive weights, which pair with
// exactly one q8_1 activation block {d8, s8, u[32]}.  A unit unpacks to
//     v[8]   : the 32 integer weight values as int8x4 words, in element order
//     wd[2]  : f32 multiplier of elements 0..15 / 16..31 (block scale x sub-scale)
//     wm     : f32 multiplier of the "min" term (q4_K only, else 0)
// and contributes   d8 * (wd[0]*dot(v[0..3],u[0..3]) + wd[1]*dot(v[4..7],u[4..7]) - wm*sum(u)).
// The integer dots are exact (dp4a); only the f32 scale/accumulate order differs from the
// reference's per-type vec_dot_*_q8_1 (ggml-cuda/vecdotq.cuh:103-122 q4_0, :231-243 q8_0,
// :295-314 mxfp4, :461-483 q4_K, :580-600 q6_K), which computes the same sums.
#pragma once
#include "common.cuh"

namespace b200 {

struct unit_w {
    int   v[8];
    float wd[2];
    float wm;
};

template <int TYPE> struct fmt;   // ::block_bytes, ::units_per_block, ::has_min, unpack()

// ---- Q4_0: 18 B block {f16 d; nibbles}, low nibble = element j, high = element j+16, value-8
template <> struct fmt<B200_Q4_0> {
    static constexpr int block_bytes = 18, units_per_block = 1; static constexpr bool has_min = false;
    static __device__ __forceinline__ void unpack(const uint8_t * row, int64_t u, unit_w & w) {
        const uint8_t * b = row + u * 18;
        const float d = __half2float(*(const __half *) b);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t q = ld_u32_a2(b + 2 + 4 * i);
            w.v[i]     = __vsubss4(q & 0x0F0F0F0Fu, 0x08080808u);
            w.v[4 + i] = __vsubss4((q >> 4) & 0x0F0F0F0Fu, 0x08080808u);
        }
        w.wd[0] = w.wd[1] = d; w.wm = 0.0f;
    }
};

// ---- Q8_0: 34 B block {f16 d; int8 qs[32]}
template <> struct fmt<B200_Q8_0> {
    static constexpr int block_bytes = 34, units_per_block = 1; static constexpr bool has_min = false;
    static __device__ __forceinline__ void unpack(const uint8_t * row, int64_t u, unit_w & w) {
        const uint8_t * b = row + u * 34;
        const float d = __half2float(*(const __half *) b);
        uint32_t q[8];
        ld_words_any<8>(b + 2, q);
#pragma unroll
        for (int i = 0; i < 8; ++i) w.v[i] = (int) q[i];
        w.wd[0] = w.wd[1] = d; w.wm = 0.0f;
    }
};

// ---- MXFP4: 17 B block {e8m0 e; nibbles}; value = kvalues[n] * 2^(e-128),
// kvalues = {0,1,2,3,4,6,8,12, 0,-1,-2,-3,-4,-6,-8,-12} (ggml-common.h:1094): bit 3 is the sign.
// PRMT looks up the 3-bit magnitude; a second PRMT in sign-replicate mode builds the 0xFF mask of
// negative lanes from bit 3 of each nibble; negation is done on the integer (mag ^ m) - m.
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;   // raw PRMT: selector bit 3 = replicate the sign of the selected byte (not exposed by __byte_perm)
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}
__device__ __forceinline__ void mxfp4_expand(uint32_t q, int & lo, int & hi) {
    constexpr uint32_t T0 = 0x03020100u, T1 = 0x0c080604u;                 // magnitudes 0..7
    const uint32_t q4 = q << 4;
    // nibble order in q: n0=elem j, n1=elem j+16, n2=elem j+1, n3=elem j+17, ...
    const uint32_t ma = prmt(T0, T1, q & 0x7777u);                  // (j, j+16, j+1, j+17)
    const uint32_t mb = prmt(T0, T1, (q >> 16) & 0x7777u);          // (j+2, j+18, j+3, j+19)
    const uint32_t sa = prmt(q4, q, 0xD9C8u);                       // 0xFF where bit 3 set
    const uint32_t sb = prmt(q4, q, 0xFBEAu);
    const uint32_t va = __vsub4(ma ^ sa, sa);
    const uint32_t vb = __vsub4(mb ^ sb, sb);
    lo = (int) prmt(va, vb, 0x6420u);                               // elements j..j+3
    hi = (int) prmt(va, vb, 0x7531u);                               // elements j+16..j+19
}
template <> struct fmt<B200_MXFP4> {
    static constexpr int block_bytes = 17, units_per_block = 1; static constexpr bool has_min = false;
    static __device__ __forceinline__ void unpack(const uint8_t * row, int64_t u, unit_w & w) {
        const uint8_t * b = row + u * 17;
        const float d = e8m0_half_to_f32(b[0]);
        uint32_t q[4];
        ld_words_any<4>(b + 1, q);
#pragma unroll
        for (int i = 0; i < 4; ++i) mxfp4_expand(q[i], w.v[i], w.v[4 + i]);
        w.wd[0] = w.wd[1] = d; w.wm = 0.0f;
    }
};

// ---- Q4_K: 144 B super-block {f16 d, dmin; 12 B packed 6-bit (scale,min) x 8; 128 B nibbles}
// unit j of a super-block = sub-block j: bytes qs[32*(j/2) .. +31], low nibbles if j even.
__device__ __forceinline__ void q4k_scale_min(const uint8_t * sc, int j, int & s, int & m) {
    if (j < 4) { s = sc[j] & 63; m = sc[j + 4] & 63; }
    else       { s = (sc[j + 4] & 0xF) | ((sc[j - 4] >> 6) << 4); m = (sc[j + 4] >> 4) | ((sc[j] >> 6) << 4); }
}
template <> struct fmt<B200_Q4_K> {
    static constexpr int block_bytes = 144, units_per_block = 8; static constexpr bool has_min = true;
    static __device__ __forceinline__ void unpack(const uint8_t * row, int64_t u, unit_w & w) {
        const uint8_t * b = row + (u >> 3) * 144;
        const int j = (int) (u & 7);
        const float2 dm = __half22float2(*(const __half2 *) b);
        int s, m; q4k_scale_min(b + 4, j, s, m);
        const uint32_t * qs = (const uint32_t *) (b + 16 + 32 * (j >> 1));
        const int sh = 4 * (j & 1);
#pragma unroll
        for (int i = 0; i < 8; ++i) w.v[i] = (int) ((qs[i] >> sh) & 0x0F0F0F0Fu);
        w.wd[0] = w.wd[1] = dm.x * (float) s; w.wm = dm.y * (float) m;
    }
};

// ---- Q6_K: 210 B super-block {ql[128]; qh[64]; int8 scales[16]; f16 d}, value = (6-bit q) - 32,
// one scale per 16 weights. unit j: half h=j/4, quarter r=j%4 (see dequantize_row_q6_K, ggml-quants.c:1762-1791).
template <> struct fmt<B200_Q6_K> {
    static constexpr int block_bytes = 210, units_per_block = 8; static constexpr bool has_min = false;
    static __device__ __forceinline__ void unpack(const uint8_t * row, int64_t u, unit_w & w) {
        const uint8_t * b = row + (u >> 3) * 210;
        const int j = (int) (u & 7), h = j >> 2, r = j & 3;
        const uint8_t * ql = b + 64 * h + 32 * (r & 1);
        const uint8_t * qh = b + 128 + 32 * h;
        const int sh4 = 4 * (r >> 1), sh2 = 2 * r;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const uint32_t l = (ld_u32_a2(ql + 4 * i) >> sh4) & 0x0F0F0F0Fu;
            const uint32_t hbits = ((ld_u32_a2(qh + 4 * i) >> sh2) & 0x03030303u) << 4;
            w.v[i] = __vsubss4(l | hbits, 0x20202020u);
        }
        const int8_t * sc = (const int8_t *) (b + 192) + 8 * h + 2 * r;
        const float d = __half2float(*(const __half *) (b + 208));
        w.wd[0] = d * (float) sc[0]; w.wd[1] = d * (float) sc[1]; w.wm = 0.0f;
    }
};

// dot of one unpacked unit with one q8_1 block
template <bool HAS_MIN>
__device__ __forceinline__ float unit_dot(const unit_w & w, const blk_q8_1 * x) {
    const int * u = (const int *) x->qs;   // 4-byte aligned (36 B blocks, 4 B header)
    int s0 = 0, s1 = 0, su = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int a = u[i], c = u[4 + i];
        s0 = __dp4a(w.v[i], a, s0);
        s1 = __dp4a(w.v[4 + i], c, s1);
        if (HAS_MIN) { su = __dp4a(0x01010101, a, su); su = __dp4a(0x01010101, c, su); }
    }
    const float d8 = __low2float(x->ds);
    float r = w.wd[0] * (float) s0 + w.wd[1] * (float) s1;
    if (HAS_MIN) r -= w.wm * (float) su;
    return d8 * r;
}

}  // namespace b200
