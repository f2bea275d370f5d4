// mmq_tc.cu — prefill GEMM  dst[M, N] = W[M, K](quantized) . x[K, N]  on the 5th-generation tensor cores.
//
// Replaces, for n > 8 columns, the reference's MMQ path (ggml-cuda/mmq.cu:71-218, mmq.cuh:3459-3698: int8 mma.sync tiles with
// per-block scaling in registers, stream-K).  B200 design:
//   * weights stay in their GGUF wire format in HBM and travel ONCE per output tile as raw bytes: one 2-D TMA box
//     (128 rows x one 256-weight super-step) per stage, no repacking, no global dequantised copy;
//   * 8 dequant warps expand the raw blocks to f16 straight into the 128B-swizzled K-major operand layout the tensor
//     core reads (half2 bit tricks: 0x6400|q is 1024+q exactly, so nibbles become halves with one PRMT and no I2F);
//   * activations are converted once per matmul group to 16-bit tiles that are already swizzled in global memory, so a
//     stage of B is one contiguous 32 KiB bulk copy;
//   * one elected thread issues tcgen05.mma (M=128, N=256, K=16, kind::f16, f32 accumulate) into TMEM; completion is
//     committed to mbarriers which release the smem stages; the dequant warps turn into the epilogue (tcgen05.ld ->
//     coalesced f32 stores, dst is column-major [M, N] like ggml's).
// Numerics: weights are expanded exactly as dequantize_row_* would (f16 rounding of d*sc products, rel. 2^-11), the
// activations are rounded to f16 (clamped to +-65504) instead of q8_1 (rel. 2^-11, far below q8_1's quantisation noise);
// accumulation is f32.  Tolerance of the reference harness for this op: NMSE <= 5e-4 (tests/test-backend-ops.cpp:3718-3720).
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "vecdot.cuh"

namespace b200 {

constexpr int TC_BM = 128, TC_BN = 256, TC_BK = 64;       // CTA tile; one k-block = 64 x 16-bit = one 128 B swizzle row
constexpr int TC_STAGES = 3, TC_RAW_STAGES = 2;
constexpr int TC_DQ_WARPS = 8, TC_DQ_THREADS = TC_DQ_WARPS * 32;
constexpr int TC_THREADS = (2 + TC_DQ_WARPS) * 32;         // warp 0 producer, warp 1 MMA issuer + TMEM owner, warps 2.. dequant/epilogue
constexpr int TC_A_BYTES = TC_BM * 128, TC_B_BYTES = TC_BN * 128;
constexpr int TC_KSTEP = 256;                              // weights per raw stage (one k-quant super-block column)
constexpr int TC_SUB = TC_KSTEP / TC_BK;

template <int TYPE> struct tc_fmt;
// gbytes: bytes of one 256-weight super-step of a row; box: bytes of the TMA box row; elem: tensor-map element size
template <> struct tc_fmt<B200_Q4_K> { static constexpr int gbytes = 144, box = 144, elem = 1; };
template <> struct tc_fmt<B200_Q6_K> { static constexpr int gbytes = 210, box = 240, elem = 1; };    // 16 B aligned window around the 210 B block
template <> struct tc_fmt<B200_Q4_0> { static constexpr int gbytes = 144, box = 144, elem = 1; };    // 8 blocks of 18 B
template <> struct tc_fmt<B200_Q8_0> { static constexpr int gbytes = 272, box = 272, elem = 2; };
template <> struct tc_fmt<B200_MXFP4> { static constexpr int gbytes = 136, box = 144, elem = 1; };  // 8 blocks of 17 B; 16 B aligned window    // 8 blocks of 34 B (tensor map in 16-bit elements:
                                                                                            // a box dimension holds at most 256 elements)

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t tc_smem(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc_mbar_init(uint64_t * b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc_smem(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void tc_mbar_arrive(uint64_t * b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc_smem(b)) : "memory"); }
__device__ __forceinline__ void tc_mbar_expect(uint64_t * b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc_smem(b)), "r"(bytes) : "memory");
}
// bounded wait: a protocol error traps instead of hanging the GPU
__device__ __forceinline__ void tc_mbar_wait(uint64_t * b, uint32_t parity) {
    const uint32_t a = tc_smem(b);
    uint32_t ok, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 24)) __trap();
    } while (!ok);
}
__device__ __forceinline__ void tc_bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tc_smem(dst)), "l"(src), "r"(bytes), "r"(tc_smem(bar)) : "memory");
}
__device__ __forceinline__ void tc_tma_2d(void * dst, const CUtensorMap * map, int c0, int c1, uint64_t * bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(tc_smem(dst)), "l"(map), "r"(tc_smem(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t * b) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc_smem(b)) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile in shared memory, 128 B rows, 128 B swizzle, 8-row groups 1024 B apart (UMMA shared-memory descriptor:
// start >> 4 in bits 0-13, LBO in 16-29 (unused for swizzled K-major), SBO >> 4 in 32-45, version 1 in 46-47, layout 2 = SWIZZLE_128B in 61-63)
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {
    return (uint64_t) ((saddr & 0x3FFFFu) >> 4) | ((uint64_t) 1 << 16) | ((uint64_t) (1024 >> 4) << 32) | ((uint64_t) 1 << 46) | ((uint64_t) 2 << 61);
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t * v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// 4 bytes (values < 64 each) -> 4 halves holding the same integers: 0x6400 | b is exactly 1024 + b
__device__ __forceinline__ void tc_bytes_to_half(uint32_t x, float bias, __half2 & lo, __half2 & hi) {
    const uint32_t a = __byte_perm(x, 0x64646464u, 0x4140), b = __byte_perm(x, 0x64646464u, 0x4342);
    const __half2 off = __float2half2_rn(bias);             // 1024 (+ the format's zero point), exact in f16
    lo = __hsub2(*(const __half2 *) &a, off);
    hi = __hsub2(*(const __half2 *) &b, off);
}
__device__ __forceinline__ uint32_t tc_h2u(__half2 h) { return *(const uint32_t *) &h; }
// 16 B chunk c (8 halves) of row r of a swizzled operand tile
__device__ __forceinline__ void tc_store_chunk(uint8_t * tile, int r, int c, uint4 v) {
    *(uint4 *) (tile + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4)) = v;
}

// 8 consecutive little-endian words from a 2-byte aligned shared-memory address
__device__ __forceinline__ void tc_lds_words8(const uint8_t * p, uint32_t * out) {
    const uint32_t a = tc_smem(p), sh = (a & 3u) * 8u;
    const uint32_t * q = (const uint32_t *) (p - (a & 3u));
    uint32_t w[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) w[i] = q[i];
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}

// ---------------------------------------------------------------- per-format expansion of one 64-weight k-block
// thread (r, h): row r of the tile, half h of the k-block (32 weights -> chunks 4h .. 4h+3); `row` = the row's raw bytes of the
// current 256-weight super-step; sub = which of its 4 k-blocks.
template <int TYPE> struct tc_expand;

template <> struct tc_expand<B200_Q4_K> {
    __device__ __forceinline__ static void run(const uint8_t * row, int sub, int r, int h, uint8_t * tile) {
        const uint8_t * sc8 = row + 4;                           // half2 dm, then 12 bytes of 6-bit scales/mins
        const int j = 2 * sub + h;                               // sub-block (get_scale_min_k4, ggml-quants.c)
        int sc, mn;
        if (j < 4) { sc = sc8[j] & 63; mn = sc8[j + 4] & 63; }
        else       { sc = (sc8[j + 4] & 0xF) | ((sc8[j - 4] >> 6) << 4); mn = (sc8[j + 4] >> 4) | ((sc8[j] >> 6) << 4); }
        const float2 dm = __half22float2(*(const __half2 *) row);
        const __half2 s2 = __float2half2_rn(dm.x * (float) sc), b2 = __float2half2_rn(-dm.y * (float) mn);
        const uint4 q0 = *(const uint4 *) (row + 16 + 32 * sub), q1 = *(const uint4 *) (row + 32 + 32 * sub);
        const uint32_t q[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const uint32_t x = h ? (q[2 * c + w] >> 4) & 0x0F0F0F0Fu : q[2 * c + w] & 0x0F0F0F0Fu;
                __half2 lo, hi;
                tc_bytes_to_half(x, 1024.0f, lo, hi);
                o[2 * w] = tc_h2u(__hfma2(lo, s2, b2)); o[2 * w + 1] = tc_h2u(__hfma2(hi, s2, b2));
            }
            tc_store_chunk(tile, r, 4 * h + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
};

template <> struct tc_expand<B200_Q6_K> {
    __device__ __forceinline__ static void run(const uint8_t * row, int sub, int r, int h, uint8_t * tile) {
        // dequantize_row_q6_K (ggml-quants.c): per 128-weight half n, y[l + 32*g] for g = 0..3 uses ql[l + 32*(g&1)] (low nibble for
        // g < 2, high for g >= 2), bits 2g..2g+1 of qh[l] and scale sc[l/16 + 2g]; this k-block holds g = 2*(sub&1) + h
        const int n = sub >> 1, g = 2 * (sub & 1) + h;
        uint32_t ql[8], qh[8];                                   // the block sits at an even byte offset of the aligned window
        tc_lds_words8(row + 64 * n + 32 * h, ql);
        tc_lds_words8(row + 128 + 32 * n, qh);
        const int8_t * scp = (const int8_t *) (row + 192 + 8 * n + 2 * g);
        const float d = __half2float(*(const __half *) (row + 208));
        const __half2 s2[2] = {__float2half2_rn(d * (float) scp[0]), __float2half2_rn(d * (float) scp[1])};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const int i = 2 * c + w;
                const uint32_t lo4 = (sub & 1) ? (ql[i] >> 4) & 0x0F0F0F0Fu : ql[i] & 0x0F0F0F0Fu;
                const uint32_t x = lo4 | (((qh[i] >> (2 * g)) & 0x03030303u) << 4);
                __half2 lo, hi;
                tc_bytes_to_half(x, 1056.0f, lo, hi);           // 1024 + the format's zero point 32
                o[2 * w] = tc_h2u(__hmul2(lo, s2[c >> 1])); o[2 * w + 1] = tc_h2u(__hmul2(hi, s2[c >> 1]));
            }
            tc_store_chunk(tile, r, 4 * h + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
};

template <> struct tc_expand<B200_Q4_0> {
    __device__ __forceinline__ static void run(const uint8_t * row, int sub, int r, int h, uint8_t * tile) {
        const uint8_t * b = row + 18 * (2 * sub + h);            // block: half d, 16 bytes: weight j = low nibble of qs[j], j+16 = high
        const uint16_t * p = (const uint16_t *) b;               // 2-byte aligned
        const __half2 s2 = __half2half2(*(const __half *) b);
        uint32_t q[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = (uint32_t) p[1 + 2 * i] | ((uint32_t) p[2 + 2 * i] << 16);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const uint32_t src = q[2 * (c & 1) + w];
                const uint32_t x = (c >> 1) ? (src >> 4) & 0x0F0F0F0Fu : src & 0x0F0F0F0Fu;
                __half2 lo, hi;
                tc_bytes_to_half(x, 1032.0f, lo, hi);           // zero point 8
                o[2 * w] = tc_h2u(__hmul2(lo, s2)); o[2 * w + 1] = tc_h2u(__hmul2(hi, s2));
            }
            tc_store_chunk(tile, r, 4 * h + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
};

template <> struct tc_expand<B200_Q8_0> {
    __device__ __forceinline__ static void run(const uint8_t * row, int sub, int r, int h, uint8_t * tile) {
        const uint8_t * b = row + 34 * (2 * sub + h);            // block: half d, 32 int8 (2-byte aligned)
        const __half2 s2 = __half2half2(*(const __half *) b);
        uint32_t q[8];
        tc_lds_words8(b + 2, q);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                __half2 lo, hi;
                tc_bytes_to_half(q[2 * c + w] ^ 0x80808080u, 1152.0f, lo, hi);      // int8 + 128 = 0..255; 1024 + 128
                o[2 * w] = tc_h2u(__hmul2(lo, s2)); o[2 * w + 1] = tc_h2u(__hmul2(hi, s2));
            }
            tc_store_chunk(tile, r, 4 * h + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
};

// MXFP4: e2m1 x 2^(e-127) does not fit f16's exponent range, so this type runs the MMA in bf16 — both
// operands bf16 (kind::f16 takes f16 x f16 or bf16 x bf16), every weight exact (a 4-bit integer times a power of two), the
// activations rounded to bf16.  The reference's B200 path (LUT -> int8, q8_1 activations, mmq.cuh:720-783) and its sm_120
// block-scaled FP4 MMA (:785-827) are replaced; activations are NOT quantised to FP4, so the 5e-4 NMSE gate applies.
template <> struct tc_expand<B200_MXFP4> {
    __device__ __forceinline__ static void run(const uint8_t * row, int sub, int r, int h, uint8_t * tile) {
        const uint8_t * b = row + 17 * (2 * sub + h);           // block: e8m0 e, 16 bytes: weight j = low nibble of qs[j], j+16 = high
        const __nv_bfloat162 s2 = __float2bfloat162_rn(e8m0_half_to_f32(b[0])), off = __float2bfloat162_rn(144.0f);
        const uint32_t a = tc_smem(b + 1), sh = (a & 3u) * 8u;
        const uint32_t * qp = (const uint32_t *) (b + 1 - (a & 3u));
        uint32_t w[5];
#pragma unroll
        for (int i = 0; i < 5; ++i) w[i] = qp[i];
        int lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) mxfp4_expand(__funnelshift_r(w[i], w[i + 1], sh), lo[i], hi[i]);      // int8: 2 x the e2m1 value
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int q = (c >> 1) ? hi[2 * (c & 1) + v] : lo[2 * (c & 1) + v];
                const uint32_t x = __vadd4((uint32_t) q, 0x10101010u);                  // 4..28 per byte; 0x4300 | x is exactly 128 + x in bf16
                const uint32_t p0 = __byte_perm(x, 0x43434343u, 0x4140), p1 = __byte_perm(x, 0x43434343u, 0x4342);
                const __nv_bfloat162 v0 = __hmul2(__hsub2(*(const __nv_bfloat162 *) &p0, off), s2), v1 = __hmul2(__hsub2(*(const __nv_bfloat162 *) &p1, off), s2);
                o[2 * v] = *(const uint32_t *) &v0; o[2 * v + 1] = *(const uint32_t *) &v1;
            }
            tc_store_chunk(tile, r, 4 * h + c, make_uint4(o[0], o[1], o[2], o[3]));
        }
    }
};

// ---------------------------------------------------------------- activations -> swizzled 16-bit tiles
// out: tiles [n_tile][k_block] of 256 rows x 128 B, each already in the shared-memory image (16 B chunk c of row r at
// (r/8)*1024 + (r%8)*128 + ((c ^ r%8) * 16)); columns beyond N are zero.  fmt 0 = f16 (clamped to +-65504), 1 = bf16.
__global__ void mmq_tc_act_kernel(const float * __restrict__ x, int64_t ldx, int N, int K, uint8_t * __restrict__ out, int fmt) {
    const int kb_n = K / TC_BK;
    const int64_t chunk = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    const int c = (int) (chunk & 7);
    const int64_t t = chunk >> 3;
    const int kb = (int) (t % kb_n);
    const int64_t n = t / kb_n;                                  // padded column index
    const int64_t n_pad = ((int64_t) N + TC_BN - 1) / TC_BN * TC_BN;
    if (n >= n_pad) return;
    float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (n < N) {
        const float4 a = *(const float4 *) (x + n * ldx + kb * TC_BK + c * 8), b = *(const float4 *) (x + n * ldx + kb * TC_BK + c * 8 + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (fmt == 0) {
            const __half2 hv = __floats2half2_rn(fminf(fmaxf(v[2 * i], -65504.0f), 65504.0f), fminf(fmaxf(v[2 * i + 1], -65504.0f), 65504.0f));
            o[i] = *(const uint32_t *) &hv;
        } else {
            const __nv_bfloat162 bv = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            o[i] = *(const uint32_t *) &bv;
        }
    }
    const int r = (int) (n % TC_BN);
    const int64_t tile = (n / TC_BN) * kb_n + kb;
    *(uint4 *) (out + tile * TC_B_BYTES + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
}

// ---------------------------------------------------------------- the GEMM kernel
struct tc_shared {
    uint64_t raw_full[TC_RAW_STAGES], raw_empty[TC_RAW_STAGES];
    uint64_t ab_full[TC_STAGES], b_full[TC_STAGES], ab_empty[TC_STAGES];
    uint64_t acc_full;
    uint32_t tmem_base;
};

// BN = columns of the CTA tile: 256, or 128 (half of a 256-column activation tile) when 256-wide tiles would leave most
// SMs without a CTA (M = 4096 at N = 512: 64 CTAs on 148 SMs)
template <int TYPE, int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
mmq_tc_kernel(const __grid_constant__ CUtensorMap tmA, const uint8_t * __restrict__ xb, float * dst,
              int M, int N, int K, int64_t ldd, uint32_t idesc, const float * add, int64_t lda) {
    using F = tc_fmt<TYPE>;
    constexpr int RAW_BYTES = TC_BM * F::box;
    extern __shared__ uint8_t smem_raw[];
    __shared__ tc_shared sh;
    uint8_t * base = smem_raw + ((1024u - (tc_smem(smem_raw) & 1023u)) & 1023u);       // swizzled tiles need 1024 B alignment
    constexpr int B_BYTES = BN * 128;
    uint8_t * sA = base, * sB = sA + TC_STAGES * TC_A_BYTES, * sR = sB + TC_STAGES * B_BYTES;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_tile = (int) blockIdx.x / (TC_BN / BN), n_half = (int) blockIdx.x % (TC_BN / BN), m0 = blockIdx.y * TC_BM, n0 = (int) blockIdx.x * BN;
    const int ksteps = K / TC_KSTEP, kb_n = K / TC_BK;

    if (tid == 0) {
        for (int s = 0; s < TC_RAW_STAGES; ++s) { tc_mbar_init(&sh.raw_full[s], 1); tc_mbar_init(&sh.raw_empty[s], TC_DQ_WARPS); }
        for (int s = 0; s < TC_STAGES; ++s) { tc_mbar_init(&sh.ab_full[s], TC_DQ_WARPS); tc_mbar_init(&sh.b_full[s], 1); tc_mbar_init(&sh.ab_empty[s], 1); }
        tc_mbar_init(&sh.acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tc_fence_async();
    }
    if (warp == 1) {                                            // TMEM: 256 columns of f32 accumulators (128 lanes x 256)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem(&sh.tmem_base)), "r"(BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sh.tmem_base;

    if (warp == 0) {
        // ================= producer: raw weight boxes (TMA 2-D) and activation tiles (bulk)
        if (lane == 0) {
            const uint8_t * xt = xb + (size_t) n_tile * kb_n * TC_B_BYTES + (size_t) n_half * B_BYTES;    // rows 128.. of a tile are its second half
            for (int ks = 0; ks < ksteps; ++ks) {
                const int rs = ks % TC_RAW_STAGES;
                tc_mbar_wait(&sh.raw_empty[rs], ((ks / TC_RAW_STAGES) & 1) ^ 1);
                tc_mbar_expect(&sh.raw_full[rs], RAW_BYTES);
                tc_tma_2d(sR + rs * RAW_BYTES, &tmA, ((ks * F::gbytes) & ~15) / F::elem, m0, &sh.raw_full[rs]);   // box starts are kept 16 B aligned (an unaligned start faulted)
                for (int sub = 0; sub < TC_SUB; ++sub) {
                    const int it = ks * TC_SUB + sub, s = it % TC_STAGES;
                    tc_mbar_wait(&sh.ab_empty[s], ((it / TC_STAGES) & 1) ^ 1);
                    tc_mbar_expect(&sh.b_full[s], B_BYTES);
                    tc_bulk_g2s(sB + s * B_BYTES, xt + (size_t) it * TC_B_BYTES, B_BYTES, &sh.b_full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer
        for (int it = 0; it < kb_n; ++it) {
            const int s = it % TC_STAGES;
            const uint32_t par = (it / TC_STAGES) & 1;
            tc_mbar_wait(&sh.ab_full[s], par);
            tc_mbar_wait(&sh.b_full[s], par);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t a0 = tc_smem(sA + s * TC_A_BYTES), b0 = tc_smem(sB + s * B_BYTES);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k)             // UMMA_K = 16 halves = 32 B inside the swizzle row
                    tc_mma_f16(tmem, tc_desc(a0 + k * 32), tc_desc(b0 + k * 32), idesc, (it | k) != 0);
                tc_commit(&sh.ab_empty[s]);                      // stage reusable once these MMAs have read it
                if (it == kb_n - 1) tc_commit(&sh.acc_full);
            }
            __syncwarp();
        }
    } else {
        // ================= dequant warps, then epilogue
        const int dt = tid - 64, r = dt & (TC_BM - 1), h = dt >> 7;
        for (int ks = 0; ks < ksteps; ++ks) {
            const int rs = ks % TC_RAW_STAGES;
            tc_mbar_wait(&sh.raw_full[rs], (ks / TC_RAW_STAGES) & 1);
            const uint8_t * row = sR + rs * RAW_BYTES + r * F::box + ((ks * F::gbytes) & 15);
            for (int sub = 0; sub < TC_SUB; ++sub) {
                const int it = ks * TC_SUB + sub, s = it % TC_STAGES;
                tc_mbar_wait(&sh.ab_empty[s], ((it / TC_STAGES) & 1) ^ 1);
                tc_expand<TYPE>::run(row, sub, r, h, sA + s * TC_A_BYTES);
                tc_fence_async();                                // generic-proxy stores -> visible to the tensor core's async reads
                __syncwarp();
                if (lane == 0) tc_mbar_arrive(&sh.ab_full[s]);
            }
            __syncwarp();
            if (lane == 0) tc_mbar_arrive(&sh.raw_empty[rs]);
        }
        // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31; the two warps of a lane quarter split the 256 columns
        tc_mbar_wait(&sh.acc_full, 0);
        tc_fence_after();
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int m = m0 + q * 32 + lane;
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
            uint32_t v[32];
            const int col = half * (BN / 2) + c * 32;
            tc_ld32(tmem + ((uint32_t) (q * 32) << 16) + (uint32_t) col, v);
            if (m < M) {
                // the addend may alias dst (in-place residual add): all its loads are issued before the first store, otherwise
                // every load would have to wait for the store in front of it (measured: 5 % of pp512)
                float r[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = n0 + col + j;
                    r[j] = add && n < N ? add[(int64_t) n * lda + m] : 0.0f;
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = n0 + col + j;
                    if (n < N) dst[(int64_t) n * ldd + m] = __uint_as_float(v[j]) + r[j];
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(BN) : "memory");
}

// ---------------------------------------------------------------- host side
typedef CUresult (*tc_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tc_encode_fn tc_encoder() {
    static tc_encode_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void * p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (tc_encode_fn) p;
        else cudaGetLastError();
    });
    return fn;
}

static int tc_box(int t) { return t == B200_Q6_K ? 240 : t == B200_Q8_0 ? 272 : 144; }
static int tc_act_fmt(int t) { return t == B200_MXFP4 ? 1 : 0; }      // 16-bit format of BOTH operands: 0 = f16, 1 = bf16
static int tc_elem(int t) { return t == B200_Q8_0 ? 2 : 1; }
static bool tc_type_ok(int t) { return t == B200_Q4_K || t == B200_Q6_K || t == B200_Q4_0 || t == B200_Q8_0 || t == B200_MXFP4; }
static size_t tc_smem_bytes(int t, int bn) { return (size_t) TC_STAGES * (TC_A_BYTES + bn * 128) + (size_t) TC_RAW_STAGES * TC_BM * tc_box(t) + 1024; }

static bool tc_args_ok(const b200_mmq_args * a) {
    if (!a) return false;
    const b200_tensor & w = a->w; const b200_tensor & x = a->x; const b200_tensor & d = a->dst;
    const int64_t K = w.ne[0], M = w.ne[1], N = x.ne[1];
    if (!tc_type_ok(w.type) || w.ne[2] != 1 || w.ne[3] != 1 || K < TC_KSTEP || K % TC_KSTEP != 0 || M < 1 || M > (1 << 30)) return false;
    if (w.nb[0] != (int64_t) b200_type_size(w.type) || w.nb[1] != (int64_t) b200_row_size(w.type, K) || w.nb[1] % 16 != 0 || ((uintptr_t) w.data) % 16 != 0) return false;
    if (x.type != B200_F32 || x.nb[0] != 4 || x.ne[0] != K || x.ne[2] != 1 || x.ne[3] != 1 || N < 1 || N > 65536 || x.nb[1] % 16 != 0 || ((uintptr_t) x.data) % 16 != 0) return false;
    if (d.type != B200_F32 || d.nb[0] != 4 || d.ne[0] != M || d.ne[1] != N || d.ne[2] != 1 || d.ne[3] != 1 || d.nb[1] % 4 != 0) return false;
    if (a->addend) {
        const b200_tensor & r = *a->addend;
        if (r.type != B200_F32 || r.nb[0] != 4 || r.ne[0] != M || r.ne[1] != N || r.ne[2] != 1 || r.ne[3] != 1 || r.nb[1] % 4 != 0) return false;
    }
    if (a->act_fmt != tc_act_fmt(w.type)) return false;   // kind::f16 rejects an f16 A with a bf16 B (illegal instruction, measured): same format on both sides
    return true;
}

template <int TYPE, int BN>
static int tc_launch_bn(const CUtensorMap & map, const b200_mmq_args * a, cudaStream_t st) {
    static bool attr_set[64] = {};
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    const size_t smem = tc_smem_bytes(TYPE, BN);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(mmq_tc_kernel<TYPE, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
        attr_set[dev] = true;
    }
    const int M = (int) a->w.ne[1], N = (int) a->x.ne[1], K = (int) a->w.ne[0];
    // instruction descriptor: D f32 (bits 4-5 = 1), A f16 (bits 7-9 = 0), B f16/bf16 (bits 10-12), both K-major, N >> 3 in 17-22, M >> 4 in 24-28
    const uint32_t idesc = (1u << 4) | ((uint32_t) a->act_fmt << 7) | ((uint32_t) a->act_fmt << 10) | ((uint32_t) (BN >> 3) << 17) | ((uint32_t) (TC_BM >> 4) << 24);
    dim3 grid((unsigned) ((N + BN - 1) / BN), (unsigned) ((M + TC_BM - 1) / TC_BM));
    mmq_tc_kernel<TYPE, BN><<<grid, TC_THREADS, smem, st>>>(map, (const uint8_t *) a->xb, (float *) a->dst.data, M, N, K, a->dst.nb[1] / 4, idesc,
                                                          a->addend ? (const float *) a->addend->data : nullptr, a->addend ? a->addend->nb[1] / 4 : 0);
    count_launch();
    return check_launch("mmq_tc_kernel");
}
template <int TYPE>
static int tc_launch(const CUtensorMap & map, const b200_mmq_args * a, cudaStream_t st) {
    // 256-column tiles unless they leave more than a third of the SMs idle and 128-column tiles fill more of them
    const int64_t M = a->w.ne[1], N = a->x.ne[1];
    const int64_t c256 = ((N + 255) / 256) * ((M + TC_BM - 1) / TC_BM), c128 = ((N + 127) / 128) * ((M + TC_BM - 1) / TC_BM);
    if (N > 128 && c256 < 100 && c128 > c256) return tc_launch_bn<TYPE, 128>(map, a, st);
    return tc_launch_bn<TYPE, 256>(map, a, st);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_mmq_act_fmt(int type) { return tc_act_fmt(type); }

extern "C" size_t b200_mmq_act_size(int64_t K, int64_t N) {
    if (K < TC_BK || N < 1) return 0;
    return (size_t) ((N + TC_BN - 1) / TC_BN) * (size_t) (K / TC_BK) * TC_B_BYTES;
}

extern "C" int b200_mmq_supported(const b200_mmq_args * a) { return tc_args_ok(a) && tc_encoder() != nullptr ? 1 : 0; }

extern "C" int b200_mmq_convert_act(const b200_tensor * x, void * xb, int act_fmt, void * stream) {
    if (!x || !xb || x->type != B200_F32 || x->nb[0] != 4 || x->ne[0] % TC_BK != 0 || x->ne[2] != 1 || x->ne[3] != 1 || x->nb[1] % 16 != 0 ||
        ((uintptr_t) x->data) % 16 != 0 || ((uintptr_t) xb) % 1024 != 0 || (act_fmt != 0 && act_fmt != 1)) {
        set_error("b200_mmq_convert_act: unsupported arguments"); return B200_ERR_UNSUPPORTED;
    }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    const int64_t K = x->ne[0], N = x->ne[1], n_pad = (N + TC_BN - 1) / TC_BN * TC_BN;
    const int64_t chunks = n_pad * (K / 8);
    mmq_tc_act_kernel<<<(unsigned) ((chunks + 255) / 256), 256, 0, (cudaStream_t) stream>>>((const float *) x->data, x->nb[1] / 4, (int) N, (int) K, (uint8_t *) xb, act_fmt);
    count_launch();
    return check_launch("mmq_tc_act_kernel");
}

extern "C" int b200_mul_mat_q(const b200_mmq_args * a, void * stream) {
    if (!tc_args_ok(a) || !a->xb || ((uintptr_t) a->xb) % 1024 != 0) { set_error("b200_mul_mat_q: unsupported arguments"); return B200_ERR_UNSUPPORTED; }
    tc_encode_fn enc = tc_encoder();
    if (!enc) { set_error("b200_mul_mat_q: cuTensorMapEncodeTiled is not available"); return B200_ERR_UNSUPPORTED; }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    // tensor map over the raw weight bytes: [M rows][row bytes], box = 128 rows x one super-step (cached per weight tensor)
    struct key { const void * p; int64_t rb, m; int t; bool operator==(const key & o) const { return p == o.p && rb == o.rb && m == o.m && t == o.t; } };
    struct hash { size_t operator()(const key & k) const { return std::hash<const void *>()(k.p) ^ (size_t) k.rb * 31 ^ (size_t) k.m * 131 ^ (size_t) k.t; } };
    static std::unordered_map<key, CUtensorMap, hash> cache;
    static std::mutex mu;
    CUtensorMap map;
    {
        std::lock_guard<std::mutex> lock(mu);
        const key k{a->w.data, a->w.nb[1], a->w.ne[1], a->w.type};
        auto it = cache.find(k);
        if (it == cache.end()) {
            const int el = tc_elem(a->w.type);
            const cuuint64_t gdim[2] = {(cuuint64_t) a->w.nb[1] / el, (cuuint64_t) a->w.ne[1]};
            const cuuint64_t gstr[1] = {(cuuint64_t) a->w.nb[1]};
            const cuuint32_t box[2] = {(cuuint32_t) (tc_box(a->w.type) / el), (cuuint32_t) TC_BM};
            const cuuint32_t estr[2] = {1, 1};
            const CUresult rc = enc(&map, el == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, a->w.data, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (rc != CUDA_SUCCESS) { set_error("b200_mul_mat_q: cuTensorMapEncodeTiled failed (%d)", (int) rc); return B200_ERR_CUDA; }
            if (cache.size() > 4096) cache.clear();
            cache.emplace(k, map);
        } else map = it->second;
    }
    switch (a->w.type) {
        case B200_Q4_K: return tc_launch<B200_Q4_K>(map, a, (cudaStream_t) stream);
        case B200_Q6_K: return tc_launch<B200_Q6_K>(map, a, (cudaStream_t) stream);
        case B200_Q4_0: return tc_launch<B200_Q4_0>(map, a, (cudaStream_t) stream);
        case B200_Q8_0: return tc_launch<B200_Q8_0>(map, a, (cudaStream_t) stream);
        case B200_MXFP4: return tc_launch<B200_MXFP4>(map, a, (cudaStream_t) stream);
    }
    return B200_ERR_UNSUPPORTED;
}
