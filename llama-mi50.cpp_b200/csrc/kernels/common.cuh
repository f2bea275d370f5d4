// common.cuh — shared device/host helpers of the B200 kernel library.
// Wire formats restate reference ggml/src/ggml-common.h (block_q4_0 :170-175, block_mxfp4
// :190-195, block_q8_0 :219-224, block_q8_1 :226-237, block_q4_K :295-306, block_q6_K :330-336).
// The structs are layout declarations of the GGUF on-disk format, written here from the
// format description; kernels consume the bytes exactly as set_tensor delivers them.
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200_kernels.h"

#define B200_WARP 32

namespace b200 {

// ---------------------------------------------------------------- block formats (packed, little endian)
struct __align__(2) blk_q4_0 { __half d; uint8_t qs[16]; };               // 18 B / 32 weights
struct __align__(2) blk_q8_0 { __half d; int8_t qs[32]; };                // 34 B / 32
struct __align__(4) blk_q8_1 { __half2 ds; int8_t qs[32]; };              // 36 B / 32 (d, d*sum(q))
struct blk_mxfp4             { uint8_t e; uint8_t qs[16]; };              // 17 B / 32
struct __align__(4) blk_q4_K { __half2 dm; uint8_t scales[12]; uint8_t qs[128]; };          // 144 B / 256
struct __align__(2) blk_q6_K { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; __half d; }; // 210 B / 256
static_assert(sizeof(blk_q4_0) == 18 && sizeof(blk_q8_0) == 34 && sizeof(blk_q8_1) == 36, "layout");
static_assert(sizeof(blk_mxfp4) == 17 && sizeof(blk_q4_K) == 144 && sizeof(blk_q6_K) == 210, "layout");

// ---------------------------------------------------------------- error handling / launch accounting
void set_error(const char * fmt, ...);
void count_launch(int n = 1);
int  check_launch(const char * what);  // cudaGetLastError -> B200_OK / B200_ERR_CUDA

#define B200_CUDA_TRY(expr)                                                             \
    do {                                                                                \
        cudaError_t e_ = (expr);                                                        \
        if (e_ != cudaSuccess) {                                                        \
            b200::set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); \
            return B200_ERR_CUDA;                                                       \
        }                                                                               \
    } while (0)

// ---------------------------------------------------------------- host helpers
static inline bool is_quantized(int t) {
    return t == B200_Q4_0 || t == B200_Q8_0 || t == B200_Q4_K || t == B200_Q6_K || t == B200_MXFP4 || t == B200_Q8_1;
}
static inline bool is_contiguous_rows(const b200_tensor & t) {   // dims 1..3 dense over rows of nb[1]
    return t.nb[2] == t.nb[1] * t.ne[1] && t.nb[3] == t.nb[2] * t.ne[2];
}
static inline bool is_contiguous(const b200_tensor & t) {
    return t.nb[0] == (int64_t) b200_type_size(t.type) && t.nb[1] == (int64_t) b200_row_size(t.type, t.ne[0]) && is_contiguous_rows(t);
}
static inline int64_t nelements(const b200_tensor & t) { return t.ne[0] * t.ne[1] * t.ne[2] * t.ne[3]; }
static inline int64_t nrows(const b200_tensor & t) { return t.ne[1] * t.ne[2] * t.ne[3]; }

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// E8M0 exponent -> 2^(e-128) ("e8m0 half": ggml's MXFP4 LUT holds 2x the e2m1 values;
// reference ggml-impl.h:467-488 ggml_e8m0_to_fp32_half)
__device__ __forceinline__ float e8m0_half_to_f32(uint8_t e) {
    const uint32_t bits = e < 2 ? (0x00200000u << e) : ((uint32_t) (e - 1) << 23);
    return __uint_as_float(bits);
}
// unaligned-safe little-endian loads
__device__ __forceinline__ uint32_t ld_u32_a2(const void * p) {   // 2-byte aligned
    const uint16_t * q = (const uint16_t *) p;
    return (uint32_t) q[0] | ((uint32_t) q[1] << 16);
}
__device__ __forceinline__ uint32_t ld_u32_a1(const void * p) {   // byte aligned
    const uint8_t * q = (const uint8_t *) p;
    return (uint32_t) q[0] | ((uint32_t) q[1] << 8) | ((uint32_t) q[2] << 16) | ((uint32_t) q[3] << 24);
}
// N consecutive little-endian u32 words from an arbitrarily aligned global address: N + 1 aligned loads and N funnel shifts
// (byte loads cost four requests per word).  Reads up to 3 bytes before and after the range: inside the row or the buffer's
// tail slack (tensors are 256-byte aligned with 256 bytes of slack).
template <int N>
__device__ __forceinline__ void ld_words_any(const uint8_t * p, uint32_t * out) {
    const uintptr_t a = (uintptr_t) p;
    const uint32_t sh = (uint32_t) (a & 3) * 8;
    const uint32_t * q = (const uint32_t *) (a & ~(uintptr_t) 3);
    uint32_t w[N + 1];
#pragma unroll
    for (int i = 0; i <= N; ++i) w[i] = q[i];
#pragma unroll
    for (int i = 0; i < N; ++i) out[i] = __funnelshift_r(w[i], w[i + 1], sh);
}
// Programmatic dependent launch (PDL) controls, sm_90+.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait()              { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

}  // namespace b200
