// quantize.cuh — q8_1 / q8_0 block quantization device functions shared by the standalone
// quantizer (quantize.cu), the fused GEMV prologues and SET_ROWS.
//
// Arithmetic follows the reference CPU routines quantize_row_q8_1_ref (ggml-quants.c:225-258)
// and quantize_row_q8_0_ref (:197-223), NOT the reference CUDA kernel (quantize.cu:4-48),
// which differs in two places:
//   d  = amax / 127            (IEEE division)
//   id = d ? 1/d : 0           (IEEE division)
//   q  = roundf(x * id)        (round half away from zero)
//   s  = fp16(d * sum(q))      (f32 d, integer sum)
// Every operation is forced to round-to-nearest single ops (__fdiv_rn/__fmul_rn) so the
// compiler cannot contract or approximate them; the result is bit-identical to the C routine.
#pragma once
#include "common.cuh"

namespace b200 {

// One warp quantizes one 32-element block; lane i holds element i. Returns q; lane 0 should
// store the scales. amax/sum reductions are warp-wide.
__device__ __forceinline__ int8_t quantize_block32_lane(float x, float & d_out, int & sum_out) {
    const float amax = warp_max(fabsf(x));
    const float d    = __fdiv_rn(amax, 127.0f);
    const float id   = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    const int   q    = (int) roundf(__fmul_rn(x, id));
    d_out   = d;
    sum_out = warp_sum_i(q);
    return (int8_t) q;
}

// Sub-warp variant: `G` consecutive lanes (G = 8: four elements per lane) own one block.
// x[4] are the lane's 4 consecutive elements; returns them packed as int8x4.
template <int G>
__device__ __forceinline__ int quantize_block32_quad(const float x[4], float & d_out, int & sum_out) {
    float amax = fmaxf(fmaxf(fabsf(x[0]), fabsf(x[1])), fmaxf(fabsf(x[2]), fabsf(x[3])));
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    int q[4], s = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) { q[i] = (int) roundf(__fmul_rn(x[i], id)); s += q[i]; }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    d_out = d; sum_out = s;
    return (q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24);
}

// Four lanes own one block, eight consecutive elements per lane; returns them packed as two int8x4 words.
__device__ __forceinline__ int2 quantize_block32_oct(const float x[8], float & d_out, int & sum_out) {
    float amax = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(x[i]));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const float d  = __fdiv_rn(amax, 127.0f);
    const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
    int q[8], s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { q[i] = (int) roundf(__fmul_rn(x[i], id)); s += q[i]; }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    d_out = d; sum_out = s;
    return make_int2((q[0] & 0xff) | ((q[1] & 0xff) << 8) | ((q[2] & 0xff) << 16) | ((q[3] & 0xff) << 24),
                     (q[4] & 0xff) | ((q[5] & 0xff) << 8) | ((q[6] & 0xff) << 16) | ((q[7] & 0xff) << 24));
}

__device__ __forceinline__ __half2 q8_1_scales(float d, int sum) {
    return __halves2half2(__float2half_rn(d), __float2half_rn(__fmul_rn((float) sum, d)));
}

}  // namespace b200
