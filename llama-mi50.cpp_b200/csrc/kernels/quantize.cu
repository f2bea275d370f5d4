// quantize.cu — K1: f32 activations -> block_q8_1 (C-ABI b200_quantize_q8_1).
// Replaces quantize_row_q8_1_cuda (reference ggml-cuda/quantize.cu:273-287).
// Layout: each thread owns 4 consecutive elements, 8 lanes own one 32-element block,
// 128-bit loads when the row is 16 B aligned; output blocks are written as 4-byte words.
#include "quantize.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) quantize_q8_1_kernel(
        const char * __restrict__ x, blk_q8_1 * __restrict__ y,
        const int64_t ne0, const int64_t ne0_padded, const int64_t ne1, const int64_t ne2,
        const int64_t nb1, const int64_t nb2, const int64_t nb3) {
    pdl_launch_dependents();
    const int64_t i0 = 4 * ((int64_t) blockIdx.x * blockDim.x + threadIdx.x);
    const bool active = i0 < ne0_padded;   // no early return: the shuffles below use the full-warp mask
    const int64_t i1 = blockIdx.y;
    const int64_t i2 = blockIdx.z % ne2;
    const int64_t i3 = blockIdx.z / ne2;
    const float * row = (const float *) (x + i1 * nb1 + i2 * nb2 + i3 * nb3);

    pdl_wait();
    float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (!active) {
    } else if (i0 + 3 < ne0 && (((uintptr_t) row) & 15) == 0) {
        const float4 t = *(const float4 *) (row + i0);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) if (i0 + i < ne0) v[i] = row[i0 + i];
    }
    float d; int sum;
    const int q = quantize_block32_quad<8>(v, d, sum);

    if (!active) return;
    const int64_t nblk_row = ne0_padded / 32;
    const int64_t ib = ((i3 * ne2 + i2) * ne1 + i1) * nblk_row + i0 / 32;
    int * out = (int *) (y + ib);          // blk_q8_1 is 36 B: 4-byte aligned words
    const int l = (int) ((i0 / 4) & 7);
    out[1 + l] = q;
    if (l == 0) *(__half2 *) out = q8_1_scales(d, sum);
}

}  // namespace b200

extern "C" size_t b200_q8_1_buffer_size(const b200_tensor * x, int64_t ne0_padded) {
    return (size_t) (ne0_padded / 32) * 36 * (size_t) (x->ne[1] * x->ne[2] * x->ne[3]);
}

extern "C" int b200_quantize_q8_1(const b200_tensor * x, void * y, int64_t ne0_padded, void * stream) {
    using namespace b200;
    if (!x || !y || x->type != B200_F32 || x->nb[0] != 4 || ne0_padded % 32 != 0 || ne0_padded < x->ne[0]) {
        set_error("b200_quantize_q8_1: bad arguments");
        return B200_ERR_ARG;
    }
    if (nelements(*x) == 0) return B200_OK;
    if (x->ne[1] > 65535 || x->ne[2] * x->ne[3] > 65535) { set_error("b200_quantize_q8_1: grid too large"); return B200_ERR_UNSUPPORTED; }
    const int64_t nthr = ne0_padded / 4;
    dim3 grid((unsigned) ((nthr + 255) / 256), (unsigned) x->ne[1], (unsigned) (x->ne[2] * x->ne[3]));
    quantize_q8_1_kernel<<<grid, 256, 0, (cudaStream_t) stream>>>(
        (const char *) x->data, (blk_q8_1 *) y, x->ne[0], ne0_padded, x->ne[1], x->ne[2], x->nb[1], x->nb[2], x->nb[3]);
    count_launch();
    return check_launch("quantize_q8_1");
}
