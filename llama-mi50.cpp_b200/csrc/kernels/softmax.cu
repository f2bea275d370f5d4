// softmax.cu — row G glue: SOFT_MAX with scale, mask (f16/f32), ALiBi slope and attention sinks.
// Replaces reference ggml-cuda/softmax.cu.  One block per row, row cached in shared memory.
#include "common.cuh"

namespace b200 {

struct sm_params {
    int64_t ne0, ne1, ne2, ne3;
    int64_t x_nb1, x_nb2, x_nb3, d_nb1, d_nb2, d_nb3;
    int64_t m_ne2, m_ne3, m_nb1, m_nb2, m_nb3;
    int32_t mask_f16;
    float scale, max_bias, m0, m1; uint32_t n_head_log2;
};

__device__ __forceinline__ float block_reduce(float v, float * red, bool is_max) {
    v = is_max ? warp_max(v) : warp_sum(v);
    const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    const int lane = threadIdx.x & 31;   // every warp reduces the per-warp partials
    v = lane < nw ? red[lane] : (is_max ? -INFINITY : 0.0f);
    return is_max ? warp_max(v) : warp_sum(v);
}

__global__ void __launch_bounds__(256) soft_max_kernel(const char * x, const char * mask, const float * sinks, char * dst, sm_params p) {
    extern __shared__ float vals[];
    __shared__ float red[32];
    pdl_launch_dependents();
    const int64_t i1 = blockIdx.x, i2 = blockIdx.y, i3 = blockIdx.z;
    const float * xr = (const float *) (x + i1 * p.x_nb1 + i2 * p.x_nb2 + i3 * p.x_nb3);
    float * dr = (float *) (dst + i1 * p.d_nb1 + i2 * p.d_nb2 + i3 * p.d_nb3);
    const char * mr = mask ? mask + i1 * p.m_nb1 + (i2 % p.m_ne2) * p.m_nb2 + (i3 % p.m_ne3) * p.m_nb3 : nullptr;
    float slope = 1.0f;
    if (p.max_bias > 0.0f) {
        const uint32_t h = (uint32_t) i2;
        slope = h < p.n_head_log2 ? powf(p.m0, (float) (h + 1)) : powf(p.m1, (float) (2 * (h - p.n_head_log2) + 1));
    }
    pdl_wait();
    float mx = sinks ? sinks[i2] : -INFINITY;
    for (int64_t i = threadIdx.x; i < p.ne0; i += blockDim.x) {
        float v = xr[i] * p.scale;
        if (mr) v += slope * (p.mask_f16 ? __half2float(((const __half *) mr)[i]) : ((const float *) mr)[i]);
        vals[i] = v;
        mx = fmaxf(mx, v);
    }
    mx = block_reduce(mx, red, true);
    float sum = 0.0f;
    for (int64_t i = threadIdx.x; i < p.ne0; i += blockDim.x) { const float e = expf(vals[i] - mx); vals[i] = e; sum += e; }
    sum = block_reduce(sum, red, false);
    if (sinks) sum += expf(sinks[i2] - mx);
    const float inv = 1.0f / sum;
    for (int64_t i = threadIdx.x; i < p.ne0; i += blockDim.x) dr[i] = vals[i] * inv;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_soft_max(const b200_tensor * x, const b200_tensor * mask, const b200_tensor * sinks,
                             const b200_tensor * dst, float scale, float max_bias, void * stream) {
    if (!x || !dst) return B200_ERR_ARG;
    if (x->type != B200_F32 || dst->type != B200_F32 || x->nb[0] != 4 || dst->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    for (int i = 0; i < 4; ++i) if (x->ne[i] != dst->ne[i]) return B200_ERR_UNSUPPORTED;
    if (x->ne[0] > 48 * 1024 / 4 || x->ne[2] > 65535 || x->ne[3] > 65535) return B200_ERR_UNSUPPORTED;
    if (mask) {
        if ((mask->type != B200_F16 && mask->type != B200_F32) || mask->nb[0] != (int64_t) b200_type_size(mask->type)) return B200_ERR_UNSUPPORTED;
        if (mask->ne[0] != x->ne[0] || mask->ne[1] < x->ne[1] || mask->ne[2] < 1 || mask->ne[3] < 1 || x->ne[2] % mask->ne[2] != 0 || x->ne[3] % mask->ne[3] != 0) return B200_ERR_UNSUPPORTED;
    }
    if (sinks && (sinks->type != B200_F32 || sinks->ne[0] != x->ne[2] || sinks->nb[0] != 4)) return B200_ERR_UNSUPPORTED;
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    if (nelements(*x) == 0) return B200_OK;
    sm_params p{};
    p.ne0 = x->ne[0]; p.ne1 = x->ne[1]; p.ne2 = x->ne[2]; p.ne3 = x->ne[3];
    p.x_nb1 = x->nb[1]; p.x_nb2 = x->nb[2]; p.x_nb3 = x->nb[3];
    p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2]; p.d_nb3 = dst->nb[3];
    if (mask) { p.m_ne2 = mask->ne[2]; p.m_ne3 = mask->ne[3]; p.m_nb1 = mask->nb[1]; p.m_nb2 = mask->nb[2]; p.m_nb3 = mask->nb[3]; p.mask_f16 = mask->type == B200_F16; }
    p.scale = scale; p.max_bias = max_bias;
    p.n_head_log2 = 1u << (uint32_t) floorf(log2f((float) x->ne[2]));
    p.m0 = powf(2.0f, -max_bias / p.n_head_log2); p.m1 = powf(2.0f, -(max_bias / 2.0f) / p.n_head_log2);
    dim3 grid((unsigned) x->ne[1], (unsigned) x->ne[2], (unsigned) x->ne[3]);
    soft_max_kernel<<<grid, 256, (size_t) x->ne[0] * 4, (cudaStream_t) stream>>>(
        (const char *) x->data, mask ? (const char *) mask->data : nullptr, sinks ? (const float *) sinks->data : nullptr, (char *) dst->data, p);
    count_launch();
    return check_launch("soft_max");
}
