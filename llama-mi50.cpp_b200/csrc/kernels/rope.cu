// rope.cu — row G glue: ROPE (NORMAL / NEOX) with optional fused SET_ROWS into the KV cache.
// Replaces reference ggml-cuda/rope.cu (:563) and the {ROPE, VIEW, SET_ROWS} fusion
// (ggml-cuda.cu:3614-3621).  YaRN scaling follows ggml_rope_yarn_corr_dims / rope_yarn
// (reference ggml/src/ggml.c and ggml-cpu/ops.cpp:5531-5546).
#include <math.h>

#include "common.cuh"

namespace b200 {

struct rope_dev_params {
    int64_t ne0, ne1, ne2, ne3;
    int64_t x_nb1, x_nb2, x_nb3, d_nb1, d_nb2, d_nb3;
    int32_t n_dims, mode;
    float theta_scale, freq_scale, ext_factor, attn_factor, corr0, corr1;
    // fused SET_ROWS
    const char * set_idx; char * set_dst; int64_t set_nb1; int32_t idx_is_i64, set_type;
};

__device__ __forceinline__ void rope_yarn(float theta_extrap, float freq_scale, float corr0, float corr1, int i0,
                                          float ext_factor, float mscale, float & c, float & s) {
    float theta_interp = freq_scale * theta_extrap, theta = theta_interp;
    if (ext_factor != 0.0f) {
        const float y = ((float) (i0 / 2) - corr0) / fmaxf(0.001f, corr1 - corr0);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * ext_factor;
        theta = theta_interp * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / freq_scale);
    }
    c = cosf(theta) * mscale;
    s = sinf(theta) * mscale;
}

// one thread per rotated pair (or per 2 pass-through elements); grid.x over ne0/2, grid.y = rows
__global__ void __launch_bounds__(128) rope_kernel(const char * x, const int32_t * pos, const float * ff, char * dst, rope_dev_params p) {
    pdl_launch_dependents();
    const int i0 = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i0 >= p.ne0) return;
    const int64_t row = blockIdx.y;
    const int64_t i1 = row % p.ne1, i2 = (row / p.ne1) % p.ne2, i3 = row / (p.ne1 * p.ne2);
    pdl_wait();
    const float * xr = (const float *) (x + i1 * p.x_nb1 + i2 * p.x_nb2 + i3 * p.x_nb3);
    float v0, v1; int a, b;
    if (i0 >= p.n_dims) { a = i0; b = i0 + 1; v0 = xr[a]; v1 = xr[b]; }
    else {
        if (p.mode == 2) { a = i0 / 2; b = a + p.n_dims / 2; } else { a = i0; b = i0 + 1; }
        const float theta_base = (float) pos[i2] * powf(p.theta_scale, (float) (i0 / 2));
        const float f = ff ? ff[i0 / 2] : 1.0f;
        float c, s;
        rope_yarn(theta_base / f, p.freq_scale, p.corr0, p.corr1, i0, p.ext_factor, p.attn_factor, c, s);
        const float x0 = xr[a], x1 = xr[b];
        v0 = x0 * c - x1 * s;
        v1 = x0 * s + x1 * c;
    }
    if (p.set_dst) {
        const int64_t r = p.idx_is_i64 ? ((const int64_t *) p.set_idx)[i2] : (int64_t) ((const int32_t *) p.set_idx)[i2];
        char * o = p.set_dst + r * p.set_nb1;
        const int64_t e = i1 * p.ne0;
        if (p.set_type == B200_F16) { ((__half *) o)[e + a] = __float2half_rn(v0); ((__half *) o)[e + b] = __float2half_rn(v1); }
        else                        { ((float *) o)[e + a] = v0; ((float *) o)[e + b] = v1; }
    } else {
        float * dr = (float *) (dst + i1 * p.d_nb1 + i2 * p.d_nb2 + i3 * p.d_nb3);
        dr[a] = v0; dr[b] = v1;
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rope(const b200_tensor * x, const b200_tensor * pos, const b200_tensor * ff, const b200_tensor * dst,
                         const b200_rope_params * rp, const b200_tensor * set_idx, const b200_tensor * set_dst, void * stream) {
    if (!x || !pos || !dst || !rp) return B200_ERR_ARG;
    if (x->type != B200_F32 || dst->type != B200_F32 || x->nb[0] != 4 || dst->nb[0] != 4 || pos->type != B200_I32) return B200_ERR_UNSUPPORTED;
    for (int i = 0; i < 4; ++i) if (x->ne[i] != dst->ne[i]) return B200_ERR_UNSUPPORTED;
    if ((rp->mode != 0 && rp->mode != 2) || rp->n_dims <= 0 || rp->n_dims % 2 != 0 || rp->n_dims > x->ne[0] || x->ne[0] % 2 != 0) return B200_ERR_UNSUPPORTED;
    if (pos->ne[0] != x->ne[2] || pos->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    if (ff && (ff->type != B200_F32 || ff->ne[0] < rp->n_dims / 2 || ff->nb[0] != 4)) return B200_ERR_UNSUPPORTED;
    const int64_t rows = x->ne[1] * x->ne[2] * x->ne[3];
    if (rows > 65535) return B200_ERR_UNSUPPORTED;
    if (set_idx || set_dst) {
        if (!set_idx || !set_dst || x->ne[3] != 1) return B200_ERR_UNSUPPORTED;
        if ((set_idx->type != B200_I64 && set_idx->type != B200_I32) || set_idx->ne[0] != x->ne[2] || set_idx->nb[0] != (int64_t) b200_type_size(set_idx->type)) return B200_ERR_UNSUPPORTED;
        if ((set_dst->type != B200_F16 && set_dst->type != B200_F32) || set_dst->ne[0] != x->ne[0] * x->ne[1] || set_dst->nb[0] != (int64_t) b200_type_size(set_dst->type)) return B200_ERR_UNSUPPORTED;
    }
    if (stream == B200_STREAM_VALIDATE) return B200_OK;
    if (rows * x->ne[0] == 0) return B200_OK;

    rope_dev_params p{};
    p.ne0 = x->ne[0]; p.ne1 = x->ne[1]; p.ne2 = x->ne[2]; p.ne3 = x->ne[3];
    p.x_nb1 = x->nb[1]; p.x_nb2 = x->nb[2]; p.x_nb3 = x->nb[3];
    p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2]; p.d_nb3 = dst->nb[3];
    p.n_dims = rp->n_dims; p.mode = rp->mode;
    p.theta_scale = powf(rp->freq_base, -2.0f / (float) rp->n_dims);
    p.freq_scale = rp->freq_scale; p.ext_factor = rp->ext_factor; p.attn_factor = rp->attn_factor;
    {   // ggml_rope_yarn_corr_dims
        const float PI = 3.14159265358979323846f;
        auto corr_dim = [&](float n_rot) { return rp->n_dims * logf(rp->n_ctx_orig / (n_rot * 2.0f * PI)) / (2.0f * logf(rp->freq_base)); };
        p.corr0 = fmaxf(0.0f, floorf(corr_dim(rp->beta_fast)));
        p.corr1 = fminf((float) rp->n_dims - 1.0f, ceilf(corr_dim(rp->beta_slow)));
    }
    if (set_dst) {
        p.set_idx = (const char *) set_idx->data; p.set_dst = (char *) set_dst->data; p.set_nb1 = set_dst->nb[1];
        p.idx_is_i64 = set_idx->type == B200_I64; p.set_type = set_dst->type;
    }
    const int64_t npairs = x->ne[0] / 2;
    dim3 grid((unsigned) ((npairs + 127) / 128), (unsigned) rows);
    rope_kernel<<<grid, 128, 0, (cudaStream_t) stream>>>((const char *) x->data, (const int32_t *) pos->data,
                                                         ff ? (const float *) ff->data : nullptr, (char *) dst->data, p);
    count_launch();
    return check_launch("rope");
}
