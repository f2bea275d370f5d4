// ops_elementwise.cu — row G glue: bin-broadcast, scale, GLU, unary, RMS norm (+mul +add).
// Replaces reference ggml-cuda/binbcast.cu, scale.cu, unary.cu (GLU :337), norm.cu (:476) and the
// {RMS_NORM, MUL[, ADD]} fusion (ggml-cuda.cu:3844-3854).  All f32 arithmetic.
#include "common.cuh"

namespace b200 {

#define VALIDATE_ONLY(stream) ((stream) == B200_STREAM_VALIDATE)

struct strides4 { int64_t ne[4]; int64_t nb[4]; };
static inline strides4 S(const b200_tensor & t) { strides4 s; for (int i = 0; i < 4; ++i) { s.ne[i] = t.ne[i]; s.nb[i] = t.nb[i]; } return s; }

template <typename T> __device__ __forceinline__ float ldf(const void * p);
template <> __device__ __forceinline__ float ldf<float>(const void * p) { return *(const float *) p; }
template <> __device__ __forceinline__ float ldf<__half>(const void * p) { return __half2float(*(const __half *) p); }
template <typename T> __device__ __forceinline__ void stf(void * p, float v);
template <> __device__ __forceinline__ void stf<float>(void * p, float v) { *(float *) p = v; }
template <> __device__ __forceinline__ void stf<__half>(void * p, float v) { *(__half *) p = __float2half_rn(v); }

// ------------------------------------------------------------------ bin broadcast
template <int OP, typename TA, typename TB, typename TD>
__global__ void __launch_bounds__(256) bin_bcast_kernel(const char * a, const char * b, char * d, strides4 sa, strides4 sb, strides4 sd, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t r = i;
    const int64_t i0 = r % sd.ne[0]; r /= sd.ne[0];
    const int64_t i1 = r % sd.ne[1]; r /= sd.ne[1];
    const int64_t i2 = r % sd.ne[2]; const int64_t i3 = r / sd.ne[2];
    pdl_wait();
    const float x = ldf<TA>(a + i0 * sa.nb[0] + i1 * sa.nb[1] + i2 * sa.nb[2] + i3 * sa.nb[3]);
    const float y = ldf<TB>(b + (i0 % sb.ne[0]) * sb.nb[0] + (i1 % sb.ne[1]) * sb.nb[1] + (i2 % sb.ne[2]) * sb.nb[2] + (i3 % sb.ne[3]) * sb.nb[3]);
    const float z = OP == B200_BIN_ADD ? x + y : OP == B200_BIN_SUB ? x - y : OP == B200_BIN_MUL ? x * y : x / y;
    stf<TD>(d + i0 * sd.nb[0] + i1 * sd.nb[1] + i2 * sd.nb[2] + i3 * sd.nb[3], z);
}

template <int OP>
static int bin_launch(const b200_tensor * a, const b200_tensor * b, const b200_tensor * d, cudaStream_t st) {
    const int64_t n = nelements(*d);
    const unsigned grid = (unsigned) ((n + 255) / 256);
    const char * pa = (const char *) a->data; const char * pb = (const char *) b->data; char * pd = (char *) d->data;
#define BIN_CASE(TA, TB, TD) bin_bcast_kernel<OP, TA, TB, TD><<<grid, 256, 0, st>>>(pa, pb, pd, S(*a), S(*b), S(*d), n)
    if (a->type == B200_F32 && b->type == B200_F32 && d->type == B200_F32) BIN_CASE(float, float, float);
    else if (a->type == B200_F16 && b->type == B200_F16 && d->type == B200_F16) BIN_CASE(__half, __half, __half);
    else if (a->type == B200_F16 && b->type == B200_F32 && d->type == B200_F16) BIN_CASE(__half, float, __half);
    else if (a->type == B200_F16 && b->type == B200_F32 && d->type == B200_F32) BIN_CASE(__half, float, float);
    else if (a->type == B200_F32 && b->type == B200_F16 && d->type == B200_F32) BIN_CASE(float, __half, float);
    else return B200_ERR_UNSUPPORTED;
#undef BIN_CASE
    count_launch();
    return check_launch("bin_bcast");
}

// ------------------------------------------------------------------ scale / unary / glu
__global__ void __launch_bounds__(256) scale_kernel(const float * x, float * y, float s, float b, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pdl_wait();
    y[i] = x[i] * s + b;
}

__device__ __forceinline__ float unary_f(int op, float x) {
    switch (op) {
        case B200_UNARY_SILU: return x / (1.0f + expf(-x));
        case B200_UNARY_GELU: return 0.5f * x * (1.0f + tanhf(0.79788456080286535587989211986876f * x * (1.0f + 0.044715f * x * x)));
        case B200_UNARY_RELU: return fmaxf(x, 0.0f);
        case B200_UNARY_SIGMOID: return 1.0f / (1.0f + expf(-x));
        case B200_UNARY_TANH: return tanhf(x);
        case B200_UNARY_GELU_QUICK: return x / (1.0f + expf(-1.702f * x));
        case B200_UNARY_NEG: return -x;
        case B200_UNARY_EXP: return expf(x);
        default: return x;
    }
}
__global__ void __launch_bounds__(256) unary_kernel(int op, const float * x, float * y, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pdl_wait();
    y[i] = unary_f(op, x[i]);
}

__device__ __forceinline__ float glu_f(int op, float g, float x, float alpha, float limit) {
    switch (op) {
        case B200_GLU_REGLU:  return fmaxf(g, 0.0f) * x;
        case B200_GLU_GEGLU:  return unary_f(B200_UNARY_GELU, g) * x;
        case B200_GLU_SWIGLU: return g / (1.0f + expf(-g)) * x;
        case B200_GLU_SWIGLU_OAI: g = fminf(g, limit); x = fmaxf(fminf(x, limit), -limit); return g / (1.0f + expf(-g * alpha)) * (1.0f + x);
        case B200_GLU_GEGLU_ERF:   return 0.5f * g * (1.0f + erff(g * 0.70710678118654752440f)) * x;
        case B200_GLU_GEGLU_QUICK: return g / (1.0f + expf(-1.702f * g)) * x;
        default: return x;
    }
}
// rows of nc outputs; gate/up rows at byte strides sg/su
__global__ void __launch_bounds__(256) glu_kernel(int op, const char * g, const char * u, float * d, int64_t nc, int64_t nrows,
                                                  int64_t sg, int64_t su, int64_t sdst, float alpha, float limit) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc * nrows) return;
    const int64_t r = i / nc, c = i % nc;
    pdl_wait();
    const float gv = ((const float *) (g + r * sg))[c], uv = ((const float *) (u + r * su))[c];
    ((float *) ((char *) d + r * sdst))[c] = glu_f(op, gv, uv, alpha, limit);
}

// ------------------------------------------------------------------ RMS norm (+mul, +add)
// one block per row; x row cached in registers when ne0 <= 8*blockDim
template <int BS>
__global__ void __launch_bounds__(BS) rms_norm_kernel(const char * x, const char * mul, const char * add, char * dst,
                                                      strides4 sx, strides4 sm, strides4 sa, strides4 sd, float eps) {
    pdl_launch_dependents();
    const int64_t i1 = blockIdx.x, i2 = blockIdx.y, i3 = blockIdx.z;
    const int64_t ne0 = sx.ne[0];
    const float * xr = (const float *) (x + i1 * sx.nb[1] + i2 * sx.nb[2] + i3 * sx.nb[3]);
    float * dr = (float *) (dst + i1 * sd.nb[1] + i2 * sd.nb[2] + i3 * sd.nb[3]);
    const float * mr = mul ? (const float *) (mul + (i1 % sm.ne[1]) * sm.nb[1] + (i2 % sm.ne[2]) * sm.nb[2] + (i3 % sm.ne[3]) * sm.nb[3]) : nullptr;
    const float * ar = add ? (const float *) (add + (i1 % sa.ne[1]) * sa.nb[1] + (i2 % sa.ne[2]) * sa.nb[2] + (i3 % sa.ne[3]) * sa.nb[3]) : nullptr;
    pdl_wait();
    float ss = 0.0f;
    for (int64_t i = threadIdx.x; i < ne0; i += BS) { const float v = xr[i]; ss += v * v; }
    ss = warp_sum(ss);
    __shared__ float red[32];
    if (BS > 32) {
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
        __syncthreads();
        ss = threadIdx.x < BS / 32 ? red[threadIdx.x] : 0.0f;
        ss = warp_sum(ss);
        if (threadIdx.x == 0) red[0] = ss;
        __syncthreads();
        ss = red[0];
    }
    const float scale = rsqrtf(ss / (float) ne0 + eps);
    for (int64_t i = threadIdx.x; i < ne0; i += BS) {
        float v = xr[i] * scale;
        if (mr) v *= mr[i % sm.ne[0]];
        if (ar) v += ar[i % sa.ne[0]];
        dr[i] = v;
    }
}

}  // namespace b200

using namespace b200;

static bool bcast_ok(const b200_tensor & big, const b200_tensor & small) {
    for (int i = 0; i < 4; ++i) if (small.ne[i] < 1 || big.ne[i] % small.ne[i] != 0) return false;
    return true;
}

extern "C" int b200_bin_bcast(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * d, void * stream) {
    if (!a || !b || !d || op < 0 || op > 3) return B200_ERR_ARG;
    for (int i = 0; i < 4; ++i) if (a->ne[i] != d->ne[i]) return B200_ERR_UNSUPPORTED;
    if (!bcast_ok(*d, *b)) return B200_ERR_UNSUPPORTED;
    const bool f32 = a->type == B200_F32 && b->type == B200_F32 && d->type == B200_F32;
    const bool mixed = (a->type == B200_F16 && (b->type == B200_F16 || b->type == B200_F32) && (d->type == B200_F16 || (d->type == B200_F32 && b->type == B200_F32))) ||
                       (a->type == B200_F32 && b->type == B200_F16 && d->type == B200_F32);
    if (!f32 && !mixed) return B200_ERR_UNSUPPORTED;
    if (a->type == B200_F16 && b->type == B200_F16 && d->type != B200_F16) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*d) == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    switch (op) {
        case B200_BIN_ADD: return bin_launch<B200_BIN_ADD>(a, b, d, st);
        case B200_BIN_SUB: return bin_launch<B200_BIN_SUB>(a, b, d, st);
        case B200_BIN_MUL: return bin_launch<B200_BIN_MUL>(a, b, d, st);
        default:           return bin_launch<B200_BIN_DIV>(a, b, d, st);
    }
}

extern "C" int b200_scale(const b200_tensor * x, const b200_tensor * d, float scale, float bias, void * stream) {
    if (!x || !d) return B200_ERR_ARG;
    if (x->type != B200_F32 || d->type != B200_F32 || !is_contiguous(*x) || !is_contiguous(*d) || nelements(*x) != nelements(*d)) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    const int64_t n = nelements(*x);
    if (n == 0) return B200_OK;
    scale_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, (cudaStream_t) stream>>>((const float *) x->data, (float *) d->data, scale, bias, n);
    count_launch();
    return check_launch("scale");
}

extern "C" int b200_unary(int op, const b200_tensor * x, const b200_tensor * d, void * stream) {
    if (!x || !d || op < 0 || op > B200_UNARY_EXP) return B200_ERR_ARG;
    if (x->type != B200_F32 || d->type != B200_F32 || !is_contiguous(*x) || !is_contiguous(*d) || nelements(*x) != nelements(*d)) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    const int64_t n = nelements(*x);
    if (n == 0) return B200_OK;
    unary_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, (cudaStream_t) stream>>>(op, (const float *) x->data, (float *) d->data, n);
    count_launch();
    return check_launch("unary");
}

extern "C" int b200_glu(int op, const b200_tensor * gate, const b200_tensor * up, const b200_tensor * d,
                        int swapped, float alpha, float limit, void * stream) {
    if (!gate || !d || op < B200_GLU_REGLU || op > B200_GLU_GEGLU_QUICK) return B200_ERR_ARG;
    if (gate->type != B200_F32 || d->type != B200_F32 || gate->nb[0] != 4 || d->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    if (!is_contiguous_rows(*gate) || !is_contiguous_rows(*d)) return B200_ERR_UNSUPPORTED;
    const int64_t nc = d->ne[0], nr = nrows(*d);
    const char * g = (const char *) gate->data; const char * u;
    int64_t su;
    if (up) {
        if (up->type != B200_F32 || up->nb[0] != 4 || !is_contiguous_rows(*up) || up->ne[0] != nc || gate->ne[0] != nc || nrows(*up) != nr || nrows(*gate) != nr) return B200_ERR_UNSUPPORTED;
        u = (const char *) up->data; su = up->nb[1];
    } else {
        if (gate->ne[0] != 2 * nc || nrows(*gate) != nr) return B200_ERR_UNSUPPORTED;
        u = g + (swapped ? 0 : nc * 4); g = g + (swapped ? nc * 4 : 0); su = gate->nb[1];
    }
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nc * nr == 0) return B200_OK;
    glu_kernel<<<(unsigned) ((nc * nr + 255) / 256), 256, 0, (cudaStream_t) stream>>>(op, g, u, (float *) d->data, nc, nr, gate->nb[1], su, d->nb[1], alpha, limit);
    count_launch();
    return check_launch("glu");
}

extern "C" int b200_rms_norm(const b200_tensor * x, const b200_tensor * mul, const b200_tensor * add,
                             const b200_tensor * d, float eps, void * stream) {
    if (!x || !d) return B200_ERR_ARG;
    if (x->type != B200_F32 || d->type != B200_F32 || x->nb[0] != 4 || d->nb[0] != 4) return B200_ERR_UNSUPPORTED;
    for (int i = 0; i < 4; ++i) if (x->ne[i] != d->ne[i]) return B200_ERR_UNSUPPORTED;
    if (mul && (mul->type != B200_F32 || mul->nb[0] != 4 || !bcast_ok(*d, *mul))) return B200_ERR_UNSUPPORTED;
    if (add && (add->type != B200_F32 || add->nb[0] != 4 || !bcast_ok(*d, *add))) return B200_ERR_UNSUPPORTED;
    if (x->ne[2] > 65535 || x->ne[3] > 65535) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*x) == 0) return B200_OK;
    dim3 grid((unsigned) x->ne[1], (unsigned) x->ne[2], (unsigned) x->ne[3]);
    strides4 sm = mul ? S(*mul) : S(*x), sa = add ? S(*add) : S(*x);
    const char * pm = mul ? (const char *) mul->data : nullptr; const char * pa = add ? (const char *) add->data : nullptr;
    if (x->ne[0] < 1024) rms_norm_kernel<256><<<grid, 256, 0, (cudaStream_t) stream>>>((const char *) x->data, pm, pa, (char *) d->data, S(*x), sm, sa, S(*d), eps);
    else                 rms_norm_kernel<1024><<<grid, 1024, 0, (cudaStream_t) stream>>>((const char *) x->data, pm, pa, (char *) d->data, S(*x), sm, sa, S(*d), eps);
    count_launch();
    return check_launch("rms_norm");
}
