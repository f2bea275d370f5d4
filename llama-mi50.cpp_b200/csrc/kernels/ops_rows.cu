// ops_rows.cu — row G glue: GET_ROWS, SET_ROWS, CPY/CONT.
// Replaces reference ggml-cuda/getrows.cu (:238), set-rows.cu (:318), cpy.cu.
#include <cuda_bf16.h>

#include "quantize.cuh"
#include "vecdot.cuh"

namespace b200 {

#define VALIDATE_ONLY(stream) ((stream) == B200_STREAM_VALIDATE)

struct st4 { int64_t ne[4]; int64_t nb[4]; };
static inline st4 S4(const b200_tensor & t) { st4 s; for (int i = 0; i < 4; ++i) { s.ne[i] = t.ne[i]; s.nb[i] = t.nb[i]; } return s; }

template <typename T> __device__ __forceinline__ float ld_as_f32(const void * p);
template <> __device__ __forceinline__ float ld_as_f32<float>(const void * p) { return *(const float *) p; }
template <> __device__ __forceinline__ float ld_as_f32<__half>(const void * p) { return __half2float(*(const __half *) p); }
template <> __device__ __forceinline__ float ld_as_f32<__nv_bfloat16>(const void * p) { return __bfloat162float(*(const __nv_bfloat16 *) p); }
template <typename T> __device__ __forceinline__ void st_from_f32(void * p, float v);
template <> __device__ __forceinline__ void st_from_f32<float>(void * p, float v) { *(float *) p = v; }
template <> __device__ __forceinline__ void st_from_f32<__half>(void * p, float v) { *(__half *) p = __float2half_rn(v); }
template <> __device__ __forceinline__ void st_from_f32<__nv_bfloat16>(void * p, float v) { *(__nv_bfloat16 *) p = __float2bfloat16_rn(v); }

// ------------------------------------------------------------------ GET_ROWS
// grid: (ne10, ne11*ne12); quantized: one thread per 32-weight unit
template <int TYPE>
__global__ void __launch_bounds__(128) get_rows_q_kernel(const char * src, const char * ids, char * dst, st4 ss, st4 si, st4 sd) {
    using F = fmt<TYPE>;
    pdl_launch_dependents();
    const int64_t i10 = blockIdx.x, i11 = blockIdx.y % si.ne[1], i12 = blockIdx.y / si.ne[1];
    pdl_wait();
    const int32_t r = *(const int32_t *) (ids + i10 * si.nb[0] + i11 * si.nb[1] + i12 * si.nb[2]);
    const uint8_t * row = (const uint8_t *) (src + (int64_t) r * ss.nb[1] + i11 * ss.nb[2] + i12 * ss.nb[3]);
    float * out = (float *) (dst + i10 * sd.nb[1] + i11 * sd.nb[2] + i12 * sd.nb[3]);
    for (int64_t u = threadIdx.x; u < ss.ne[0] / 32; u += blockDim.x) {
        unit_w w;
        F::unpack(row, u, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float s = w.wd[i >> 2];
            const int v = w.v[i];
            float4 o;
            o.x = s * (float) (int8_t) (v) - w.wm;       o.y = s * (float) (int8_t) (v >> 8) - w.wm;
            o.z = s * (float) (int8_t) (v >> 16) - w.wm; o.w = s * (float) (int8_t) (v >> 24) - w.wm;
            *(float4 *) (out + u * 32 + i * 4) = o;
        }
    }
}
template <typename T>
__global__ void __launch_bounds__(256) get_rows_f_kernel(const char * src, const char * ids, char * dst, st4 ss, st4 si, st4 sd) {
    pdl_launch_dependents();
    const int64_t i10 = blockIdx.x, i11 = blockIdx.y % si.ne[1], i12 = blockIdx.y / si.ne[1];
    pdl_wait();
    const int32_t r = *(const int32_t *) (ids + i10 * si.nb[0] + i11 * si.nb[1] + i12 * si.nb[2]);
    const char * row = src + (int64_t) r * ss.nb[1] + i11 * ss.nb[2] + i12 * ss.nb[3];
    float * out = (float *) (dst + i10 * sd.nb[1] + i11 * sd.nb[2] + i12 * sd.nb[3]);
    for (int64_t i = threadIdx.x; i < ss.ne[0]; i += blockDim.x) out[i] = ld_as_f32<T>(row + i * sizeof(T));
}

// ------------------------------------------------------------------ SET_ROWS
template <typename TI, typename TD>
__global__ void __launch_bounds__(256) set_rows_kernel(const char * src, const char * idx, char * dst, st4 ss, st4 si, st4 sd) {
    pdl_launch_dependents();
    const int64_t i01 = blockIdx.x, i02 = blockIdx.y % ss.ne[2], i03 = blockIdx.y / ss.ne[2];
    pdl_wait();
    const int64_t r = (int64_t) *(const TI *) (idx + i01 * si.nb[0] + (i02 % si.ne[1]) * si.nb[1] + (i03 % si.ne[2]) * si.nb[2]);
    const float * in = (const float *) (src + i01 * ss.nb[1] + i02 * ss.nb[2] + i03 * ss.nb[3]);
    char * out = dst + r * sd.nb[1] + i02 * sd.nb[2] + i03 * sd.nb[3];
    for (int64_t i = threadIdx.x; i < ss.ne[0]; i += blockDim.x) st_from_f32<TD>(out + i * sizeof(TD), in[i]);
}
// q8_0 destination: one warp per 32-element block, blocks bit-exact vs quantize_row_q8_0_ref (ggml-quants.c:197-223)
template <typename TI>
__global__ void __launch_bounds__(256) set_rows_q8_0_kernel(const char * src, const char * idx, char * dst, st4 ss, st4 si, st4 sd) {
    pdl_launch_dependents();
    const int64_t i01 = blockIdx.x, i02 = blockIdx.y % ss.ne[2], i03 = blockIdx.y / ss.ne[2];
    pdl_wait();
    const int64_t r = (int64_t) *(const TI *) (idx + i01 * si.nb[0] + (i02 % si.ne[1]) * si.nb[1] + (i03 % si.ne[2]) * si.nb[2]);
    const float * in = (const float *) (src + i01 * ss.nb[1] + i02 * ss.nb[2] + i03 * ss.nb[3]);
    blk_q8_0 * out = (blk_q8_0 *) (dst + r * sd.nb[1] + i02 * sd.nb[2] + i03 * sd.nb[3]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    for (int64_t b = warp; b < ss.ne[0] / 32; b += nwarp) {
        float d; int sum;
        const int8_t q = quantize_block32_lane(in[b * 32 + lane], d, sum);
        out[b].qs[lane] = q;
        if (lane == 0) out[b].d = __float2half_rn(d);
    }
}

// ------------------------------------------------------------------ CPY (generic strided, same element count)
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) cpy_kernel(const char * src, char * dst, st4 ss, st4 sd, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t r = i;
    const int64_t s0 = r % ss.ne[0]; r /= ss.ne[0];
    const int64_t s1 = r % ss.ne[1]; r /= ss.ne[1];
    const int64_t s2 = r % ss.ne[2]; const int64_t s3 = r / ss.ne[2];
    r = i;
    const int64_t d0 = r % sd.ne[0]; r /= sd.ne[0];
    const int64_t d1 = r % sd.ne[1]; r /= sd.ne[1];
    const int64_t d2 = r % sd.ne[2]; const int64_t d3 = r / sd.ne[2];
    pdl_wait();
    const float v = ld_as_f32<TS>(src + s0 * ss.nb[0] + s1 * ss.nb[1] + s2 * ss.nb[2] + s3 * ss.nb[3]);
    st_from_f32<TD>(dst + d0 * sd.nb[0] + d1 * sd.nb[1] + d2 * sd.nb[2] + d3 * sd.nb[3], v);
}
// bit-preserving 4-byte copy (i32 and f32->f32)
__global__ void __launch_bounds__(256) cpy_b32_kernel(const char * src, char * dst, st4 ss, st4 sd, int64_t n) {
    pdl_launch_dependents();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t r = i;
    const int64_t s0 = r % ss.ne[0]; r /= ss.ne[0];
    const int64_t s1 = r % ss.ne[1]; r /= ss.ne[1];
    const int64_t s2 = r % ss.ne[2]; const int64_t s3 = r / ss.ne[2];
    r = i;
    const int64_t d0 = r % sd.ne[0]; r /= sd.ne[0];
    const int64_t d1 = r % sd.ne[1]; r /= sd.ne[1];
    const int64_t d2 = r % sd.ne[2]; const int64_t d3 = r / sd.ne[2];
    pdl_wait();
    *(uint32_t *) (dst + d0 * sd.nb[0] + d1 * sd.nb[1] + d2 * sd.nb[2] + d3 * sd.nb[3]) =
        *(const uint32_t *) (src + s0 * ss.nb[0] + s1 * ss.nb[1] + s2 * ss.nb[2] + s3 * ss.nb[3]);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_get_rows(const b200_tensor * src, const b200_tensor * ids, const b200_tensor * dst, void * stream) {
    if (!src || !ids || !dst) return B200_ERR_ARG;
    if (ids->type != B200_I32 || dst->type != B200_F32 || dst->nb[0] != 4 || dst->ne[0] != src->ne[0]) return B200_ERR_UNSUPPORTED;
    if (src->nb[0] != (int64_t) b200_type_size(src->type) || ids->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (dst->ne[1] != ids->ne[0] || dst->ne[2] != ids->ne[1] || dst->ne[3] != ids->ne[2]) return B200_ERR_UNSUPPORTED;
    if (src->ne[2] != ids->ne[1] || src->ne[3] != ids->ne[2]) return B200_ERR_UNSUPPORTED;
    if (ids->ne[1] * ids->ne[2] > 65535) return B200_ERR_UNSUPPORTED;
    const bool q = src->type == B200_Q4_0 || src->type == B200_Q8_0 || src->type == B200_Q4_K || src->type == B200_Q6_K || src->type == B200_MXFP4;
    if (q && (src->ne[0] % b200_blck_size(src->type) != 0 || dst->nb[1] % 16 != 0 || ((uintptr_t) dst->data) % 16 != 0 || dst->nb[2] % 16 != 0 || dst->nb[3] % 16 != 0)) return B200_ERR_UNSUPPORTED;
    if (!q && src->type != B200_F32 && src->type != B200_F16 && src->type != B200_BF16) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*dst) == 0) return B200_OK;
    dim3 grid((unsigned) ids->ne[0], (unsigned) (ids->ne[1] * ids->ne[2]));
    cudaStream_t st = (cudaStream_t) stream;
    const char * s = (const char *) src->data; const char * i = (const char *) ids->data; char * d = (char *) dst->data;
    switch (src->type) {
        case B200_Q4_0:  get_rows_q_kernel<B200_Q4_0><<<grid, 128, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_Q8_0:  get_rows_q_kernel<B200_Q8_0><<<grid, 128, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_Q4_K:  get_rows_q_kernel<B200_Q4_K><<<grid, 128, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_Q6_K:  get_rows_q_kernel<B200_Q6_K><<<grid, 128, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_MXFP4: get_rows_q_kernel<B200_MXFP4><<<grid, 128, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_F32:   get_rows_f_kernel<float><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        case B200_F16:   get_rows_f_kernel<__half><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
        default:         get_rows_f_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*ids), S4(*dst)); break;
    }
    count_launch();
    return check_launch("get_rows");
}

extern "C" int b200_set_rows(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst, void * stream) {
    if (!src || !idx || !dst) return B200_ERR_ARG;
    if (src->type != B200_F32 || src->nb[0] != 4 || (idx->type != B200_I64 && idx->type != B200_I32)) return B200_ERR_UNSUPPORTED;
    if (dst->ne[0] != src->ne[0] || dst->ne[2] != src->ne[2] || dst->ne[3] != src->ne[3] || idx->ne[0] != src->ne[1] || idx->ne[3] != 1) return B200_ERR_UNSUPPORTED;
    if (idx->ne[1] < 1 || idx->ne[2] < 1 || src->ne[2] % idx->ne[1] != 0 || src->ne[3] % idx->ne[2] != 0) return B200_ERR_UNSUPPORTED;
    if (dst->nb[0] != (int64_t) b200_type_size(dst->type) || src->ne[2] * src->ne[3] > 65535) return B200_ERR_UNSUPPORTED;
    switch (dst->type) { case B200_F32: case B200_F16: case B200_BF16: break;
        case B200_Q8_0: if (src->ne[0] % 32 != 0) return B200_ERR_UNSUPPORTED; break;
        default: return B200_ERR_UNSUPPORTED; }
    if (VALIDATE_ONLY(stream)) return B200_OK;
    if (nelements(*src) == 0) return B200_OK;
    dim3 grid((unsigned) src->ne[1], (unsigned) (src->ne[2] * src->ne[3]));
    cudaStream_t st = (cudaStream_t) stream;
    const char * s = (const char *) src->data; const char * i = (const char *) idx->data; char * d = (char *) dst->data;
    const bool i64 = idx->type == B200_I64;
#define SR(TD) do { if (i64) set_rows_kernel<int64_t, TD><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*idx), S4(*dst)); \
                    else     set_rows_kernel<int32_t, TD><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*idx), S4(*dst)); } while (0)
    switch (dst->type) {
        case B200_F32: SR(float); break;
        case B200_F16: SR(__half); break;
        case B200_BF16: SR(__nv_bfloat16); break;
        default:
            if (i64) set_rows_q8_0_kernel<int64_t><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*idx), S4(*dst));
            else     set_rows_q8_0_kernel<int32_t><<<grid, 256, 0, st>>>(s, i, d, S4(*src), S4(*idx), S4(*dst));
    }
#undef SR
    count_launch();
    return check_launch("set_rows");
}

extern "C" int b200_cpy(const b200_tensor * src, const b200_tensor * dst, void * stream) {
    if (!src || !dst) return B200_ERR_ARG;
    if (nelements(*src) != nelements(*dst)) return B200_ERR_UNSUPPORTED;
    const int ts = src->type, td = dst->type;
    auto flt = [](int t) { return t == B200_F32 || t == B200_F16 || t == B200_BF16; };
    const bool b32 = (ts == B200_I32 && td == B200_I32) || (ts == B200_F32 && td == B200_F32);
    if (!b32 && !(flt(ts) && flt(td))) return B200_ERR_UNSUPPORTED;
    if (VALIDATE_ONLY(stream)) return B200_OK;
    const int64_t n = nelements(*src);
    if (n == 0) return B200_OK;
    cudaStream_t st = (cudaStream_t) stream;
    const char * s = (const char *) src->data; char * d = (char *) dst->data;
    if (is_contiguous(*src) && is_contiguous(*dst) && ts == td) {
        if (cudaMemcpyAsync(d, s, (size_t) n * b200_type_size(ts), cudaMemcpyDeviceToDevice, st) != cudaSuccess) { set_error("cpy: memcpy failed"); return B200_ERR_CUDA; }
        return B200_OK;
    }
    const unsigned grid = (unsigned) ((n + 255) / 256);
#define CP(TS, TD) cpy_kernel<TS, TD><<<grid, 256, 0, st>>>(s, d, S4(*src), S4(*dst), n)
    if (b32) cpy_b32_kernel<<<grid, 256, 0, st>>>(s, d, S4(*src), S4(*dst), n);
    else if (ts == B200_F32 && td == B200_F16) CP(float, __half);
    else if (ts == B200_F32 && td == B200_BF16) CP(float, __nv_bfloat16);
    else if (ts == B200_F16 && td == B200_F32) CP(__half, float);
    else if (ts == B200_F16 && td == B200_F16) CP(__half, __half);
    else if (ts == B200_F16 && td == B200_BF16) CP(__half, __nv_bfloat16);
    else if (ts == B200_BF16 && td == B200_F32) CP(__nv_bfloat16, float);
    else if (ts == B200_BF16 && td == B200_F16) CP(__nv_bfloat16, __half);
    else CP(__nv_bfloat16, __nv_bfloat16);
#undef CP
    count_launch();
    return check_launch("cpy");
}
