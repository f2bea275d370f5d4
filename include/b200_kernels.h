/* b200_kernels.h — C ABI of the B200 (sm_100a) kernel library (libb200_kernels.so).
 *
 * Plain pointers, sizes and strides only: no ggml, torch or C++ types cross this boundary.
 * The ggml backend plugin (libggml-b200.so, include/b200_backend.h) translates ggml tensors
 * into b200_tensor descriptors and calls these entry points; tests call them through ctypes.
 *
 * Each entry point names the reference interface it replaces (paths under the reference's
 * ggml/src/ggml-cuda/).  All functions are asynchronous on `stream` (a cudaStream_t passed as
 * void*), return B200_OK (0) or a negative error, and never fall back to the CPU.
 * Every *_supported() predicate is pure host code (no CUDA call) so the plugin's supports_op
 * and the launchers share one definition of "claimed".
 */
#ifndef B200_KERNELS_H
#define B200_KERNELS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_API __attribute__((visibility("default")))

enum b200_status {
    B200_OK              = 0,
    B200_ERR_UNSUPPORTED = -1, /* shape/type/stride combination not claimed */
    B200_ERR_CUDA        = -2, /* a CUDA runtime call failed; see b200_last_error() */
    B200_ERR_ARG         = -3
};

/* Element types; numeric values equal ggml_type (reference ggml/include/ggml.h:390-430) */
enum b200_type {
    B200_F32 = 0, B200_F16 = 1, B200_Q4_0 = 2, B200_Q8_0 = 8, B200_Q8_1 = 9,
    B200_Q4_K = 12, B200_Q6_K = 14, B200_I32 = 26, B200_I64 = 27, B200_BF16 = 30, B200_MXFP4 = 39
};

/* Tensor descriptor: same meaning as ggml_tensor {data, type, ne[4], nb[4]}:
 * ne[0] is the contiguous dimension, nb[i] are byte strides, quantized rows are block arrays. */
typedef struct b200_tensor {
    void *  data;
    int32_t type;
    int32_t _pad;
    int64_t ne[4];
    int64_t nb[4];
} b200_tensor;

/* ---- library ------------------------------------------------------------------------- */
B200_API const char * b200_last_error(void);
B200_API int          b200_device_count(void);
B200_API uint64_t     b200_launch_count(void);          /* kernels launched by this library so far */
B200_API size_t       b200_type_size(int type);         /* bytes per block */
B200_API int64_t      b200_blck_size(int type);         /* elements per block */
B200_API size_t       b200_row_size(int type, int64_t ne0);

/* ---- K1: activation quantizer --------------------------------------------------------
 * replaces quantize_row_q8_1_cuda (quantize.cu:273-287) / kernel quantize_q8_1 (:4-48).
 * x: f32 [ne0, ne1, ne2, ne3] (nb[0]==4). y: block_q8_1 rows of ne0_padded/32 blocks,
 * laid out contiguously as [ne0_padded/32, ne1, ne2, ne3]; columns >= ne0 quantize as zeros.
 * Arithmetic is that of quantize_row_q8_1_ref (ggml-quants.c:225-258): d = amax/127,
 * q = roundf(x * (1/d)), s = fp16(d * sum(q))  — bit-exact, see DESIGN.md. */
B200_API int b200_quantize_q8_1(const b200_tensor * x, void * y, int64_t ne0_padded, void * stream);
B200_API size_t b200_q8_1_buffer_size(const b200_tensor * x, int64_t ne0_padded);

/* ---- K4: quantized GEMV (decode MUL_MAT / MUL_MAT_ID) -------------------------------
 * replaces ggml_cuda_mul_mat_vec_q (mmvq.cu:638-736) and kernel mul_mat_vec_q (:140-356). */
enum b200_glu_op { B200_GLU_NONE = -1, B200_GLU_REGLU = 0, B200_GLU_GEGLU = 1, B200_GLU_SWIGLU = 2,
                   B200_GLU_SWIGLU_OAI = 3, B200_GLU_GEGLU_ERF = 4, B200_GLU_GEGLU_QUICK = 5 };

typedef struct b200_mmvq_args {
    b200_tensor w;          /* quantized weights [K, M, ne02, ne03] (ne02 = n_expert for MUL_MAT_ID) */
    const void * xq;        /* q8_1 activations from b200_quantize_q8_1 */
    int64_t xq_ne0_padded;  /* K padded (multiple of 32; the plugin uses GGML_PAD(K,512)) */
    int64_t x_ne[4];        /* logical shape of the f32 activation tensor [K, ne11, ne12, ne13] */
    b200_tensor dst;        /* f32 [M, ne1, ne2, ne3] */
    const b200_tensor * ids;    /* NULL, or i32 [n_expert_used, n_tokens] (MUL_MAT_ID) */
    /* optional fused epilogue (mmvq.cu:195-232, 321-350) */
    const b200_tensor * bias;       /* f32 [M,...] added to w.x            (or NULL) */
    const b200_tensor * gate_w;     /* second weight matrix, same shape/type as w (or NULL) */
    const b200_tensor * gate_bias;  /* f32 (or NULL) */
    int32_t glu_op;                 /* b200_glu_op, used iff gate_w != NULL */
    float   glu_alpha, glu_limit;   /* SWIGLU_OAI parameters */
} b200_mmvq_args;

B200_API int b200_mmvq_supported(const b200_mmvq_args * a);
B200_API int b200_mul_mat_vec_q(const b200_mmvq_args * a, void * stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_KERNELS_H */
