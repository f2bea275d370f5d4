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
    /* optional column window [col_begin, col_begin+col_count) of x/dst dim 1 (0,0 = all columns);
     * lets a caller walk an activation batch wider than 8 columns in chunks (MUL_MAT only) */
    int32_t col_begin, col_count;
} b200_mmvq_args;

B200_API int b200_mmvq_supported(const b200_mmvq_args * a);
B200_API int b200_mul_mat_vec_q(const b200_mmvq_args * a, void * stream);
/* MUL_MAT_ID with many tokens (prefill): the (expert slot, token) pairs are bucketed by expert on the device (reference
 * mm_ids_helper, mmid.cu:28-116) and multiplied in windows of up to 8 pairs of one expert, so an expert matrix is read once per
 * 8 columns instead of once per column (reference: mmq over expert_bounds, mmq.cu:160-217).  `a` as for b200_mul_mat_vec_q with
 * ids set and no fused epilogue; scratch: device memory of b200_mul_mat_id_grouped_scratch_size() bytes, 16-byte aligned.
 * Results are bit-identical to the one-column-per-pair path. */
/* MUL_MAT with 2-D weights and ANY number of columns in one launch (every 8-column window; no ids, no epilogue, no column
 * window): the fallback of the tensor-core GEMM for rows that are not 16-byte multiples (reference: mul_mat_vec_q / mmq router,
 * ggml-cuda.cu:2183-2266) */
B200_API int    b200_mul_mat_vec_q_wide(const b200_mmvq_args * a, void * stream);
B200_API size_t b200_mul_mat_id_grouped_scratch_size(const b200_mmvq_args * a);
B200_API int    b200_mul_mat_id_grouped(const b200_mmvq_args * a, void * scratch, size_t scratch_size, void * stream);


/* ---- K4s: streaming fused GEMV (the decode hot path) ---------------------------------------------
 * One launch computes up to B200_GEMV_MAX_SEG matrix-vector products that share ONE activation
 * column (wq/wk/wv of a layer, or ffn_gate/ffn_up), replacing the reference's sequence
 *   rms_norm (norm.cu:476) -> mul -> quantize_q8_1 (quantize.cu:4-48) -> mul_mat_vec_q (mmvq.cu:140-356) x n
 *   [-> add / glu (unary.cu:337)]
 * with
 *   prologue : x f32 [K] (optionally RMS-normalised and multiplied by norm_w) is quantised to q8_1 in
 *              shared memory with the arithmetic of quantize_row_q8_1_ref (ggml-quants.c:225-258);
 *   stream   : weight rows travel HBM -> shared memory as 1-D TMA bulk copies (cp.async.bulk) through
 *              an mbarrier ring; consumer warps do the dp4a dot products out of shared memory;
 *   epilogue : + bias / residual, or GLU pairing (segment 0 = gate, segment 1 = up, one dst).
 * Weight tensors must be row-contiguous 2-D (nb[1] == row size) with a 16-byte aligned base, and the
 * allocation must extend >= 16 bytes past the last row (the plugin's buffers and torch do).
 * Kernels are chained with programmatic dependent launch: weight prefetch of launch N+1 overlaps
 * the tail of launch N. */
#define B200_GEMV_MAX_SEG 4
typedef struct b200_gemv_seg {
    b200_tensor w;              /* quantized [K, M] */
    const b200_tensor * bias;   /* f32 [M] added to the product (bias or residual stream), or NULL */
    b200_tensor dst;            /* f32 [M] (ignored for segment 0 of a GLU pair) */
} b200_gemv_seg;
typedef struct b200_gemv_fused_args {
    b200_tensor x;                  /* f32 [K] activation column */
    const b200_tensor * norm_w;     /* f32 [K]: x := rms_norm(x, eps) * norm_w before quantisation; NULL = none */
    float   eps;
    int32_t n_seg;
    b200_gemv_seg seg[B200_GEMV_MAX_SEG];
    int32_t glu_op;                 /* b200_glu_op; != NONE requires n_seg == 2: dst = glu(seg0) * seg1 -> seg[1].dst */
    float   glu_alpha, glu_limit;
    void *  xq_out;                 /* optional: the q8_1 blocks of the (normalised) activation, K/32 blocks (tests) */
    int32_t n_sm;                   /* CTAs to launch; 0 = one per SM of the current device */
    int32_t flags;                  /* 0. Tuning: bits 4-7 ring depth, bits 8-15 stage KiB, bit 0 bandwidth probe (no dot products) */
} b200_gemv_fused_args;
B200_API int b200_gemv_fused_supported(const b200_gemv_fused_args * a);
B200_API int b200_gemv_fused(const b200_gemv_fused_args * a, void * stream);

typedef struct b200_rope_params {
    int32_t n_dims, mode, n_ctx_orig;
    float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
} b200_rope_params;

/* ---- decode chain: every matmul of a token in ONE cooperative launch -------------------------------
 * A chain is a list of phases run by the persistent streaming-GEMV kernel: GEMV phases (exactly the
 * arguments of b200_gemv_fused) and attention phases (one query position: RoPE(q), RoPE(k), the KV-cache
 * write and flash attention over an f16 cache — reference rope.cu:563, set-rows.cu:318,
 * fattn-vec.cuh:19-507).  Phase i may read anything phase i-1 wrote: phases are separated by in-kernel grid
 * barriers which only the consuming warps wait on, while the producer warp keeps streaming the next
 * phase's weights, so HBM does not drain at matmul boundaries (the reference pays one launch + one
 * quantize launch per matmul).  All pointers are captured at create time. */
enum b200_phase_kind { B200_PHASE_GEMV = 0, B200_PHASE_ATTN = 1,
                       B200_PHASE_NOP = 2 /* nothing to compute (a rank that owns no rows of a matrix); takes part in the barrier */ };
typedef struct b200_attn_phase {
    b200_tensor q, k_new, v_new;     /* f32 raw projections of the current token: [D*H], [D*H_kv], [D*H_kv] */
    b200_tensor k_cache, v_cache;    /* f16 or q8_0 [D, n_kv, H_kv] views of the layer's cache (nb[1] = position stride) */
    const b200_tensor * pos;         /* i32 [1]: position of the token (RoPE) */
    const b200_tensor * slot;        /* i64 [1]: cache row the token is written to */
    const b200_tensor * mask;        /* f16 [n_kv] (0 / -inf) or NULL */
    b200_rope_params rope;           /* modes NORMAL(0) / NEOX(2); YaRN as in ggml_rope_ext (ext_factor, attn_factor, beta_fast / beta_slow) */
    float scale;
    b200_tensor dst;                 /* f32 [D*H] */
    const b200_tensor * slot_v;      /* i64 [1]: cache row of V (NULL: same as slot; ggml passes k_idxs / v_idxs separately) */
    const b200_tensor * sinks;       /* f32 [H] attention sinks (one extra logit per head with no value: gpt-oss), or NULL */
} b200_attn_phase;
typedef struct b200_phase {
    int32_t kind;                    /* b200_phase_kind */
    int32_t _pad;
    b200_gemv_fused_args gemv;
    b200_attn_phase attn;
} b200_phase;
typedef struct b200_chain b200_chain;
B200_API b200_chain * b200_chain_create(const b200_phase * phases, int n_phases, int flags);   /* NULL: see b200_last_error() */
B200_API int          b200_chain_launch(b200_chain * chain, void * stream);
B200_API void         b200_chain_destroy(b200_chain * chain);
/* flags at create time (tuning / debug): bit 0 consumers skip the dot products (bandwidth probe), bit 2 per-phase globaltimer
 * stamps of the first and last CTA (tools/timeline_chain.py), bit 3 publish a phase with __threadfence in every thread before
 * the CTA barrier instead of one release RED by thread 0 after it, bits 4-7 ring depth, bits 8-15 stage KiB, bits 24-27 / 28-30 share
 * withheld from the CTAs that ran an attention phase in the next / next-but-one phase (steps of 5 % / 1 %; 0 = default 20 % / 4 %,
 * bits 24-27 = 0xF = even split) */
B200_API int          b200_chain_debug_read(b200_chain * chain, unsigned long long * out, int n_phases);
/* same flag: %globaltimer of EVERY CTA when it entered a phase and when its rows were consumed: out[n_phases][b200_chain_grid()][2] */
B200_API int          b200_chain_debug_read_ctas(b200_chain * chain, unsigned long long * out, int n_phases);

/* ---- row split across GPUs: the decode chain over NVLink peer memory --------------------------------
 * Replaces the reference's row-split matmul (ggml-cuda.cu:791-804 get_row_split, :1497-1700 ggml_cuda_op_mul_mat:
 * per matmul, the main device quantizes the activation, copies it to every device, each device runs its row slice,
 * and the slices are copied back behind events).  Here one process per GPU runs the SAME chain over its own row
 * slices (each rank passes its slice of every weight, bias and dst; a slice may have 0 rows).  The GEMV epilogue
 * stores every output value to ALL other ranks over NVLink as ONE 64-bit word {f32 bits, tag of (launch, phase)} in a
 * shadow of the activation arena; the consuming phase reads its own rows from plain memory behind the local grid
 * barrier and spins on the tags of the others' rows — one NVLink traversal per dependency, no fence round trip, no
 * counter (the idea of NCCL's LL protocol, inside the matmul).  Outputs nobody in the chain consumes (the logits) go
 * out as plain stores followed by one system-scope fence and an arrival on every rank, which the launch waits for
 * before it ends.  No collective, no extra launch.  Attention phases are computed by every rank (replicated KV cache).
 * Epilogue operands (bias / residual) must be rows the rank computed itself.
 * Why buffers can be reused across layers and launches: a rank enters phase p+1 only after it has read, with matching
 * tags, the rows every CTA of every other rank produced in phase p, so no rank is more than one phase ahead of the
 * slowest CTA anywhere; a buffer and its shadow are rewritten five phases later at the earliest, and the next launch
 * starts only after the final arrival of all ranks.  Tags differ per (launch, phase): a stale word never matches.
 * Every spin is bounded by wall time (20 s) and traps.
 *
 * Requirements: every rank allocates a "symmetric arena" of the same size with the same layout (b200_peer_alloc),
 * maps the arenas of the others (b200_peer_export / b200_peer_open) and passes delta[r] = arena of rank r - own arena
 * (as mapped in this process).  Every GEMV dst and the barrier counters `sync` (>= n_phases u32, zero before the first
 * launch of any rank) must lie inside the arena.  All ranks create the same list of phases (kinds in the same order)
 * and launch the chain the same number of times. */
#define B200_MAX_RANKS 8
typedef struct b200_peer_group {
    int32_t rank, n_ranks;
    int32_t total_ctas;              /* sum over ranks of the SM count (the grid of the chain on that rank) */
    int32_t _pad;
    void *  arena; size_t arena_bytes;
    int64_t delta[B200_MAX_RANKS];   /* bytes; delta[rank] == 0 */
    unsigned int * sync;
    size_t  act_bytes;               /* [arena, arena + act_bytes): activations (every GEMV dst and every x / q / k / v a row-split matmul produced) */
    size_t  ll_offset;               /* [arena + ll_offset, + 2 * act_bytes): shadow of tagged 64-bit words, zero before the first launch */
} b200_peer_group;
B200_API b200_chain * b200_chain_create_peer(const b200_phase * phases, int n_phases, int flags, const b200_peer_group * pg);
B200_API int  b200_chain_grid(void);                                       /* CTAs the chain uses on the current device */
B200_API int  b200_peer_alloc(size_t bytes, void ** ptr);                  /* zero-filled device memory suitable for export */
B200_API int  b200_peer_free(void * ptr);
B200_API int  b200_peer_export(void * ptr, unsigned char handle[64]);      /* handle: send to the other processes */
B200_API int  b200_peer_open(const unsigned char handle[64], void ** ptr); /* maps another rank's arena (enables peer access) */
B200_API int  b200_peer_close(void * ptr);

/* ---- K5: prefill GEMM on the tensor cores (tcgen05 + TMEM) ------------------------------------------
 * dst[M, N] = W[M, K] . x[K, N] for N > 8 columns.  Replaces ggml_cuda_mul_mat_q (reference ggml-cuda/mmq.cu:71-218,
 * mmq.cuh:3459-3698): the raw GGUF blocks travel by TMA, are expanded to f16 in shared memory by dequant warps and
 * multiplied with 16-bit activations by tcgen05.mma (f32 accumulation in TMEM).  Two steps so that matmuls sharing an
 * activation (q/k/v, gate/up) convert it once:
 *   b200_mmq_convert_act : x f32 [K, N] -> pre-swizzled 16-bit tiles in `xb` (b200_mmq_act_size bytes, 1024 B aligned);
 *                          act_fmt 0 = f16 (clamped to +-65504); 1 = bf16 tiles can be produced, but the GEMM only takes 0
 *                          (tcgen05 kind::f16 traps on an f16 A with a bf16 B)
 *   b200_mul_mat_q       : the GEMM; weights Q4_K / Q6_K / Q4_0, K % 256 == 0, 2-D, 16 B aligned rows */
typedef struct b200_mmq_args {
    b200_tensor w;        /* quantized [K, M] */
    b200_tensor x;        /* f32 [K, N] (shape / validation; the kernel reads xb) */
    b200_tensor dst;      /* f32 [M, N] */
    const void * xb;      /* output of b200_mmq_convert_act for x */
    int32_t act_fmt;      /* must match the conversion AND b200_mmq_act_fmt(w.type) */
    int32_t _pad;
    const b200_tensor * addend;   /* optional f32 [M, N] added in the epilogue (the residual stream: replaces the ADD node that
                                     follows the matmul, reference ggml-cuda/binbcast.cu); may alias dst */
} b200_mmq_args;
B200_API int    b200_mmq_act_fmt(int weight_type);   /* 16-bit format of both MMA operands for this weight type: 0 = f16, 1 = bf16 (MXFP4) */
B200_API size_t b200_mmq_act_size(int64_t K, int64_t N);
B200_API int    b200_mmq_supported(const b200_mmq_args * a);
B200_API int    b200_mmq_convert_act(const b200_tensor * x, void * xb, int act_fmt, void * stream);
B200_API int    b200_mul_mat_q(const b200_mmq_args * a, void * stream);

/* ---- validation mode ------------------------------------------------------------------
 * Passing B200_STREAM_VALIDATE as `stream` to any op below makes it check its arguments and
 * return B200_OK / B200_ERR_UNSUPPORTED without touching the GPU.  The plugin's supports_op uses
 * this so "claimed" and "implemented" cannot drift apart. */
#define B200_STREAM_VALIDATE ((void *) (intptr_t) -1)

/* ---- row G: decode glue ops (reference ggml-cuda/{getrows,norm,binbcast,rope,set-rows,cpy,unary,
 * scale,softmax}.cu).  All f32 compute; shapes/strides as in ggml. -------------------------- */

/* GET_ROWS (getrows.cu:238): dst[:, i, b, c] = dequant(src[:, ids[i,b,c], b, c]); src f32/f16/q4_0/q8_0/q4_K/q6_K/mxfp4, dst f32 */
B200_API int b200_get_rows(const b200_tensor * src, const b200_tensor * ids, const b200_tensor * dst, void * stream);

/* RMS_NORM (+MUL (+ADD)) (norm.cu:476 and the fused forms ggml-cuda.cu:3844-3854):
 * dst = x * rsqrt(mean(x^2) + eps) [* mul] [+ add]; mul/add broadcast over rows (may be NULL) */
B200_API int b200_rms_norm(const b200_tensor * x, const b200_tensor * mul, const b200_tensor * add,
                           const b200_tensor * dst, float eps, void * stream);

/* ADD/SUB/MUL/DIV with ggml broadcasting of b over a (binbcast.cu); f32 (dst may be f16 for ADD) */
enum b200_bin_op { B200_BIN_ADD = 0, B200_BIN_SUB = 1, B200_BIN_MUL = 2, B200_BIN_DIV = 3 };
B200_API int b200_bin_bcast(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, void * stream);

/* SCALE (scale.cu): dst = x * scale + bias */
B200_API int b200_scale(const b200_tensor * x, const b200_tensor * dst, float scale, float bias, void * stream);

/* GLU (unary.cu:337): dst = act(gate) * up. `up` NULL => single tensor split in halves along ne0
 * (first half is the activated one unless swapped) */
B200_API int b200_glu(int glu_op, const b200_tensor * gate, const b200_tensor * up, const b200_tensor * dst,
                      int swapped, float alpha, float limit, void * stream);

/* unary ops on contiguous f32 (unary.cu) */
enum b200_unary_op { B200_UNARY_SILU = 0, B200_UNARY_GELU = 1, B200_UNARY_RELU = 2, B200_UNARY_SIGMOID = 3,
                     B200_UNARY_TANH = 4, B200_UNARY_GELU_QUICK = 5, B200_UNARY_NEG = 6, B200_UNARY_EXP = 7 };
B200_API int b200_unary(int op, const b200_tensor * x, const b200_tensor * dst, void * stream);

/* CPY / CONT / DUP (cpy.cu): same element count, arbitrary strides; f32<->f16/bf16, f32->q8_0 (contiguous dst), same-type copies */
B200_API int b200_cpy(const b200_tensor * src, const b200_tensor * dst, void * stream);

/* SET_ROWS (set-rows.cu:318): dst[:, idx[i,b%ne11,c%ne12], b, c] = convert(src[:, i, b, c]);
 * src f32, idx i64/i32, dst f32/f16/bf16/q8_0/q4_0 (q8_0/q4_0 blocks bit-exact vs quantize_row_*_ref) */
B200_API int b200_set_rows(const b200_tensor * src, const b200_tensor * idx, const b200_tensor * dst, void * stream);

/* ROPE (rope.cu:563): modes NORMAL(0) and NEOX(2); YaRN parameters as in ggml_rope_ext */
/* b200_rope_params: declared above (decode chain section) */
/* x f32 [head_dim, n_head, n_tok, 1]; pos i32 [n_tok]; freq_factors f32 [n_dims/2] or NULL.
 * Optional fused SET_ROWS (ggml-cuda.cu:3614-3621): when set_idx != NULL the rotated rows are written
 * to set_dst (f16/f32 rows of head_dim*n_head elements, row index from set_idx) instead of dst. */
B200_API int b200_rope(const b200_tensor * x, const b200_tensor * pos, const b200_tensor * freq_factors,
                       const b200_tensor * dst, const b200_rope_params * p,
                       const b200_tensor * set_idx, const b200_tensor * set_dst, void * stream);

/* SOFT_MAX (softmax.cu): dst = softmax(x*scale + slope*mask [, sinks]) over ne0; mask f16/f32 broadcast */
B200_API int b200_soft_max(const b200_tensor * x, const b200_tensor * mask, const b200_tensor * sinks,
                           const b200_tensor * dst, float scale, float max_bias, void * stream);

/* ---- MoE routing glue (reference ggml-cuda/argsort.cu, top-k.cu, add-id.cu:28, sumrows.cu, clamp.cu) and the float-weight
 * matmul of the router matrix (ggml-cuda.cu:2183-2266 / mmvf.cu).  Graph: reference src/llama-graph.cpp build_moe_ffn. */
/* ARGSORT: dst i32 [ne0, rows] = indices that sort each f32 row (order 0 ascending, 1 descending); ne0 <= 32768 */
B200_API int b200_argsort(const b200_tensor * x, const b200_tensor * dst, int order, void * stream);
/* TOP_K: dst i32 [k, rows] = indices of the k largest values of each row (descending) */
B200_API int b200_top_k(const b200_tensor * x, const b200_tensor * dst, void * stream);
/* fused routing (reference ggml-cuda/topk-moe.cu:314, softmax over the selected logits): logits f32 [n_expert <= 512, n_tok];
 * ids i32 [n_used, n_tok] (row stride free: the first n_used slots of an ARGSORT row) = the n_used largest logits in
 * descending order (lowest index first on ties); weights f32 [n_used, n_tok] = softmax of those logits */
B200_API int b200_topk_moe(const b200_tensor * logits, const b200_tensor * ids, const b200_tensor * weights, void * stream);
/* weighted sum over the selected experts (the MUL + VIEWs + ADDs at the end of build_moe_ffn): x f32 [n, n_used, n_tok],
 * w f32 [1, n_used, n_tok], dst f32 [n, n_tok] = sum_e x[:, e, t] * w[e, t] (products rounded, added left to right)
 * [+ residual f32 [n, n_tok], or NULL: the ADD of the residual stream that follows] */
B200_API int b200_moe_weighted_sum(const b200_tensor * x, const b200_tensor * w, const b200_tensor * residual, const b200_tensor * dst, void * stream);
/* ADD_ID: dst[:, e, t] = a[:, e, t] + b[:, ids[e, t]]; a f32 [n, n_used, n_tok], b f32 [n, n_expert], ids i32 [n_used, n_tok] (strided) */
B200_API int b200_add_id(const b200_tensor * a, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * dst, void * stream);
/* SUM_ROWS: dst[0, i1, i2, i3] = sum over ne0; CLAMP: dst = min(max(x, lo), hi) (contiguous) */
B200_API int b200_sum_rows(const b200_tensor * x, const b200_tensor * dst, void * stream);
B200_API int b200_clamp(const b200_tensor * x, const b200_tensor * dst, float lo, float hi, void * stream);
/* MUL_MAT with f32 / f16 / bf16 weights [K, M] (2-D) and f32 activations [K, N]: dst f32 [M, N] (+ bias f32 [M] per row, or NULL) */
B200_API int b200_mul_mat_f(const b200_tensor * w, const b200_tensor * x, const b200_tensor * bias, const b200_tensor * dst, void * stream);

/* ---- K7: FLASH_ATTN_EXT decode (fattn-vec.cuh:19-507, fattn-common.cuh:728-779, launch :781-1036)
 * q f32 [D, n_q, H, n_seq]; k,v f16 or q8_0 [D, n_kv, H_kv, n_seq] (row stride = cache stride);
 * mask f16 [n_kv, n_q_pad, 1|.., 1|n_seq] or NULL; sinks f32 [H] or NULL; dst f32 [D, H, n_q, n_seq].
 * Split-KV across SMs with an in-kernel last-block combine; scratch must hold b200_fattn_scratch_size(). */
typedef struct b200_fattn_args {
    b200_tensor q, k, v, dst;
    const b200_tensor * mask;
    const b200_tensor * sinks;
    float scale, max_bias, logit_softcap;
    void * scratch; size_t scratch_size;
} b200_fattn_args;
B200_API size_t b200_fattn_scratch_size(const b200_fattn_args * a);
B200_API int    b200_flash_attn_ext(const b200_fattn_args * a, void * stream);

/* FLASH_ATTN_EXT for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Same arguments, kernels and results as b200_flash_attn_ext, but a->scratch / a->scratch_size are IGNORED: the split-KV
 * scratch is a file-scope `static __device__` array of B200_FATTN_STATIC_SCRATCH_BYTES inside the library (every GPU that
 * loads the library holds it until exit).  The call performs no cudaMalloc / cudaFree / allocating library routine.
 * Shapes whose b200_fattn_scratch_size() exceeds the array are rejected with B200_ERR_ARG (b200_last_error() names both
 * sizes).  All calls on a device share the one array: calls that overlap on the GPU (different streams) race — serialise
 * them on one stream.  The array keeps the previous call's contents; the arrival counters are cleared on `stream` by every
 * call that uses them. */
#define B200_FATTN_STATIC_SCRATCH_BYTES (8u << 20)
B200_API int    b200_flash_attn_ext_static(const b200_fattn_args * a, void * stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_KERNELS_H */
