// b200_plugin.h — internal declarations of the ggml backend plugin (libggml-b200.so).
// Host side in C++ (the reference's language); compute goes through include/b200_kernels.h only.
#pragma once

#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <string>
#include <vector>

#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-impl.h"
#include "ggml-backend-impl.h"

#include "b200_backend.h"
#include "b200_kernels.h"

#define B200_MAX_DEVICES 16
#define B200_MATRIX_ROW_PADDING 512   // q8_1 activation rows are padded to this many elements (reference MATRIX_ROW_PADDING)

[[noreturn]] void b200_abort(const char * file, int line, const char * what, const char * detail);

#define B200_CUDA_CHECK(expr)                                                              \
    do {                                                                                   \
        cudaError_t err_ = (expr);                                                         \
        if (err_ != cudaSuccess) b200_abort(__FILE__, __LINE__, #expr, cudaGetErrorString(err_)); \
    } while (0)

#define B200_KCHECK(expr)                                                                  \
    do {                                                                                   \
        int rc_ = (expr);                                                                  \
        if (rc_ != B200_OK) b200_abort(__FILE__, __LINE__, #expr, b200_last_error());      \
    } while (0)

struct b200_device_info {
    int         id;
    std::string name;         // "B200_0"
    std::string description;  // cudaDeviceProp.name
    std::string pci_bus_id;
    size_t      total_mem;
    int         cc_major, cc_minor, n_sm;
};

int                      b200_n_devices();
const b200_device_info & b200_device(int id);
void                     b200_set_device(int id);

// ---- buffer types (buffer.cpp)
ggml_backend_buffer_type_t b200_buffer_type(int device);
ggml_backend_buffer_type_t b200_host_buffer_type();   // pinned host memory (cudaMallocHost)
bool                       b200_buft_is_b200(ggml_backend_buffer_type_t buft);
bool                       b200_buft_is_split(ggml_backend_buffer_type_t buft);
int                        b200_buft_device(ggml_backend_buffer_type_t buft);

// ---- row-split tensors (split_buffer.cpp): tensor->extra of a tensor in a split buffer
struct b200_split_extra {
    void *  data[B200_MAX_DEVICES];                            // rows [row_lo, row_hi) of the tensor on each device (null: no rows)
    int64_t row_lo[B200_MAX_DEVICES], row_hi[B200_MAX_DEVICES];
};
void b200_split_rows(ggml_backend_buffer_type_t split_buft, int64_t nrows, int device, int64_t & lo, int64_t & hi);
void b200_enable_peer_access();
// what a backend context keeps on each OTHER device to run its slice of a row-split matmul
struct b200_split_dev {
    cudaStream_t stream = nullptr; cudaEvent_t done = nullptr;
    char * x = nullptr;  size_t x_size = 0;                    // copy of the f32 activations
    char * q8 = nullptr; size_t q8_size = 0;                   // their q8_1 form (GEMV)
    char * xb = nullptr; size_t xb_size = 0;                   // their 16-bit tiles (tensor-core GEMM)
};

// ---- backend context (exec.cpp)
struct b200_graph_cache;
struct b200_context {
    int          device = 0;
    cudaStream_t stream = nullptr;
    std::string  name;
    // scratch arenas (stable addresses across CUDA-graph replays; growing one drops the graph cache)
    char * q8_scratch = nullptr; size_t q8_size = 0;      // q8_1 activations
    char * fa_scratch = nullptr; size_t fa_size = 0;      // flash-attn counters + partials
    char * xb_scratch = nullptr; size_t xb_size = 0;      // 16-bit activation tiles of the tensor-core prefill GEMM
    b200_graph_cache * cache = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    bool timing = false, ev_pending = false;
    b200_split_dev split[B200_MAX_DEVICES];               // row split: streams / scratch on the other devices
    cudaEvent_t    split_fork = nullptr;
};

ggml_status b200_graph_compute(b200_context * ctx, ggml_cgraph * cgraph);
bool        b200_supports_op(int device, const ggml_tensor * op);
void        b200_context_free(b200_context * ctx);
void        b200_timing_collect(b200_context * ctx);

// ---- decode-chain plan (chain_plan.cpp): the graph as a list of chain phases and plain nodes
enum b200_item_kind { B200_ITEM_NODE = 0, B200_ITEM_PHASE = 1 };
struct b200_plan_item {
    int        kind = B200_ITEM_NODE;
    int        first_node = 0;      // index into the graph
    int        n_nodes = 0;
    bool       scratch_rel = false; // phase tensors whose data pointer is an offset into the chain's private scratch
    bool       hoisted = false;     // plain node moved out of an attention block: runs alone, out of index order
    b200_phase phase;
};
struct b200_plan {
    std::vector<b200_plan_item> items;
    std::deque<b200_tensor>     pool;            // descriptors the phases point to (norm weights, biases, positions, masks)
    int64_t                     scratch_floats = 0;  // per-chain scratch of the attention blocks (raw q/k/v of the current layer)
    // Phase outputs whose ggml buffer cannot be used (ggml-alloc gave them the memory of a tensor the phase still reads)
    // and whose consumers are all chain phases live in a plan-private buffer instead.  The planner hands out addresses
    // from a virtual range; the executor maps them onto the real allocation when it builds the chains.
    int64_t                     redirect_bytes = 0;
};
#define B200_REDIRECT_VBASE ((uintptr_t) 0x7e0000000000ull)
static inline bool b200_is_redirect(const void * p) { return (uintptr_t) p >= B200_REDIRECT_VBASE && (uintptr_t) p < B200_REDIRECT_VBASE + ((uintptr_t) 1 << 32); }
void b200_plan_build(ggml_cgraph * g, b200_plan & plan);
bool b200_gemv_fused_supported_rel(const b200_gemv_fused_args & a);

// ---- tensor translation
static inline b200_tensor b200_desc(const ggml_tensor * t) {
    b200_tensor d;
    d.data = t->data; d.type = (int32_t) t->type; d._pad = 0;
    for (int i = 0; i < 4; ++i) { d.ne[i] = t->ne[i]; d.nb[i] = (int64_t) t->nb[i]; }
    return d;
}

// ---- stats (device.cpp)
struct b200_stats_t {
    std::mutex mu;
    ggml_b200_stats s{};
};
b200_stats_t & b200_stats();
