// split_buffer.cpp — row-split (tensor-parallel) buffer type: SURVEY section 8(e).
// Implements the interface of ggml_backend_cuda_split_buffer_type (reference ggml-cuda/ggml-cuda.cu:776-1107) for the
// B200 devices of this process: libllama asks the registry for "ggml_backend_split_buffer_type" under --split-mode row
// (src/llama-model.cpp:391-409) and places every 2-D matmul weight there.
//
// A tensor in a split buffer has no single address: `tensor->data` is an offset from a dummy base that is never
// dereferenced, and `tensor->extra` points to a b200_split_extra with one allocation per device that holds the rows
// [row_lo, row_hi) of that device, in the unchanged GGUF wire format.  Same partition rule as the reference
// (get_row_split, ggml-cuda.cu:791-804): cumulative fractions, boundaries rounded DOWN to 128 rows, the last device takes
// the remainder, so a device may own no rows of a small matrix.  Execution: exec.cpp (run_mul_mat_split).
#include "b200_plugin.h"

#include <algorithm>
#include <array>
#include <cstring>
#include <map>

namespace {

constexpr int64_t kRowRounding = 128;
constexpr size_t  kTailSlack = 256;       // like buffer.cpp: a 16 B vector load of the last row stays inside the allocation

struct split_buft_ctx {
    int main_device;
    std::array<float, B200_MAX_DEVICES> split;     // cumulative start fraction per device
    std::string name;
};
struct split_buf_ctx {
    std::vector<b200_split_extra *> extras;
};

const char * split_buft_name(ggml_backend_buffer_type_t t) { return ((split_buft_ctx *) t->context)->name.c_str(); }

// rows [lo, hi) of device `id` out of n devices with cumulative start fractions `split` (reference get_row_split,
// ggml-cuda.cu:791-804: boundaries rounded down to the row rounding, the last device takes the remainder)
void row_range_n(const float * split, int n, int64_t nrows, int id, int64_t & lo, int64_t & hi) {
    lo = id == 0 ? 0 : (int64_t) (nrows * split[id]);
    lo -= lo % kRowRounding;
    if (id == n - 1) hi = nrows;
    else { hi = (int64_t) (nrows * split[id + 1]); hi -= hi % kRowRounding; }
    if (hi < lo) hi = lo;
}
void row_range(const split_buft_ctx * c, int64_t nrows, int id, int64_t & lo, int64_t & hi) {
    row_range_n(c->split.data(), b200_n_devices(), nrows, id, lo, hi);
}
// user-facing tensor_split (relative sizes, all zero = default) -> cumulative start fractions (reference ggml-cuda.cu:1078-1090)
void cumulative_split(const float * tensor_split, int n, const double * default_weight, float * out) {
    bool all_zero = true;
    for (int i = 0; tensor_split && i < n; ++i) all_zero = all_zero && tensor_split[i] == 0.0f;
    if (all_zero) {
        double total = 0, acc = 0;
        for (int i = 0; i < n; ++i) total += default_weight[i];
        for (int i = 0; i < n; ++i) { out[i] = (float) (acc / total); acc += default_weight[i]; }
    } else {
        float sum = 0.0f;
        for (int i = 0; i < n; ++i) { out[i] = sum; sum += tensor_split[i]; }
        for (int i = 0; i < n; ++i) out[i] /= sum;
    }
}

size_t slice_bytes(const ggml_tensor * t, int64_t rows) {
    size_t n = (size_t) rows * ggml_row_size(t->type, t->ne[0]);
    if (t->ne[0] % B200_MATRIX_ROW_PADDING != 0)      // last row padded to 512 elements, as the reference does (ggml-cuda.cu:863-896)
        n += ggml_row_size(t->type, B200_MATRIX_ROW_PADDING - t->ne[0] % B200_MATRIX_ROW_PADDING);
    return n;
}

// ---------------------------------------------------------------- buffer
void split_buf_free(ggml_backend_buffer_t b) {
    split_buf_ctx * c = (split_buf_ctx *) b->context;
    for (b200_split_extra * e : c->extras) {
        for (int id = 0; id < B200_MAX_DEVICES; ++id)
            if (e->data[id]) { b200_set_device(id); cudaFree(e->data[id]); }
        delete e;
    }
    delete c;
}
void * split_buf_base(ggml_backend_buffer_t) { return (void *) 0x1000; }      // never dereferenced

ggml_status split_buf_init_tensor(ggml_backend_buffer_t b, ggml_tensor * t) {
    GGML_ASSERT(t->view_src == nullptr);                // views of split tensors are not supported (reference :841)
    GGML_ASSERT(ggml_is_contiguous(t) && "split buffers only hold contiguous tensors");
    split_buf_ctx * c = (split_buf_ctx *) b->context;
    const split_buft_ctx * bt = (const split_buft_ctx *) b->buft->context;
    b200_split_extra * e = new b200_split_extra{};
    c->extras.push_back(e);
    const int64_t nrows = ggml_nrows(t);
    for (int id = 0; id < b200_n_devices(); ++id) {
        row_range(bt, nrows, id, e->row_lo[id], e->row_hi[id]);
        const int64_t rows = e->row_hi[id] - e->row_lo[id];
        if (rows == 0) continue;
        const size_t bytes = slice_bytes(t, rows) + kTailSlack;
        b200_set_device(id);
        B200_CUDA_CHECK(cudaMalloc(&e->data[id], bytes));
        B200_CUDA_CHECK(cudaMemset(e->data[id], 0, bytes));       // the padding must not hold NaN bit patterns
    }
    B200_CUDA_CHECK(cudaDeviceSynchronize());
    t->extra = e;
    return GGML_STATUS_SUCCESS;
}

void split_buf_set_tensor(ggml_backend_buffer_t, ggml_tensor * t, const void * data, size_t off, size_t size) {
    GGML_ASSERT(off == 0 && size == ggml_nbytes(t) && "split tensors are written whole");       // reference :901-903
    const b200_split_extra * e = (const b200_split_extra *) t->extra;
    const size_t nb1 = ggml_row_size(t->type, t->ne[0]);
    for (int id = 0; id < b200_n_devices(); ++id) {
        const int64_t rows = e->row_hi[id] - e->row_lo[id];
        if (rows == 0) continue;
        b200_set_device(id);
        B200_CUDA_CHECK(cudaMemcpy(e->data[id], (const char *) data + (size_t) e->row_lo[id] * nb1, (size_t) rows * nb1, cudaMemcpyHostToDevice));
    }
}
void split_buf_get_tensor(ggml_backend_buffer_t, const ggml_tensor * t, void * data, size_t off, size_t size) {
    GGML_ASSERT(off == 0 && size == ggml_nbytes(t) && "split tensors are read whole");
    const b200_split_extra * e = (const b200_split_extra *) t->extra;
    const size_t nb1 = ggml_row_size(t->type, t->ne[0]);
    for (int id = 0; id < b200_n_devices(); ++id) {
        const int64_t rows = e->row_hi[id] - e->row_lo[id];
        if (rows == 0) continue;
        b200_set_device(id);
        B200_CUDA_CHECK(cudaMemcpy((char *) data + (size_t) e->row_lo[id] * nb1, e->data[id], (size_t) rows * nb1, cudaMemcpyDeviceToHost));
    }
}
void split_buf_clear(ggml_backend_buffer_t, uint8_t) {}

const ggml_backend_buffer_i kSplitBufIface = {
    /* free_buffer   */ split_buf_free,
    /* get_base      */ split_buf_base,
    /* init_tensor   */ split_buf_init_tensor,
    /* memset_tensor */ nullptr,
    /* set_tensor    */ split_buf_set_tensor,
    /* get_tensor    */ split_buf_get_tensor,
    /* cpy_tensor    */ nullptr,
    /* clear         */ split_buf_clear,
    /* reset         */ nullptr,
};

// ---------------------------------------------------------------- buffer type
ggml_backend_buffer_t split_buft_alloc(ggml_backend_buffer_type_t t, size_t size) {
    // the per-device slices are allocated tensor by tensor in init_tensor; `size` only sums get_alloc_size
    return ggml_backend_buffer_init(t, kSplitBufIface, new split_buf_ctx{}, size);
}
size_t split_buft_alignment(ggml_backend_buffer_type_t) { return 128; }
size_t split_buft_alloc_size(ggml_backend_buffer_type_t t, const ggml_tensor * tensor) {
    const split_buft_ctx * c = (const split_buft_ctx *) t->context;
    GGML_ASSERT(ggml_is_contiguous(tensor) && "split buffers only hold contiguous tensors");
    size_t total = 0;
    const int64_t nrows = ggml_nrows(tensor);
    for (int id = 0; id < b200_n_devices(); ++id) {
        int64_t lo, hi; row_range(c, nrows, id, lo, hi);
        if (hi > lo) total += slice_bytes(tensor, hi - lo);
    }
    return total;
}
bool split_buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i kSplitBuftIface = {
    /* get_name       */ split_buft_name,
    /* alloc_buffer   */ split_buft_alloc,
    /* get_alignment  */ split_buft_alignment,
    /* get_max_size   */ nullptr,
    /* get_alloc_size */ split_buft_alloc_size,
    /* is_host        */ split_buft_is_host,
};

}  // namespace

ggml_backend_dev_t b200_dev(int device);   // device.cpp

bool b200_buft_is_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == split_buft_name; }
void b200_split_rows(ggml_backend_buffer_type_t split_buft, int64_t nrows, int device, int64_t & lo, int64_t & hi) {
    row_range((const split_buft_ctx *) split_buft->context, nrows, device, lo, hi);
}

// every device can read and write every other device's memory (kernels of a slice store straight into the main device's dst)
void b200_enable_peer_access() {
    static std::once_flag once;
    std::call_once(once, [] {
        const int n = b200_n_devices();
        for (int a = 0; a < n; ++a) {
            b200_set_device(a);
            for (int b = 0; b < n; ++b) {
                if (a == b) continue;
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, b200_device(a).id, b200_device(b).id) != cudaSuccess || !can) {
                    cudaGetLastError();
                    GGML_LOG_ERROR("ggml-b200: no peer access from device %d to device %d: row split needs NVLink/PCIe peer access\n", a, b);
                    GGML_ABORT("ggml-b200: peer access unavailable");
                }
                const cudaError_t e = cudaDeviceEnablePeerAccess(b200_device(b).id, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) B200_CUDA_CHECK(e);
                cudaGetLastError();
            }
        }
    });
}

extern "C" __attribute__((visibility("default")))
ggml_backend_buffer_type_t ggml_backend_b200_split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mu;
    static std::map<std::pair<int, std::array<float, B200_MAX_DEVICES>>, ggml_backend_buffer_type> types;
    std::lock_guard<std::mutex> lock(mu);
    const int n = b200_n_devices();
    if (n < 1 || main_device < 0 || main_device >= n) return nullptr;
    std::array<float, B200_MAX_DEVICES> split{};
    double mem[B200_MAX_DEVICES];      // default: proportional to device memory (reference default_tensor_split, ggml-cuda.cu:231-232, 304-306)
    for (int i = 0; i < n; ++i) mem[i] = (double) b200_device(i).total_mem;
    cumulative_split(tensor_split, n, mem, split.data());
    auto key = std::make_pair(main_device, split);
    auto it = types.find(key);
    if (it != types.end()) return &it->second;
    if (n > 1) b200_enable_peer_access();
    ggml_backend_buffer_type t{ kSplitBuftIface, b200_dev(main_device), new split_buft_ctx{main_device, split, b200_device(main_device).name + "_Split"} };
    return &types.emplace(key, t).first->second;
}

// The partition arithmetic alone (no device needed): rows [*row_lo, *row_hi) of `device` for a tensor of `nrows` rows
// split over `n_devices` equal devices with the user's `tensor_split` (NULL / all zero: equal shares).  Returns 0, or -1
// on bad arguments.  Used by the CPU tests and by hosts that want to place data before creating buffers.
extern "C" __attribute__((visibility("default")))
int ggml_backend_b200_split_rows(const float * tensor_split, int n_devices, long long nrows, int device, long long * row_lo, long long * row_hi) {
    if (n_devices < 1 || n_devices > B200_MAX_DEVICES || device < 0 || device >= n_devices || nrows < 0 || !row_lo || !row_hi) return -1;
    float split[B200_MAX_DEVICES];
    double w[B200_MAX_DEVICES];
    for (int i = 0; i < n_devices; ++i) w[i] = 1.0;
    cumulative_split(tensor_split, n_devices, w, split);
    int64_t lo, hi;
    row_range_n(split, n_devices, nrows, device, lo, hi);
    *row_lo = lo; *row_hi = hi;
    return 0;
}
