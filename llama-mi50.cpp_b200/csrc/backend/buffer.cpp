// buffer.cpp — device buffer type and buffer of the B200 backend.
// Implements ggml_backend_buffer_type_i / ggml_backend_buffer_i (reference interface:
// ggml/src/ggml-backend-impl.h:17-72; the reference's CUDA implementation of the same interface is
// ggml-cuda/ggml-cuda.cu:569-774).  Weights are stored exactly as set_tensor delivers them
// (GGUF wire format); the kernels consume those bytes directly, so get_tensor(set_tensor(x)) == x.
#include "b200_plugin.h"

#include <cstring>

namespace {

struct buft_ctx { int device; std::string name; };
struct buf_ctx  { int device; void * base; size_t size; };

constexpr size_t kAlign = 256;       // tensor alignment inside a buffer (>= 16 B for 128-bit / bulk loads)
constexpr size_t kTailSlack = 256;   // allocation slack so a 16 B vector load of the last row never leaves the allocation

// ---------------------------------------------------------------- buffer
void buf_free(ggml_backend_buffer_t b) {
    buf_ctx * c = (buf_ctx *) b->context;
    b200_set_device(c->device);
    cudaFree(c->base);
    delete c;
}
void * buf_get_base(ggml_backend_buffer_t b) { return ((buf_ctx *) b->context)->base; }

ggml_status buf_init_tensor(ggml_backend_buffer_t, ggml_tensor *) { return GGML_STATUS_SUCCESS; }

void buf_memset_tensor(ggml_backend_buffer_t b, ggml_tensor * t, uint8_t v, size_t off, size_t size) {
    b200_set_device(((buf_ctx *) b->context)->device);
    B200_CUDA_CHECK(cudaMemset((char *) t->data + off, v, size));
    B200_CUDA_CHECK(cudaStreamSynchronize(0));
}
void buf_set_tensor(ggml_backend_buffer_t b, ggml_tensor * t, const void * data, size_t off, size_t size) {
    b200_set_device(((buf_ctx *) b->context)->device);
    B200_CUDA_CHECK(cudaMemcpy((char *) t->data + off, data, size, cudaMemcpyHostToDevice));
}
void buf_get_tensor(ggml_backend_buffer_t b, const ggml_tensor * t, void * data, size_t off, size_t size) {
    b200_set_device(((buf_ctx *) b->context)->device);
    B200_CUDA_CHECK(cudaMemcpy(data, (const char *) t->data + off, size, cudaMemcpyDeviceToHost));
}
bool buf_cpy_tensor(ggml_backend_buffer_t b, const ggml_tensor * src, ggml_tensor * dst) {
    ggml_backend_buffer_t sb = src->view_src ? src->view_src->buffer : src->buffer;
    if (!sb || !b200_buft_is_b200(sb->buft) || b200_buft_is_split(sb->buft)) return false;
    if (!ggml_is_contiguous(src) || !ggml_is_contiguous(dst) || ggml_nbytes(src) != ggml_nbytes(dst)) return false;
    const int sd = ((buf_ctx *) sb->context)->device, dd = ((buf_ctx *) b->context)->device;
    b200_set_device(dd);
    // device-to-device copies return before they are done and the legacy stream they would run on does not order with
    // the backend's non-blocking compute streams: copy on the per-thread stream and wait for it (found on a 2-GPU box:
    // layer split fed the second device the previous token's activations once graph replay made the launch fast enough)
    if (sd == dd) B200_CUDA_CHECK(cudaMemcpyAsync(dst->data, src->data, ggml_nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    else          B200_CUDA_CHECK(cudaMemcpyPeerAsync(dst->data, b200_device(dd).id, src->data, b200_device(sd).id, ggml_nbytes(src), cudaStreamPerThread));
    B200_CUDA_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}
void buf_clear(ggml_backend_buffer_t b, uint8_t v) {
    buf_ctx * c = (buf_ctx *) b->context;
    b200_set_device(c->device);
    B200_CUDA_CHECK(cudaMemset(c->base, v, c->size));
    B200_CUDA_CHECK(cudaStreamSynchronize(0));
}

const ggml_backend_buffer_i kBufIface = {
    /* free_buffer   */ buf_free,
    /* get_base      */ buf_get_base,
    /* init_tensor   */ buf_init_tensor,
    /* memset_tensor */ buf_memset_tensor,
    /* set_tensor    */ buf_set_tensor,
    /* get_tensor    */ buf_get_tensor,
    /* cpy_tensor    */ buf_cpy_tensor,
    /* clear         */ buf_clear,
    /* reset         */ nullptr,
};

// ---------------------------------------------------------------- buffer type
const char * buft_name(ggml_backend_buffer_type_t t) { return ((buft_ctx *) t->context)->name.c_str(); }

ggml_backend_buffer_t buft_alloc(ggml_backend_buffer_type_t t, size_t size) {
    buft_ctx * c = (buft_ctx *) t->context;
    b200_set_device(c->device);
    void * p = nullptr;
    const size_t n = (size ? size : 1) + kTailSlack;
    cudaError_t e = cudaMalloc(&p, n);
    if (e != cudaSuccess) {   // OOM convention: clear the error, return NULL (reference ggml-cuda.cu:706-712)
        cudaGetLastError();
        GGML_LOG_ERROR("%s: allocating %.2f MiB on device %d failed: %s\n", __func__, n / 1048576.0, c->device, cudaGetErrorString(e));
        return nullptr;
    }
    return ggml_backend_buffer_init(t, kBufIface, new buf_ctx{c->device, p, size}, size);
}
size_t buft_alignment(ggml_backend_buffer_type_t) { return kAlign; }
size_t buft_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * t) { return ggml_nbytes(t); }
bool   buft_is_host(ggml_backend_buffer_type_t) { return false; }

const ggml_backend_buffer_type_i kBuftIface = {
    /* get_name       */ buft_name,
    /* alloc_buffer   */ buft_alloc,
    /* get_alignment  */ buft_alignment,
    /* get_max_size   */ nullptr,
    /* get_alloc_size */ buft_alloc_size,
    /* is_host        */ buft_is_host,
};

}  // namespace

ggml_backend_dev_t b200_dev(int device);   // device.cpp

ggml_backend_buffer_type_t b200_buffer_type(int device) {
    static std::mutex mu;
    static ggml_backend_buffer_type types[B200_MAX_DEVICES];
    static bool init[B200_MAX_DEVICES] = {};
    std::lock_guard<std::mutex> lock(mu);
    if (device < 0 || device >= b200_n_devices()) return nullptr;
    if (!init[device]) {
        types[device].iface = kBuftIface;
        types[device].device = b200_dev(device);
        types[device].context = new buft_ctx{device, b200_device(device).name};
        init[device] = true;
    }
    return &types[device];
}

bool b200_buft_is_split(ggml_backend_buffer_type_t buft);   // split_buffer.cpp

bool b200_buft_is_b200(ggml_backend_buffer_type_t buft) {
    return buft && (buft->iface.get_name == buft_name || b200_buft_is_split(buft));
}
int b200_buft_device(ggml_backend_buffer_type_t buft) {
    return buft && buft->iface.get_name == buft_name ? ((buft_ctx *) buft->context)->device : -1;
}

// ---------------------------------------------------------------- pinned host buffer type
// Replaces ggml_backend_cuda_host_buffer_type (reference ggml-cuda.cu:1111-1176): libllama places the logits output buffer
// and CPU-side tensors here, so the per-token logits D2H and the input H2D copies run at PCIe speed instead of being staged
// through pageable memory.  The buffer behaves like a CPU buffer (ggml-base's ggml_backend_cpu_buffer_from_ptr) whose
// memory is cudaMallocHost'ed; allocation failure falls back to the caller's default (nullptr).
namespace {
const char * host_buft_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
void host_buffer_free(ggml_backend_buffer_t buffer) { cudaFreeHost(buffer->context); }
ggml_backend_buffer_t host_buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    if (getenv("GGML_B200_NO_PINNED")) return nullptr;
    void * ptr = nullptr;
    if (cudaMallocHost(&ptr, size ? size : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    ggml_backend_buffer_t buffer = ggml_backend_cpu_buffer_from_ptr(ptr, size);
    buffer->buft = buft;
    buffer->iface.free_buffer = host_buffer_free;
    return buffer;
}
size_t host_buft_alignment(ggml_backend_buffer_type_t) { return 64; }
bool host_buft_is_host(ggml_backend_buffer_type_t) { return true; }
}  // namespace

ggml_backend_buffer_type_t b200_host_buffer_type() {
    static ggml_backend_buffer_type type = {
        /* iface   */ { host_buft_name, host_buft_alloc, host_buft_alignment, nullptr, nullptr, host_buft_is_host },
        /* device  */ nullptr,
        /* context */ nullptr,
    };
    static std::once_flag once;
    std::call_once(once, [] { type.device = b200_dev(0); });
    return &type;
}
