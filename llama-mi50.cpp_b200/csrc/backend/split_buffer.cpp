// split_buffer.cpp — row-split (tensor-parallel) buffer type: SURVEY section 8(e).
// Replaces ggml_backend_cuda_split_buffer_type (reference ggml-cuda/ggml-cuda.cu:776-1107).
#include "b200_plugin.h"

bool b200_buft_is_split(ggml_backend_buffer_type_t) { return false; }

extern "C" __attribute__((visibility("default")))
ggml_backend_buffer_type_t ggml_backend_b200_split_buffer_type(int, const float *) {
    return nullptr;   // not built yet: libllama then uses the per-device buffer type (src/llama-model.cpp:397-409)
}
