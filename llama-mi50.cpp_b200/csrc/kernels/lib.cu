// lib.cu — library-level entry points: error string, launch counter, type tables.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char * fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add((uint64_t) n, std::memory_order_relaxed); }
int check_launch(const char * what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: launch failed: %s", what, cudaGetErrorString(e));
        return B200_ERR_CUDA;
    }
    return B200_OK;
}

}  // namespace b200

extern "C" {

const char * b200_last_error(void) { return b200::g_err; }

int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

uint64_t b200_launch_count(void) { return b200::g_launches.load(std::memory_order_relaxed); }

size_t b200_type_size(int type) {
    switch (type) {
        case B200_F32: case B200_I32: return 4;
        case B200_F16: case B200_BF16: return 2;
        case B200_I64: return 8;
        case B200_Q4_0: return 18;
        case B200_Q8_0: return 34;
        case B200_Q8_1: return 36;
        case B200_MXFP4: return 17;
        case B200_Q4_K: return 144;
        case B200_Q6_K: return 210;
        default: return 0;
    }
}

int64_t b200_blck_size(int type) {
    switch (type) {
        case B200_F32: case B200_I32: case B200_F16: case B200_BF16: case B200_I64: return 1;
        case B200_Q4_0: case B200_Q8_0: case B200_Q8_1: case B200_MXFP4: return 32;
        case B200_Q4_K: case B200_Q6_K: return 256;
        default: return 0;
    }
}

size_t b200_row_size(int type, int64_t ne0) {
    const int64_t b = b200_blck_size(type);
    return b ? (size_t) (ne0 / b) * b200_type_size(type) : 0;
}

}  // extern "C"
