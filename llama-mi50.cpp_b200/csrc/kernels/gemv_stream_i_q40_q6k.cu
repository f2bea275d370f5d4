// explicit instantiations of the streaming GEMV / decode-chain kernel for the weight-type pair (q40, q6k)
#include "gemv_stream.cuh"

namespace b200 {
template int gs_launch<2, 14, 0>(const gs_params &, size_t, int, cudaStream_t);
template int gs_launch<2, 14, 1>(const gs_params &, size_t, int, cudaStream_t);
template int gs_launch<2, 14, 2>(const gs_params &, size_t, int, cudaStream_t);
}  // namespace b200
