// explicit instantiations of the streaming GEMV / decode-chain kernel for the weight-type pair (mxfp4, mxfp4)
#include "gemv_stream.cuh"

namespace b200 {
template int gs_launch<39, 39, 0>(const gs_params &, size_t, int, cudaStream_t);
template int gs_launch<39, 39, 1>(const gs_params &, size_t, int, cudaStream_t);
template int gs_launch<39, 39, 2>(const gs_params &, size_t, int, cudaStream_t);
}  // namespace b200
