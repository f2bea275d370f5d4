// mmvq_stream.cu — streaming fast path of K4 for contiguous decode weights.
#include "vecdot.cuh"

namespace b200 {

int mmvq_stream_try(const b200_mmvq_args * a, void * stream) {
    (void) a; (void) stream;
    return B200_ERR_UNSUPPORTED;
}

}  // namespace b200
