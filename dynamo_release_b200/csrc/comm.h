// internal: communicator handle shared between comm.cu and the EM loop
#pragma once
#include <cstdint>
#include <cstring>
#include <cuda_runtime.h>

namespace dvf {
struct dvf_nccl_uid { char internal[128]; };   // layout of ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)
struct comm_t {
    void* nccl = nullptr;
    int rank = 0, world = 1;
};
// in-place allreduce of `count` doubles (SUM, or MIN if is_min) on the stream; no-op for world == 1 / NULL comm
int comm_allreduce(void* comm, double* buf, int64_t count, bool is_min, cudaStream_t st);
}  // namespace dvf
