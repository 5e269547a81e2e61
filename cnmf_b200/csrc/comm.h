// Internal view of the communicator handle (include/cnmf_b200.h keeps it opaque).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

struct cnmf_comm {
    void *nccl = nullptr;  // ncclComm_t
    int world_size = 1;
    int rank = 0;
};

namespace cnmf {
// in-place sum of `count` doubles over the communicator; no-op for world_size 1
int comm_allreduce_f64(cnmf_comm *c, double *buf, size_t count, cudaStream_t st);
}  // namespace cnmf
