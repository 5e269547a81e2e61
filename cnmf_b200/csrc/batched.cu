// K4 -- batched small-problem solver (one CTA per problem).  Placeholder until
// the kernel lands: reports "not supported" so callers fail loudly.
#include "launch.h"

namespace cnmf {
int batched_smem_bytes(int, int, long long, int *threads_out) {
    if (threads_out) *threads_out = 0;
    return -1;
}
int launch_batched(const BatchedArgs &, cudaStream_t) { return (int)cudaErrorNotSupported; }
}  // namespace cnmf
