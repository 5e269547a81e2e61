// Internal view of the communicator handle (include/cnmf_b200.h keeps it opaque).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace cnmf {
constexpr int kMaxPeers = 16;
constexpr int kPeerCap = 1 << 16;  // doubles per (slot, sender): exchanges larger than this go through NCCL

// One rank's exchange window (device memory, exported to the peers through CUDA IPC).  Senders
// PUSH: rank r stores its partial sums into slot[seq & 1][r] of every peer's window (its own
// included) and then releases flag[seq & 1][r] = seq there; the receiver only reads local memory.
struct PeerWindow {
    unsigned int flag[2][kMaxPeers];
    unsigned int pad[64 - 2 * kMaxPeers];
    // double slot[2][world][kPeerCap] follows this 256-byte header
};
static_assert(sizeof(PeerWindow) == 256, "window header");
__host__ __device__ inline double *peer_slot(PeerWindow *w, int parity, int world, int sender) {
    return reinterpret_cast<double *>(reinterpret_cast<char *>(w) + sizeof(PeerWindow)) +
           ((size_t)parity * world + sender) * kPeerCap;
}

// what the exchange kernel needs, passed by value
struct PeerView {
    PeerWindow *win[kMaxPeers];  // win[rank] is the local window
    unsigned int *seq;           // local exchange counter (device)
    int world, rank;
    unsigned long long timeout_ns;  // bound on the wait for the peers' flags (cnmf_comm_set_timeout)
};
}  // namespace cnmf

struct cnmf_comm {
    void *nccl = nullptr;  // ncclComm_t
    int world_size = 1;
    int rank = 0;
    // peer-memory exchange (cnmf_comm_peer_export / _import); NCCL stays the fallback
    cnmf::PeerWindow *window = nullptr;  // this rank's window
    size_t window_bytes = 0;
    void *peer[cnmf::kMaxPeers] = {nullptr};  // opened IPC mappings (nullptr for self)
    unsigned int *d_seq = nullptr;
    bool peer_mapped = false;  // every peer window is opened
    bool peer_ready = false;   // ... and the host layer switched the exchange on (all ranks or none)
    // cnmf_comm_create_local: the ranks of ONE process (one per device, or several on one device); the
    // peers' windows are plain device pointers (cudaDeviceEnablePeerAccess), no NCCL and no IPC
    bool local = false;
    int device = -1;                                      // device of this rank (local groups)
    unsigned long long timeout_ns = 120ull * 1000000000;  // exchange_finalize_kernel gives up after this
};

namespace cnmf {
// in-place sum of `count` doubles over the communicator; no-op for world_size 1
int comm_allreduce_f64(cnmf_comm *c, double *buf, size_t count, cudaStream_t st);
// peer view of a communicator whose windows are attached (comm_peer_ok), for launch_exchange
bool comm_peer_ok(const cnmf_comm *c, size_t count);
PeerView comm_peer_view(const cnmf_comm *c);
}  // namespace cnmf
