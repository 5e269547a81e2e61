// NCCL plumbing for the row-sharded mode.  NCCL is dlopen-ed on first use so the
// library loads (and single-GPU solves run) on a box without it; inside a
// torch process the already-loaded bundled libnccl.so.2 is picked up by soname.
// Only three calls are used: one ncclAllReduce(sum, fp64) of the
// [2*k*ld + 1] partial buffer per iteration.
#include "comm.h"

#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/cnmf_b200.h"

namespace cnmf {

void set_error(const char *fmt, ...);

namespace {
// minimal NCCL ABI (stable since 2.x): opaque comm, 128-byte id, enums as ints
typedef struct { char internal[128]; } nccl_uid_t;
typedef int (*fn_get_uid)(nccl_uid_t *);
typedef int (*fn_init_rank)(void **, int, nccl_uid_t, int);
typedef int (*fn_destroy)(void *);
typedef int (*fn_allreduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef const char *(*fn_errstr)(int);
typedef int (*fn_version)(int *);
constexpr int kNcclFloat64 = 8;
constexpr int kNcclSum = 0;

struct Api {
    void *lib = nullptr;
    fn_get_uid get_uid = nullptr;
    fn_init_rank init_rank = nullptr;
    fn_destroy destroy = nullptr;
    fn_allreduce allreduce = nullptr;
    fn_errstr errstr = nullptr;
    fn_version version = nullptr;
    bool ok = false;
};
Api g_api;
std::once_flag g_once;

void load_api() {
    const char *names[] = {getenv("CNMF_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        if (!n || !*n) continue;
        g_api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_api.lib) break;
    }
    if (!g_api.lib) return;
    g_api.get_uid = (fn_get_uid)dlsym(g_api.lib, "ncclGetUniqueId");
    g_api.init_rank = (fn_init_rank)dlsym(g_api.lib, "ncclCommInitRank");
    g_api.destroy = (fn_destroy)dlsym(g_api.lib, "ncclCommDestroy");
    g_api.allreduce = (fn_allreduce)dlsym(g_api.lib, "ncclAllReduce");
    g_api.errstr = (fn_errstr)dlsym(g_api.lib, "ncclGetErrorString");
    g_api.version = (fn_version)dlsym(g_api.lib, "ncclGetVersion");
    g_api.ok = g_api.get_uid && g_api.init_rank && g_api.destroy && g_api.allreduce;
}

int need_api() {
    std::call_once(g_once, load_api);
    if (!g_api.ok) {
        set_error("NCCL could not be loaded (libnccl.so.2; set CNMF_NCCL_LIB): %s", dlerror() ? dlerror() : "missing symbols");
        return CNMF_ERR_NCCL;
    }
    return 0;
}

int wrap(int r, const char *what) {
    if (r == 0) return 0;
    set_error("%s failed: NCCL error %d (%s)", what, r, g_api.errstr ? g_api.errstr(r) : "?");
    return 2000 + r;
}
}  // namespace

int comm_allreduce_f64(cnmf_comm *c, double *buf, size_t count, cudaStream_t st) {
    if (!c || c->world_size <= 1) return 0;
    if (!c->nccl) {
        set_error("exchange of %zu doubles exceeds the peer windows (%d) and this communicator has no NCCL", count, kPeerCap);
        return CNMF_ERR_UNSUPPORTED;
    }
    return wrap(g_api.allreduce(buf, buf, count, kNcclFloat64, kNcclSum, c->nccl, st), "ncclAllReduce");
}

bool comm_peer_ok(const cnmf_comm *c, size_t count) {
    return c && c->world_size > 1 && c->peer_ready && count <= (size_t)kPeerCap;
}

PeerView comm_peer_view(const cnmf_comm *c) {
    PeerView v;
    for (int i = 0; i < kMaxPeers; ++i)
        v.win[i] = i < c->world_size ? (i == c->rank ? c->window : (PeerWindow *)c->peer[i]) : nullptr;
    v.seq = c->d_seq;
    v.world = c->world_size;
    v.rank = c->rank;
    v.timeout_ns = c->timeout_ns;
    return v;
}

}  // namespace cnmf

using namespace cnmf;

static int cuda_err(cudaError_t e, const char *what) {
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    cudaGetLastError();
    return 1000 + (int)e;
}

extern "C" int cnmf_comm_unique_id(uint8_t id_out[128]) {
    if (!id_out) return CNMF_ERR_ARG;
    if (int e = need_api()) return e;
    nccl_uid_t id;
    if (int e = wrap(g_api.get_uid(&id), "ncclGetUniqueId")) return e;
    memcpy(id_out, id.internal, 128);
    return 0;
}

extern "C" int cnmf_comm_create(cnmf_comm_t **comm, const uint8_t id[128], int32_t world_size, int32_t rank) {
    if (!comm || !id || world_size < 1 || rank < 0 || rank >= world_size) {
        set_error("cnmf_comm_create: bad arguments");
        return CNMF_ERR_ARG;
    }
    if (int e = need_api()) return e;
    nccl_uid_t uid;
    memcpy(uid.internal, id, 128);
    void *nc = nullptr;
    if (int e = wrap(g_api.init_rank(&nc, world_size, uid, rank), "ncclCommInitRank")) return e;
    cnmf_comm *c = new cnmf_comm();
    c->nccl = nc;
    c->world_size = world_size;
    c->rank = rank;
    *comm = c;
    return 0;
}

// ---- peer-memory exchange: every rank exports one window, the host layer gathers the 64-byte
// handles (as it broadcasts the NCCL id), every rank opens the others'.  From then on the
// per-iteration sum of the H partials is one kernel that stores into the peers' windows over
// NVLink (reduce.cu: exchange_finalize_kernel) and ncclAllReduce is off the iteration path.
extern "C" int cnmf_comm_peer_export(cnmf_comm_t *comm, uint8_t handle_out[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!comm || !handle_out) {
        set_error("cnmf_comm_peer_export: bad arguments");
        return CNMF_ERR_ARG;
    }
    if (comm->world_size > kMaxPeers) {
        set_error("cnmf_comm_peer_export: at most %d ranks", kMaxPeers);
        return CNMF_ERR_UNSUPPORTED;
    }
    if (!comm->window) {
        comm->window_bytes = sizeof(PeerWindow) + 2 * (size_t)comm->world_size * kPeerCap * sizeof(double);
        cudaError_t e = cudaMalloc((void **)&comm->window, comm->window_bytes);
        if (e != cudaSuccess) return cuda_err(e, "cudaMalloc(peer window)");
        if ((e = cudaMemset(comm->window, 0, comm->window_bytes)) != cudaSuccess) return cuda_err(e, "cudaMemset(peer window)");
        if ((e = cudaMalloc((void **)&comm->d_seq, 64)) != cudaSuccess) return cuda_err(e, "cudaMalloc(seq)");
        if ((e = cudaMemset(comm->d_seq, 0, 64)) != cudaSuccess) return cuda_err(e, "cudaMemset(seq)");
        if ((e = cudaDeviceSynchronize()) != cudaSuccess) return cuda_err(e, "cudaDeviceSynchronize");
    }
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, comm->window);
    if (e != cudaSuccess) return cuda_err(e, "cudaIpcGetMemHandle");
    memcpy(handle_out, &h, 64);
    return 0;
}

extern "C" int cnmf_comm_peer_import(cnmf_comm_t *comm, const uint8_t *handles) {
    if (!comm || !handles || !comm->window) {
        set_error("cnmf_comm_peer_import: export this rank's window first");
        return CNMF_ERR_ARG;
    }
    for (int r = 0; r < comm->world_size; ++r) {
        if (r == comm->rank || comm->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * 64, 64);
        cudaError_t e = cudaIpcOpenMemHandle(&comm->peer[r], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            comm->peer[r] = nullptr;
            return cuda_err(e, "cudaIpcOpenMemHandle (ranks must share a node with peer access)");
        }
    }
    comm->peer_mapped = true;
    return 0;
}

// Collective decision of the host layer: every rank enables (all imports succeeded) or none does --
// a rank left on ncclAllReduce while the others wait on their windows would deadlock.
extern "C" int cnmf_comm_peer_enable(cnmf_comm_t *comm, int32_t on) {
    if (!comm || (on && !comm->peer_mapped)) {
        set_error("cnmf_comm_peer_enable: windows are not imported");
        return CNMF_ERR_ARG;
    }
    comm->peer_ready = on != 0;
    return 0;
}

extern "C" int cnmf_comm_peer_active(const cnmf_comm_t *comm) { return comm && comm->peer_ready ? 1 : 0; }

// ---- single-process group: rank r lives on devices[r] (the same device may appear more than once:
// "virtual shards", which exercise the sharded schedule and the exchange kernel on one GPU).  Every
// rank's window is ordinary device memory the other devices reach through peer access, so neither
// NCCL nor CUDA IPC is involved.  One host thread per rank must drive the solves (a rank's final
// synchronise waits for its peers' kernels).
extern "C" int cnmf_comm_create_local(cnmf_comm_t **comms, int32_t world_size, const int32_t *devices) {
    if (!comms || !devices || world_size < 1 || world_size > kMaxPeers) {
        set_error("cnmf_comm_create_local: need 1..%d ranks", kMaxPeers);
        return CNMF_ERR_ARG;
    }
    int ndev = 0, prev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libcnmf_b200 has no CPU path");
        return CNMF_ERR_NODEVICE;
    }
    for (int r = 0; r < world_size; ++r) {
        comms[r] = nullptr;
        if (devices[r] < 0 || devices[r] >= ndev) {
            set_error("cnmf_comm_create_local: device %d of rank %d does not exist (%d devices)", devices[r], r, ndev);
            return CNMF_ERR_ARG;
        }
    }
    cudaGetDevice(&prev);
    auto fail = [&](int rc) {
        for (int r = 0; r < world_size; ++r) {
            if (comms[r]) cnmf_comm_destroy(comms[r]);
            comms[r] = nullptr;
        }
        cudaSetDevice(prev);
        return rc;
    };
    for (int r = 0; r < world_size; ++r) {
        cudaError_t e = cudaSetDevice(devices[r]);
        if (e != cudaSuccess) return fail(cuda_err(e, "cudaSetDevice"));
        for (int q = 0; q < world_size; ++q) {
            if (devices[q] == devices[r]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[r], devices[q]);
            if (!can) {
                set_error("cnmf_comm_create_local: device %d cannot access device %d", devices[r], devices[q]);
                return fail(CNMF_ERR_UNSUPPORTED);
            }
            e = cudaDeviceEnablePeerAccess(devices[q], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(cuda_err(e, "cudaDeviceEnablePeerAccess"));
            cudaGetLastError();
        }
        cnmf_comm *c = new cnmf_comm();
        comms[r] = c;
        c->world_size = world_size;
        c->rank = r;
        c->local = true;
        c->device = devices[r];
        c->window_bytes = sizeof(PeerWindow) + 2 * (size_t)world_size * kPeerCap * sizeof(double);
        if ((e = cudaMalloc((void **)&c->window, c->window_bytes)) != cudaSuccess) return fail(cuda_err(e, "cudaMalloc(peer window)"));
        if ((e = cudaMemset(c->window, 0, c->window_bytes)) != cudaSuccess) return fail(cuda_err(e, "cudaMemset(peer window)"));
        if ((e = cudaMalloc((void **)&c->d_seq, 64)) != cudaSuccess) return fail(cuda_err(e, "cudaMalloc(seq)"));
        if ((e = cudaMemset(c->d_seq, 0, 64)) != cudaSuccess) return fail(cuda_err(e, "cudaMemset(seq)"));
        if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(cuda_err(e, "cudaDeviceSynchronize"));
    }
    for (int r = 0; r < world_size; ++r) {
        for (int q = 0; q < world_size; ++q) comms[r]->peer[q] = q == r ? nullptr : (void *)comms[q]->window;
        comms[r]->peer_mapped = comms[r]->peer_ready = world_size > 1;
    }
    cudaSetDevice(prev);
    return 0;
}

extern "C" int cnmf_comm_set_timeout(cnmf_comm_t *comm, double seconds) {
    if (!comm || !(seconds > 0)) {
        set_error("cnmf_comm_set_timeout: need a communicator and a positive time");
        return CNMF_ERR_ARG;
    }
    comm->timeout_ns = (unsigned long long)(seconds * 1e9);
    return 0;
}

extern "C" int cnmf_comm_destroy(cnmf_comm_t *comm) {
    if (!comm) return 0;
    int e = 0;
    int prev = -1;
    if (comm->local) {
        cudaGetDevice(&prev);
        cudaSetDevice(comm->device);
    } else {
        for (int r = 0; r < kMaxPeers; ++r)
            if (comm->peer[r]) cudaIpcCloseMemHandle(comm->peer[r]);
    }
    if (comm->window) cudaFree(comm->window);
    if (comm->d_seq) cudaFree(comm->d_seq);
    if (prev >= 0) cudaSetDevice(prev);
    if (comm->nccl && g_api.ok) e = wrap(g_api.destroy(comm->nccl), "ncclCommDestroy");
    delete comm;
    return e;
}

extern "C" int cnmf_comm_info(const cnmf_comm_t *comm, int32_t *world_size, int32_t *rank) {
    if (!comm) return CNMF_ERR_ARG;
    if (world_size) *world_size = comm->world_size;
    if (rank) *rank = comm->rank;
    return 0;
}
