// cnmf_nmf_host: the host-buffer entry point (what cnmf.nmf calls for NumPy
// input, and what bench.py times as the end-to-end number).  Uploads X /
// weights / mask in row chunks through two staging buffers, folds and pads them
// on the device (cnmf_prepare), runs the plan, downloads W, H and the history.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <chrono>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/cnmf_b200.h"

namespace cnmf {
void set_error(const char *fmt, ...);
// sum of `count` doubles over the ranks of `comm`, in place on the device (api.cu); d_state: 32 zeroed bytes
int comm_sum_small(cnmf_comm_t *comm, double *d_buf, int count, void *d_state, cudaStream_t st);
}
using cnmf::set_error;

namespace {
struct DevBufs {
    void *base = nullptr;  // the one device allocation everything below points into
    void *x = nullptr, *v = nullptr, *w = nullptr, *h = nullptr, *sx = nullptr, *sw = nullptr;
    uint8_t *sm = nullptr;
    double *hist = nullptr;
    int *status = nullptr;
    cudaStream_t st = nullptr, cp = nullptr;
    cudaEvent_t copied[2] = {nullptr, nullptr}, folded[2] = {nullptr, nullptr};
    cnmf_plan_t *plan = nullptr;
    ~DevBufs() {
        if (plan) cnmf_plan_destroy(plan);
        for (int i = 0; i < 2; ++i) {
            if (copied[i]) cudaEventDestroy(copied[i]);
            if (folded[i]) cudaEventDestroy(folded[i]);
        }
        if (cp) cudaStreamDestroy(cp);
        cudaFree(base);
        if (st) cudaStreamDestroy(st);
    }
};
// wall-clock phases of the last cnmf_nmf_host call of this thread (cnmf_host_timing)
thread_local double g_phase[CNMF_HOST_PHASES] = {0};
double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int fail(cudaError_t e, const char *what) {
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return 1000 + (int)e;
}
// NVTX range around a phase of the call (visible in nsys / ncu timelines; a no-op without a tool attached)
struct Range {
    explicit Range(const char *name) { nvtxRangePushA(name); }
    ~Range() { nvtxRangePop(); }
};

// ---- pageable input: a ring of pinned slots filled by a few host threads -------------------------
// cudaMemcpyAsync from pageable memory goes through the driver's own single-threaded staging (measured
// 2-3x slower than a pinned source on this host).  For large pageable arrays the upload instead runs
// kStageThreads workers: each copies a piece of the caller's array into one of its pinned slots with
// memcpy and issues the slot's asynchronous copy on its own stream, so the host-side copies of several
// pieces overlap each other and the DMA transfers.  Slots are pooled across calls (pinning costs
// ~0.3 ms per MB).
constexpr size_t kSlotBytes = 8u << 20;
constexpr int kSlotsPerThread = 2;
constexpr int kStageThreads = 6;
constexpr size_t kStageMinBytes = 32u << 20;  // smaller uploads take the plain path

struct StageSet {  // the slots, stream and events of one worker
    void *slot[kSlotsPerThread] = {nullptr, nullptr};
    cudaEvent_t done[kSlotsPerThread] = {nullptr, nullptr};
    cudaStream_t st = nullptr;
    int device = -1;
};
std::mutex g_stage_mu;
std::vector<StageSet *> g_stage_free;

StageSet *stage_get(int device) {
    {
        std::lock_guard<std::mutex> lk(g_stage_mu);
        for (size_t i = 0; i < g_stage_free.size(); ++i)
            if (g_stage_free[i]->device == device) {
                StageSet *s = g_stage_free[i];
                g_stage_free.erase(g_stage_free.begin() + i);
                return s;
            }
    }
    StageSet *s = new StageSet();
    s->device = device;
    bool ok = cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < kSlotsPerThread; ++i)
        ok = cudaHostAlloc(&s->slot[i], kSlotBytes, cudaHostAllocPortable) == cudaSuccess &&
             cudaEventCreateWithFlags(&s->done[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        for (int i = 0; i < kSlotsPerThread; ++i) {
            if (s->slot[i]) cudaFreeHost(s->slot[i]);
            if (s->done[i]) cudaEventDestroy(s->done[i]);
        }
        if (s->st) cudaStreamDestroy(s->st);
        delete s;
        return nullptr;
    }
    return s;
}
void stage_put(StageSet *s) {
    std::lock_guard<std::mutex> lk(g_stage_mu);
    g_stage_free.push_back(s);
}

bool is_pageable(const void *p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

// One piece of the upload: `bytes` from src (host, pageable) to dst (device staging set `set` of chunk `chunk`).
struct Piece {
    const char *src;
    char *dst;
    size_t bytes;
    int64_t chunk;
};

struct Stager {
    std::vector<Piece> pieces;
    std::vector<int64_t> chunk_first;            // index of the first piece of each chunk (+ sentinel)
    std::atomic<size_t> next{0};
    std::vector<std::atomic<int>> issued;         // per piece: 1 once its device copy has been enqueued
    std::vector<cudaEvent_t> piece_event;         // the slot event that covers the piece (worker streams are in order)
    std::atomic<int64_t> folded_upto{-1};         // chunks whose fold has been enqueued (folded[] recorded) by the main thread
    cudaEvent_t *folded = nullptr;                // folded[chunk & 1]
    std::atomic<int> failed{0};
    std::vector<StageSet *> sets;
    std::vector<std::thread> threads;
    int device = 0;

    explicit Stager(size_t n) : issued(n), piece_event(n, nullptr) {
        for (auto &f : issued) f.store(0, std::memory_order_relaxed);
    }
    void worker(StageSet *ss) {
        cudaSetDevice(device);
        int turn = 0;
        for (;;) {
            const size_t i = next.fetch_add(1, std::memory_order_relaxed);
            if (i >= pieces.size() || failed.load(std::memory_order_relaxed)) break;
            const Piece &pc = pieces[i];
            const int sl = turn++ % kSlotsPerThread;
            if (cudaEventSynchronize(ss->done[sl]) != cudaSuccess) { failed = 1; break; }  // the slot's previous copy has left it
            memcpy(ss->slot[sl], pc.src, pc.bytes);
            // the device staging set of this chunk is free once the fold of chunk - 2 has been enqueued and run
            if (pc.chunk >= 2) {
                while (folded_upto.load(std::memory_order_acquire) < pc.chunk - 2 && !failed.load(std::memory_order_relaxed))
                    std::this_thread::yield();
                if (cudaStreamWaitEvent(ss->st, folded[pc.chunk & 1], 0) != cudaSuccess) { failed = 1; break; }
            }
            if (cudaMemcpyAsync(pc.dst, ss->slot[sl], pc.bytes, cudaMemcpyHostToDevice, ss->st) != cudaSuccess ||
                cudaEventRecord(ss->done[sl], ss->st) != cudaSuccess) {
                failed = 1;
                break;
            }
            piece_event[i] = ss->done[sl];
            issued[i].store(1, std::memory_order_release);
        }
    }
    bool start(int dev) {
        device = dev;
        for (int t = 0; t < kStageThreads; ++t) {
            StageSet *ss = stage_get(dev);
            if (!ss) break;
            sets.push_back(ss);
        }
        if (sets.empty()) return false;
        for (StageSet *ss : sets) threads.emplace_back([this, ss] { worker(ss); });
        return true;
    }
    // make stream `st` wait for every piece of `chunk`; returns false on failure
    bool wait_chunk(int64_t chunk, cudaStream_t st) {
        for (int64_t i = chunk_first[chunk]; i < chunk_first[chunk + 1]; ++i) {
            while (!issued[i].load(std::memory_order_acquire)) {
                if (failed.load(std::memory_order_relaxed)) return false;
                std::this_thread::yield();
            }
            if (cudaStreamWaitEvent(st, piece_event[i], 0) != cudaSuccess) return false;
        }
        return true;
    }
    ~Stager() {
        failed.store(failed.load() | 2);  // stops workers that have not finished (error paths)
        next.store(pieces.size());
        for (auto &t : threads) t.join();
        for (StageSet *ss : sets) {
            cudaStreamSynchronize(ss->st);
            stage_put(ss);
        }
    }
};
}  // namespace

#define HCU(call)                                        \
    do {                                                 \
        cudaError_t e__ = (call);                        \
        if (e__ != cudaSuccess) return fail(e__, #call); \
    } while (0)

extern "C" int cnmf_nmf_host(int32_t dtype, int64_t n_rows, int32_t m, int32_t k, const void *h_x,
                             const void *h_weights, const uint8_t *h_mask, void *h_w, void *h_h, double *h_hist,
                             int32_t max_iter, double tol, double eps, int32_t update, cnmf_comm_t *comm,
                             int32_t *n_iter_out, int64_t *h2d_bytes, int64_t *d2h_bytes) {
    if ((dtype != CNMF_F32 && dtype != CNMF_F64) || n_rows < 1 || m < 1 || k < 1 || k > CNMF_MAX_K || !h_x || !h_w ||
        !h_h || !h_hist || max_iter < 0) {
        set_error("cnmf_nmf_host: bad arguments");
        return CNMF_ERR_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device: libcnmf_b200 has no CPU path");
        return CNMF_ERR_NODEVICE;
    }
    const size_t sz = dtype == CNMF_F32 ? 4 : 8;
    const int64_t ld = cnmf_padded_ld(m, dtype);
    for (double &x : g_phase) x = 0;
    const double t_begin = now();
    double t0 = t_begin;
    auto lap = [&](int i) {
        const double t = now();
        g_phase[i] += t - t0;
        t0 = t;
    };
    struct Total {
        double t;
        ~Total() { g_phase[CNMF_HOST_T_TOTAL] = now() - t; }  // includes the buffers' release below
    } total{t_begin};
    DevBufs b;
    int64_t up = 0, down = 0;
    HCU(cudaStreamCreateWithFlags(&b.st, cudaStreamNonBlocking));
    // staging: two sets of ~64 MB of raw rows per operand; the copies of chunk c + 1 (copy
    // stream) run under the fold-and-pad of chunk c (compute stream)
    int64_t chunk = (64LL << 20) / ((int64_t)m * sz);
    if (chunk < 1) chunk = 1;
    if (chunk > n_rows) chunk = n_rows;
    const size_t cb = (size_t)chunk * m * sz;
    {
        // one allocation (a dozen cudaMalloc calls cost more than a small solve), 256-byte carved
        auto up = [](size_t v) { return (v + 255) / 256 * 256; };
        const size_t xb = up((size_t)n_rows * ld * sz), wb = up((size_t)n_rows * k * sz), hb = up((size_t)k * m * sz),
                     histb = up((size_t)(max_iter + 1) * sizeof(double)), stb = 512, sxb = up(2 * cb),
                     swb = h_weights ? up(2 * cb) : 0, smb = h_mask ? up(2 * (size_t)chunk * m) : 0;
        HCU(cudaMalloc(&b.base, 2 * xb + wb + hb + histb + stb + sxb + swb + smb));
        char *q = (char *)b.base;
        b.x = q; q += xb;
        b.v = q; q += xb;
        b.w = q; q += wb;
        b.h = q; q += hb;
        b.hist = (double *)q; q += histb;
        b.status = (int *)q; q += stb;
        b.sx = q; q += sxb;
        if (h_weights) { b.sw = q; q += swb; }
        if (h_mask) { b.sm = (uint8_t *)q; q += smb; }
    }
    HCU(cudaMemsetAsync(b.status, 0, 512, b.st));
    // the history starts as NaN: entries a solve never reached (early stop, error) cannot pass for values
    HCU(cudaMemsetAsync(b.hist, 0xff, (size_t)(max_iter + 1) * sizeof(double), b.st));
    lap(CNMF_HOST_T_ALLOC);
    HCU(cudaStreamCreateWithFlags(&b.cp, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        HCU(cudaEventCreateWithFlags(&b.copied[i], cudaEventDisableTiming));
        HCU(cudaEventCreateWithFlags(&b.folded[i], cudaEventDisableTiming));
    }
    // Pageable sources of a large upload go through the pinned ring (see Stager); pinned / registered
    // sources are copied directly.  Both paths fold chunk c under the copies of chunk c + 1.
    int dev = 0;
    HCU(cudaGetDevice(&dev));
    const int64_t n_chunks = (n_rows + chunk - 1) / chunk;
    const size_t total_in = (size_t)n_rows * m * (sz * (h_weights ? 2 : 1) + (h_mask ? 1 : 0));
    const bool px = is_pageable(h_x), pw = h_weights && is_pageable(h_weights), pm = h_mask && is_pageable(h_mask);
    std::unique_ptr<Stager> stager;
    if ((px || pw || pm) && total_in >= kStageMinBytes) {
        Range r("cnmf: plan pageable upload");
        size_t n_pieces = 0;
        auto count = [&](size_t nb) { n_pieces += (nb + kSlotBytes - 1) / kSlotBytes; };
        for (int64_t c = 0; c < n_chunks; ++c) {
            const int64_t rows = (n_rows - c * chunk) < chunk ? (n_rows - c * chunk) : chunk;
            if (px) count((size_t)rows * m * sz);
            if (pw) count((size_t)rows * m * sz);
            if (pm) count((size_t)rows * m);
        }
        stager.reset(new Stager(n_pieces));
        stager->folded = b.folded;
        stager->pieces.reserve(n_pieces);
        auto add = [&](const char *src, char *dst, size_t nb, int64_t c) {
            for (size_t o = 0; o < nb; o += kSlotBytes)
                stager->pieces.push_back(Piece{src + o, dst + o, nb - o < kSlotBytes ? nb - o : kSlotBytes, c});
        };
        for (int64_t c = 0; c < n_chunks; ++c) {
            const int64_t r0 = c * chunk, rows = (n_rows - r0) < chunk ? (n_rows - r0) : chunk;
            const int q = (int)(c & 1);
            stager->chunk_first.push_back((int64_t)stager->pieces.size());
            if (px) add((const char *)h_x + (size_t)r0 * m * sz, (char *)b.sx + q * cb, (size_t)rows * m * sz, c);
            if (pw) add((const char *)h_weights + (size_t)r0 * m * sz, (char *)b.sw + q * cb, (size_t)rows * m * sz, c);
            if (pm) add((const char *)h_mask + (size_t)r0 * m, (char *)b.sm + q * (size_t)chunk * m, (size_t)rows * m, c);
        }
        stager->chunk_first.push_back((int64_t)stager->pieces.size());
        if (!stager->start(dev)) stager.reset();  // no pinned memory to be had: the plain path below
    }
    const bool sx_staged = stager && px, sw_staged = stager && pw, sm_staged = stager && pm;
    {
        Range r("cnmf: upload + prepare");
        int64_t c = 0;
        for (int64_t r0 = 0; r0 < n_rows; r0 += chunk, ++c) {
            const int64_t rows = (n_rows - r0) < chunk ? (n_rows - r0) : chunk;
            const size_t nb = (size_t)rows * m * sz;
            const int q = (int)(c & 1);
            char *sx = (char *)b.sx + q * cb;
            char *sw = h_weights ? (char *)b.sw + q * cb : nullptr;
            uint8_t *sm = h_mask ? b.sm + q * (size_t)chunk * m : nullptr;
            if (c >= 2) HCU(cudaStreamWaitEvent(b.cp, b.folded[q], 0));  // set q is free again
            if (!sx_staged) HCU(cudaMemcpyAsync(sx, (const char *)h_x + (size_t)r0 * m * sz, nb, cudaMemcpyHostToDevice, b.cp));
            up += (int64_t)nb;
            if (h_weights) {
                if (!sw_staged)
                    HCU(cudaMemcpyAsync(sw, (const char *)h_weights + (size_t)r0 * m * sz, nb, cudaMemcpyHostToDevice, b.cp));
                up += (int64_t)nb;
            }
            if (h_mask) {
                if (!sm_staged) HCU(cudaMemcpyAsync(sm, h_mask + (size_t)r0 * m, (size_t)rows * m, cudaMemcpyHostToDevice, b.cp));
                up += rows * m;
            }
            HCU(cudaEventRecord(b.copied[q], b.cp));
            HCU(cudaStreamWaitEvent(b.st, b.copied[q], 0));
            if (stager && !stager->wait_chunk(c, b.st)) {
                set_error("upload through the pinned staging ring failed (%s)", cudaGetErrorString(cudaGetLastError()));
                return CNMF_ERR_NOMEM;
            }
            int e = cnmf_prepare(dtype, sx, sw, sm, m, m, (char *)b.x + (size_t)r0 * ld * sz,
                                 (char *)b.v + (size_t)r0 * ld * sz, ld, rows, m, b.status, b.st);
            if (e) return e;
            HCU(cudaEventRecord(b.folded[q], b.st));
            if (stager) stager->folded_upto.store(c, std::memory_order_release);
        }
        stager.reset();  // joins the workers (all pieces are issued by now)
    }
    HCU(cudaMemcpyAsync(b.w, h_w, (size_t)n_rows * k * sz, cudaMemcpyHostToDevice, b.cp));
    HCU(cudaMemcpyAsync(b.h, h_h, (size_t)k * m * sz, cudaMemcpyHostToDevice, b.cp));
    HCU(cudaEventRecord(b.copied[0], b.cp));
    HCU(cudaStreamWaitEvent(b.st, b.copied[0], 0));
    up += (int64_t)((size_t)n_rows * k * sz + (size_t)k * m * sz);
    int status = 0;
    HCU(cudaMemcpyAsync(&status, b.status, sizeof(int), cudaMemcpyDeviceToHost, b.st));
    HCU(cudaStreamSynchronize(b.st));
    if (comm) {
        // The verdict on the input is collective: a rank that left here alone would leave its peers
        // waiting in the first exchange.  Every rank contributes "my shard is bad" and all of them
        // return CNMF_ERR_DATA together.  (b.status + 64: one double, + 128: a zeroed State block.)
        double *d_flag = (double *)((char *)b.status + 64);
        const double mine = status ? 1.0 : 0.0;
        double any = 0.0;
        HCU(cudaMemcpyAsync(d_flag, &mine, sizeof(double), cudaMemcpyHostToDevice, b.st));
        if (int e = cnmf::comm_sum_small(comm, d_flag, 1, (char *)b.status + 128, b.st)) return e;
        HCU(cudaMemcpyAsync(&any, d_flag, sizeof(double), cudaMemcpyDeviceToHost, b.st));
        HCU(cudaStreamSynchronize(b.st));
        if (any != 0.0 && !status) {
            set_error("invalid input on another rank of the communicator (its shard failed validation)");
            return CNMF_ERR_DATA;
        }
    }
    lap(CNMF_HOST_T_UPLOAD);
    if (status) {
        set_error("invalid input:%s%s%s%s", (status & CNMF_BAD_X_NEGATIVE) ? " negative X under positive weight;" : "",
                  (status & CNMF_BAD_X_NONFINITE) ? " non-finite X under positive weight;" : "",
                  (status & CNMF_BAD_W_NEGATIVE) ? " negative weights;" : "",
                  (status & CNMF_BAD_W_NONFINITE) ? " non-finite weights;" : "");
        return CNMF_ERR_DATA;
    }
    cnmf_desc_t d;
    d.n_rows = n_rows;
    d.ld = ld;
    d.m = m;
    d.k = k;
    d.dtype = dtype;
    d.batch = 1;
    d.eps = eps;
    d.comm = comm;
    d.sm_limit = 0;
    d.variant = 0;
    // A single problem small enough for a thread-block cluster's shared memory (1 000 x 81 fp64 = 8 CTAs)
    // runs on the resident cluster kernel: all iterations in one launch, 7 us per iteration where the
    // streaming plan's three launches need 16.  CNMF_ONE_CTA=0 keeps the streaming plan; CNMF_ONE_CTA=1
    // sends every single problem through the batched solver (resident where it fits, else one CTA
    // streaming its tiles -- slow, tests only).
    bool one_cta = false;
    int want = comm ? 0 : CNMF_VARIANT_RESIDENT;
    if (const char *e = getenv("CNMF_ONE_CTA")) want = comm ? 0 : (e[0] == '1' ? CNMF_VARIANT_ONE_CTA : (e[0] == '0' ? 0 : want));
    int n_iter = max_iter;
    int rc = 0;
    if (want) {
        d.variant = want;
        one_cta = cnmf_plan_create(&b.plan, &d) == 0;  // fails when the problem does not fit shared memory
        if (!one_cta) d.variant = 0;
    }
    nvtxRangePushA("cnmf: solve");
    struct PopAtExit {
        bool armed = true;
        ~PopAtExit() { if (armed) nvtxRangePop(); }
    } solve_range;
    if (one_cta) {
        lap(CNMF_HOST_T_PLAN);
        int *d_niter = b.status;  // free again: the data check above has read it
        rc = cnmf_solve_batched(b.plan, b.x, b.v, b.w, b.h, b.hist, d_niter, max_iter, tol, update, b.st);
        if (rc) return rc;
        HCU(cudaMemcpyAsync(&n_iter, d_niter, sizeof(int), cudaMemcpyDeviceToHost, b.st));
        HCU(cudaStreamSynchronize(b.st));
        double last = 0;
        HCU(cudaMemcpy(&last, b.hist + n_iter, sizeof(double), cudaMemcpyDeviceToHost));
        if (!(last == last) || last > 1.7e308) {
            set_error("chi2 became non-finite at iteration %d", n_iter);
            rc = CNMF_ERR_NONFINITE;
        }
    } else {
        if (int e = cnmf_plan_create(&b.plan, &d)) return e;
        lap(CNMF_HOST_T_PLAN);
        rc = cnmf_solve(b.plan, b.x, b.v, b.w, b.h, b.hist, max_iter, tol, 0, update, &n_iter, b.st);
        if (rc && rc != CNMF_ERR_NONFINITE) return rc;
        HCU(cudaStreamSynchronize(b.st));
    }
    lap(CNMF_HOST_T_SOLVE);
    nvtxRangePop();
    solve_range.armed = false;
    Range download_range("cnmf: download");
    HCU(cudaMemcpyAsync(h_w, b.w, (size_t)n_rows * k * sz, cudaMemcpyDeviceToHost, b.st));
    HCU(cudaMemcpyAsync(h_h, b.h, (size_t)k * m * sz, cudaMemcpyDeviceToHost, b.st));
    HCU(cudaMemcpyAsync(h_hist, b.hist, (size_t)(n_iter + 1) * sizeof(double), cudaMemcpyDeviceToHost, b.st));
    HCU(cudaStreamSynchronize(b.st));
    lap(CNMF_HOST_T_DOWNLOAD);
    down += (int64_t)((size_t)n_rows * k * sz + (size_t)k * m * sz + (size_t)(n_iter + 1) * 8);
    if (n_iter_out) *n_iter_out = n_iter;
    if (h2d_bytes) *h2d_bytes = up;
    if (d2h_bytes) *d2h_bytes = down;
    return rc;
}

extern "C" int cnmf_host_timing(double *seconds, int32_t n) {
    if (!seconds || n < 0) return CNMF_ERR_ARG;
    for (int i = 0; i < n && i < CNMF_HOST_PHASES; ++i) seconds[i] = g_phase[i];
    return 0;
}
