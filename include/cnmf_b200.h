/*
 * cnmf_b200 -- C ABI of the B200-native weighted / masked multiplicative-update
 * NMF iteration (libcnmf_b200.so, hand-written sm_100a CUDA).
 *
 * Boundary: this is the interface the `cnmf` Python package binds for its NMF
 * entry point.  The upstream snapshot pins only the import name
 * (/root/reference/setup.py:9, /root/reference/cnmf/__init__.py is 0 bytes);
 * no function, signature or error convention exists there, so the entry points
 * below follow SURVEY.md section 8(b) and BASELINE.json `north_star` (k,
 * weights/mask, initial W/H, max iterations, tolerance -> W, H, chi2 history).
 * INTEGRATION.md shows the ctypes stub that `cnmf/__init__.py` would add.
 *
 * Conventions
 *  - every function returns 0 on success, a negative CNMF_ERR_* on argument
 *    errors, or a positive value wrapping a cudaError_t (1000 + code) or an
 *    ncclResult_t (2000 + code); cnmf_last_error() gives the thread-local text.
 *    No exception crosses the ABI.
 *  - all device buffers are caller-owned raw device pointers; a plan owns only
 *    its scratch (partial blocks, reduce buffer, graphs).
 *  - streams are passed as void* (a cudaStream_t); NULL is the default stream.
 *  - a plan is not thread-safe: one plan per host thread per device.
 *
 * Mathematics (oracle/cnmf_oracle.py is the CPU restatement):
 *    chi2 = sum V (X - WH)^2
 *    per iteration:  H <- H * [W^T (V*X)] / max(W^T (V*WH), eps)
 *                    W <- W * [(V*X) H^T] / max((V*WH) H^T, eps)   (new H)
 *                    hist[it] = chi2(W, H)
 *    hist[0] = chi2(W0, H0); stop at max_iter or, for tol > 0, when
 *    hist[it-1] - hist[it] <= tol * hist[it-1].
 */
#ifndef CNMF_B200_H
#define CNMF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CNMF_API __attribute__((visibility("default")))
#else
#define CNMF_API
#endif

#define CNMF_VERSION 100 /* 0.1.0 */

#define CNMF_F32 0
#define CNMF_F64 1

#define CNMF_MAX_K 16
#define CNMF_VARIANT_ONE_CTA 1
#define CNMF_VARIANT_RESIDENT 2 /* one problem per thread-block cluster, X and V resident in shared memory */

/* argument errors */
#define CNMF_ERR_ARG (-1)       /* NULL pointer, bad enum, k/m/n out of range */
#define CNMF_ERR_ALIGN (-2)     /* ld or a base pointer not 16-byte aligned   */
#define CNMF_ERR_UNSUPPORTED (-3)
#define CNMF_ERR_NODEVICE (-4)  /* no CUDA device / not sm_100                */
#define CNMF_ERR_NCCL (-5)      /* NCCL library could not be loaded           */
#define CNMF_ERR_DATA (-6)      /* negative / non-finite X or weights          */
#define CNMF_ERR_NONFINITE (-7) /* chi2 became non-finite                     */
#define CNMF_ERR_NOMEM (-8)

/* cnmf_solve update flags */
#define CNMF_UPDATE_W 1
#define CNMF_UPDATE_H 2
#define CNMF_UPDATE_BOTH 3
/* OR-ed into `update`: continue the previous cnmf_solve on the same plan and buffers
 * (no state reset, no pass over (W0, H0)); hist indices keep counting, so d_hist must
 * have room for the accumulated iteration count + 1. */
#define CNMF_SOLVE_RESUME 4
/* OR-ed into `update`: bracket every fused-kernel launch with CUDA events on the launching
 * stream (disables graph replay for this call); read the sum with cnmf_plan_kernel_time. */
#define CNMF_SOLVE_TIME_KERNELS 8
/* OR-ed into `update`: do the plan-time work of a coming cnmf_solve with exactly these buffers, max_iter,
 * tol, check_every and stream (capture, instantiate and upload the CUDA graph of its iteration chunks) and
 * return without running an iteration. */
#define CNMF_SOLVE_PREPARE 16

/* cnmf_synth V modes (cnmf_b200/synth.py VMODES) */
#define CNMF_V_ONES 0
#define CNMF_V_IVAR 1
#define CNMF_V_MASK 2
#define CNMF_V_IVAR_MASK 3

/* cnmf_prepare status bits (device int, OR-ed) */
#define CNMF_BAD_X_NEGATIVE 1
#define CNMF_BAD_X_NONFINITE 2
#define CNMF_BAD_W_NEGATIVE 4
#define CNMF_BAD_W_NONFINITE 8

typedef struct cnmf_plan cnmf_plan_t;
typedef struct cnmf_comm cnmf_comm_t;

/* Problem description.  X and V are row-major [n_rows][ld] in the compute dtype
 * with ld >= m, ld * sizeof(T) a multiple of 16 and 16-byte aligned bases; pad
 * columns m..ld-1 must hold X = 0 and V = 0 (cnmf_prepare writes them so).
 * W is [n_rows][k], H is [k][m], both dense row-major in the compute dtype.
 * batch > 1: X,V are [batch][n_rows][ld], W [batch][n_rows][k], H [batch][k][m]. */
typedef struct cnmf_desc {
    int64_t n_rows;     /* rows held by THIS rank */
    int64_t ld;         /* leading dimension of X and V, in elements */
    int32_t m;          /* wavelengths */
    int32_t k;          /* components, 1..CNMF_MAX_K */
    int32_t dtype;      /* CNMF_F32 | CNMF_F64 */
    int32_t batch;      /* 0 or 1 = one problem; > 1 = independent problems */
    double eps;         /* denominator guard; <= 0 selects 1e-16 (f64) / 1e-12 (f32) */
    cnmf_comm_t *comm;  /* NULL = single GPU; else rows are sharded over the comm */
    int32_t sm_limit;   /* 0 = all SMs; > 0 caps the persistent grid (tests) */
    int32_t variant;    /* 0 = default kernel choice; CNMF_VARIANT_ONE_CTA = the batched (one CTA per
                         * problem, all iterations in-kernel) solver also for batch <= 1: solve such a
                         * plan with cnmf_solve_batched.  (Not a fast path: one CTA walks its tiles serially;
                         * the streaming plan is quicker even at 1 000 x 81.)
                         * CNMF_VARIANT_RESIDENT = a batched plan that must run on the resident cluster
                         * kernel (X and V of a problem stay in the shared memory of 1..8 CTAs for all
                         * iterations); plan creation fails with CNMF_ERR_UNSUPPORTED when the problem does
                         * not fit (k > 8, or more than 8 SMs' worth of X and V).  The fast path for ONE
                         * small problem: cnmf_nmf_host takes it by default (CNMF_ONE_CTA=0 turns that off). */
} cnmf_desc_t;

/* What a plan decided; bytes_per_iter is the ALGORITHMIC figure
 * n_rows * m * 2 * sizeof(T) that roofline numbers are quoted on.
 * variant: < 100000 the scalar stream kernel (lanes per row * 1000 + packs per thread * 100 + rows per lane * 10
 * + row-owner warps); 7000cs the resident cluster kernel (cs CTAs per problem); 90m000 + 100 * column parts +
 * row-owner warps the fp64 tensor-path kernel (mma_kernel.cuh; m = 3: 8 column-owner warps, m = 1: 4). */
typedef struct cnmf_info {
    int32_t grid, threads, smem_bytes, stages, tile_rows, lanes_per_row;
    int32_t h_streams, partial_blocks, variant, launches_per_iter;
    int64_t bytes_per_iter;
    int64_t scratch_bytes;
} cnmf_info_t;

CNMF_API int cnmf_version(void);
CNMF_API const char *cnmf_last_error(void);

/* leading dimension (elements) cnmf_prepare / the Python layer use for m */
CNMF_API int64_t cnmf_padded_ld(int32_t m, int32_t dtype);

CNMF_API int cnmf_plan_create(cnmf_plan_t **plan, const cnmf_desc_t *desc);
CNMF_API int cnmf_plan_destroy(cnmf_plan_t *plan);
CNMF_API int cnmf_plan_info(const cnmf_plan_t *plan, cnmf_info_t *info);
/* The launch configuration cnmf_plan_create would choose on a B200 (148 SMs,
 * 227 KB shared memory) -- host-only, needs no device; used by the CPU tests. */
CNMF_API int cnmf_plan_dryrun(const cnmf_desc_t *desc, cnmf_info_t *info);

/* chi2 of (W, H) summed over all ranks of the plan's comm -> one double on the device. */
CNMF_API int cnmf_chi2(cnmf_plan_t *plan, const void *dX, const void *dV, const void *dW,
              const void *dH, double *d_out, void *stream);

/* Run up to max_iter iterations in place on dW, dH.
 * d_hist: device double[max_iter + 1].  tol <= 0: fixed count, the iterations are
 * enqueued without a host round trip.  tol > 0: the host reads the convergence flag
 * every check_every iterations (<= 0 selects 10).  Either way the call ends by reading
 * the 32-byte device state (one stream synchronise): a peer that never answered or a
 * non-finite chi2 is reported (CNMF_ERR_NCCL / CNMF_ERR_NONFINITE) and *n_iter_out is
 * the number of iterations actually completed.
 * update: CNMF_UPDATE_* (W-only / H-only fits keep the other factor fixed). */
CNMF_API int cnmf_solve(cnmf_plan_t *plan, const void *dX, const void *dV, void *dW, void *dH,
               double *d_hist, int32_t max_iter, double tol, int32_t check_every,
               int32_t update, int32_t *n_iter_out, void *stream);

/* chi2 of (W, H) on ONE GPU for callers that may not allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc) -> one double at d_out (device).  No counterpart in the reference.
 * Needs no plan; calls no cudaMalloc / cudaFree or any
 * other allocating routine: its scratch (one partial sum per 1024 rows + an arrival counter, 262 400
 * bytes) is a file-scope device array that every GPU loading the library holds until exit.  Every call
 * shares that array: calls that overlap on one GPU race -- keep them on one stream, or serialise them.
 * Limits: n_rows <= CNMF_CHI2_STATIC_MAX_ROWS and k * m * sizeof(T) <= CNMF_CHI2_STATIC_MAX_H_BYTES,
 * beyond which it returns CNMF_ERR_UNSUPPORTED (use cnmf_chi2 on a plan).  X, V: [n_rows][ld] with any
 * ld >= m (no alignment or padding requirement), W [n_rows][k], H [k][m].  The work is enqueued on
 * `stream`; d_out is valid once the stream has reached that point. */
#define CNMF_CHI2_STATIC_MAX_ROWS 33554432 /* 2^25 */
#define CNMF_CHI2_STATIC_MAX_H_BYTES 49152
CNMF_API int cnmf_chi2_static(int32_t dtype, const void *dX, const void *dV, const void *dW, const void *dH,
                     int64_t n_rows, int64_t ld, int32_t m, int32_t k, double *d_out, void *stream);

/* Device time of the fused row-streaming kernel launches of the last cnmf_solve that
 * carried CNMF_SOLVE_TIME_KERNELS: total milliseconds and number of launches.
 * Synchronises the events it recorded. */
CNMF_API int cnmf_plan_kernel_time(cnmf_plan_t *plan, double *ms_total, int32_t *n_launches);

/* Batched plans (desc.batch > 1): one CTA per problem, all iterations in-kernel.
 * d_hist: double[batch][max_iter + 1] (unused tail = NaN), d_niter: int32[batch]. */
CNMF_API int cnmf_solve_batched(cnmf_plan_t *plan, const void *dX, const void *dV, void *dW, void *dH,
                       double *d_hist, int32_t *d_niter, int32_t max_iter, double tol,
                       int32_t update, void *stream);

/* Device-side operand preparation: dst X,V [n_rows][dst_ld] from src X
 * [n_rows][src_ld] (+ optional weights of the same pitch, optional uint8 mask
 * [n_rows][mask_ld], nonzero = good).  V = weights * mask, X zero-filled where
 * V == 0, pad columns zeroed, *d_status |= CNMF_BAD_* for offending entries.
 * src and dst may not overlap unless identical with src_ld == dst_ld. */
CNMF_API int cnmf_prepare(int32_t dtype, const void *src_x, const void *src_w, const uint8_t *src_mask,
                 int64_t src_ld, int64_t mask_ld, void *dst_x, void *dst_v, int64_t dst_ld,
                 int64_t n_rows, int32_t m, int32_t *d_status, void *stream);

/* Host-buffer entry point (what cnmf.nmf calls for NumPy input): uploads
 * X / weights / mask in row chunks, prepares them on the device, solves and
 * downloads W, H, hist.  Host arrays are dense row-major ([n][m], [n][k],
 * [k][m]); weights and mask may be NULL.  hist: host double[max_iter + 1]. */
CNMF_API int cnmf_nmf_host(int32_t dtype, int64_t n_rows, int32_t m, int32_t k, const void *h_x,
                  const void *h_weights, const uint8_t *h_mask, void *h_w, void *h_h,
                  double *h_hist, int32_t max_iter, double tol, double eps, int32_t update,
                  cnmf_comm_t *comm, int32_t *n_iter_out, int64_t *h2d_bytes, int64_t *d2h_bytes);

/* Wall-clock seconds of the phases of this thread's last cnmf_nmf_host call (the end-to-end
 * number's breakdown): device allocation, upload + prepare, plan creation, solve, download,
 * whole call including the release of the device buffers.  Fills min(n, CNMF_HOST_PHASES). */
enum { CNMF_HOST_T_ALLOC = 0, CNMF_HOST_T_UPLOAD = 1, CNMF_HOST_T_PLAN = 2, CNMF_HOST_T_SOLVE = 3,
       CNMF_HOST_T_DOWNLOAD = 4, CNMF_HOST_T_TOTAL = 5, CNMF_HOST_PHASES = 6 };
CNMF_API int cnmf_host_timing(double *seconds, int32_t n);

/* Synthetic spectra on the device (rows row0 .. row0+n_rows of the set that
 * cnmf_b200/synth.py defines).  d_hstar: device double[k][m] basis. */
CNMF_API int cnmf_synth(int32_t dtype, void *dX, void *dV, int64_t n_rows, int64_t row0, int32_t m,
               int64_t ld, int32_t k, const double *d_hstar, uint64_t seed, int32_t vmode,
               double p_missing, void *stream);

/* NCCL communicator owned by the library (NCCL is dlopen-ed on first use).
 * Rank 0 makes the 128-byte id, the host layer broadcasts it (torch.distributed). */
CNMF_API int cnmf_comm_unique_id(uint8_t id_out[128]);
CNMF_API int cnmf_comm_create(cnmf_comm_t **comm, const uint8_t id[128], int32_t world_size, int32_t rank);
CNMF_API int cnmf_comm_destroy(cnmf_comm_t *comm);
CNMF_API int cnmf_comm_info(const cnmf_comm_t *comm, int32_t *world_size, int32_t *rank);

/* Peer-memory exchange: with the ranks of one NVLink / NVSwitch node,
 * the per-iteration sum of the 2*k*ld + 1 partials leaves NCCL -- one kernel stores each rank's
 * partials into every peer's window, waits for the peers' flags, sums in rank order and applies
 * the H update.  cnmf_comm_peer_export allocates this rank's window and returns its 64-byte CUDA
 * IPC handle; the host layer gathers the handles of all ranks (rank order, world_size x 64 bytes)
 * and passes them to cnmf_comm_peer_import; cnmf_comm_peer_enable(1) must then be called on
 * EVERY rank or on none (the host layer agrees on it), and before the first plan uses the
 * communicator.  Exchanges larger than 65536 doubles keep using ncclAllReduce. */
CNMF_API int cnmf_comm_peer_export(cnmf_comm_t *comm, uint8_t handle_out[64]);
CNMF_API int cnmf_comm_peer_import(cnmf_comm_t *comm, const uint8_t *handles);
CNMF_API int cnmf_comm_peer_enable(cnmf_comm_t *comm, int32_t on);
CNMF_API int cnmf_comm_peer_active(const cnmf_comm_t *comm);

/* Single-process group (the `devices=` keyword of cnmf.nmf): comms[r] is rank r of world_size ranks,
 * living on devices[r].  The peers' windows are reached through cudaDeviceEnablePeerAccess -- no NCCL,
 * no IPC -- and the exchange is switched on from the start.  A device may be listed more than once
 * ("virtual shards": the sharded schedule and the exchange kernel on one GPU; tests).  Each rank needs
 * its own host thread (with its device current) for plan creation and solves; destroy every rank with
 * cnmf_comm_destroy after all of them finished. */
CNMF_API int cnmf_comm_create_local(cnmf_comm_t **comms, int32_t world_size, const int32_t *devices);
/* The sharded schedule with ALL `world` ranks of a local group on the CURRENT device ("virtual shards":
 * how a one-GPU box tests the exchange arithmetic).  plans[r] was created with comms[r] of
 * cnmf_comm_create_local(comms, world, {dev, dev, ...}); the arrays hold each rank's operands.  One host
 * thread, one stream: per iteration every rank's partial reduction, ONE cooperative launch in which CTA r
 * runs rank r's exchange + H update (the CTAs that wait on each other are co-resident by construction),
 * then every rank's fused kernel.  n_iter_out: int32[world].  Separate per-rank solves that spin on one
 * another must not share a GPU -- cnmf.nmf(devices=...) therefore wants distinct devices. */
CNMF_API int cnmf_solve_group(cnmf_plan_t **plans, int32_t world, const void *const *dX, const void *const *dV,
                     void *const *dW, void *const *dH, double *const *d_hist, int32_t max_iter, double tol,
                     int32_t update, int32_t *n_iter_out, void *stream);
/* Upper bound on the wait for a peer's partial sums inside the exchange kernel (default 120 s).  When it
 * expires the solve ends with CNMF_ERR_NCCL on that rank and the communicator must be destroyed. */
CNMF_API int cnmf_comm_set_timeout(cnmf_comm_t *comm, double seconds);

/* number of kernels this library has launched in this process (bench's gpu_launches) */
CNMF_API int64_t cnmf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CNMF_B200_H */
