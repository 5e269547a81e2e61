// Shared device helpers: 8-byte "packs" (two fp32 columns handled by one
// FFMA2, or one fp64 column), 16-byte vectors of two packs, mbarrier and
// bulk-copy (TMA engine, SASS UBLKCP) wrappers.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace cnmf {

// ----------------------------------------------------------------------------
// Pack arithmetic.  A pack is 8 bytes of one row: float2 (columns j, j+1) or one
// double.  fp32 packs go through the packed f32x2 PTX ops (SASS FFMA2), which
// halve the issue slots of the 6k+7 FMAs per element this path needs.
// ----------------------------------------------------------------------------
template <typename T> struct Pk;

template <> struct Pk<float> {
    typedef float2 P;
    static constexpr int W = 2;  // columns per pack
    static __device__ __forceinline__ unsigned long long u(const P &a) {
        return *reinterpret_cast<const unsigned long long *>(&a);
    }
    static __device__ __forceinline__ P f(unsigned long long a) {
        return *reinterpret_cast<P *>(&a);
    }
    static __device__ __forceinline__ P fma(const P &a, const P &b, const P &c) {
        unsigned long long r;
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(u(a)), "l"(u(b)), "l"(u(c)));
        return f(r);
    }
    static __device__ __forceinline__ P mul(const P &a, const P &b) {
        unsigned long long r;
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(u(a)), "l"(u(b)));
        return f(r);
    }
    static __device__ __forceinline__ P add(const P &a, const P &b) {
        unsigned long long r;
        asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(u(a)), "l"(u(b)));
        return f(r);
    }
    static __device__ __forceinline__ P sub(const P &a, const P &b) {
        unsigned long long r;
        asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(u(a)), "l"(u(b)));
        return f(r);
    }
    static __device__ __forceinline__ P dup(float s) { return make_float2(s, s); }
    static __device__ __forceinline__ P zero() { return make_float2(0.f, 0.f); }
    static __device__ __forceinline__ float hsum(const P &a) { return a.x + a.y; }
    static __device__ __forceinline__ double dsum(const P &a) { return (double)a.x + (double)a.y; }
    static __device__ __forceinline__ float lane(const P &a, int i) { return i ? a.y : a.x; }
    static __device__ __forceinline__ float first(const P &a) { return a.x; }
};

template <> struct Pk<double> {
    typedef double P;
    static constexpr int W = 1;
    static __device__ __forceinline__ P fma(P a, P b, P c) { return ::fma(a, b, c); }
    static __device__ __forceinline__ P mul(P a, P b) { return a * b; }
    static __device__ __forceinline__ P add(P a, P b) { return a + b; }
    static __device__ __forceinline__ P sub(P a, P b) { return a - b; }
    static __device__ __forceinline__ P dup(double s) { return s; }
    static __device__ __forceinline__ P zero() { return 0.0; }
    static __device__ __forceinline__ double hsum(P a) { return a; }
    static __device__ __forceinline__ double dsum(P a) { return a; }
    static __device__ __forceinline__ double lane(P a, int) { return a; }
    static __device__ __forceinline__ double first(P a) { return a; }
};

// 16 bytes = two packs; the unit of every shared-memory load on the hot path.
template <typename T> struct alignas(16) Vec {
    typename Pk<T>::P p[2];
};

template <typename T> __device__ __forceinline__ Vec<T> lds_vec(const void *smem) {
    Vec<T> v;
    const uint4 r = *reinterpret_cast<const uint4 *>(smem);
    *reinterpret_cast<uint4 *>(&v) = r;
    return v;
}

// ----------------------------------------------------------------------------
// mbarrier + bulk async copy (global -> shared through the TMA engine)
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// same, on precomputed 32-bit shared addresses (keeps the S2R/cvta out of the tile loops)
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_a(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!ok);
}
// Whole-warp wait that leaves the warp CONVERGED: the lanes' try_wait results can differ, and
// lanes that leave a per-lane loop at different times keep running as separate partial warps
// (__syncwarp orders them but does not merge them: ncu showed lane 0 walking 37 % of the row
// chunks on its own, every instruction issued twice).  The loop exit is a warp vote.
__device__ __forceinline__ void mbar_wait_warp_a(uint32_t bar, uint32_t parity, unsigned mask = 0xffffffffu) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!__all_sync(mask, ok != 0));
}
// Same, with a suspend-time hint (ns): the hardware parks the warp until the phase completes or the hint runs out, instead
// of returning after its short default limit.  A column-owner warp that waits 15-40 % of its time on W' otherwise spends
// that time in a six-instruction poll loop -- 25 % of all instructions the tensor-path kernel issued at the bench size
// (profiles/r2b_ncu_bench_f64_mma.txt), issue slots and power taken from the warps that have work.
__device__ __forceinline__ void mbar_wait_warp_hint_a(uint32_t bar, uint32_t parity, uint32_t hint_ns) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity), "r"(hint_ns)
            : "memory");
    } while (!__all_sync(0xffffffffu, ok != 0));
}
// one try_wait (suspends up to the hardware's time limit, then reports)
__device__ __forceinline__ bool mbar_try_a(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// non-blocking probe of a phase (issued early, consumed after the arithmetic it overlaps)
__device__ __forceinline__ bool mbar_test_a(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// release-increment / acquire-load of a 32-bit shared-memory counter (MEMBAR.ALL.CTA + ATOMS / LDS)
__device__ __forceinline__ void red_release_add_a(uint32_t addr) {
    asm volatile("red.release.cta.shared::cta.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}
__device__ __forceinline__ unsigned int atom_add_acqrel_a(uint32_t addr) {
    unsigned int old;
    asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(addr) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t lds_acquire_u32_a(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void bulk_g2s_a(uint32_t dst_smem, const void *src_gmem, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_smem),
                 "l"(src_gmem), "r"(bytes), "r"(bar)
                 : "memory");
}
// bulk prefetch into L2 (no shared memory involved): shortens the time a stage sits reserved
// for an in-flight load, which is what bounds the rows-in-flight pipeline
__device__ __forceinline__ void bulk_prefetch_l2(const void *src_gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src_gmem), "r"(bytes) : "memory");
}
// 1-D bulk copy; dst, src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Device-resident solver state shared by the per-iteration kernels.
struct State {
    int k1_count;   // fused passes launched so far (incremented by the stream kernel)
    int done;       // set by finalize when the tolerance test fires
    int n_iter;     // iterations completed when done was set
    int nonfinite;  // chi2 was NaN/Inf at some point
    int peer_timeout;  // a peer's partial sums did not arrive (exchange_finalize_kernel)
};

}  // namespace cnmf
