// The two per-tile phases of the fused iteration as device functions on a tile
// that sits in shared memory (used by the batched kernel; the streaming kernel
// carries the same arithmetic inline in its warp-specialised roles).
#pragma once
#include "common.cuh"

namespace cnmf {

// Phase 1 for one row shared by G lanes (lane index `sub`): returns the new W row
// in wn[] on every one of the G lanes.  xrow / vrow / Hs are shared-memory
// pointers; ldv = 16-byte vectors per row; Hs is [K][ldv] vectors.
template <typename T, int K, int G>
__device__ __forceinline__ void phase1_row(const unsigned char *xrow, const unsigned char *vrow,
                                           const unsigned char *Hs, int ldv, int sub, const T (&w)[K], T eps,
                                           T (&wn)[K]) {
    typedef Pk<T> O;
    typedef typename O::P P;
    P wd[K], num[K], den[K];
#pragma unroll
    for (int a = 0; a < K; ++a) {
        wd[a] = O::dup(w[a]);
        num[a] = O::zero();
        den[a] = O::zero();
    }
#pragma unroll 1
    for (int q = sub; q < ldv; q += G) {
        const Vec<T> x = lds_vec<T>(xrow + q * 16);
        const Vec<T> v = lds_vec<T>(vrow + q * 16);
        Vec<T> h[K];
#pragma unroll
        for (int a = 0; a < K; ++a) h[a] = lds_vec<T>(Hs + ((long long)a * ldv + q) * 16);
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const P vx = O::mul(v.p[e], x.p[e]);
            P wh = O::mul(wd[0], h[0].p[e]);
#pragma unroll
            for (int a = 1; a < K; ++a) wh = O::fma(wd[a], h[a].p[e], wh);
            const P vwh = O::mul(v.p[e], wh);
#pragma unroll
            for (int a = 0; a < K; ++a) {
                num[a] = O::fma(vx, h[a].p[e], num[a]);
                den[a] = O::fma(vwh, h[a].p[e], den[a]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < K; ++a) {
        T nu = O::hsum(num[a]), de = O::hsum(den[a]);
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) {
            nu += __shfl_xor_sync(0xffffffffu, nu, o);
            de += __shfl_xor_sync(0xffffffffu, de, o);
        }
        wn[a] = (w[a] * nu) / (de > eps ? de : eps);
    }
}

// Phase 1 for NR rows at once on the same G lanes: every H vector a lane loads serves all of them
// (the resident kernel's form: (2 NR + K) / NR shared-memory vectors per X vector instead of 2 + K).
template <typename T, int K, int G, int NR>
__device__ __forceinline__ void phase1_rows(const unsigned char *const (&xrow)[NR], const unsigned char *const (&vrow)[NR],
                                            const unsigned char *Hs, int ldv, int sub, const T (&w)[NR][K], T eps,
                                            T (&wn)[NR][K]) {
    typedef Pk<T> O;
    typedef typename O::P P;
    P wd[NR][K], num[NR][K], den[NR][K];
#pragma unroll
    for (int j = 0; j < NR; ++j)
#pragma unroll
        for (int a = 0; a < K; ++a) {
            wd[j][a] = O::dup(w[j][a]);
            num[j][a] = O::zero();
            den[j][a] = O::zero();
        }
#pragma unroll 1
    for (int q = sub; q < ldv; q += G) {
        Vec<T> h[K];
#pragma unroll
        for (int a = 0; a < K; ++a) h[a] = lds_vec<T>(Hs + ((long long)a * ldv + q) * 16);
#pragma unroll
        for (int j = 0; j < NR; ++j) {
            const Vec<T> x = lds_vec<T>(xrow[j] + q * 16);
            const Vec<T> v = lds_vec<T>(vrow[j] + q * 16);
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const P vx = O::mul(v.p[e], x.p[e]);
                P wh = O::mul(wd[j][0], h[0].p[e]);
#pragma unroll
                for (int a = 1; a < K; ++a) wh = O::fma(wd[j][a], h[a].p[e], wh);
                const P vwh = O::mul(v.p[e], wh);
#pragma unroll
                for (int a = 0; a < K; ++a) {
                    num[j][a] = O::fma(vx, h[a].p[e], num[j][a]);
                    den[j][a] = O::fma(vwh, h[a].p[e], den[j][a]);
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NR; ++j)
#pragma unroll
        for (int a = 0; a < K; ++a) {
            T nu = O::hsum(num[j][a]), de = O::hsum(den[j][a]);
#pragma unroll
            for (int o = G / 2; o > 0; o >>= 1) {
                nu += __shfl_xor_sync(0xffffffffu, nu, o);
                de += __shfl_xor_sync(0xffffffffu, de, o);
            }
            wn[j][a] = (w[j][a] * nu) / (de > eps ? de : eps);
        }
}

// Phase 2 for one row and one pack of columns: accumulates chi2 and the k
// numerator / denominator partials of the H update.  wd[] are W' packs
// (duplicated halves in fp32), Hr[] this thread's H packs.
template <typename T, int K>
__device__ __forceinline__ void phase2_pack(typename Pk<T>::P x, typename Pk<T>::P v,
                                            const typename Pk<T>::P (&wd)[K], const typename Pk<T>::P (&Hr)[K],
                                            typename Pk<T>::P (&nH)[K], typename Pk<T>::P (&dH)[K],
                                            typename Pk<T>::P &chi) {
    typedef Pk<T> O;
    typedef typename O::P P;
    const P vx = O::mul(v, x);
    P wh = O::mul(wd[0], Hr[0]);
#pragma unroll
    for (int a = 1; a < K; ++a) wh = O::fma(wd[a], Hr[a], wh);
    const P res = O::sub(x, wh);
    const P vr = O::mul(v, res);
    chi = O::fma(vr, res, chi);
    const P vwh = O::mul(v, wh);
#pragma unroll
    for (int a = 0; a < K; ++a) {
        nH[a] = O::fma(wd[a], vx, nH[a]);
        dH[a] = O::fma(wd[a], vwh, dH[a]);
    }
}

}  // namespace cnmf
