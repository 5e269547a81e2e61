// Round-2 microbenchmark C: tcgen05.mma kind::tf32 issue rate at the tile shapes the NMF contractions
// would use (M = 128 rows, N = 8 / 16 components or N = 64..256 columns, K = 8 per instruction), operands
// from shared-memory descriptors, accumulator in TMEM.  The data is garbage on purpose -- only the
// instruction's legality and its rate matter.  Prints cycles per MMA and MAC per cycle per SM, next to
// what the operand fetch from shared memory alone would allow (A: 128 x 8 x 4 B = 4 KB per instruction).
//
// Encodings follow the CUTLASS descriptors shipped in the image (cute/arch/mma_sm100_desc.hpp):
//   instruction descriptor: c_format[4,6) = 1 (F32), a_format[7,10) = b_format[10,13) = 2 (TF32),
//                           a_major[15] = b_major[16] = 0 (K-major), n_dim[17,23) = N >> 3, m_dim[24,29) = M >> 4
//   shared-memory descriptor: start[0,14) = addr >> 4, LBO[16,30) = bytes >> 4, SBO[32,46) = bytes >> 4,
//                             version[46,48) = 1, layout_type[61,64) = 0 (no swizzle)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

__global__ void __launch_bounds__(128) mma_rate(long long *cyc, int n_mma, int N, int same_a) {
    extern __shared__ __align__(1024) unsigned char sm[];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) ((float *)sm)[i] = 1.0f + (i & 7) * 0.125f;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const uint32_t tm = tmem_base;
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        // A: 128 rows x 8 tf32 (K-major, core matrices of 8 rows x 16 B): 16 row groups x 2 K chunks x 128 B
        // B: N rows x 8 tf32, same layout, behind A's 16 KB of slots
        const uint32_t a0 = smem_u32(sm), b0 = smem_u32(sm) + 16 * 1024;
        t0 = clock64();
        for (int i = 0; i < n_mma; ++i) {
            // walk A through 4 slots of 4 KB unless same_a (then every instruction re-reads one A tile)
            const uint32_t aoff = same_a ? 0u : (uint32_t)(i & 3) * 4096u;
            const uint64_t da = make_desc(a0 + aoff, 128, 256);
            const uint64_t db = make_desc(b0, 128, 256);
            const uint32_t acc = i > 0;
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.b64 [%0];" ::"l"((uint64_t)smem_u32(&bar)) : "memory");
        uint32_t ok;
        do {
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(ok)
                : "r"(smem_u32(&bar))
                : "memory");
        } while (!ok);
        t1 = clock64();
        cyc[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    printf("%s, %d SMs\n", p.name, sms);
    long long *cyc;
    cudaMalloc(&cyc, sms * 8);
    const int smem = 48 * 1024;
    cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int n_mma = 20000;
    for (int same_a = 0; same_a < 2; ++same_a)
        for (int N : {8, 16, 32, 64, 128, 256}) {
            mma_rate<<<sms, 128, smem>>>(cyc, n_mma, N, same_a);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
                printf("N=%d: %s\n", N, cudaGetErrorString(e));
                return 1;
            }
            long long h[256];
            cudaMemcpy(h, cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
            double mean = 0;
            for (int i = 0; i < sms; ++i) mean += (double)h[i];
            mean /= sms;
            const double per = mean / n_mma;
            printf("tcgen05.mma kind::tf32 M=128 N=%3d K=8 (%s): %7.2f cycles per MMA, %7.1f MAC/cycle/SM; "
                   "operand bytes per MMA %5d -> %6.1f B/cycle from shared memory\n",
                   N, same_a ? "one A tile re-read" : "A walks 4 tiles  ", per, 128.0 * N * 8 / per, (128 + N) * 32,
                   (128 + N) * 32 / per);
        }
    return 0;
}
