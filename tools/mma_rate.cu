// mma_rate.cu -- how fast does tcgen05.mma kind::i8 run on a B200 when the B
// operand is N-major (the digit planes as they are) compared with K-major?
// One CTA per SM, operands resident in shared memory (no loads in the timed
// loop), one thread issues ITER back-to-back 128 x N x 32 MMAs and waits for
// the commit; cycles per MMA from clock64.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/bin/mma_rate tools/mma_rate.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e = (x);                                                          \
        if (e != cudaSuccess) {                                                       \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                  \
        }                                                                             \
    } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t it = 0; it < (1u << 26); ++it) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (done) return;
    }
    __trap();
}

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61; /* SWIZZLE_128B */
    return d;
}

// mode 0: B N-major (as dense_tc.cu), mode 1: B K-major
template <int N, int MODE>
__global__ void __launch_bounds__(128) k_rate(long long* cycles, int iters) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int warp = threadIdx.x >> 5;
    /* defined operand bytes (values do not matter for the rate) */
    for (int i = threadIdx.x; i < (128 * 128 + 128 * N) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * (i & 3);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(N));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;");
        asm volatile("fence.proxy.async.shared::cta;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem_base = tmem_base_s;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (2u << 4) | ((MODE == 0 ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t aA = smem_u32(smem), aB = aA + 128 * 128;
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                const uint64_t da = make_desc(aA + k4 * 32, 16, 1024);
                const uint64_t db = MODE == 0 ? make_desc(aB + k4 * 32 * 128, N > 128 ? 16384 : 16, 1024) : make_desc(aB + k4 * 32, 16, 1024);
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}" ::"r"(tmem_base),
                    "l"(da), "l"(db), "r"(idesc), "r"(it | k4), "r"(0), "r"(0), "r"(0), "r"(0)
                    : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        mbar_wait(&bar, 0);
        const long long t1 = clock64();
        cycles[blockIdx.x] = t1 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(N));
}

template <int N, int MODE>
static void run(int ctas, int iters, long long* d_cycles) {
    const int smem = 128 * 128 + 128 * N + 1024;
    CK(cudaFuncSetAttribute(k_rate<N, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    k_rate<N, MODE><<<ctas, 128, smem>>>(d_cycles, iters);
    CK(cudaDeviceSynchronize());
    k_rate<N, MODE><<<ctas, 128, smem>>>(d_cycles, iters);
    CK(cudaDeviceSynchronize());
    long long* h = (long long*)malloc(sizeof(long long) * ctas);
    CK(cudaMemcpy(h, d_cycles, sizeof(long long) * ctas, cudaMemcpyDeviceToHost));
    long long mx = 0;
    for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
    const double per = (double)mx / (iters * 4.0);
    printf("{\"N\": %d, \"b_major\": \"%s\", \"ctas\": %d, \"cycles_per_mma_128xNx32\": %.1f, \"mac_per_clk_per_sm\": %.0f}\n", N,
           MODE == 0 ? "N" : "K", ctas, per, 128.0 * N * 32 / per);
    free(h);
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int ctas = prop.multiProcessorCount;
    long long* d;
    CK(cudaMalloc(&d, sizeof(long long) * ctas));
    run<128, 0>(ctas, 2000, d);
    run<128, 1>(ctas, 2000, d);
    run<256, 0>(ctas, 2000, d);
    run<256, 1>(ctas, 2000, d);
    run<64, 0>(ctas, 2000, d);
    run<64, 1>(ctas, 2000, d);
    run<128, 0>(1, 2000, d);
    run<128, 1>(1, 2000, d);
    cudaFree(d);
    return 0;
}
