// tmem_a_rate.cu -- how fast does tcgen05.mma.kind::i8 M128 N128 K32 retire when A comes from TENSOR memory instead of shared memory?
// One issuing thread per SM, back-to-back MMAs into rotating accumulators, +1/-1 operands, no loads, no epilogue (as
// mslam_b200_measure_tensor_peak).   nvcc -gencode arch=compute_100a,code=sm_100a -o build/tmem_a_rate tools/tmem_a_rate.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

constexpr int M = 128, ROW = 256;
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr)
{
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(2048 >> 4) << 32) | (1ull << 46);
}
template <int N, bool TA>
__global__ void __launch_bounds__(128, 1) k_rate(int iters)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* sA = smem;
    uint8_t* sB = smem + M * ROW;
    __shared__ __align__(8) unsigned long long s_bar;
    __shared__ uint32_t s_tmem;
    for(int i = threadIdx.x; i < (M + N) * ROW / 4; i += blockDim.x)
    {
        unsigned h = (i + blockIdx.x * 7919u) * 2654435761u;
        h ^= h >> 15;
        reinterpret_cast<unsigned*>(smem)[i] = 0x01010101u + ((h & 1u) * 0xfeu) + (((h >> 1) & 1u) * 0xfe00u) + (((h >> 2) & 1u) * 0xfe0000u) +
                                               (((h >> 3) & 1u) * 0xfe000000u);
    }
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if(threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(threadIdx.x < 32)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    constexpr int A_COL = 384; // 64 columns of A (K = 256 bytes per row)
    if(TA)
    {
        const uint32_t la = tmem + ((uint32_t)((threadIdx.x >> 5) * 32) << 16) + A_COL;
        for(int c0 = 0; c0 < 64; c0 += 8)
        {
            const unsigned v = 0x01ff01ffu; // +1 / -1 bytes
            asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(la + c0), "r"(v) : "memory");
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if(threadIdx.x == 0)
    {
        constexpr uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t a_addr = (uint32_t)__cvta_generic_to_shared(sA), b_addr = (uint32_t)__cvta_generic_to_shared(sB);
        constexpr int NBUF = TA ? (384 / N) : (512 / N);
        for(int it = 0; it < iters; ++it)
        {
            const uint32_t d = tmem + (uint32_t)(it % NBUF) * N;
#pragma unroll
            for(int k = 0; k < ROW / 32; ++k)
            {
                if(TA)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                                 "r"(tmem + A_COL + k * 8), "l"(smem_desc(b_addr + k * 256)), "r"(idesc), "r"((uint32_t)(k > 0))
                                 : "memory");
                else
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                                 "l"(smem_desc(a_addr + k * 256)), "l"(smem_desc(b_addr + k * 256)), "r"(idesc), "r"((uint32_t)(k > 0))
                                 : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
        for(unsigned spins = 0;; ++spins)
        {
            uint32_t ok;
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(0) : "memory");
            if(ok || spins > (1u << 28))
                break;
        }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if(threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

template <int N, bool TA>
static void run(const char* what, int sms)
{
    const int smem = (M + N) * ROW;
    cudaFuncSetAttribute(k_rate<N, TA>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    const int iters = 8192;
    k_rate<N, TA><<<sms, 128, smem>>>(64);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for(int r = 0; r < 5; ++r)
        k_rate<N, TA><<<sms, 128, smem>>>(iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = 5.0 * sms * iters * (ROW / 32) * 2.0 * M * N * 32;
    printf("%-44s %s  %.0f TOP/s  (%.1f ms)\n", what, cudaGetErrorName(e), ops / (ms * 1e-3) / 1e12, ms);
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    run<128, false>("M128 N128 K32, A and B from shared memory", p.multiProcessorCount);
    run<128, true>("M128 N128 K32, A from tensor memory", p.multiProcessorCount);
    run<256, false>("M128 N256 K32, A and B from shared memory", p.multiProcessorCount);
    return 0;
}
