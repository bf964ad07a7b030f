// tmem_a_probe.cu -- tcgen05.mma kind::i8 with the A operand in TENSOR MEMORY: which layout does it expect?  One M128 N64 K32 MMA;
// A written by tcgen05.st (lane = row, column c = bytes 4c .. 4c + 3 of the row's 32 K bytes, little endian), B in shared memory
// in the matcher's no-swizzle K-major core-matrix layout; D compared with the integer dot products on the host.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o build/tmem_a_probe tools/tmem_a_probe.cu && build/tmem_a_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

constexpr int M = 128, N = 64, KB = 256; // KB: bytes of K the smem layout is sized for (the matcher's rows); the MMA uses 32

__device__ __forceinline__ uint64_t smem_desc(uint32_t addr)
{
    return (uint64_t)((addr >> 4) & 0x3fffu) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(2048 >> 4) << 32) | (1ull << 46);
}

__global__ void k_probe(const int8_t* __restrict__ a, const uint8_t* __restrict__ b_canon, int* __restrict__ out, int kstep)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) unsigned long long s_bar;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for(int i = threadIdx.x; i < N * KB / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(b_canon)[i];
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    if(threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if(warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(128) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy smem writes -> visible to the tensor core's reads
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
    // A: row r = 32 * warp + lane, its 32 K bytes of this K step as 8 columns at column 64
    {
        const int r = 32 * warp + lane;
        uint32_t v[8];
        for(int c = 0; c < 8; ++c)
        {
            uint32_t w = 0;
            for(int j = 0; j < 4; ++j)
                w |= (uint32_t)(uint8_t)a[r * KB + kstep * 32 + 4 * c + j] << (8 * j);
            v[c] = w;
        }
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(lane_addr + 64), "r"(v[0]), "r"(v[1]),
                     "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                     : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if(threadIdx.x == 0)
    {
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint64_t bdesc = smem_desc((uint32_t)__cvta_generic_to_shared(smem) + kstep * 256);
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem),
                     "r"(tmem + 64), "l"(bdesc), "r"(idesc), "r"(0u)
                     : "memory");
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    }
    for(unsigned spins = 0;; ++spins)
    {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(0) : "memory");
        if(ok || spins > (1u << 22))
            break;
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t d[32];
    for(int half = 0; half < 2; ++half)
    {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]), "=r"(d[9]),
                       "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15]), "=r"(d[16]), "=r"(d[17]), "=r"(d[18]),
                       "=r"(d[19]), "=r"(d[20]), "=r"(d[21]), "=r"(d[22]), "=r"(d[23]), "=r"(d[24]), "=r"(d[25]), "=r"(d[26]), "=r"(d[27]),
                       "=r"(d[28]), "=r"(d[29]), "=r"(d[30]), "=r"(d[31])
                     : "r"(lane_addr + 32 * half));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for(int c = 0; c < 32; ++c)
            out[(32 * warp + lane) * N + 32 * half + c] = (int)d[c];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if(warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128) : "memory");
}

int main()
{
    static int8_t a[M * KB], b[N * KB];
    static uint8_t bc[N * KB];
    for(int r = 0; r < M; ++r)
        for(int k = 0; k < KB; ++k)
            a[r * KB + k] = (int8_t)(((r * 7 + k * 3) % 5) - 2);
    for(int n = 0; n < N; ++n)
        for(int k = 0; k < KB; ++k)
        {
            b[n * KB + k] = (int8_t)(((n * 5 + k) % 7) - 3);
            bc[(n / 8) * 2048 + (k / 16) * 128 + (n % 8) * 16 + (k % 16)] = (uint8_t)b[n * KB + k];
        }
    int8_t* d_a;
    uint8_t* d_b;
    int* d_o;
    cudaMalloc(&d_a, sizeof(a)), cudaMalloc(&d_b, sizeof(bc)), cudaMalloc(&d_o, M * N * 4);
    cudaMemcpy(d_a, a, sizeof(a), cudaMemcpyHostToDevice), cudaMemcpy(d_b, bc, sizeof(bc), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, N * KB);
    for(int kstep : {0, 3})
    {
        cudaMemset(d_o, 0xee, M * N * 4);
        k_probe<<<1, 128, N * KB>>>(d_a, d_b, d_o, kstep);
        cudaError_t e = cudaDeviceSynchronize();
        static int h[M * N];
        cudaMemcpy(h, d_o, sizeof(h), cudaMemcpyDeviceToHost);
        int bad = 0, first = -1;
        for(int r = 0; r < M; ++r)
            for(int n = 0; n < N; ++n)
            {
                int want = 0;
                for(int k = 0; k < 32; ++k)
                    want += a[r * KB + kstep * 32 + k] * b[n * KB + kstep * 32 + k];
                if(h[r * N + n] != want)
                {
                    if(first < 0)
                        first = r * N + n;
                    ++bad;
                }
            }
        printf("k step %d: run %s, %d of %d wrong", kstep, cudaGetErrorName(e), bad, M * N);
        if(first >= 0)
            printf(" (first: row %d col %d got %d)", first / N, first % N, h[first]);
        printf("\n");
        if(e != cudaSuccess)
            return 1;
    }
    return 0;
}
