// tmem_pack_probe.cu -- what does tcgen05.ld ... .pack::16b return for 32-bit accumulator columns?  (design aid for the database
// scoring epilogue: if it returns the LOW 16 bits of two adjacent columns in one register, the epilogue's packing IMADs and half
// of its tensor-memory read traffic go away)
//   nvcc -gencode arch=compute_100a,code=sm_100a -o build/tmem_pack_probe tools/tmem_pack_probe.cu && build/tmem_pack_probe
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__global__ void k_probe(uint32_t* out)
{
    __shared__ uint32_t s_tmem;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if(warp == 0)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = s_tmem;
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
    // column c of lane L holds 0xA000'0000 + (c << 20) | ((L + 32 * warp) << 8) | c   (high half differs from low half)
    uint32_t v[8];
    for(int c0 = 0; c0 < 32; c0 += 8)
    {
        for(int i = 0; i < 8; ++i)
            v[i] = 0xA0000000u + ((uint32_t)(c0 + i) << 20) + ((uint32_t)(lane + 32 * warp) << 8) + (uint32_t)(c0 + i);
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr + c0), "r"(v[0]), "r"(v[1]),
                     "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                     : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.pack::16b.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for(int i = 0; i < 8; ++i)
        out[threadIdx.x * 8 + i] = r[i];
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if(warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
}

int main()
{
    uint32_t* d;
    cudaMalloc(&d, 128 * 8 * 4);
    cudaMemset(d, 0, 128 * 8 * 4);
    k_probe<<<1, 128>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("run: %s\n", cudaGetErrorName(e));
    uint32_t h[128 * 8];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    for(int t : {0, 1, 37})
    {
        printf("thread %3d:", t);
        for(int i = 0; i < 8; ++i)
            printf(" %08x", h[t * 8 + i]);
        printf("\n");
    }
    return 0;
}
