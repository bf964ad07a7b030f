// alu_probe.cu -- issue rates of the instructions the matcher's epilogue is made of (register-only loops, all SMs):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/alu_probe tools/alu_probe.cu && build/alu_probe
#include <cuda_runtime.h>
#include <stdio.h>

template <int OP>
__global__ void __launch_bounds__(256) k(unsigned* out, int iters, unsigned seed, unsigned mul)
{
    unsigned v[8], w[8];
#pragma unroll
    for(int i = 0; i < 8; ++i)
        v[i] = threadIdx.x * 2654435761u + i * 40503u + seed, w[i] = v[i] ^ 0x9e3779b9u;
#pragma unroll 4
    for(int it = 0; it < iters; ++it)
    {
#pragma unroll
        for(int i = 0; i < 8; ++i)
        {
            if(OP == 0) v[i] = __vminu2(v[i], w[i] + it);                 // VIMNMX.U16x2 (+ an IADD feeding it)
            if(OP == 1) v[i] = __vimin3_u16x2(v[i], w[i], w[(i + 1) & 7]) ^ it; // VIMNMX3.U16x2
            if(OP == 2) v[i] = min(v[i], w[i] + it);                     // VIMNMX.U32
            if(OP == 3) v[i] = __vimin3_u32(v[i], w[i], w[(i + 1) & 7]) ^ it;   // VIMNMX3.U32
            if(OP == 4) v[i] = v[i] * mul + w[i];                        // IMAD
            if(OP == 5) v[i] = (v[i] ^ w[i]) + it;                       // LOP3 + IADD reference
        }
    }
    unsigned s = 0;
#pragma unroll
    for(int i = 0; i < 8; ++i)
        s ^= v[i];
    if(s == 0x12345678u)
        out[0] = s;
}

template <int OP>
double run(int sms, const char* name, double extra)
{
    unsigned* d;
    cudaMalloc(&d, 64);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 8192, grid = sms * 8;
    float ms = 0;
    for(int r = 0; r < 3; ++r)
    {
        cudaEventRecord(e0);
        k<OP><<<grid, 256>>>(d, iters, 17u, 3u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
    }
    const double per_sm_clk = (double)grid * 256 * iters * 8 / (ms * 1e-3) / sms / 1.9e9;
    printf("%-34s %8.3f ms  %6.1f lane-results / clk / SM (at 1.9 GHz)  [loop also issues %.0f other op(s) per result]\n", name, ms, per_sm_clk, extra);
    cudaFree(d);
    return per_sm_clk;
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    run<5>(sms, "LOP3 + IADD (reference)", 1);
    run<0>(sms, "VIMNMX.U16x2 (+ IADD)", 1);
    run<1>(sms, "VIMNMX3.U16x2 (+ LOP)", 1);
    run<2>(sms, "VIMNMX.U32 (+ IADD)", 1);
    run<3>(sms, "VIMNMX3.U32 (+ LOP)", 1);
    run<4>(sms, "IMAD", 0);
    return 0;
}
