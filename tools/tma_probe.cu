// tma_probe.cu -- which tiled-TMA tile loads of a u8 image plane stack does this GPU accept?  (debugging aid for k_fast's staging)
//   nvcc -gencode arch=compute_100a,code=sm_100a -o build/tma_probe tools/tma_probe.cu && build/tma_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

struct alignas(64) Map
{
    unsigned char b[128];
};

template <int RANK>
__global__ void k_probe(const __grid_constant__ Map M, int boxw, int boxh, int x0, int y0, int z0, uint8_t* out)
{
    extern __shared__ __align__(128) uint8_t tile[];
    __shared__ __align__(8) unsigned long long bar;
    const unsigned sb = (unsigned)__cvta_generic_to_shared(&bar), st = (unsigned)__cvta_generic_to_shared(tile);
    if(threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(sb) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(sb), "r"(boxw * boxh) : "memory");
        if(RANK == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(st),
                         "l"(&M), "r"(x0), "r"(y0), "r"(z0), "r"(sb)
                         : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(st),
                         "l"(&M), "r"(x0), "r"(y0), "r"(sb)
                         : "memory");
    }
    __syncthreads();
    for(unsigned spins = 0;; ++spins)
    {
        unsigned ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(sb), "r"(0) : "memory");
        if(ok || spins > (1u << 20))
            break;
    }
    for(int i = threadIdx.x; i < boxw * boxh; i += blockDim.x)
        out[i] = tile[i];
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv)
{
    const int only_rank = argc > 1 ? atoi(argv[1]) : 0, only_boxw = argc > 2 ? atoi(argv[2]) : 0, only_x0 = argc > 3 ? atoi(argv[3]) : -1;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    EncodeFn encode = (EncodeFn)fn;
    const int pitch = 704, h = 480, frames = 2;
    const size_t fs = (size_t)pitch * h;
    std::vector<uint8_t> img(fs * frames);
    for(size_t i = 0; i < img.size(); ++i)
        img[i] = (uint8_t)(i * 131 + (i >> 9));
    uint8_t *d_img, *d_out;
    cudaMalloc(&d_img, img.size());
    cudaMalloc(&d_out, 256 * 256);
    cudaMemcpy(d_img, img.data(), img.size(), cudaMemcpyHostToDevice);
    const int ranks[] = {2, 3}, boxws[] = {128, 144, 160}, x0s[] = {32, 24}, promos[] = {0, 2};
    for(int rank : ranks)
        for(int boxw : boxws)
            for(int x0 : x0s)
                for(int promo : promos)
                {
                    if((only_rank && rank != only_rank) || (only_boxw && boxw != only_boxw) || (only_x0 >= 0 && x0 != only_x0))
                        continue;
                    const int boxh = 70, y0 = 27, z0 = 1;
                    Map m;
                    memset(&m, 0, sizeof(m));
                    const cuuint64_t dims[3] = {(cuuint64_t)pitch, (cuuint64_t)h, (cuuint64_t)frames};
                    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)fs};
                    const cuuint32_t box[3] = {(cuuint32_t)boxw, (cuuint32_t)boxh, 1u}, es[3] = {1, 1, 1};
                    const CUresult r = encode((CUtensorMap*)&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d_img + (rank == 2 ? fs * z0 : 0), dims, strides, box,
                                              es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo,
                                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    cudaMemset(d_out, 0xee, 256 * 256);
                    cudaError_t e = cudaSuccess;
                    if(r == CUDA_SUCCESS)
                    {
                        cudaFuncSetAttribute(k_probe<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
                        cudaFuncSetAttribute(k_probe<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
                        if(rank == 2)
                            k_probe<2><<<1, 256, boxw * boxh + 128>>>(m, boxw, boxh, x0, y0, z0, d_out);
                        else
                            k_probe<3><<<1, 256, boxw * boxh + 128>>>(m, boxw, boxh, x0, y0, z0, d_out);
                        e = cudaDeviceSynchronize();
                    }
                    int bad = -1;
                    if(r == CUDA_SUCCESS && e == cudaSuccess)
                    {
                        std::vector<uint8_t> out((size_t)boxw * boxh);
                        cudaMemcpy(out.data(), d_out, out.size(), cudaMemcpyDeviceToHost);
                        bad = 0;
                        for(int y = 0; y < boxh; ++y)
                            for(int x = 0; x < boxw; ++x)
                            {
                                const int gx = x0 + x, gy = y0 + y;
                                const uint8_t want = (gx < pitch && gy < h) ? img[fs * z0 + (size_t)gy * pitch + gx] : 0;
                                bad += out[(size_t)y * boxw + x] != want;
                            }
                    }
                    printf("rank %d boxw %3d x0 %2d l2promo %d : encode %d, run %s, %d wrong bytes\n", rank, boxw, x0, promo, (int)r,
                           cudaGetErrorName(e), bad);
                    if(e != cudaSuccess)
                    {
                        printf("(sticky error: stopping)\n");
                        return 1;
                    }
                }
    return 0;
}
