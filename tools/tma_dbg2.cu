// bisect: v0 = bulk (non-tensor) copy; v1 = 2-D tensor aligned coords; v2 = 2-D unaligned x; v3 = 3-D aligned; v4 = 3-D unaligned
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap tm, const uint8_t* img, int mode, int x, int y, int z, uint32_t* out)
{
    __shared__ __align__(128) uint32_t tile[40 * 32];
    __shared__ __align__(8) unsigned long long mbar;
    const unsigned mb = (unsigned)__cvta_generic_to_shared(&mbar), dst = (unsigned)__cvta_generic_to_shared(tile);
    const unsigned bytes = mode == 0 ? 4096 : 40 * 128;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes) : "memory");
        if (mode == 0)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(img), "r"(bytes), "r"(mb) : "memory");
        else if (mode == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(&tm)), "r"(x), "r"(y), "r"(mb) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(&tm)), "r"(x), "r"(y), "r"(z), "r"(mb) : "memory");
    }
    __syncthreads();
    unsigned done = 0;
    for (int spin = 0; spin < (1 << 22) && !done; spin++)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mb) : "memory");
    for (int i = threadIdx.x; i < 40 * 32; i += blockDim.x) out[i] = done ? tile[i] : 0xdeadbeefu;
}
int main(int argc, char** argv)
{
    const int v = argc > 1 ? atoi(argv[1]) : 0;
    const int W = 640, H = 480, F = 3, P = 640;
    uint8_t* img; cudaMalloc(&img, (size_t)P * H * F);
    uint8_t* h = (uint8_t*)malloc((size_t)P * H * F);
    for (size_t i = 0; i < (size_t)P * H * F; i++) h[i] = (uint8_t)(i * 7 + (i >> 9));
    cudaMemcpy(img, h, (size_t)P * H * F, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    CUtensorMap tm; memset(&tm, 0, sizeof tm);
    const int rank = (v == 1 || v == 2) ? 2 : 3;
    const cuuint64_t dims[3] = {W, H, F}; const cuuint64_t str[2] = {P, (cuuint64_t)P * H};
    const cuuint32_t box[3] = {128, 40, 1}, es[3] = {1, 1, 1};
    CUresult r = ((enc_fn)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, img, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("v=%d encode rc=%d rank=%d\n", v, (int)r, rank);
    uint32_t* out; cudaMalloc(&out, 40 * 128);
    uint32_t* ho = (uint32_t*)malloc(40 * 128);
    const int x = (v == 2 || v == 4) ? 116 : 128, y = 32, z = rank == 3 ? 1 : 0;
    k<<<1, 128>>>(tm, img, v == 0 ? 0 : rank, x, y, z, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("v=%d: %s\n", v, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(ho, out, 40 * 128, cudaMemcpyDeviceToHost);
    int bad = 0; const uint8_t* ob = (const uint8_t*)ho;
    if (v == 0) { for (int i = 0; i < 4096; i++) bad += ob[i] != h[i]; }
    else for (int r2 = 0; r2 < 40; r2++) for (int cc = 0; cc < 128; cc++) {
        const int gx = x + cc, gy = y + r2;
        const uint8_t want = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? h[(size_t)z * P * H + (size_t)gy * P + gx] : 0;
        bad += ob[r2 * 128 + cc] != want;
    }
    printf("v=%d first word %08x mismatches %d\n", v, ho[0], bad);
    return 0;
}
