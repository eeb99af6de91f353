// minimal TMA experiment: which way of passing a CUtensorMap works here?   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/tma_dbg tools/tma_dbg.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
typedef CUresult (*enc_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                           CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
struct Args { CUtensorMap tm; int x, y, z; uint32_t* out; };
__device__ void body(const CUtensorMap* tm, int x, int y, int z, uint32_t* out)
{
    __shared__ __align__(128) uint32_t tile[40 * 32];
    __shared__ __align__(8) unsigned long long mbar;
    const unsigned mb = (unsigned)__cvta_generic_to_shared(&mbar), dst = (unsigned)__cvta_generic_to_shared(tile);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(40 * 128) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(dst), "l"(reinterpret_cast<unsigned long long>(tm)), "r"(x), "r"(y), "r"(z), "r"(mb) : "memory");
    }
    __syncthreads();
    unsigned done = 0;
    for (int spin = 0; spin < (1 << 22) && !done; spin++)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mb) : "memory");
    for (int i = threadIdx.x; i < 40 * 32; i += blockDim.x) out[i] = done ? tile[i] : 0xdeadbeefu;
}
__global__ void k_nested(const __grid_constant__ Args a) { body(&a.tm, a.x, a.y, a.z, a.out); }
__global__ void k_top(const __grid_constant__ CUtensorMap tm, int x, int y, int z, uint32_t* out) { body(&tm, x, y, z, out); }
__global__ void k_glob(const CUtensorMap* tm, int x, int y, int z, uint32_t* out) { body(tm, x, y, z, out); }
int main(int argc, char** argv)
{
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    const int W = 640, H = 480, F = 3, P = 640;
    uint8_t* img; cudaMalloc(&img, (size_t)P * H * F);
    uint8_t* h = (uint8_t*)malloc((size_t)P * H * F);
    for (size_t i = 0; i < (size_t)P * H * F; i++) h[i] = (uint8_t)(i * 7 + (i >> 9));
    cudaMemcpy(img, h, (size_t)P * H * F, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    printf("entry point: %s q=%d p=%p\n", cudaGetErrorString(e), (int)q, p);
    Args a; memset(&a, 0, sizeof a);
    const cuuint64_t dims[3] = {W, H, F}; const cuuint64_t str[2] = {P, (cuuint64_t)P * H};
    const cuuint32_t box[3] = {128, 40, 1}, es[3] = {1, 1, 1};
    CUresult r = ((enc_fn)p)(&a.tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, img, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d sizeof(Args)=%zu align=%zu\n", (int)r, sizeof(Args), alignof(Args));
    uint32_t* out; cudaMalloc(&out, 40 * 128);
    uint32_t* ho = (uint32_t*)malloc(40 * 128);
    CUtensorMap* dtm; cudaMalloc(&dtm, sizeof(CUtensorMap)); cudaMemcpy(dtm, &a.tm, sizeof(CUtensorMap), cudaMemcpyHostToDevice);
    const int coords[4][3] = {{116, 28, 1}, {-4, -4, 0}, {596, 444, 2}, {117, 3, 1}};
    for (int variant = 0; variant < 3; variant++) {
        if (only >= 0 && variant != only) continue;
        for (int c = 0; c < 4; c++) {
            const int x = coords[c][0], y = coords[c][1], z = coords[c][2];
            a.x = x; a.y = y; a.z = z; a.out = out;
            cudaMemset(out, 0, 40 * 128);
            if (variant == 0) k_nested<<<1, 128>>>(a);
            else if (variant == 1) k_top<<<1, 128>>>(a.tm, x, y, z, out);
            else k_glob<<<1, 128>>>(dtm, x, y, z, out);
            e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("variant %d coord %d: %s\n", variant, c, cudaGetErrorString(e)); return 1; }
            cudaMemcpy(ho, out, 40 * 128, cudaMemcpyDeviceToHost);
            int bad = 0; const uint8_t* ob = (const uint8_t*)ho;
            for (int r2 = 0; r2 < 40; r2++) for (int cc = 0; cc < 128; cc++) {
                const int gx = x + cc, gy = y + r2;
                const uint8_t want = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? h[(size_t)z * P * H + (size_t)gy * P + gx] : 0;
                bad += ob[r2 * 128 + cc] != want;
            }
            printf("variant %d coord (%d,%d,%d): first word %08x mismatches %d\n", variant, x, y, z, ho[0], bad);
        }
    }
    return 0;
}
