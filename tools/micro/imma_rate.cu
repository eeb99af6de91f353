// Microbenchmark: legacy mma.sync m16n8k32 u8 (IMMA.16832) issue rate on sm_100a, register operands only.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k(int* out, int iters)
{
    unsigned a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = threadIdx.x * 5, a3 = threadIdx.x * 7, b0 = threadIdx.x * 11, b1 = threadIdx.x * 13;
    int c[8][4] = {};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[j][0]), "+r"(c[j][1]), "+r"(c[j][2]), "+r"(c[j][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main()
{
    int* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int wpb = 1; wpb <= 8; wpb *= 2) {
        const int blocks = 148 * wpb, iters = 20000;
        k<<<blocks, 256>>>(d, 100); cudaDeviceSynchronize();
        cudaEventRecord(e0); k<<<blocks, 256>>>(d, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double mmas = (double)blocks * 8 /*warps*/ * iters * 8;
        printf("blocks/SM %d: %.3f ms, %.1f G mma/s, %.1f int8 TOPS (2*16*8*32 per mma), %.2f T hamming-cmp/s equiv (8 mma per 128 cmp)\n", wpb, ms,
               mmas / ms / 1e6, mmas * 2 * 16 * 8 * 32 / ms / 1e9, mmas / 8 * 128 / ms / 1e9);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
