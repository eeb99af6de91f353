#include "../orb-slam2-annotation_b200/csrc/extract_kernels.cuh"
#include <cstdio>
#include <cstdlib>
using namespace orbb;
__global__ void dbg(const uint8_t* in, int* out, int* out2)
{
    __shared__ __align__(16) uint8_t tile[SMH][SMW];
    for (int i = threadIdx.x; i < SMH * SMW; i += blockDim.x) (&tile[0][0])[i] = in[i];
    __syncthreads();
    for (int i = threadIdx.x; i < 34 * 66; i += blockDim.x) {
        int r = i / 66, c = i % 66;
        out[i] = fast_arc_score(tile, r + 3, c + 3);
        // brute force
        const int rdx[16] = {0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}, rdy[16] = {3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
        int v = tile[r+3][c+3], d[16];
        for (int k = 0; k < 16; k++) d[k] = v - tile[r+3+rdy[k]][c+3+rdx[k]];
        int best = -999;
        for (int k = 0; k < 16; k++) { int mn = 999, mx = -999; for (int j = 0; j < 9; j++) { mn = min(mn, d[(k+j)&15]); mx = max(mx, d[(k+j)&15]); } best = max(best, max(mn, -mx)); }
        out2[i] = best;
    }
}
int main()
{
    uint8_t h[SMH * SMW]; srand(3);
    for (auto& p : h) p = rand() % 256;
    uint8_t* d; int *o, *o2; cudaMalloc(&d, sizeof h); cudaMalloc(&o, 34*66*4); cudaMalloc(&o2, 34*66*4);
    cudaMemcpy(d, h, sizeof h, cudaMemcpyHostToDevice);
    dbg<<<1, 256>>>(d, o, o2);
    static int a[34*66], b[34*66];
    cudaMemcpy(a, o, sizeof a, cudaMemcpyDeviceToHost); cudaMemcpy(b, o2, sizeof b, cudaMemcpyDeviceToHost);
    int bad = 0; for (int i = 0; i < 34*66; i++) if (a[i] != b[i]) { if (bad < 10) printf("i=%d fast=%d brute=%d\n", i, a[i], b[i]); bad++; }
    printf("bad=%d err=%s\n", bad, cudaGetErrorString(cudaGetLastError()));
}
