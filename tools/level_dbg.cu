#include "../orb-slam2-annotation_b200/csrc/extract_kernels.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
using namespace orbb;
namespace orbb { std::atomic<uint64_t> g_launches{0}; void set_cuda_error(cudaError_t, const char*) {} }
int main()
{
    const int W = 640, H = 480;
    std::vector<uint8_t> img(W * H); srand(5);
    // smooth-ish random texture
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) img[y * W + x] = (uint8_t)((((x / 7) * 37 + (y / 5) * 91) % 200) + rand() % 9);
    uint8_t *d_img, *d_blur; uint32_t* d_rec; int* d_cnt;
    cudaMalloc(&d_img, W * H); cudaMalloc(&d_blur, W * H); cudaMalloc(&d_rec, 4 * 100000); cudaMalloc(&d_cnt, 4);
    cudaMemcpy(d_img, img.data(), W * H, cudaMemcpyHostToDevice); cudaMemset(d_cnt, 0, 4);
    LevelArgs a; memset(&a, 0, sizeof a);
    a.src = d_img; a.src_pitch = W; a.w = W; a.h = H; a.blur = d_blur; a.blur_pitch = W;
    a.fast_on = 1; a.wCell = 31; a.hCell = 32; a.minTh = 7; a.magicW = (unsigned)((0x100000000ull + 30) / 31); a.magicH = (unsigned)((0x100000000ull + 31) / 32); a.rec = d_rec; a.rec_cap = 100000; a.rec_count = d_cnt; a.rec_count_stride = 1;
    const int Q[7] = {18, 34, 48, 56, 48, 34, 18}; for (int i = 0; i < 7; i++) a.q[i] = Q[i];
    level_kernel<<<dim3((W + TW - 1) / TW, (H + TH - 1) / TH, 1), NT>>>(a);
    int n = 0; cudaMemcpy(&n, d_cnt, 4, cudaMemcpyDeviceToHost);
    std::vector<uint32_t> rec(n); cudaMemcpy(rec.data(), d_rec, 4 * n, cudaMemcpyDeviceToHost);
    printf("err=%s n=%d\n", cudaGetErrorString(cudaGetLastError()), n);
    // CPU m map
    const int rdx[16] = {0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}, rdy[16] = {3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
    std::vector<int> M(W * H, 0);
    for (int y = 19; y <= H - 20; y++) for (int x = 19; x <= W - 20; x++) {
        int v = img[y * W + x], d[16]; for (int k = 0; k < 16; k++) d[k] = v - img[(y + rdy[k]) * W + x + rdx[k]];
        int best = -999; for (int k = 0; k < 16; k++) { int mn = 999, mx = -999; for (int j = 0; j < 9; j++) { mn = std::min(mn, d[(k+j)&15]); mx = std::max(mx, d[(k+j)&15]); } best = std::max(best, std::max(mn, -mx)); }
        M[y * W + x] = best > 7 ? best : 0;
    }
    printf("CPU around (599,450):"); for (int yy = -3; yy <= 3; yy++) { printf("\n   "); for (int xx = -3; xx <= 3; xx++) printf("%4d", img[(450 + yy) * W + 599 + xx]); } printf("\n M=%d\n", M[450 * W + 599]);
    int expect = 0;
    for (int y = 19; y <= H - 20; y++) for (int x = 19; x <= W - 20; x++) {
        int m = M[y * W + x]; if (!m) continue; bool keep = true; int cx = (x - 19) / 31, cy = (y - 19) / 32;
        for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) { if (!dx && !dy) continue; int xx = x + dx, yy = y + dy;
            if (M[yy * W + xx] >= m && xx >= 19 && yy >= 19 && (xx - 19) / 31 == cx && (yy - 19) / 32 == cy) keep = false; }
        expect += keep;
    }
    int bad = 0;
    for (int i = 0; i < n; i++) { int m = rec[i] & 255, x = (rec[i] >> 8) & 0xfff, y = rec[i] >> 20; if (x >= W || y >= H || M[y * W + x] != m) { if (bad < 8) printf("rec x=%d y=%d m=%d true=%d\n", x, y, m, (x < W && y < H) ? M[y * W + x] : -1); bad++; } }
    printf("expected %d records, got %d, bad %d\n", expect, n, bad);
}
