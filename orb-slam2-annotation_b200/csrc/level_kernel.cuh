// level_kernel.cuh — the fused per-pyramid-level kernel (resize + 7x7 blur + FAST/NMS), sm_100a.
//
// One launch per level, grid = (tiles_x, tiles_y, frames).  A CTA stages a 64x32 tile of level L (+4 px halo) in
// shared memory ONCE (32-bit loads) and produces from it
//   (a) the blurred level        cv::GaussianBlur 7x7 sigma 2      ORBextractor.cpp:1097-1098
//   (b) FAST-9-16 candidates     cv::FAST per 30-px cell + NMS     ORBextractor.cpp:797-838
//   (c) the tile of level L+1    cv::resize INTER_LINEAR           ORBextractor.cpp:1136
// so level L is read from HBM once.  All pixel math is byte-SIMD inside 32-bit words (4 pixels per thread):
//   blur rows    : IDP.4A (dp4a) on funnel-shifted words, columns: IMAD on the symmetric Q8 kernel
//   FAST reject  : VABSDIFF4 on the 4 axis + 4 diagonal ring pixels (any 9-arc holds two adjacent ones of each set)
//   FAST score   : both polarities packed as biased u16x2, min/max networks on VIMNMX(3).U16x2
// Arithmetic is exact integer throughout; results are bit-identical to the oracle.
#pragma once
#include "common.cuh"

namespace orbb {

constexpr int TW = 64, TH = 32, HALO = 4, NT = 256;
constexpr int TWW = 20;                      // shared tile pitch in 32-bit words (18 used = 72 pixels)
constexpr int SMW = TWW * 4;                 // ... in bytes
constexpr int SMH = TH + 2 * HALO;           // 40 rows
constexpr int SCW = TW + 2, SCH = TH + 2;    // score region: owned pixels + 1 (for the 3x3 NMS)
constexpr int SCP = SCW + 2;                 // score map pitch (68 bytes)
constexpr int EMIT_CAP = 1024;
constexpr int QCAP = SCH * 72;

struct LevelArgs {
    const uint8_t* src; long long src_pitch, src_fstride;   // level L (external input image when L == 0)
    int w, h;
    uint8_t* blur; long long blur_pitch, blur_fstride;      // blurred level L out
    uint8_t* dst; long long dst_pitch, dst_fstride;         // level L+1 out (null on the last level)
    const int2* xt; const int2* yt; const int* dxs; const int* dys;
    int fast_on, wCell, hCell, minTh;
    unsigned magicW, magicH;                                // ceil(2^32 / wCell), ceil(2^32 / hCell)
    uint32_t* rec; int rec_cap; long long rec_fstride;      // raw FAST records  y<<20 | x<<8 | m
    int* rec_count; int rec_count_stride;
    int q[7];                                               // Q8 blur kernel
};

__device__ __forceinline__ int reflect101(int p, int n)
{
    if (p < 0) p = -p;
    if (p >= n) p = 2 * (n - 1) - p;
    return min(max(p, 0), n - 1);
}

// per byte: 0x80 where x > t (0 <= t <= 126), else 0
__device__ __forceinline__ unsigned gt_bytes(unsigned x, unsigned t1 /* (t+1) * 0x01010101 */)
{
    return ((((x & 0x7f7f7f7fu) | 0x80808080u) - t1) | x) & 0x80808080u;
}

// Threshold-independent FAST arc score m = max over the 16 arcs of 9 contiguous ring pixels of
// max(min(v - p_k), min(p_k - v)); cv::FAST's response is m - 1 and a pixel is a corner at threshold t iff m > t.
// Both polarities ride in one register as biased u16 halves: lo = 256 + v - p_k, hi = 256 - v + p_k.
__device__ __forceinline__ int fast_arc_score(const uint8_t* tb, int ty, int tx)
{
    const uint8_t* c = tb + ty * SMW + tx;
    const unsigned v = c[0];
    const unsigned av = (256u + v) + ((256u - v) << 16);
    unsigned X[16];
    X[0] = c[3 * SMW] * 65535u + av;       X[1] = c[3 * SMW + 1] * 65535u + av;   X[2] = c[2 * SMW + 2] * 65535u + av;
    X[3] = c[SMW + 3] * 65535u + av;       X[4] = c[3] * 65535u + av;             X[5] = c[-SMW + 3] * 65535u + av;
    X[6] = c[-2 * SMW + 2] * 65535u + av;  X[7] = c[-3 * SMW + 1] * 65535u + av;  X[8] = c[-3 * SMW] * 65535u + av;
    X[9] = c[-3 * SMW - 1] * 65535u + av;  X[10] = c[-2 * SMW - 2] * 65535u + av; X[11] = c[-SMW - 3] * 65535u + av;
    X[12] = c[-3] * 65535u + av;           X[13] = c[SMW - 3] * 65535u + av;      X[14] = c[2 * SMW - 2] * 65535u + av;
    X[15] = c[3 * SMW - 1] * 65535u + av;
    unsigned a2[16], a4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) a2[k] = __vminu2(X[k], X[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) a4[k] = __vminu2(a2[k], a2[(k + 2) & 15]);
    unsigned best = 0;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        const unsigned a9 = __vimin3_u16x2(a4[k], a4[(k + 4) & 15], X[(k + 8) & 15]);
        const unsigned b9 = __vimin3_u16x2(a4[k + 1], a4[(k + 5) & 15], X[(k + 9) & 15]);
        best = __vimax3_u16x2(best, a9, b9);
    }
    return (int)max(best & 0xffffu, best >> 16) - 256;
}

__global__ void __launch_bounds__(NT) level_kernel(const __grid_constant__ LevelArgs a)
{
    __shared__ __align__(16) uint32_t tw[SMH][TWW];      // tile + halo, 4 pixels per word
    __shared__ __align__(16) uint32_t hb[TH + 6][TW];    // horizontal blur sums (Q8.8)
    __shared__ __align__(4) uint8_t sc[SCH * SCP];       // arc scores of the candidate pixels
    __shared__ uint16_t queue[QCAP];
    __shared__ uint32_t emit[EMIT_CAP];
    __shared__ int s_qn, s_en, s_ebase;

    const int tid = threadIdx.x;
    const int f = blockIdx.z;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const uint8_t* __restrict__ src = a.src + (long long)f * a.src_fstride;
    uint8_t* tb = reinterpret_cast<uint8_t*>(&tw[0][0]);
    if (tid == 0) { s_qn = 0; s_en = 0; }

    // ---- 1. stage tile + halo (REFLECT_101 at the image border, as the blur of a cloned level sees it) ----
    const bool word_load = ((a.src_pitch & 3) == 0) && ((reinterpret_cast<unsigned long long>(src) & 3) == 0) &&
                           x0 >= HALO && x0 + TW + HALO <= a.w;
    if (word_load) {
        for (int i = tid; i < SMH * 18; i += NT) {
            const int r = i / 18, c = i - r * 18;
            const int gy = reflect101(y0 - HALO + r, a.h);
            tw[r][c] = *reinterpret_cast<const uint32_t*>(src + (long long)gy * a.src_pitch + (x0 - HALO + 4 * c));
        }
    } else {
        for (int r = tid >> 5; r < SMH; r += NT / 32) {
            const int gy = reflect101(y0 - HALO + r, a.h);
            const uint8_t* row = src + (long long)gy * a.src_pitch;
            for (int c = tid & 31; c < TW + 2 * HALO; c += 32) tb[r * SMW + c] = row[reflect101(x0 - HALO + c, a.w)];
        }
    }
    for (int i = tid; i < SCH * SCP / 4; i += NT) reinterpret_cast<uint32_t*>(sc)[i] = 0;
    __syncthreads();

    // ---- 2. blur, horizontal pass: 4 outputs per task, two dp4a each on funnel-shifted words ----
    {
        const unsigned qlo = (unsigned)a.q[0] | ((unsigned)a.q[1] << 8) | ((unsigned)a.q[2] << 16) | ((unsigned)a.q[3] << 24);
        const unsigned qhi = (unsigned)a.q[4] | ((unsigned)a.q[5] << 8) | ((unsigned)a.q[6] << 16);
        for (int i = tid; i < (TH + 6) * 16; i += NT) {
            const int r = i >> 4, cw = i & 15;
            const unsigned w0 = tw[r + 1][cw], w1 = tw[r + 1][cw + 1], w2 = tw[r + 1][cw + 2];
            uint4 o;   // output j covers tile bytes 4cw+1+j .. 4cw+7+j
            o.x = __dp4a(__funnelshift_r(w0, w1, 8), qlo, __dp4a(__funnelshift_r(w1, w2, 8), qhi, 0u));
            o.y = __dp4a(__funnelshift_r(w0, w1, 16), qlo, __dp4a(__funnelshift_r(w1, w2, 16), qhi, 0u));
            o.z = __dp4a(__funnelshift_r(w0, w1, 24), qlo, __dp4a(__funnelshift_r(w1, w2, 24), qhi, 0u));
            o.w = __dp4a(w1, qlo, __dp4a(w2, qhi, 0u));
            *reinterpret_cast<uint4*>(&hb[r][4 * cw]) = o;
        }
    }

    // ---- 3. FAST quick reject on words (4 pixels at a time) -> candidate queue ----
    const bool fast_tile = a.fast_on && (x0 + TW >= 19) && (x0 - 1 <= a.w - 20) && (y0 + TH >= 19) && (y0 - 1 <= a.h - 20);
    if (fast_tile) {
        const unsigned t1 = (unsigned)(min(a.minTh, 126) + 1) * 0x01010101u;
        const int vx_lo = max(19, x0 - 1) - (x0 - HALO), vx_hi = min(a.w - 20, x0 + TW) - (x0 - HALO);   // tile byte coords
        for (int i = tid; i < SCH * 18; i += NT) {
            const int r = i / 18, c = i - r * 18;
            const int gy = y0 - 1 + r;
            if (gy < 19 || gy > a.h - 20) continue;
            const int lo = vx_lo - 4 * c, hi = vx_hi - 4 * c;
            if (lo > 3 || hi < 0) continue;
            unsigned valid = 0x80808080u;
            if (lo > 0) valid &= 0xffffffffu << (8 * lo);
            if (hi < 3) valid &= 0xffffffffu >> (8 * (3 - hi));
            const int ty = r + 3;
            const unsigned V = tw[ty][c], L = tw[ty][c - 1], R = tw[ty][c + 1];    // c-1 / c+1 may be pad words: masked out
            // ring pixels 0 (0,+3), 8 (0,-3), 4 (+3,0), 12 (-3,0): any 9-arc contains two ADJACENT ones of these four
            const unsigned m0 = gt_bytes(__vabsdiffu4(V, tw[ty + 3][c]), t1), m8 = gt_bytes(__vabsdiffu4(V, tw[ty - 3][c]), t1);
            const unsigned m4 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(V, R, 24)), t1);
            const unsigned m12 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(L, V, 8)), t1);
            unsigned cand = (m0 | m8) & (m4 | m12) & valid;
            if (!cand) continue;
            // ... and two adjacent ones of the diagonal four: 2 (+2,+2), 6 (+2,-2), 10 (-2,-2), 14 (-2,+2)
            const unsigned Vd = tw[ty + 2][c], Ld = tw[ty + 2][c - 1], Rd = tw[ty + 2][c + 1];
            const unsigned Vu = tw[ty - 2][c], Lu = tw[ty - 2][c - 1], Ru = tw[ty - 2][c + 1];
            const unsigned m2 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(Vd, Rd, 16)), t1);
            const unsigned m14 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(Ld, Vd, 16)), t1);
            const unsigned m6 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(Vu, Ru, 16)), t1);
            const unsigned m10 = gt_bytes(__vabsdiffu4(V, __funnelshift_r(Lu, Vu, 16)), t1);
            cand &= (m2 | m10) & (m6 | m14);
            if (!cand) continue;
            int slot = atomicAdd(&s_qn, __popc(cand));
#pragma unroll
            for (int b = 0; b < 4; b++)
                if (cand & (0x80u << (8 * b))) queue[slot++] = (uint16_t)(ty * 72 + 4 * c + b);
        }
    }
    __syncthreads();

    // ---- 4. blur, vertical pass (symmetric kernel, Q16.16), 4 pixels per task, word store ----
    {
        uint8_t* __restrict__ bl = a.blur + (long long)f * a.blur_fstride;
        const unsigned q0 = a.q[0], q1 = a.q[1], q2 = a.q[2], q3 = a.q[3];
        const bool sym = a.q[4] == a.q[2] && a.q[5] == a.q[1] && a.q[6] == a.q[0];
        for (int i = tid; i < TH * 16; i += NT) {
            const int r = i >> 4, cw = i & 15;
            const int gy = y0 + r, gx = x0 + 4 * cw;
            if (gy >= a.h || gx >= a.w) continue;
            uint4 hrow[7];
#pragma unroll
            for (int k = 0; k < 7; k++) hrow[k] = *reinterpret_cast<const uint4*>(&hb[r + k][4 * cw]);
            unsigned s[4];
            if (sym) {
                s[0] = q0 * (hrow[0].x + hrow[6].x) + q1 * (hrow[1].x + hrow[5].x) + q2 * (hrow[2].x + hrow[4].x) + q3 * hrow[3].x;
                s[1] = q0 * (hrow[0].y + hrow[6].y) + q1 * (hrow[1].y + hrow[5].y) + q2 * (hrow[2].y + hrow[4].y) + q3 * hrow[3].y;
                s[2] = q0 * (hrow[0].z + hrow[6].z) + q1 * (hrow[1].z + hrow[5].z) + q2 * (hrow[2].z + hrow[4].z) + q3 * hrow[3].z;
                s[3] = q0 * (hrow[0].w + hrow[6].w) + q1 * (hrow[1].w + hrow[5].w) + q2 * (hrow[2].w + hrow[4].w) + q3 * hrow[3].w;
            } else {
                s[0] = s[1] = s[2] = s[3] = 0;
#pragma unroll
                for (int k = 0; k < 7; k++) { const unsigned qk = a.q[k]; s[0] += qk * hrow[k].x; s[1] += qk * hrow[k].y; s[2] += qk * hrow[k].z; s[3] += qk * hrow[k].w; }
            }
            unsigned o = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) o |= min((s[j] + 32768u) >> 16, 255u) << (8 * j);
            uint8_t* dstp = bl + (long long)gy * a.blur_pitch + gx;
            if (gx + 3 < a.w && (a.blur_pitch & 3) == 0) *reinterpret_cast<uint32_t*>(dstp) = o;
            else for (int j = 0; j < 4 && gx + j < a.w; j++) dstp[j] = (uint8_t)(o >> (8 * j));
        }
    }

    // ---- 5. bilinear resize of the destination pixels whose source origin lies in this tile ----
    if (a.dst) {
        uint8_t* __restrict__ dst = a.dst + (long long)f * a.dst_fstride;
        const int dx0 = a.dxs[blockIdx.x], dx1 = a.dxs[blockIdx.x + 1];
        const int dy0 = a.dys[blockIdx.y], dy1 = a.dys[blockIdx.y + 1];
        const int cx_i = tid & 63;
        if (dx0 + cx_i < dx1) {
            const int dx = dx0 + cx_i;
            const int2 cx = a.xt[dx];
            const int tx = cx.x - (x0 - HALO);
            const int a0 = (short)(cx.y & 0xffff), a1 = cx.y >> 16;
            for (int dy = dy0 + (tid >> 6); dy < dy1; dy += NT / 64) {
                const int2 cy = a.yt[dy];
                const uint8_t* p = tb + (cy.x - (y0 - HALO)) * SMW + tx;
                const int b0 = (short)(cy.y & 0xffff), b1 = cy.y >> 16;
                const int h0 = p[0] * a0 + p[1] * a1;
                const int h1 = p[SMW] * a0 + p[SMW + 1] * a1;
                const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                dst[(long long)dy * a.dst_pitch + dx] = (uint8_t)min(max(v, 0), 255);
            }
        }
    }

    if (!fast_tile) return;

    // ---- 6. exact arc score for the queued pixels ----
    const int qn = s_qn;
    for (int j = tid; j < qn; j += NT) {
        const int pos = queue[j];
        const int ty = pos / 72, tx = pos - ty * 72;
        const int m = fast_arc_score(tb, ty, tx);
        sc[(ty - 3) * SCP + (tx - 3)] = (uint8_t)(m > a.minTh ? m : 0);
    }
    __syncthreads();

    // ---- 7. cell-local 3x3 non-max suppression (strict '>') and emission ----
    for (int j = tid; j < qn; j += NT) {
        const int pos = queue[j];
        const int ty = pos / 72, tx = pos - ty * 72;
        const int r = ty - 3, c = tx - 3;
        if (r < 1 || r > TH || c < 1 || c > TW) continue;       // halo ring: scored for its neighbours only
        const uint8_t* sp = sc + r * SCP + c;
        const int m = sp[0];
        if (!m) continue;
        const int gx = x0 - 1 + c, gy = y0 - 1 + r;
        // position inside the 30-px FAST cell: a neighbour in ANOTHER cell never suppresses (cv::FAST ran per cell)
        const int cellx = __umulhi((unsigned)(gx - 19), a.magicW), celly = __umulhi((unsigned)(gy - 19), a.magicH);
        const int rx = gx - 19 - cellx * a.wCell, ry = gy - 19 - celly * a.hCell;
        const bool left = rx > 0, right = rx < a.wCell - 1, up = ry > 0, down = ry < a.hCell - 1;
        bool keep = true;
        if (up) { if ((left && sp[-SCP - 1] >= m) || sp[-SCP] >= m || (right && sp[-SCP + 1] >= m)) keep = false; }
        if ((left && sp[-1] >= m) || (right && sp[1] >= m)) keep = false;
        if (down) { if ((left && sp[SCP - 1] >= m) || sp[SCP] >= m || (right && sp[SCP + 1] >= m)) keep = false; }
        if (keep) {
            const int e = atomicAdd(&s_en, 1);
            if (e < EMIT_CAP) emit[e] = ((unsigned)gy << 20) | ((unsigned)gx << 8) | (unsigned)m;
        }
    }
    __syncthreads();
    const int en = s_en;
    if (en == 0) return;
    if (tid == 0) s_ebase = atomicAdd(a.rec_count + (long long)f * a.rec_count_stride, en);
    __syncthreads();
    const int ebase = s_ebase;
    uint32_t* __restrict__ rec = a.rec + (long long)f * a.rec_fstride;
    for (int j = tid; j < min(en, EMIT_CAP); j += NT)
        if (ebase + j < a.rec_cap) rec[ebase + j] = emit[j];
}

}  // namespace orbb
