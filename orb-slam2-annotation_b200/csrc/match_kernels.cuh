// match_kernels.cuh — sm_100a kernels of the Hamming-matching half of the hot path.
//
//   hamming_pairs_kernel   : ORBmatcher::DescriptorDistance (ORBmatcher.cpp:1838-1854) on n pairs.
//   grid_*                 : Frame::AssignFeaturesToGrid / PosInGrid (Frame.cpp:241-257, :434-444) as a device CSR grid.
//   window_count/fill      : Frame::GetFeaturesInArea (Frame.cpp:379-432) + DescriptorDistance, one warp per query,
//                            candidates emitted in the reference's enumeration order.
//   stereo_*               : Frame::ComputeStereoMatches (Frame.cpp:540-748).
//   bf_top2_kernel         : brute-force top-2 (SearchByBoW update rule, ORBmatcher.cpp:249-275) over a DB shard;
//                            queries live in registers, DB rows stream through shared memory, XOR + POPC on the integer
//                            pipe; top-2 kept as packed (distance, row) keys so ties resolve to the first row.
#pragma once
#include "common.cuh"

namespace orbb {

constexpr int GRID_COLS = 64, GRID_ROWS = 48;      // Frame.h:37-38
constexpr int TH_HIGH = 100, TH_LOW = 50;          // ORBmatcher.cpp:37-38


__global__ void hamming_pairs_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n, int* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = hamming256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

// ------------------------------------------------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cpp:302-380), batched over map points: one warp per point.
// Point p owns rows [off[p], off[p+1]) (its observations).  For row i the reference sorts the N distances of row i
// (d(i,i) = 0 included) and takes element int(0.5*(N-1)); the row with the smallest such median wins, first row on ties.
// The k-th smallest of a row is found by bisection on the value (9 counting passes over the point's rows, which sit in
// shared memory for N <= DD_SMEM_ROWS); rows are spread over the lanes, the winner is a packed-key min reduction.
// ------------------------------------------------------------------------------------------------------
constexpr int DD_WARPS = 8, DD_SMEM_ROWS = 64;

__global__ void __launch_bounds__(32 * DD_WARPS) distinctive_kernel(const uint4* __restrict__ desc, const long long* __restrict__ off, int npoints,
                                                                    int* __restrict__ best_row, int* __restrict__ best_median)
{
    __shared__ uint4 rows_s[DD_WARPS][DD_SMEM_ROWS * 2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int p = blockIdx.x * DD_WARPS + warp;
    if (p >= npoints) return;
    const long long b = off[p];
    const long long N64 = off[p + 1] - b;
    if (N64 <= 0) { if (lane == 0) { best_row[p] = -1; if (best_median) best_median[p] = -1; } return; }
    const int N = (int)N64;
    const uint4* rows = desc + 2 * b;
    const bool staged = N <= DD_SMEM_ROWS;
    if (staged) {
        for (int i = lane; i < 2 * N; i += 32) rows_s[warp][i] = rows[i];
        __syncwarp();
        rows = rows_s[warp];
    }
    const int k = (int)(0.5 * (N - 1));                      // vDists[0.5*(N-1)]
    unsigned long long best = ~0ull;                         // median << 32 | row
    for (int i = lane; i < N; i += 32) {
        const uint4 a0 = rows[2 * i], a1 = rows[2 * i + 1];
        int lo = 0, hi = 256;                                // smallest v with #{j : d(i,j) <= v} >= k + 1
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int c = 0;
            for (int j = 0; j < N; j++) c += hamming256(a0, a1, rows[2 * j], rows[2 * j + 1]) <= mid;
            if (c >= k + 1) hi = mid; else lo = mid + 1;
        }
        const unsigned long long key = ((unsigned long long)lo << 32) | (unsigned)i;
        best = key < best ? key : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, best, o);
        best = t < best ? t : best;
    }
    if (lane == 0) { best_row[p] = (int)(best & 0xffffffffu); if (best_median) best_median[p] = (int)(best >> 32); }
}

// ------------------------------------------------------------------------------------------------------
// Node join of ORBmatcher::SearchByBoW (ORBmatcher.cpp:224-262): features of two frames that fell into the same vocabulary
// node are compared all-pairs.  One warp per (common node, A feature) entry: lanes run over the node's B features.
// entry = {iA, first B slot (index into featB), number of B features, output offset}.
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) bow_join_kernel(const uint4* __restrict__ descA, const uint4* __restrict__ descB, const int* __restrict__ featB,
                                                       const int4* __restrict__ entries, int nentries, uint16_t* __restrict__ dist)
{
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (e >= nentries) return;
    const int4 en = entries[e];
    const uint4 a0 = descA[2 * en.x], a1 = descA[2 * en.x + 1];
    for (int j = lane; j < en.z; j += 32) {
        const int iB = featB[en.y + j];
        dist[en.w + j] = (uint16_t)hamming256(a0, a1, descB[2 * iB], descB[2 * iB + 1]);
    }
}

// ------------------------------------------------------------------------------------------------------
// Frame::UndistortKeyPoints (Frame.cpp:462-496) = cv::undistortPoints(pts, K, D, noArray(), K) on the keypoint coordinates.
// OpenCV 4.x (cvUndistortPointsInternal): double arithmetic on float inputs, five fixed-point iterations of the Brown model
// (TermCriteria(MAX_ITER, 5, 0.01) of the 6-argument overload), the icdist < 0 escape, then re-projection with P = K.  Every
// operation is an explicit round-to-nearest intrinsic so that no multiply-add is contracted: bit-equal to cv2 4.13.
// ------------------------------------------------------------------------------------------------------
struct UndistortArgs {
    double fx, fy, cx, cy, ifx, ify;
    double k[12];
    int identity;          // mDistCoef(0) == 0: mvKeysUn = mvKeys
};

__global__ void __launch_bounds__(256) undistort_kernel(const orbx_keypoint* in, orbx_keypoint* out, int n,   // in == out allowed
                                                         const __grid_constant__ UndistortArgs a)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orbx_keypoint kp = in[i];
    if (!a.identity) {
#define DM(p, q) __dmul_rn((p), (q))
#define DA(p, q) __dadd_rn((p), (q))
#define DS(p, q) __dsub_rn((p), (q))
        const double u = (double)kp.x, v = (double)kp.y;
        double x = DM(DS(u, a.cx), a.ifx), y = DM(DS(v, a.cy), a.ify);
        const double x0 = x, y0 = y;
        const double* k = a.k;
#pragma unroll 1
        for (int j = 0; j < 5; j++) {
            const double r2 = DA(DM(x, x), DM(y, y));
            const double num = DA(1.0, DM(DA(DM(DA(DM(k[7], r2), k[6]), r2), k[5]), r2));
            const double den = DA(1.0, DM(DA(DM(DA(DM(k[4], r2), k[1]), r2), k[0]), r2));
            const double icdist = __ddiv_rn(num, den);
            if (icdist < 0) { x = DM(DS(u, a.cx), a.ifx); y = DM(DS(v, a.cy), a.ify); break; }
            const double deltaX = DA(DA(DA(DM(DM(DM(2.0, k[2]), x), y), DM(k[3], DA(r2, DM(DM(2.0, x), x)))), DM(k[8], r2)), DM(DM(k[9], r2), r2));
            const double deltaY = DA(DA(DA(DM(k[2], DA(r2, DM(DM(2.0, y), y))), DM(DM(DM(2.0, k[3]), x), y)), DM(k[10], r2)), DM(DM(k[11], r2), r2));
            x = DM(DS(x0, deltaX), icdist);
            y = DM(DS(y0, deltaY), icdist);
        }
        const double xx = DA(DA(DM(a.fx, x), DM(0.0, y)), a.cx);
        const double yy = DA(DA(DM(0.0, x), DM(a.fy, y)), a.cy);
        const double ww = __ddiv_rn(1.0, DA(DA(DM(0.0, x), DM(0.0, y)), 1.0));
        kp.x = (float)DM(xx, ww);
        kp.y = (float)DM(yy, ww);
#undef DM
#undef DA
#undef DS
    }
    out[i] = kp;
}

// ------------------------------------------------------------------------------------------------------
// Frame grid
// ------------------------------------------------------------------------------------------------------
struct FrameDev {
    const orbx_keypoint* kps; const uint4* desc; int n;
    float minX, minY, invW, invH;
    int* cell_of;        // [n]  cell id (ix * GRID_ROWS + iy) or -1
    int* cell_start;     // [GRID_COLS*GRID_ROWS + 1]
    int* cell_items;     // [n]  keypoint indices, ascending inside each cell (insertion order)
};

__global__ void grid_assign_kernel(FrameDev f, int* cell_count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= f.n) return;
    // PosInGrid (Frame.cpp:434-444): round(), not floor()
    const int px = (int)roundf(__fmul_rn(__fsub_rn(f.kps[i].x, f.minX), f.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(f.kps[i].y, f.minY), f.invH));
    int c = -1;
    if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) { c = px * GRID_ROWS + py; atomicAdd(&cell_count[c], 1); }
    f.cell_of[i] = c;
}

// single block: exclusive scan of the 3072 cell counts -> cell_start; cell_count is reset to the cursor
__global__ void grid_scan_kernel(int* cell_count, int* cell_start)
{
    __shared__ int s[1024];
    const int NC = GRID_COLS * GRID_ROWS;       // 3072 = 3 * 1024
    int v[3], sum = 0;
    for (int j = 0; j < 3; j++) { v[j] = cell_count[threadIdx.x * 3 + j]; sum += v[j]; }
    s[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
        __syncthreads();
        s[threadIdx.x] += t;
        __syncthreads();
    }
    int acc = s[threadIdx.x] - sum;
    for (int j = 0; j < 3; j++) { cell_start[threadIdx.x * 3 + j] = acc; cell_count[threadIdx.x * 3 + j] = acc; acc += v[j]; }
    if (threadIdx.x == 1023) cell_start[NC] = acc;
}

__global__ void grid_fill_kernel(FrameDev f, int* cursor)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= f.n) return;
    const int c = f.cell_of[i];
    if (c >= 0) f.cell_items[atomicAdd(&cursor[c], 1)] = i;
}

// one thread per cell: insertion sort => ascending keypoint index == push_back order of AssignFeaturesToGrid
__global__ void grid_sort_kernel(FrameDev f)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= GRID_COLS * GRID_ROWS) return;
    const int b = f.cell_start[c], e = f.cell_start[c + 1];
    for (int i = b + 1; i < e; i++) {
        const int v = f.cell_items[i];
        int j = i - 1;
        while (j >= b && f.cell_items[j] > v) { f.cell_items[j + 1] = f.cell_items[j]; j--; }
        f.cell_items[j + 1] = v;
    }
}

// ------------------------------------------------------------------------------------------------------
// GetFeaturesInArea + DescriptorDistance.  One warp per query; FILL=false counts, FILL=true writes.
// ------------------------------------------------------------------------------------------------------
// FILTER (the device resolve): records are orbm_proj_query; candidates that can never be chosen — the Frame slot already holds a
// MapPoint with observations (ORBmatcher.cpp:112-114, :1590-1592) or the stereo gate fails (:116-121, :1594-1601) — are dropped
// here, where the lanes of the warp test them in parallel, and cand_dist carries the candidate's octave in bits 16+.
// frame_mp_obs == nullptr: every slot that already holds a MapPoint is skipped (relocalisation / loop-closing overloads, ORBmatcher.cpp
// :1727-1728, :438-439); else only slots whose MapPoint has observations (:1590-1592, :107-109)
struct WindowFilter { const int* frame_mp; const int* frame_mp_obs; const float* uright; };

// One walk over the query's cell window.  FILL = false: count the candidates; FILL = true: write them from cand[obase] on (GetFeaturesInArea
// order) with their distances.
template <bool FILL, bool FILTER>
__device__ __forceinline__ int window_walk(const FrameDev& f, const orbm_query& q, int lane, int obase, int* __restrict__ cand_idx,
                                           int* __restrict__ cand_dist, int cap, const WindowFilter& flt)
{
    const float x = q.u, y = q.v, r = q.r;
    const int minLevel = q.min_level, maxLevel = q.max_level;
    int total = 0;
    // cell window (Frame.cpp:384-398)
    const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, f.minX), r), f.invW)));
    const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, f.minX), r), f.invW)));
    const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, f.minY), r), f.invH)));
    const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, f.minY), r), f.invH)));
    const bool empty = nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0;
    if (empty) return 0;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    uint4 q0, q1;
    if (FILL) {   // orbm_query::desc sits at byte offset 20 of a 52-byte record: 4-byte aligned only
        const unsigned* qd = reinterpret_cast<const unsigned*>(q.desc);
        q0 = make_uint4(qd[0], qd[1], qd[2], qd[3]); q1 = make_uint4(qd[4], qd[5], qd[6], qd[7]);
    }
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        // cells (ix, nMinCellY..nMaxCellY) are contiguous in the CSR (cell id = ix*GRID_ROWS + iy)
        const int b = f.cell_start[ix * GRID_ROWS + nMinCellY], e = f.cell_start[ix * GRID_ROWS + nMaxCellY + 1];
        for (int base = b; base < e; base += 32) {
            const int j = base + lane;
            bool ok = false; int idx = -1, oct = 0;
            if (j < e) {
                idx = f.cell_items[j];
                const orbx_keypoint kp = f.kps[idx];
                ok = true;
                if (bCheckLevels) {
                    if (kp.octave < minLevel) ok = false;
                    if (maxLevel >= 0 && kp.octave > maxLevel) ok = false;
                }
                const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                if (!(fabsf(dx) < r && fabsf(dy) < r)) ok = false;
                oct = kp.octave;
                if (FILTER && ok) {
                    const orbm_proj_query& pq = reinterpret_cast<const orbm_proj_query&>(q);
                    if (flt.frame_mp[idx] >= 0 && (!flt.frame_mp_obs || flt.frame_mp_obs[idx] > 0)) ok = false;
                    if (flt.uright) {
                        const float urk = flt.uright[idx];
                        if (urk > 0 && fabsf(__fsub_rn(pq.u_right, urk)) > pq.max_err_right) ok = false;
                    }
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (FILL && ok) {
                const int pos = obase + total + __popc(bal & ((1u << lane) - 1));
                if (pos < cap) {
                    cand_idx[pos] = idx;
                    const int d = hamming256(q0, q1, f.desc[2 * idx], f.desc[2 * idx + 1]);
                    cand_dist[pos] = FILTER ? (d | (oct << 16)) : d;
                }
            }
            total += __popc(bal);
        }
    }
    return total;
}

template <bool FILL, bool FILTER = false>
__global__ void window_search_kernel(FrameDev f, const void* __restrict__ qs, size_t qstride, int nq,   // records that START with an orbm_query
                                     int* __restrict__ counts, const int* __restrict__ offsets,
                                     int* __restrict__ cand_idx, int* __restrict__ cand_dist, int cap, WindowFilter flt = WindowFilter{})
{
    const int lane = threadIdx.x & 31;
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (qi >= nq) return;
    const orbm_query& q = *reinterpret_cast<const orbm_query*>(static_cast<const char*>(qs) + (size_t)qi * qstride);
    const int total = window_walk<FILL, FILTER>(f, q, lane, FILL ? offsets[qi] : 0, cand_idx, cand_dist, cap, flt);
    if (!FILL && lane == 0) counts[qi] = total;
}

// Count, reserve and fill in ONE launch (the device resolve does not need the candidate lists in query order): the warp counts its
// window, lane 0 reserves that many slots with one atomicAdd on *total (zeroed by the caller) and the warp walks the window again to
// fill them - the second walk hits the lines the first one brought in.  begin[qi] / end[qi] delimit the query's candidates;
// *total > cap afterwards means overflow (slots past cap were not written), as offs[nq] > cap does in the three-launch form.
template <bool FILTER>
__global__ void window_search_onepass_kernel(FrameDev f, const void* __restrict__ qs, size_t qstride, int nq, int* __restrict__ begin, int* __restrict__ end,
                                             int* __restrict__ total, int* __restrict__ cand_idx, int* __restrict__ cand_dist, int cap, WindowFilter flt)
{
    const int lane = threadIdx.x & 31;
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (qi >= nq) return;
    const orbm_query& q = *reinterpret_cast<const orbm_query*>(static_cast<const char*>(qs) + (size_t)qi * qstride);
    const int n = window_walk<false, FILTER>(f, q, lane, 0, nullptr, nullptr, 0, flt);
    int obase = 0;
    if (lane == 0) { obase = n > 0 ? atomicAdd(total, n) : 0; begin[qi] = obase; end[qi] = obase + n; }
    obase = __shfl_sync(0xffffffffu, obase, 0);
    if (n > 0) window_walk<true, FILTER>(f, q, lane, obase, cand_idx, cand_dist, cap, flt);
}

// single-block exclusive scan (n up to a few hundred thousand): offsets[0..n], offsets[n] = total
__global__ void excl_scan_kernel(const int* __restrict__ in, int* __restrict__ out, int n)
{
    __shared__ int s[1024];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n ? in[i] : 0;
        s[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {
            int t = threadIdx.x >= o ? s[threadIdx.x - o] : 0;
            __syncthreads();
            s[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < n) out[i] = carry + s[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry += s[1023];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[n] = carry;
}

// ------------------------------------------------------------------------------------------------------
// Device resolve of ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>, th) (MODE 0, ORBmatcher.cpp:63-155) and
// SearchByProjection(Frame& Current, const Frame& Last, th, bMono) (MODE 1, :1506-1659) on the candidate CSR of
// window_search_kernel (no candidate list travels to the host).
//
// The reference walks the queries in order; query i skips every candidate whose Frame slot already holds a MapPoint with
// Observations() > 0 (:112-114, :1590-1592), and its own match, written into the slot, is such a point for the queries after
// it.  So choice(i) depends only on the choices of queries j < i: the sequence of choices is the unique fixed point of
//     choice(i) = best candidate k of i with  !(initially taken(k))  and  !(exists j < i: choice(j) == k and has_obs(j)).
// One CTA iterates it in parallel (thread per query; taker[k] = min such j by atomicMin) until no choice changes; after round
// r the first r queries are final, and dependency chains are a handful long in practice.  Only displaced queries are evaluated again
// (see the loop).  Slot k finally holds the LAST query
// that chose it (a point without observations does not block and is overwritten, as in the reference), nmatches counts every
// accepted query (:150, :1618), and MODE 1's rotation histogram (:1620-1656, ComputeThreeMaxima :1792-1833) then clears the
// slots of the queries outside the three dominant bins.
// ------------------------------------------------------------------------------------------------------
struct ResolveArgs {
    FrameDev f;
    const orbm_proj_query* q; int nq;
    const int* offs; const int* ends; const int* total;            // candidates of query i: [offs[i], ends[i]); *total > cap: overflow, nothing is written
    const int* idx; const int* dist; int cap;
    const float* uright;                                           // mvuRight or nullptr
    float nnratio; int check_ori;
    int max_dist;                                                  // accept best <= max_dist: TH_HIGH (a10, a11), ORBdist (reloc), TH_LOW (loop)
    int* frame_mp; int* frame_mp_obs;                              // [f.n] in/out (frame_mp_obs null in MODE 2)
    int* choice;                                                   // [nq] scratch
    int* taker;                                                    // [f.n] scratch
    int* nmatches; int* status; int* rounds;                       // device scalars (rounds may be nullptr)
};

constexpr int RESOLVE_SMEM_SLOTS = 10 * 1024;       // taker table in shared memory up to this many Frame features (40 KB dynamic + static < the 48 KB default limit)

template <int MODE>
__device__ __forceinline__ int resolve_choose(const ResolveArgs& a, const int* __restrict__ taker, int i)
{
    int bestDist = 256, bestDist2 = 256, bestLevel = -1, bestLevel2 = -1, bestIdx = -1;
    const int e = a.ends[i];
    for (int c0 = a.offs[i]; c0 < e; c0 += 4) {          // four candidates' loads in flight, then the reference's sequential update
        int k[4], pd[4], tk[4];
#pragma unroll
        for (int j = 0; j < 4; j++) { const int c = min(c0 + j, e - 1); k[j] = a.idx[c]; pd[j] = a.dist[c]; }
#pragma unroll
        for (int j = 0; j < 4; j++) tk[j] = taker[k[j]];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (c0 + j >= e || tk[j] < i) continue;
            const int d = pd[j] & 0xffff, oct = pd[j] >> 16;
            if (MODE == 0) {
                if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = k[j]; }
                else if (d < bestDist2) { bestLevel2 = oct; bestDist2 = d; }
            } else if (d < bestDist) { bestDist = d; bestIdx = k[j]; }
        }
    }
    if (bestIdx < 0 || bestDist > a.max_dist) return -1;
    if (MODE == 0 && bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(a.nnratio, (float)bestDist2)) return -1;
    return bestIdx;
}

#ifdef RESOLVE_TIMING
__device__ unsigned long long g_resolve_t[8];
#define RT_STAMP(k) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_resolve_t[k] = t_; } } while (0)
#else
#define RT_STAMP(k)
#endif
template <int MODE>
__global__ void __launch_bounds__(1024) resolve_projection_kernel(const __grid_constant__ ResolveArgs a)
{
    RT_STAMP(0);
    extern __shared__ int taker_s[];
    __shared__ int nm_s, hist_s[30], ind_s[3];
    const int tid = threadIdx.x, nt = blockDim.x;
    if (*a.total > a.cap) { if (tid == 0) { *a.status = ORBB_E_CAPACITY; *a.nmatches = 0; } return; }
    int* taker = a.f.n <= RESOLVE_SMEM_SLOTS ? taker_s : a.taker;
    for (int i = tid; i < a.nq; i += nt) a.choice[i] = -2;
    for (int k = tid; k < a.f.n; k += nt) taker[k] = 0x7fffffff;
    if (tid == 0) nm_s = 0;
    if (tid < 30) hist_s[tid] = 0;
    __syncthreads();
    RT_STAMP(1);
    int rounds = 0;
    if (MODE == 0) {
        // The ratio test makes a query's verdict depend on its best AND second-best unblocked candidate, and a query that loses its
        // second best may turn from accepted to rejected, which frees its slot again: every round rebuilds taker[] and evaluates all.
        for (;;) {
            bool changed = false;
            for (int i = tid; i < a.nq; i += nt) {
                const int c = resolve_choose<MODE>(a, taker, i);
                if (c != a.choice[i]) { a.choice[i] = c; changed = true; }
            }
            rounds++;
            const int any = __syncthreads_or(changed);
            for (int k = tid; k < a.f.n; k += nt) taker[k] = 0x7fffffff;
            __syncthreads();
            if (!any) break;
            for (int i = tid; i < a.nq; i += nt) {
                const int c = a.choice[i];
                if (c >= 0 && a.q[i].has_obs > 0) atomicMin(&taker[c], i);
            }
            __syncthreads();
        }
    } else {
        // MODE 1 / 2 accept the best unblocked candidate on its distance alone, so blocked sets only grow: slot k is blocked for query i
        // by taker[k] = the smallest j < i (with observations) that chose it, and that j leaves k only when an even smaller index takes
        // k.  So (1) taker[] accumulates by atomicMin without being rebuilt - the minimum over every query that ever chose k is a query
        // that stays on k - (2) a rejected query stays rejected, and (3) a query is evaluated again only when its current choice became
        // blocked: after the first round a round costs one shared-memory load per query plus the few displaced queries.
        for (;;) {
            bool changed = false;
            for (int i = tid; i < a.nq; i += nt) {
                const int c = a.choice[i];
                if (c == -2 || (c >= 0 && taker[c] < i)) {
                    const int nc = resolve_choose<MODE>(a, taker, i);
                    a.choice[i] = nc; changed = true;
                    if (nc >= 0 && (MODE == 2 || a.q[i].has_obs > 0)) atomicMin(&taker[nc], i);      // MODE 2: any MapPoint blocks the slot for later queries
                }
            }
            rounds++;
            if (!__syncthreads_or(changed)) break;
        }
        for (int k = tid; k < a.f.n; k += nt) taker[k] = 0x7fffffff;
        __syncthreads();
    }
    RT_STAMP(2);
    // slot k <- the last query that chose it; taker is reused as "last chooser" (reset to INT_MAX above: store -i-1 as min)
    for (int i = tid; i < a.nq; i += nt) {
        const int c = a.choice[i];
        if (c >= 0) atomicMin(&taker[c], -i - 1);
    }
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < a.nq; i += nt) {
        const int c = a.choice[i];
        if (c < 0) continue;
        mine++;
        if (taker[c] == -i - 1) { a.frame_mp[c] = a.q[i].id; if (MODE != 2) a.frame_mp_obs[c] = a.q[i].has_obs; }
        if (MODE >= 1 && a.check_ori) {
            float rot = __fsub_rn(a.q[i].angle, a.f.kps[c].angle);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, 30 / 360.0f));
            if (bin == 30 || bin < 0 || bin > 30) bin = 0;           // (beyond 30: angles outside [0, 360), outside the reference's domain; stay inside the histogram)
            atomicAdd(&hist_s[bin], 1);
        }
    }
    if (mine) atomicAdd(&nm_s, mine);
    __syncthreads();
    RT_STAMP(3);
    if (MODE >= 1 && a.check_ori) {
        if (tid == 0) {                                     // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int b = 0; b < 30; b++) {
                const int s = hist_s[b];
                if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = b; }
                else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = b; }
                else if (s > max3) { max3 = s; ind3 = b; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
            ind_s[0] = ind1; ind_s[1] = ind2; ind_s[2] = ind3;
        }
        __syncthreads();
        int gone = 0;
        for (int i = tid; i < a.nq; i += nt) {
            const int c = a.choice[i];
            if (c < 0) continue;
            float rot = __fsub_rn(a.q[i].angle, a.f.kps[c].angle);
            if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
            int bin = (int)roundf(__fmul_rn(rot, 30 / 360.0f));
            if (bin == 30 || bin < 0 || bin > 30) bin = 0;           // (beyond 30: angles outside [0, 360), outside the reference's domain; stay inside the histogram)
            if (bin != ind_s[0] && bin != ind_s[1] && bin != ind_s[2]) { a.frame_mp[c] = -1; gone++; }
        }
        if (gone) atomicSub(&nm_s, gone);
        __syncthreads();
    }
    RT_STAMP(4);
    if (tid == 0) { *a.nmatches = nm_s; *a.status = ORBB_OK; if (a.rounds) *a.rounds = rounds; }
#ifdef RESOLVE_TIMING
    if (tid == 0) g_resolve_t[5] = (unsigned long long)rounds;
#endif
}

// The per-point search of Fuse / SearchBySim3 (ORBmatcher.cpp:962-1115, :1119-1251, :1253-1500) on the candidate CSR: queries do not
// interact, a thread owns one projected MapPoint.  First candidate (in GetFeaturesInArea order) with the smallest distance among those that
// pass the chi-square reprojection gate when inv_level_sigma2 is given (:1053-1078: three degrees of freedom, 7.8, when the feature has a
// right coordinate, else two, 5.99; float products, double comparison as in the reference); -1 unless best <= max_dist.
__global__ void fuse_best_kernel(FrameDev f, const orbm_proj_query* __restrict__ q, int nq, const int* __restrict__ offs, const int* __restrict__ idx,
                                 const int* __restrict__ dist, int cap, const float* __restrict__ uright, const float* __restrict__ inv_level_sigma2,
                                 int max_dist, int* __restrict__ best_idx, int* __restrict__ best_dist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq || offs[nq] > cap) return;                 // overflow: the host repeats the call with the exact capacity
    const float u = q[i].q.u, v = q[i].q.v, ur = q[i].u_right;
    int bestDist = 256, bestIdx = -1;
    for (int c = offs[i]; c < offs[i + 1]; c++) {
        const int k = idx[c];
        if (inv_level_sigma2) {
            const orbx_keypoint kp = f.kps[k];
            const float ex = __fsub_rn(u, kp.x), ey = __fsub_rn(v, kp.y);
            if (uright && uright[k] >= 0) {
                const float er = __fsub_rn(ur, uright[k]);
                const float e2 = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(er, er));
                if ((double)__fmul_rn(e2, inv_level_sigma2[kp.octave]) > 7.8) continue;
            } else {
                const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                if ((double)__fmul_rn(e2, inv_level_sigma2[kp.octave]) > 5.99) continue;
            }
        }
        const int d = dist[c];
        if (d < bestDist) { bestDist = d; bestIdx = k; }
    }
    const bool ok = bestIdx >= 0 && bestDist <= max_dist;
    best_idx[i] = ok ? bestIdx : -1;
    if (best_dist) best_dist[i] = ok ? bestDist : -1;
}

// ------------------------------------------------------------------------------------------------------
// Stereo (Frame.cpp:540-748)
// ------------------------------------------------------------------------------------------------------
struct StereoArgs {
    const orbx_keypoint* kL; const uint4* dL; int nL;
    const orbx_keypoint* kR; const uint4* dR; int nR;
    const LevelGeom* geom; int nlevels; int nRows;
    const uint8_t* src0L; long long src0L_pitch; const uint8_t* pyrL;   // level 0 / levels >= 1 of the left extractor
    const uint8_t* src0R; long long src0R_pitch; const uint8_t* pyrR;
    float scale[ORBB_MAX_LEVELS], inv_scale[ORBB_MAX_LEVELS];
    float mbf, minD, maxD;
    float* uRight; float* depth; int* sad; int* bestIdx; int* bestDist;
};

__global__ void stereo_match_kernel(const __grid_constant__ StereoArgs a)
{
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (iL >= a.nL) return;
    const orbx_keypoint kpL = a.kL[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int rowL = (int)vL;                          // vRowIndices[vL] (Frame.cpp:602)
    const float minU = __fsub_rn(uL, a.maxD), maxU = __fsub_rn(uL, a.minD);
    const uint4 l0 = a.dL[2 * iL], l1 = a.dL[2 * iL + 1];
    // best right candidate in the row band: min (dist, iR) with dist < TH_HIGH (strict '<' in ascending iR)
    unsigned best = ((unsigned)TH_HIGH << 16);         // key = dist << 16 | iR  (iR < 65536 checked by the host)
    if (maxU >= 0) {
        for (int iR = lane; iR < a.nR; iR += 32) {
            const orbx_keypoint kpR = a.kR[iR];
            const float r = __fmul_rn(2.0f, a.scale[kpR.octave]);
            const int maxr = (int)ceilf(__fadd_rn(kpR.y, r)), minr = (int)floorf(__fsub_rn(kpR.y, r));
            if (rowL < minr || rowL > maxr) continue;
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            if (!(kpR.x >= minU && kpR.x <= maxU)) continue;
            const unsigned d = (unsigned)hamming256(l0, l1, a.dR[2 * iR], a.dR[2 * iR + 1]);
            if (d < (unsigned)TH_HIGH) best = min(best, (d << 16) | (unsigned)iR);
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    const int bestDist = (int)(best >> 16), bestIdxR = (int)(best & 0xffff);
    float out_u = -1.f, out_d = -1.f; int out_sad = -1;
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    if (lane == 0) { a.bestIdx[iL] = bestDist < thOrbDist ? bestIdxR : -1; a.bestDist[iL] = bestDist < thOrbDist ? bestDist : -1; }
    if (bestDist < thOrbDist) {
        const float uR0 = a.kR[bestIdxR].x;
        const float sf = a.inv_scale[levelL];
        const float scaleduL = roundf(__fmul_rn(kpL.x, sf)), scaledvL = roundf(__fmul_rn(kpL.y, sf));
        const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
        const int w = 5, Lw = 5;
        const LevelGeom g = a.geom[levelL];
        const float iniu = scaleduR0 + Lw - w, endu = scaleduR0 + Lw + w + 1;
        if (!(iniu < 0 || endu >= (float)g.w)) {
            const uint8_t* imL; const uint8_t* imR; long long pL, pR;
            if (levelL == 0) { imL = a.src0L; pL = a.src0L_pitch; imR = a.src0R; pR = a.src0R_pitch; }
            else { imL = a.pyrL + g.off; pL = g.pitch; imR = a.pyrR + g.off; pR = g.pitch; }
            const int cxL = (int)scaleduL, cy = (int)scaledvL, cxR0 = (int)scaleduR0;
            const int cL = imL[(long long)cy * pL + cxL];
            int sums[11];
#pragma unroll
            for (int s = 0; s < 11; s++) sums[s] = 0;
            for (int p = lane; p < 121; p += 32) {
                const int dy = p / 11 - w, dx = p % 11 - w;
                const int vl = (int)imL[(long long)(cy + dy) * pL + cxL + dx] - cL;
                const uint8_t* rr = imR + (long long)(cy + dy) * pR + cxR0 + dx;
#pragma unroll
                for (int s = 0; s < 11; s++) {
                    const int cR = imR[(long long)cy * pR + cxR0 + s - Lw];
                    sums[s] += abs(vl - ((int)rr[s - Lw] - cR));
                }
            }
#pragma unroll
            for (int s = 0; s < 11; s++)
#pragma unroll
                for (int o = 16; o; o >>= 1) sums[s] += __shfl_xor_sync(0xffffffffu, sums[s], o);
            int sadBest = 0x7fffffff, bestinc = 0;
#pragma unroll
            for (int s = 0; s < 11; s++) if (sums[s] < sadBest) { sadBest = sums[s]; bestinc = s - Lw; }
            if (!(bestinc == -Lw || bestinc == Lw)) {
                float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
                for (int s = 1; s < 10; s++) if (s - Lw == bestinc) { dist1 = (float)sums[s - 1]; dist2 = (float)sums[s]; dist3 = (float)sums[s + 1]; }
                const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                               __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));
                if (!(deltaR < -1 || deltaR > 1)) {
                    float bestuR = __fmul_rn(a.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= a.minD && disparity < a.maxD) {
                        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
                        out_d = __fdiv_rn(a.mbf, disparity);
                        out_u = bestuR;
                        out_sad = sadBest;
                    }
                }
            }
        }
    }
    if (lane == 0) { a.uRight[iL] = out_u; a.depth[iL] = out_d; a.sad[iL] = out_sad; }
}

// single block: median of the accepted SADs (element size()/2 of the sorted list) and the 1.5*1.4*median cut
// (Frame.cpp:734-747).  SAD <= 121*510 < 2^16: two-pass radix select.
__global__ void stereo_median_filter_kernel(int n, const int* __restrict__ sad, float* uRight, float* depth, float k21)
{
    __shared__ int hist[256];
    __shared__ int s_m, s_bin, s_rem, s_val;
    const int tid = threadIdx.x;
    if (tid == 0) s_m = 0;
    for (int i = tid; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < n; i += blockDim.x) if (sad[i] >= 0) { cnt++; atomicAdd(&hist[sad[i] >> 8], 1); }
    atomicAdd(&s_m, cnt);
    __syncthreads();
    const int M = s_m;
    if (M == 0) return;
    if (tid == 0) {
        int k = M / 2, b = 0;
        while (k >= hist[b]) { k -= hist[b]; b++; }
        s_bin = b; s_rem = k;
    }
    __syncthreads();
    const int bin = s_bin;
    __syncthreads();
    for (int i = tid; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) if (sad[i] >= 0 && (sad[i] >> 8) == bin) atomicAdd(&hist[sad[i] & 255], 1);
    __syncthreads();
    if (tid == 0) {
        int k = s_rem, b = 0;
        while (k >= hist[b]) { k -= hist[b]; b++; }
        s_val = (bin << 8) | b;
    }
    __syncthreads();
    const float median = (float)s_val;
    const float thDist = __fmul_rn(k21, median);
    for (int i = tid; i < n; i += blockDim.x)
        if (sad[i] >= 0 && !((float)sad[i] < thDist)) { uRight[i] = -1.f; depth[i] = -1.f; }
}

// ------------------------------------------------------------------------------------------------------
// Brute-force top-2
// ------------------------------------------------------------------------------------------------------
constexpr int BF_THREADS = 128;
constexpr int BF_TQ = 4;                       // queries per thread (registers)
constexpr int BF_QTILE = BF_THREADS * BF_TQ;   // 512 queries per CTA
constexpr int BF_ROWS = 128;                   // DB rows per shared-memory stage (4 KB)
constexpr int BF_IDX_BITS = 23;                // packed key = dist << 23 | row-in-chunk

// distance of one query (8 words in registers) to one DB row (8 words), variant 0: 8 XOR + 8 POPC.
__device__ __forceinline__ unsigned ham_plain(const unsigned (&q)[8], const uint4& r0, const uint4& r1)
{
    return __popc(q[0] ^ r0.x) + __popc(q[1] ^ r0.y) + __popc(q[2] ^ r0.z) + __popc(q[3] ^ r0.w) +
           __popc(q[4] ^ r1.x) + __popc(q[5] ^ r1.y) + __popc(q[6] ^ r1.z) + __popc(q[7] ^ r1.w);
}
// variant 1: three carry-save full adders (LOP3 pairs) fold 7 of the 8 XOR words into {ones, twos x3}: 5 POPC
// instead of 8, trading the quarter-rate POPC pipe for full-rate LOP3s.
__device__ __forceinline__ unsigned ham_csa(const unsigned (&q)[8], const uint4& r0, const uint4& r1)
{
    const unsigned x0 = q[0] ^ r0.x, x1 = q[1] ^ r0.y, x2 = q[2] ^ r0.z, x3 = q[3] ^ r0.w;
    const unsigned x4 = q[4] ^ r1.x, x5 = q[5] ^ r1.y, x6 = q[6] ^ r1.z, x7 = q[7] ^ r1.w;
    const unsigned s0 = x0 ^ x1 ^ x2, c0 = (x0 & x1) | (x2 & (x0 ^ x1));
    const unsigned s1 = x3 ^ x4 ^ x5, c1 = (x3 & x4) | (x5 & (x3 ^ x4));
    const unsigned s2 = s0 ^ s1 ^ x6, c2 = (s0 & s1) | (x6 & (s0 ^ s1));
    return __popc(s2) + __popc(x7) + 2 * (__popc(c0) + __popc(c1) + __popc(c2));
}

struct BfArgs {
    const uint4* db; long long nrows;          // shard rows
    const uint4* q; int nq;
    long long rows_per_chunk; int nchunks;
    unsigned* part_k1; unsigned* part_k2;      // [nchunks][nq] packed keys
    int* bound_d2;                             // [nq] or null: smallest second-best distance any chunk has published (bf_top2_tc2_kernel)
};

template <int VARIANT>
__global__ void __launch_bounds__(BF_THREADS) bf_top2_kernel(const __grid_constant__ BfArgs a)
{
    __shared__ uint4 rows[2][BF_ROWS * 2];
    const int tid = threadIdx.x;
    const int chunk = blockIdx.y;
    const long long r_begin = (long long)chunk * a.rows_per_chunk;
    const long long r_end = min(a.nrows, r_begin + a.rows_per_chunk);
    // queries -> registers
    unsigned q[BF_TQ][8];
    unsigned k1[BF_TQ], k2[BF_TQ];
#pragma unroll
    for (int j = 0; j < BF_TQ; j++) {
        const int qi = blockIdx.x * BF_QTILE + j * BF_THREADS + tid;
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
        if (qi < a.nq) { v0 = a.q[2 * qi]; v1 = a.q[2 * qi + 1]; }
        q[j][0] = v0.x; q[j][1] = v0.y; q[j][2] = v0.z; q[j][3] = v0.w;
        q[j][4] = v1.x; q[j][5] = v1.y; q[j][6] = v1.z; q[j][7] = v1.w;
        k1[j] = 0xffffffffu; k2[j] = 0xffffffffu;
    }
    const long long nrows = r_end - r_begin;
    const int nstages = (int)((nrows + BF_ROWS - 1) / BF_ROWS);
    auto load_stage = [&](int s, int buf) {
        // 128 rows x 32 B = 256 uint4: two per thread, 16-byte cp.async (LDGSTS), zero-fill past the end
        const long long base = r_begin + (long long)s * BF_ROWS;
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int e = tid + j * BF_THREADS;                  // uint4 index inside the stage
            const long long row = base + (e >> 1);
            const uint4* src = a.db + (row < r_end ? row * 2 + (e & 1) : 0);
            const unsigned dst = (unsigned)__cvta_generic_to_shared(&rows[buf][e]);
            const int bytes = row < r_end ? 16 : 0;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes));
        }
        asm volatile("cp.async.commit_group;");
    };
    if (nstages > 0) load_stage(0, 0);
    for (int s = 0; s < nstages; s++) {
        const int buf = s & 1;
        if (s + 1 < nstages) { load_stage(s + 1, buf ^ 1); asm volatile("cp.async.wait_group 1;"); }
        else asm volatile("cp.async.wait_group 0;");
        __syncthreads();
        const int nr = (int)min((long long)BF_ROWS, nrows - (long long)s * BF_ROWS);
        const unsigned rbase = (unsigned)(s * BF_ROWS);
#pragma unroll 2
        for (int r = 0; r < nr; r++) {
            const uint4 r0 = rows[buf][2 * r], r1 = rows[buf][2 * r + 1];
#pragma unroll
            for (int j = 0; j < BF_TQ; j++) {
                const unsigned d = VARIANT == 0 ? ham_plain(q[j], r0, r1) : ham_csa(q[j], r0, r1);
                const unsigned key = (d << BF_IDX_BITS) + (rbase + (unsigned)r);
                k2[j] = min(k2[j], max(k1[j], key));
                k1[j] = min(k1[j], key);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < BF_TQ; j++) {
        const int qi = blockIdx.x * BF_QTILE + j * BF_THREADS + tid;
        if (qi < a.nq) {
            a.part_k1[(long long)chunk * a.nq + qi] = k1[j];
            a.part_k2[(long long)chunk * a.nq + qi] = k2[j];
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Brute-force top-2 on the tensor cores (variant 2).  For descriptors q, d of 256 bits:
//     hamming(q, d) = |q| - sum_k a_k * b_k      with a_k = +1 / -1 for query bit 1 / 0 and b_k = database bit (0 / 1),
// so a 16 x 8 block of distances is eight mma.sync.m16n8k32.s8.u8 (IMMA) instructions.  The b1 XOR-POPC mma the reference
// algorithm suggests does not exist on sm_100a (ptxas emulates it with ~170 instructions around IMMA), the integer pipe tops
// out at ~0.65 T comparisons/s (ham_csa above), and the legacy IMMA path measures 1144 int8 TOPS on B200 = 2.2 T comparisons/s
// (tools/micro/imma_rate.cu).
//   * a CTA owns 256 queries: each warp keeps its 32 queries as A fragments in registers (+1 / -1 bytes, all eight k-steps);
//   * database rows arrive 128 at a time (cp.async, double buffered), are expanded bit -> byte in shared memory (bit p of a
//     row = byte p: the k order of the instruction is irrelevant as long as A and B agree) and read as 8-byte B fragments;
//   * the accumulators are filtered against the per-row threshold (max of the lane's two columns >= |q| - current second
//     distance); only the rare pair that can enter the top-2 builds its key (dist << 23 | row) and updates (k1, k2);
//   * the four lanes that share a query row merge their partial top-2 with shuffles at the end.
// Output: the same per-(chunk, query) packed keys as bf_top2_kernel, merged by bf_merge_chunks_kernel.
// ------------------------------------------------------------------------------------------------------
constexpr int BM_THREADS = 256;
constexpr int BM_QTILE = 256;                  // queries per CTA: 8 warps x 2 m-tiles x 16 rows
constexpr int BM_ROWS = 128;                   // database rows per stage
constexpr int BM_PITCH = 288;                  // bytes per expanded row in shared memory (256 + 32: conflict-free 8-byte fragment loads)

__device__ __forceinline__ unsigned bm_spread01(unsigned nib)       // 4 bits -> 4 bytes of 0 / 1
{
    return (nib * 0x00204081u) & 0x01010101u;
}
__device__ __forceinline__ unsigned bm_spread_pm1(unsigned nib)     // 4 bits -> 4 bytes of +1 (0x01) / -1 (0xff)
{
    return ~(bm_spread01(nib) * 0xfeu);
}
__device__ __forceinline__ void bm_mma(int (&c)[4], const unsigned (&a)[4], const uint2& b)
{
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

__global__ void __launch_bounds__(BM_THREADS, 2) bf_top2_mma_kernel(const __grid_constant__ BfArgs a)
{
    __shared__ __align__(16) uint8_t Bs[BM_ROWS * BM_PITCH];      // expanded stage
    __shared__ uint4 raw[2][BM_ROWS * 2];                         // packed rows in flight
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int chunk = blockIdx.y;
    const long long r_begin = (long long)chunk * a.rows_per_chunk;
    const long long r_end = min(a.nrows, r_begin + a.rows_per_chunk);
    const int nrows = (int)(r_end - r_begin);

    // ---- queries -> A fragments: row (i, h) = blockIdx.x * 256 + warp * 32 + 16 i + g + 8 h ----
    unsigned A[2][8][4];
    int pq[2][2], thr[2][2];
    unsigned k1[2][2], k2[2][2];
#pragma unroll
    for (int i = 0; i < 2; i++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int qi = blockIdx.x * BM_QTILE + warp * 32 + 16 * i + g + 8 * h;
            uint4 v0 = make_uint4(0, 0, 0, 0), v1 = v0;
            if (qi < a.nq) { v0 = a.q[2 * qi]; v1 = a.q[2 * qi + 1]; }
            const unsigned w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
            int pc = 0;
#pragma unroll
            for (int s = 0; s < 8; s++) {
                pc += __popc(w[s]);
                const unsigned byte = (w[s] >> (8 * t)) & 0xffu;            // bits s*32 + 8 t .. + 7
                A[i][s][h] = qi < a.nq ? bm_spread_pm1(byte & 15u) : 0u;        // k = 4 t + 0..3 of the k-step
                A[i][s][2 + h] = qi < a.nq ? bm_spread_pm1(byte >> 4) : 0u;     // k = 16 + 4 t + 0..3
            }
            pq[i][h] = pc;
            k1[i][h] = 0xffffffffu; k2[i][h] = 0xffffffffu;
            thr[i][h] = qi < a.nq ? pc - 511 : 0x7fffffff;                  // nothing yet: every row passes / padding rows: none
        }
    }

    const int nstages = (nrows + BM_ROWS - 1) / BM_ROWS;
    auto load_stage = [&](int s, int buf) {                       // 128 rows x 32 B = 256 uint4: one 16-byte cp.async per thread
        const long long row = r_begin + (long long)s * BM_ROWS + (tid >> 1);
        const uint4* src = a.db + (row < r_end ? row * 2 + (tid & 1) : 0);
        const unsigned dst = (unsigned)__cvta_generic_to_shared(&raw[buf][tid]);
        const int bytes = row < r_end ? 16 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes));
        asm volatile("cp.async.commit_group;");
    };
    if (nstages > 0) load_stage(0, 0);
    for (int s = 0; s < nstages; s++) {
        const int buf = s & 1;
        if (s + 1 < nstages) { load_stage(s + 1, buf ^ 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();                                          // stage s has landed; the previous stage's fragments are consumed
        // ---- expand: packed word (row, wd) -> 32 bytes at row * PITCH + 32 wd ----
        {
            const unsigned* rw = reinterpret_cast<const unsigned*>(raw[buf]);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int e = tid + BM_THREADS * j;
                const unsigned wv = rw[e];
                uint4* dst = reinterpret_cast<uint4*>(Bs + (e >> 3) * BM_PITCH + (e & 7) * 32);
                dst[0] = make_uint4(bm_spread01(wv & 15u), bm_spread01((wv >> 4) & 15u), bm_spread01((wv >> 8) & 15u), bm_spread01((wv >> 12) & 15u));
                dst[1] = make_uint4(bm_spread01((wv >> 16) & 15u), bm_spread01((wv >> 20) & 15u), bm_spread01((wv >> 24) & 15u), bm_spread01(wv >> 28));
            }
        }
        __syncthreads();
        // ---- 16 n-tiles of 8 rows: 2 x 8 IMMA each, threshold filter, rare exact update ----
        const int sbase = s * BM_ROWS + 2 * t;                    // chunk row of this lane's first column in n-tile 0
#pragma unroll 2
        for (int j = 0; j < BM_ROWS / 8; j++) {
            const uint8_t* bp = Bs + (8 * j + g) * BM_PITCH + 8 * t;
            uint2 b[8];
#pragma unroll
            for (int ks = 0; ks < 8; ks++) b[ks] = *reinterpret_cast<const uint2*>(bp + 32 * ks);
            int c[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
#pragma unroll
            for (int ks = 0; ks < 8; ks++) {
                bm_mma(c[0], A[0][ks], b[ks]);
                bm_mma(c[1], A[1][ks], b[ks]);
            }
#pragma unroll
            for (int i = 0; i < 2; i++) {
                if (max(c[i][0], c[i][1]) >= thr[i][0] || max(c[i][2], c[i][3]) >= thr[i][1]) {
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        if (thr[i][h] == 0x7fffffff) continue;                     // padding query row
#pragma unroll
                        for (int jj = 0; jj < 2; jj++) {
                            const int idx = sbase + 8 * j + jj;
                            if (idx < nrows) {
                                const unsigned key = ((unsigned)(pq[i][h] - c[i][2 * h + jj]) << BF_IDX_BITS) | (unsigned)idx;
                                k2[i][h] = min(k2[i][h], max(k1[i][h], key));
                                k1[i][h] = min(k1[i][h], key);
                            }
                        }
                        thr[i][h] = pq[i][h] - (int)(k2[i][h] >> BF_IDX_BITS);
                    }
                }
            }
        }
    }
    // ---- the four lanes of a query row merge their partial top-2 ----
#pragma unroll
    for (int i = 0; i < 2; i++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            unsigned x1 = k1[i][h], x2 = k2[i][h];
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const unsigned y1 = __shfl_xor_sync(0xffffffffu, x1, o), y2 = __shfl_xor_sync(0xffffffffu, x2, o);
                x2 = min(max(x1, y1), min(x2, y2));
                x1 = min(x1, y1);
            }
            const int qi = blockIdx.x * BM_QTILE + warp * 32 + 16 * i + g + 8 * h;
            if (t == 0 && qi < a.nq) {
                a.part_k1[(long long)chunk * a.nq + qi] = x1;
                a.part_k2[(long long)chunk * a.nq + qi] = x2;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Brute-force top-2 on the 5th-generation tensor cores (default).  The same restatement as bf_top2_mma_kernel (hamming = |q| - sum a.b,
// a = +-1 query bits, b = 0 / 1 database bits) issued as tcgen05.mma.kind::i8, M = 128 queries x N = 256 rows x K = 32 bits per
// instruction, both operands in shared memory (K-major, no swizzle: 8-row x 16-byte core matrices, LBO = 128 B between the two K chunks
// of an instruction, SBO = 2048 B between 8-row groups), int32 accumulators in TMEM.
//   * a CTA owns 256 queries = two A tiles of 32 KB (+1 / -1 bytes), expanded once; one CTA per SM (all 512 TMEM columns);
//   * a stage is 256 database rows = one 64 KB B tile (ring of two).  Each query tile has ONE 256-column accumulator, so the two tiles of
//     a stage double-buffer each other: tile 1 is multiplied while tile 0 is consumed.  (N = 256 rather than 128 because with both
//     operands in shared memory the kernel sits at the 128 B/clk shared-memory port before the tensor pipe: 768 instead of 1024 bytes
//     read per row and CTA.)
// The three activities of a stage run as independent pipelines coupled only by mbarriers; no block barrier inside the scan:
//   * 8 PRODUCER warps: thread t copies packed row t of the stage (cp.async, own slots) and expands it bit -> byte into the B tile;
//     full[s] <- one arrive per producer warp;  a tile is reused when empty[s] <- tcgen05.commit of the MMAs that read it;
//   * 1 MMA warp (one lane) waits full[s] and accempty[t], issues the 8 k-steps of tile t and commits them to accfull[t] (and to
//     empty[s] after tile 1);
//   * 16 CONSUMER warps (4 per TMEM lane quarter, 64 columns each, a thread = one query row) wait accfull[t], pull the accumulators
//     into registers (tcgen05.ld), release the buffer (accempty[t] <- one arrive per consumer warp) and filter them against the row
//     threshold |q| - (current second distance); only a pair that can still enter the top-2 builds its key and updates (k1, k2).  The
//     four threads of a row publish their threshold with a shared-memory atomicMax: an element above ANY thread's second distance
//     cannot enter the row's top-2.
// Output: the same per-(chunk, query) packed keys as bf_top2_kernel, merged by bf_merge_chunks_kernel.
// ------------------------------------------------------------------------------------------------------
constexpr int BT_QTILE = 256;                            // queries per CTA
constexpr int BT_TILE_BYTES = 128 * 256;                 // one 128-row x 256-byte operand tile

__device__ __forceinline__ unsigned long long bt_desc(unsigned smem_addr)
{
    // start address >> 4 | LBO (128 B) >> 4 at bit 16 | SBO (2048 B) >> 4 at bit 32 | descriptor version 1 at bit 46 | no swizzle
    return (unsigned long long)((smem_addr & 0x3ffffu) >> 4) | (8ull << 16) | (128ull << 32) | (1ull << 46);
}
// canonical K-major byte offset of (row r, byte p) inside a 128-row x 256-byte tile
__device__ __forceinline__ unsigned bt_off(unsigned r, unsigned p)
{
    return (r >> 3) * 2048u + (p >> 4) * 128u + (r & 7u) * 16u + (p & 15u);
}
__device__ __forceinline__ void bt_wait(unsigned mbar, unsigned parity)
{
    unsigned done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
}
#define BT_LD32(r, taddr)                                                                                                                  \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23," \
                 "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"                                                                                \
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),  \
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),    \
                   "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),    \
                   "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                                                                      \
                 : "r"(taddr))

constexpr int BW_PROD_WARPS = 8, BW_CONS_WARPS = 16;
constexpr int BW_THREADS = 32 * (BW_PROD_WARPS + BW_CONS_WARPS + 1);                 // 800

__device__ __forceinline__ void bw_arrive(unsigned mbar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}
constexpr int BN_ROWS = 256, BN_STAGES = 2;
constexpr int BN_BTILE = BN_ROWS * 256;
constexpr int BN_SMEM = 2 * BT_TILE_BYTES + BN_STAGES * BN_BTILE + 2 * BN_ROWS * 32;     // 212992 B

__global__ void __launch_bounds__(BW_THREADS, 1) bf_top2_tcn_kernel(const __grid_constant__ BfArgs a)
{
    extern __shared__ __align__(128) uint8_t bt_smem[];
    __shared__ __align__(8) unsigned long long bars[2 * BN_STAGES + 4];      // full[S] | empty[S] | accfull[2] | accempty[2]
    __shared__ unsigned tmem_base_s;
    __shared__ int thr_s[2 * 128];                                          // per query row: tightest threshold any of its four consumer threads holds
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk = blockIdx.y;
    const long long r_begin = (long long)chunk * a.rows_per_chunk;
    const long long r_end = min(a.nrows, r_begin + a.rows_per_chunk);
    const int nrows = (int)(r_end - r_begin);
    const int nstages = (nrows + BN_ROWS - 1) / BN_ROWS;
    uint8_t* const As = bt_smem;
    uint8_t* const Bs = bt_smem + 2 * BT_TILE_BYTES;
    uint4* const raw = reinterpret_cast<uint4*>(bt_smem + 2 * BT_TILE_BYTES + BN_STAGES * BN_BTILE);      // [2][256][2], a producer thread's own slots
    const unsigned As_addr = (unsigned)__cvta_generic_to_shared(As), Bs_addr = (unsigned)__cvta_generic_to_shared(Bs);
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
    const unsigned full0 = bar0, empty0 = bar0 + 8 * BN_STAGES, accfull0 = bar0 + 16 * BN_STAGES, accempty0 = accfull0 + 16;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((unsigned)__cvta_generic_to_shared(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        for (int s = 0; s < BN_STAGES; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full0 + 8 * s), "r"(BW_PROD_WARPS));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty0 + 8 * s));
        }
        for (int b = 0; b < 2; b++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(accfull0 + 8 * b));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(accempty0 + 8 * b), "r"(BW_CONS_WARPS));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 256) thr_s[tid] = -100000;
    // ---- A tiles: 256 queries x 256 bytes of +1 / -1, canonical K-major layout; consecutive lanes write consecutive rows ----
    for (int i = tid; i < 2048; i += BW_THREADS) {
        const int r8 = i & 7, wd = (i >> 3) & 7, row256 = (i >> 6) * 8 + r8;
        const int qi = blockIdx.x * BT_QTILE + row256;
        const bool ok = qi < a.nq;
        const unsigned w = ok ? reinterpret_cast<const unsigned*>(a.q)[(long long)qi * 8 + wd] : 0u;
        uint4 lo, hi;
        lo.x = bm_spread_pm1(w & 15u); lo.y = bm_spread_pm1((w >> 4) & 15u); lo.z = bm_spread_pm1((w >> 8) & 15u); lo.w = bm_spread_pm1((w >> 12) & 15u);
        hi.x = bm_spread_pm1((w >> 16) & 15u); hi.y = bm_spread_pm1((w >> 20) & 15u); hi.z = bm_spread_pm1((w >> 24) & 15u); hi.w = bm_spread_pm1(w >> 28);
        if (!ok) { lo = make_uint4(0, 0, 0, 0); hi = lo; }
        uint8_t* dst = As + (row256 >> 7) * BT_TILE_BYTES + bt_off(row256 & 127, wd * 32);
        *reinterpret_cast<uint4*>(dst) = lo;
        *reinterpret_cast<uint4*>(dst + 128) = hi;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");                          // the A tiles are read by the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = tmem_base_s;
    constexpr unsigned IDESC = (2u << 4) | (1u << 7) | (32u << 17) | (8u << 24);      // N = 256

    int pq[2] = {0, 0}, thr[2] = {0x7fffffff, 0x7fffffff};
    unsigned k1[2] = {0xffffffffu, 0xffffffffu}, k2[2] = {0xffffffffu, 0xffffffffu};
    const int quarter = warp & 3, cg = (warp - BW_PROD_WARPS) >> 2, row = 32 * quarter + lane;

    if (warp < BW_PROD_WARPS) {
        // ================= producers: 256 packed rows -> B ring; thread t owns row t of every stage (two 16-byte copies) =================
        auto load_stage = [&](int s, int buf) {
            const long long grow = r_begin + (long long)s * BN_ROWS + tid;
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const uint4* src = a.db + (grow < r_end ? grow * 2 + u : 0);
                const unsigned dst = (unsigned)__cvta_generic_to_shared(&raw[(buf * 256 + tid) * 2 + u]);
                const int bytes = grow < r_end ? 16 : 0;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes));
            }
            asm volatile("cp.async.commit_group;");
        };
        if (nstages > 0) load_stage(0, 0);
        for (int it = 0; it < nstages; it++) {
            const int s = it % BN_STAGES, n = it / BN_STAGES;
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            uint4 rv[2];
            rv[0] = raw[((it & 1) * 256 + tid) * 2]; rv[1] = raw[((it & 1) * 256 + tid) * 2 + 1];
            if (it + 1 < nstages) load_stage(it + 1, (it + 1) & 1);
            if (it >= BN_STAGES) bt_wait(empty0 + 8 * s, (unsigned)((n - 1) & 1));      // the MMAs that read this tile have completed
            uint8_t* dst = Bs + s * BN_BTILE + bt_off(tid, 0);
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const unsigned wv[4] = {rv[u].x, rv[u].y, rv[u].z, rv[u].w};
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const unsigned w = wv[j];
                    *reinterpret_cast<uint4*>(dst + 1024 * u + 256 * j) = make_uint4(bm_spread01(w & 15u), bm_spread01((w >> 4) & 15u), bm_spread01((w >> 8) & 15u), bm_spread01((w >> 12) & 15u));
                    *reinterpret_cast<uint4*>(dst + 1024 * u + 256 * j + 128) = make_uint4(bm_spread01((w >> 16) & 15u), bm_spread01((w >> 20) & 15u), bm_spread01((w >> 24) & 15u), bm_spread01(w >> 28));
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) bw_arrive(full0 + 8 * s);
        }
    } else if (warp == BW_PROD_WARPS + BW_CONS_WARPS) {
        // ================= MMA issuer =================
        if (lane == 0) {
            // accumulator buffer = query tile: 256 columns each; tile 1 of a stage is multiplied while tile 0 is being consumed
            const unsigned long long adesc0 = bt_desc(As_addr), bdesc0 = bt_desc(Bs_addr);
            for (int it = 0; it < nstages; it++) {
                const int s = it % BN_STAGES, n = it / BN_STAGES;
                bt_wait(full0 + 8 * s, (unsigned)(n & 1));
                const unsigned long long bd0 = bdesc0 + (unsigned long long)(s * (BN_BTILE >> 4));
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    if (it >= 1) bt_wait(accempty0 + 8 * mt, (unsigned)((it - 1) & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) {
                        const unsigned long long ad = adesc0 + (unsigned long long)(mt * (BT_TILE_BYTES >> 4) + ks * 16), bd = bd0 + (unsigned long long)(ks * 16);
                        if (ks == 0)
                            asm volatile("{ .reg .pred p; setp.ne.b32 p, 0, 0; tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%4, %4, %4, %4}, p; }"
                                         ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC), "r"(0u) : "memory");
                        else
                            asm volatile("{ .reg .pred p; setp.eq.b32 p, 0, 0; tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%4, %4, %4, %4}, p; }"
                                         ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC), "r"(0u) : "memory");
                    }
                    if (mt == 1) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(empty0 + 8u * s) : "memory");
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(accfull0 + 8u * mt) : "memory");
                }
            }
        }
    } else {
        // ================= consumers: accumulators -> top-2 =================
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            const int qi = blockIdx.x * BT_QTILE + 128 * mt + row;
            if (qi < a.nq) {
                const uint4 v0 = a.q[2 * qi], v1 = a.q[2 * qi + 1];
                pq[mt] = __popc(v0.x) + __popc(v0.y) + __popc(v0.z) + __popc(v0.w) + __popc(v1.x) + __popc(v1.y) + __popc(v1.z) + __popc(v1.w);
                thr[mt] = pq[mt] - 511;
            }
        }
        for (int e = 0; e < nstages; e++) {
            if (a.bound_d2 && (e & 15) == 0 && cg == 0) {              // chunks share their second-best bounds: see bf_top2_tc2_kernel
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    const int qi = blockIdx.x * BT_QTILE + 128 * mt + row;
                    if (qi < a.nq) {
                        const int g = __ldcg(a.bound_d2 + qi);
                        if (g < 512) atomicMax(&thr_s[mt * 128 + row], pq[mt] - g);
                    }
                }
            }
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                bt_wait(accfull0 + 8 * mt, (unsigned)(e & 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                thr[mt] = max(thr[mt], thr_s[mt * 128 + row]);
#pragma unroll
                for (int part = 0; part < 2; part++) {                // this warp's 64 columns of the tile, 32 at a time
                    int v[32];
                    BT_LD32(v, tmem + ((unsigned)(32 * quarter) << 16) + (unsigned)(mt * 256 + 64 * cg + 32 * part));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (part == 1) {
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) bw_arrive(accempty0 + 8 * mt);     // the tile's buffer can be overwritten: the values are in registers
                    }
                    const int cbase = e * BN_ROWS + 64 * cg + 32 * part;
                    int gm[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        int t3 = __vimax3_s32(v[8 * q], v[8 * q + 1], v[8 * q + 2]);
                        t3 = __vimax3_s32(t3, v[8 * q + 3], v[8 * q + 4]);
                        t3 = __vimax3_s32(t3, v[8 * q + 5], v[8 * q + 6]);
                        gm[q] = max(t3, v[8 * q + 7]);
                    }
                    if (max(__vimax3_s32(gm[0], gm[1], gm[2]), gm[3]) >= thr[mt]) {
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            if (gm[q] < thr[mt]) continue;
#pragma unroll
                            for (int c = 8 * q; c < 8 * q + 8; c++) {
                                const int idx = cbase + c;
                                if (v[c] >= thr[mt] && idx < nrows) {
                                    const unsigned key = ((unsigned)(pq[mt] - v[c]) << BF_IDX_BITS) | (unsigned)idx;
                                    k2[mt] = min(k2[mt], max(k1[mt], key));
                                    k1[mt] = min(k1[mt], key);
                                    thr[mt] = pq[mt] - (int)(k2[mt] >> BF_IDX_BITS);
                                }
                            }
                        }
                        if (thr[mt] != 0x7fffffff) {
                            atomicMax(&thr_s[mt * 128 + row], thr[mt]);
                            if (a.bound_d2 && k2[mt] != 0xffffffffu) atomicMin(a.bound_d2 + blockIdx.x * BT_QTILE + 128 * mt + row, (int)(k2[mt] >> BF_IDX_BITS));
                        }
                    }
                }
            }
        }
    }
    // ---- merge the four column groups of a query row through shared memory (the B ring is free: every MMA has been consumed) ----
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    unsigned* const scratch = reinterpret_cast<unsigned*>(Bs);         // [4][2][128][2]
    const bool consumer = warp >= BW_PROD_WARPS && warp < BW_PROD_WARPS + BW_CONS_WARPS;
    if (consumer) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++) { scratch[((cg * 2 + mt) * 128 + row) * 2] = k1[mt]; scratch[((cg * 2 + mt) * 128 + row) * 2 + 1] = k2[mt]; }
    }
    __syncthreads();
    if (consumer && cg == 0) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            unsigned x1 = k1[mt], x2 = k2[mt];
            for (int g = 1; g < 4; g++) {
                const unsigned y1 = scratch[((g * 2 + mt) * 128 + row) * 2], y2 = scratch[((g * 2 + mt) * 128 + row) * 2 + 1];
                x2 = min(max(x1, y1), min(x2, y2));
                x1 = min(x1, y1);
            }
            const int qi = blockIdx.x * BT_QTILE + 128 * mt + row;
            if (qi < a.nq) {
                a.part_k1[(long long)chunk * a.nq + qi] = x1;
                a.part_k2[(long long)chunk * a.nq + qi] = x2;
            }
        }
    }
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------------
// Brute-force top-2 on a CTA PAIR (thread-block cluster of 2, tcgen05 cta_group::2): one MMA is 256 queries x 256 rows x 32 bits, each
// CTA holding 128 of its query rows (A) and staging 128 of the stage's 256 database rows (B); the tensor cores exchange the B
// halves themselves.  A pair keeps two such query sets (512 queries) resident, so every expanded row feeds 512 queries instead of
// 256: per multiply-add, half the bit -> byte expansion work and a third of the shared-memory operand reads of bf_top2_tcn_kernel
// (which is level between the shared-memory port and the tensor pipe).  Roles per CTA are those of bf_top2_tcn_kernel; only the
// leader CTA's MMA thread issues.  Cross-CTA protocol: producers and consumers of BOTH CTAs arrive on the LEADER's full / accempty
// barriers (mapa + mbarrier.arrive.release.cluster); tcgen05.commit multicasts to the empty / accfull barriers of both CTAs.
// Output: the same per-(chunk, query) packed keys as the other kernels.
// ------------------------------------------------------------------------------------------------------
#ifndef B2_REFRESH
#define B2_REFRESH 15
#endif
constexpr int B2_QTILE = 512;                            // queries per CTA pair
constexpr int B2_STAGES = 4;
constexpr int B2_BTILE = 128 * 256;                      // one CTA's half of a 256-row stage, expanded
constexpr int B2_SMEM = 2 * BT_TILE_BYTES + B2_STAGES * B2_BTILE + 2 * 256 * 16;      // 204800 B
constexpr int B2_PROD_WARPS = 2;                         // 64 producer threads: two database rows each per stage
constexpr int B2_THREADS = 32 * (B2_PROD_WARPS + BW_CONS_WARPS + 1);               // 608: leaves 96 registers per thread (the consumers hold 64 accumulator columns)

// Wait on a local barrier.  CTA-scope acquire on purpose: what the barriers of this kernel order is shared memory written for / TMEM
// read by the tensor cores, never global data of the peer, and a cluster-scope acquire makes ptxas invalidate the whole L1
// (CCTL.IVALL) after every successful wait, twice on the accumulator hand-over path.  A protocol bug traps instead of hanging the GPU.
__device__ __forceinline__ void b2_wait(unsigned mbar, unsigned parity)
{
    unsigned done = 0;
    for (unsigned spin = 0; !done; spin++) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
        if (spin > (1u << 27)) __trap();
    }
}
__device__ __forceinline__ void b2_arrive_leader(unsigned mbar_local)              // arrive on the barrier at this offset in CTA 0 of the cluster
{
    unsigned remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(remote) : "r"(mbar_local));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

__device__ __forceinline__ void b2_arrive_leader_relaxed(unsigned mbar_local)     // no data to publish (the TMEM reads have completed)
{
    unsigned remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(remote) : "r"(mbar_local));
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

#ifdef B2_TIMING
// diagnostic build only: clock64 stamps of the hand-over chain in the leader CTA of cluster (0, 0), stages 64..95
__device__ long long g_b2_t[32][2][8];
__device__ __forceinline__ long long b2_gt() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ long long g_b2_g[2][32][2][8];       // [cta rank] globaltimer stamps (ns)
#define B2_STAMP(it, mt, slot) do { if (blockIdx.x < 2 && blockIdx.y == 0 && (it) >= 64 && (it) < 96) { if (blockIdx.x == 0) g_b2_t[(it) - 64][mt][slot] = clock64(); g_b2_g[blockIdx.x][(it) - 64][mt][slot] = b2_gt(); } } while (0)
#else
#define B2_STAMP(it, mt, slot) do {} while (0)
#endif
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(B2_THREADS, 1) bf_top2_tc2_kernel(const __grid_constant__ BfArgs a)
{
    extern __shared__ __align__(128) uint8_t bt_smem[];
    __shared__ __align__(8) unsigned long long bars[2 * B2_STAGES + 4];      // full[S] | empty[S] | accfull[2] | accempty[2]
    __shared__ unsigned tmem_base_s;
    __shared__ int thr_s[2 * 128];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int chunk = blockIdx.y;
    const int qbase = (blockIdx.x >> 1) * B2_QTILE + 128 * (int)rank;          // query of (set mt, row) = qbase + 256 * mt + row
    const long long r_begin = (long long)chunk * a.rows_per_chunk;
    const long long r_end = min(a.nrows, r_begin + a.rows_per_chunk);
    const int nrows = (int)(r_end - r_begin);
    const int nstages = (nrows + 255) / 256;
    uint8_t* const As = bt_smem;
    uint8_t* const Bs = bt_smem + 2 * BT_TILE_BYTES;
    uint4* const raw = reinterpret_cast<uint4*>(bt_smem + 2 * BT_TILE_BYTES + B2_STAGES * B2_BTILE);      // [2][256]
    const unsigned As_addr = (unsigned)__cvta_generic_to_shared(As), Bs_addr = (unsigned)__cvta_generic_to_shared(Bs);
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
    const unsigned full0 = bar0, empty0 = bar0 + 8 * B2_STAGES, accfull0 = bar0 + 16 * B2_STAGES, accempty0 = accfull0 + 16;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((unsigned)__cvta_generic_to_shared(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    if (tid == 0) {
        for (int s = 0; s < B2_STAGES; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full0 + 8 * s), "r"(2 * B2_PROD_WARPS));     // both CTAs' producer warps
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty0 + 8 * s));
        }
        for (int b = 0; b < 2; b++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(accfull0 + 8 * b));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(accempty0 + 8 * b), "r"(2 * BW_CONS_WARPS));  // both CTAs' consumer warps
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 256) thr_s[tid] = -100000;
    // ---- A tiles: this CTA's 128 rows of both query sets, +1 / -1 bytes, canonical K-major layout ----
    for (int i = tid; i < 2048; i += B2_THREADS) {
        const int r8 = i & 7, wd = (i >> 3) & 7, row256 = (i >> 6) * 8 + r8;
        const int qi = qbase + 256 * (row256 >> 7) + (row256 & 127);
        const bool ok = qi < a.nq;
        const unsigned w = ok ? reinterpret_cast<const unsigned*>(a.q)[(long long)qi * 8 + wd] : 0u;
        uint4 lo, hi;
        lo.x = bm_spread_pm1(w & 15u); lo.y = bm_spread_pm1((w >> 4) & 15u); lo.z = bm_spread_pm1((w >> 8) & 15u); lo.w = bm_spread_pm1((w >> 12) & 15u);
        hi.x = bm_spread_pm1((w >> 16) & 15u); hi.y = bm_spread_pm1((w >> 20) & 15u); hi.z = bm_spread_pm1((w >> 24) & 15u); hi.w = bm_spread_pm1(w >> 28);
        if (!ok) { lo = make_uint4(0, 0, 0, 0); hi = lo; }
        uint8_t* dst = As + (row256 >> 7) * BT_TILE_BYTES + bt_off(row256 & 127, wd * 32);
        *reinterpret_cast<uint4*>(dst) = lo;
        *reinterpret_cast<uint4*>(dst + 128) = hi;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    // both CTAs' barriers, A tiles and TMEM exist before anyone arrives remotely or the leader issues
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = tmem_base_s;
    constexpr unsigned IDESC = (2u << 4) | (1u << 7) | (32u << 17) | (16u << 24);     // s32 accumulators, A signed / B unsigned 8 bit, N = 256, M = 256

    int pq[2] = {0, 0}, thr[2] = {0x7fffffff, 0x7fffffff};
    unsigned k1[2] = {0xffffffffu, 0xffffffffu}, k2[2] = {0xffffffffu, 0xffffffffu};
    const int quarter = warp & 3, cg = (warp - B2_PROD_WARPS) >> 2, row = 32 * quarter + lane;

    if (warp < B2_PROD_WARPS) {
        // ================= producers: this CTA's 128 rows of every stage; thread t owns rows t and t + 64 (four 16-byte copies) =================
        auto load_stage = [&](int s, int buf) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const long long grow = r_begin + (long long)s * 256 + 128 * (int)rank + tid + 64 * h;
#pragma unroll
                for (int u = 0; u < 2; u++) {
                    const uint4* src = a.db + (grow < r_end ? grow * 2 + u : 0);
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(&raw[(buf * 128 + tid + 64 * h) * 2 + u]);
                    const int bytes = grow < r_end ? 16 : 0;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes));
                }
            }
            asm volatile("cp.async.commit_group;");
        };
        if (nstages > 0) load_stage(0, 0);
        for (int it = 0; it < nstages; it++) {
            const int s = it % B2_STAGES, n = it / B2_STAGES;
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            uint4 rv[4];
#pragma unroll
            for (int h = 0; h < 2; h++) { rv[2 * h] = raw[((it & 1) * 128 + tid + 64 * h) * 2]; rv[2 * h + 1] = raw[((it & 1) * 128 + tid + 64 * h) * 2 + 1]; }
            if (it + 1 < nstages) load_stage(it + 1, (it + 1) & 1);
            if (it >= B2_STAGES) b2_wait(empty0 + 8 * s, (unsigned)((n - 1) & 1));      // the MMAs that read this tile have completed
#pragma unroll
            for (int h = 0; h < 2; h++) {
                uint8_t* dst = Bs + s * B2_BTILE + bt_off(tid + 64 * h, 0);
#pragma unroll
                for (int u = 0; u < 2; u++) {
                    const unsigned wv[4] = {rv[2 * h + u].x, rv[2 * h + u].y, rv[2 * h + u].z, rv[2 * h + u].w};
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const unsigned w = wv[j];
                        *reinterpret_cast<uint4*>(dst + 1024 * u + 256 * j) = make_uint4(bm_spread01(w & 15u), bm_spread01((w >> 4) & 15u), bm_spread01((w >> 8) & 15u), bm_spread01((w >> 12) & 15u));
                        *reinterpret_cast<uint4*>(dst + 1024 * u + 256 * j + 128) = make_uint4(bm_spread01((w >> 16) & 15u), bm_spread01((w >> 20) & 15u), bm_spread01((w >> 24) & 15u), bm_spread01(w >> 28));
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) b2_arrive_leader(full0 + 8 * s);
        }
    } else if (warp == B2_PROD_WARPS + BW_CONS_WARPS) {
        // ================= MMA issuer: leader CTA only =================
        if (rank == 0 && lane == 0) {
            const unsigned long long adesc0 = bt_desc(As_addr), bdesc0 = bt_desc(Bs_addr);
            for (int it = 0; it < nstages; it++) {
                const int s = it % B2_STAGES, n = it / B2_STAGES;
                b2_wait(full0 + 8 * s, (unsigned)(n & 1));
                const unsigned long long bd0 = bdesc0 + (unsigned long long)(s * (B2_BTILE >> 4));
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    B2_STAMP(it, mt, 0);                                   // issuer reaches the accempty wait
                    if (it >= 1) b2_wait(accempty0 + 8 * mt, (unsigned)((it - 1) & 1));
                    B2_STAMP(it, mt, 1);                                   // buffer free
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) {
                        const unsigned long long ad = adesc0 + (unsigned long long)(mt * (BT_TILE_BYTES >> 4) + ks * 16), bd = bd0 + (unsigned long long)(ks * 16);
                        if (ks == 0)
                            asm volatile("{ .reg .pred p; setp.ne.b32 p, 0, 0; tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p; }"
                                         ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC) : "memory");
                        else
                            asm volatile("{ .reg .pred p; setp.eq.b32 p, 0, 0; tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p; }"
                                         ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC) : "memory");
                    }
                    // (the last uses of a stage are not followed by a refill: no empty signal, so nothing arrives at a CTA that is leaving)
                    if (mt == 1 && it + B2_STAGES < nstages)
                        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(empty0 + 8u * s), "h"((unsigned short)3) : "memory");
                    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(accfull0 + 8u * mt), "h"((unsigned short)3) : "memory");
                    B2_STAMP(it, mt, 2);                                   // the set's MMAs and commits are issued
                }
            }
        }
    } else {
        // ================= consumers: this CTA's 128 query rows of both sets =================
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            const int qi = qbase + 256 * mt + row;
            if (qi < a.nq) {
                const uint4 v0 = a.q[2 * qi], v1 = a.q[2 * qi + 1];
                pq[mt] = __popc(v0.x) + __popc(v0.y) + __popc(v0.z) + __popc(v0.w) + __popc(v1.x) + __popc(v1.y) + __popc(v1.z) + __popc(v1.w);
                thr[mt] = pq[mt] - 511;
            }
        }
        for (int e = 0; e < nstages; e++) {
            // Chunks share their progress: a row farther from the query than a second-best distance ANY chunk has already found cannot be in
            // the merged top-2 (strictly farther: ties stay, so "first index wins" is untouched), so every chunk starts from - and keeps
            // refreshing - the tightest published bound instead of re-learning its thresholds from scratch (with 270 k-row chunks 15 % of
            // the warp-level filter steps otherwise take the divergent update path, and one late warp delays the accumulator hand-over).
            if (a.bound_d2 && (e & B2_REFRESH) == 0 && cg == 0) {
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    const int qi = qbase + 256 * mt + row;
                    if (qi < a.nq) {
                        const int g = __ldcg(a.bound_d2 + qi);
                        if (g < 512) atomicMax(&thr_s[mt * 128 + row], pq[mt] - g);
                    }
                }
            }
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                if (warp == B2_PROD_WARPS && lane == 0) B2_STAMP(e, mt, 3);       // consumer warp 0 starts waiting for the set
                b2_wait(accfull0 + 8 * mt, (unsigned)(e & 1));
                if (warp == B2_PROD_WARPS && lane == 0) B2_STAMP(e, mt, 4);       // ... sees it complete
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                // this warp's 64 columns of the set: both loads in flight, one wait, and the buffer is handed back before any filtering
                // (.pack::16b loads, two columns per register, were tried: the TMEM read gets slower, not faster)
                int vv[2][32];
                BT_LD32(vv[0], tmem + ((unsigned)(32 * quarter) << 16) + (unsigned)(mt * 256 + 64 * cg));
                BT_LD32(vv[1], tmem + ((unsigned)(32 * quarter) << 16) + (unsigned)(mt * 256 + 64 * cg + 32));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (warp == B2_PROD_WARPS && lane == 0) B2_STAMP(e, mt, 5);       // ... has its 64 columns in registers
                if (warp == B2_PROD_WARPS + 15 && lane == 0) B2_STAMP(e, mt, 6);  // the last consumer warp likewise
                if (lane == 0) b2_arrive_leader_relaxed(accempty0 + 8 * mt);       // nothing to publish: the TMEM reads have completed
                thr[mt] = max(thr[mt], thr_s[mt * 128 + row]);
#pragma unroll
                for (int part = 0; part < 2; part++) {
                    const int (&v)[32] = vv[part];
                    const int cbase = e * 256 + 64 * cg + 32 * part;
                    int gm[4];
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        int t3 = __vimax3_s32(v[8 * q], v[8 * q + 1], v[8 * q + 2]);
                        t3 = __vimax3_s32(t3, v[8 * q + 3], v[8 * q + 4]);
                        t3 = __vimax3_s32(t3, v[8 * q + 5], v[8 * q + 6]);
                        gm[q] = max(t3, v[8 * q + 7]);
                    }
                    if (max(__vimax3_s32(gm[0], gm[1], gm[2]), gm[3]) >= thr[mt]) {
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            if (gm[q] < thr[mt]) continue;
#pragma unroll
                            for (int c = 8 * q; c < 8 * q + 8; c++) {
                                const int idx = cbase + c;
                                if (v[c] >= thr[mt] && idx < nrows) {
                                    const unsigned key = ((unsigned)(pq[mt] - v[c]) << BF_IDX_BITS) | (unsigned)idx;
                                    k2[mt] = min(k2[mt], max(k1[mt], key));
                                    k1[mt] = min(k1[mt], key);
                                    thr[mt] = pq[mt] - (int)(k2[mt] >> BF_IDX_BITS);
                                }
                            }
                        }
                        if (thr[mt] != 0x7fffffff) {
                            atomicMax(&thr_s[mt * 128 + row], thr[mt]);
                            if (a.bound_d2 && k2[mt] != 0xffffffffu) atomicMin(a.bound_d2 + qbase + 256 * mt + row, (int)(k2[mt] >> BF_IDX_BITS));
                        }
                    }
                }
            }
        }
    }
    // ---- merge the four column groups of a query row through shared memory.  The B ring is free only once BOTH CTAs have consumed
    //      every MMA (the peer's tensor core reads this CTA's half): cluster barrier first ----
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    unsigned* const scratch = reinterpret_cast<unsigned*>(Bs);         // [4][2][128][2]
    const bool consumer = warp >= B2_PROD_WARPS && warp < B2_PROD_WARPS + BW_CONS_WARPS;
    if (consumer) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++) { scratch[((cg * 2 + mt) * 128 + row) * 2] = k1[mt]; scratch[((cg * 2 + mt) * 128 + row) * 2 + 1] = k2[mt]; }
    }
    __syncthreads();
    if (consumer && cg == 0) {
#pragma unroll
        for (int mt = 0; mt < 2; mt++) {
            unsigned x1 = k1[mt], x2 = k2[mt];
            for (int g = 1; g < 4; g++) {
                const unsigned y1 = scratch[((g * 2 + mt) * 128 + row) * 2], y2 = scratch[((g * 2 + mt) * 128 + row) * 2 + 1];
                x2 = min(max(x1, y1), min(x2, y2));
                x1 = min(x1, y1);
            }
            const int qi = qbase + 256 * mt + row;
            if (qi < a.nq) {
                a.part_k1[(long long)chunk * a.nq + qi] = x1;
                a.part_k2[(long long)chunk * a.nq + qi] = x2;
            }
        }
    }
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// Roofline probe: tcgen05.mma kind::i8 128 x 256 x 32 back to back from static shared-memory operands into the two accumulators the
// scan kernel uses — no producers, no consumers, no global memory.  One CTA per SM, one issuing thread; groups of 16 MMAs are
// committed to an mbarrier and the issuer stays two groups ahead of the tensor pipe.
constexpr int TCP_SMEM = BT_TILE_BYTES + BN_BTILE;          // one A tile + one 256-row B tile = 96 KB
__global__ void __launch_bounds__(128, 1) tc_i8_peak_kernel(int groups, int* sink)
{
    extern __shared__ __align__(128) uint8_t tp_smem[];
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ unsigned tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < TCP_SMEM / 4; i += 128) reinterpret_cast<unsigned*>(tp_smem)[i] = 0x01010101u;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((unsigned)__cvta_generic_to_shared(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = tmem_base_s;
    constexpr unsigned IDESC = (2u << 4) | (1u << 7) | (32u << 17) | (8u << 24);      // s32 accumulate, u8 x s8 ... as bf_top2_tcn_kernel, N = 256
    if (tid == 0) {
        const unsigned long long adesc0 = bt_desc((unsigned)__cvta_generic_to_shared(tp_smem));
        const unsigned long long bdesc0 = bt_desc((unsigned)__cvta_generic_to_shared(tp_smem + BT_TILE_BYTES));
        for (int g = 0; g < groups; g++) {
            if (g >= 2) bt_wait(bar0 + 8 * (g & 1), (unsigned)(((g >> 1) - 1) & 1));
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
#pragma unroll
                for (int ks = 0; ks < 8; ks++) {
                    const unsigned long long ad = adesc0 + (unsigned long long)(ks * 16), bd = bdesc0 + (unsigned long long)(ks * 16);
                    if (ks == 0)
                        asm volatile("{ .reg .pred p; setp.ne.b32 p, 0, 0; tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%4, %4, %4, %4}, p; }"
                                     ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC), "r"(0u) : "memory");
                    else
                        asm volatile("{ .reg .pred p; setp.eq.b32 p, 0, 0; tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%4, %4, %4, %4}, p; }"
                                     ::"r"(tmem + (unsigned)(mt * 256)), "l"(ad), "l"(bd), "r"(IDESC), "r"(0u) : "memory");
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar0 + 8u * (g & 1)) : "memory");
        }
        for (int g = max(groups - 2, 0); g < groups; g++) bt_wait(bar0 + 8 * (g & 1), (unsigned)((g >> 1) & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (sink && groups < 0) sink[blockIdx.x] = (int)tmem;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// The same probe on CTA pairs: tcgen05.mma cta_group::2 kind::i8 256 x 256 x 32 (each CTA 128 A rows and 128 B rows in shared memory),
// issued by the leader, completion multicast to both CTAs.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) tc2_i8_peak_kernel(int groups, int ncols)
{
    extern __shared__ __align__(128) uint8_t tp_smem[];
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ unsigned tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    unsigned rank;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    for (int i = tid; i < (BT_TILE_BYTES + B2_BTILE) / 4; i += 128) reinterpret_cast<unsigned*>(tp_smem)[i] = 0x01010101u;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((unsigned)__cvta_generic_to_shared(&tmem_base_s)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    const unsigned bar0 = (unsigned)__cvta_generic_to_shared(&bars[0]);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const unsigned tmem = tmem_base_s;
    const unsigned IDESC = (2u << 4) | (1u << 7) | ((unsigned)(ncols >> 3) << 17) | (16u << 24);
    const int nbuf = 512 / ncols;
    if (tid == 0 && rank == 0) {
        const unsigned long long adesc0 = bt_desc((unsigned)__cvta_generic_to_shared(tp_smem));
        const unsigned long long bdesc0 = bt_desc((unsigned)__cvta_generic_to_shared(tp_smem + BT_TILE_BYTES));
        for (int g = 0; g < groups; g++) {
            if (g >= 2) b2_wait(bar0 + 8 * (g & 1), (unsigned)(((g >> 1) - 1) & 1));
#pragma unroll
            for (int mt = 0; mt < 2; mt++) {
                const unsigned dcol = (unsigned)(((2 * g + mt) % nbuf) * ncols);
#pragma unroll
                for (int ks = 0; ks < 8; ks++) {
                    const unsigned long long ad = adesc0 + (unsigned long long)(ks * 16), bd = bdesc0 + (unsigned long long)(ks * 16);
                    if (ks == 0)
                        asm volatile("{ .reg .pred p; setp.ne.b32 p, 0, 0; tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p; }"
                                     ::"r"(tmem + dcol), "l"(ad), "l"(bd), "r"(IDESC) : "memory");
                    else
                        asm volatile("{ .reg .pred p; setp.eq.b32 p, 0, 0; tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p; }"
                                     ::"r"(tmem + dcol), "l"(ad), "l"(bd), "r"(IDESC) : "memory");
                }
            }
            asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar0 + 8u * (g & 1)), "h"((unsigned short)1) : "memory");
        }
        for (int g = max(groups - 2, 0); g < groups; g++) b2_wait(bar0 + 8 * (g & 1), (unsigned)((g >> 1) & 1));
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
    }
}

// Merge the per-chunk keys in ascending chunk (= ascending row) order: sequential-scan semantics.
__global__ void bf_merge_chunks_kernel(const unsigned* __restrict__ k1, const unsigned* __restrict__ k2, int nchunks, int nq,
                                       long long rows_per_chunk, long long row_offset,
                                       int* __restrict__ d1, long long* __restrict__ i1, int* __restrict__ d2)
{
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    int b1 = 256, b2 = 256; long long bi = -1;
    for (int c = 0; c < nchunks; c++) {
        const unsigned a1 = k1[(long long)c * nq + qi], a2 = k2[(long long)c * nq + qi];
        if (a1 == 0xffffffffu) continue;
        const int c1 = (int)(a1 >> BF_IDX_BITS);
        const int c2 = a2 == 0xffffffffu ? 256 : (int)(a2 >> BF_IDX_BITS);
        if (c1 < b1) { b2 = min(b1, c2); b1 = c1; bi = row_offset + (long long)c * rows_per_chunk + (a1 & ((1u << BF_IDX_BITS) - 1)); }
        else b2 = min(b2, c1);
    }
    d1[qi] = b1; i1[qi] = bi; d2[qi] = b2;
}

// Merge G shard results [g][nq] (ascending shard row ranges) — the step after the all-gather.
__global__ void bf_merge_shards_kernel(const int* __restrict__ d1a, const long long* __restrict__ i1a, const int* __restrict__ d2a,
                                       int G, int nq, int* __restrict__ d1, long long* __restrict__ i1, int* __restrict__ d2)
{
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    int b1 = 256, b2 = 256; long long bi = -1;
    for (int g = 0; g < G; g++) {
        const int c1 = d1a[(long long)g * nq + qi], c2 = d2a[(long long)g * nq + qi];
        const long long ci = i1a[(long long)g * nq + qi];
        if (ci < 0) continue;
        if (c1 < b1) { b2 = min(b1, c2); b1 = c1; bi = ci; }
        else b2 = min(b2, c1);
    }
    d1[qi] = b1; i1[qi] = bi; d2[qi] = b2;
}

// register-only POPC throughput probe (roofline denominator for matching)
__global__ void popc_peak_kernel(unsigned* out, int iters, unsigned seed)
{
    unsigned x[8], acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { x[j] = seed * (threadIdx.x + 1) + j * 0x9e3779b9u; acc[j] = 0; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) { acc[j] += __popc(x[j]); x[j] ^= acc[j]; }
    }
    unsigned s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += acc[j];
    if (s == 0x12345678u) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}


// register-only IMMA (mma.sync.m16n8k32 int8) throughput probe: the tensor-pipe denominator of bf_top2_mma_kernel
__global__ void __launch_bounds__(256) imma_peak_kernel(int* out, int iters, unsigned seed)
{
    unsigned fa[4] = {seed * (threadIdx.x + 1), seed * 3u + threadIdx.x, seed * 5u ^ threadIdx.x, seed * 7u + 2u * threadIdx.x};
    const uint2 fb = make_uint2(seed * 11u + threadIdx.x, seed * 13u ^ threadIdx.x);
    int c[8][4] = {};
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) bm_mma(c[j], fa, fb);
    }
    int s = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    if (s == 0x12345678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace orbb
