// match.cu — host side of the matcher half of liborbb200 (C ABI of include/orbb200.h, section "matcher").
// Replaces the Hamming core of ORB_SLAM2::ORBmatcher / Frame (reference src/ORBmatcher.cpp, src/Frame.cpp).
#include "match_kernels.cuh"
#include <stdlib.h>
#include <atomic>

using namespace orbb;

// ---- internal accessors into orbx_handle (defined in extract.cu) -------------------------------------
extern "C" int orbx__pyramid_view(orbx_handle* h, const LevelGeom** geom_host, const LevelGeom** geom_dev, int* nlevels,
                                  const uint8_t** src0, long long* src0_pitch, const uint8_t** pyr, int* device,
                                  const float** scale, const float** inv_scale, void** stream);

extern "C" int orbx__extract_resident(orbx_handle* h, const uint8_t* const* imgs_host, int nframes, int w, int ht, size_t stride,
                                      int32_t* counts_host, orbx_resident_view* v);

namespace {

// Device scratch from the library's block cache (common.cuh): no cudaMalloc / cudaFree on the per-frame paths after warm-up.
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    ~DevBuf() { if (p) scratch_free(p, cap); }
    int alloc(size_t bytes) { p = scratch_alloc(bytes ? bytes : 16, &cap); if (!p) { cap = 0; return ORBB_E_NOMEM; } return ORBB_OK; }
    int reserve(size_t bytes)           // grow-only; contents are not kept
    {
        if (bytes <= cap) return ORBB_OK;
        if (p) {                        // queued work may still use the old block, and the cache may hand it to another stream: drain first (rare)
            ORBB_CUDA(cudaDeviceSynchronize());
            scratch_free(p, cap); p = nullptr; cap = 0;
        }
        return alloc(bytes + bytes / 4);
    }
    template <typename T> T* as() { return (T*)p; }
};

// Pinned staging of the host-pointer entry points, one block per calling thread (grown on demand, never shrunk): a call packs its
// uploads here, queues asynchronous copies around its kernels and synchronises once, instead of one blocking pageable copy per array.
struct PinnedArena {
    void* p = nullptr; size_t cap = 0;
    ~PinnedArena() { if (p) cudaFreeHost(p); }
    uint8_t* get(size_t bytes)           // only at the start of a call: nothing of this thread is in flight then
    {
        if (bytes <= cap) return (uint8_t*)p;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 2;
        if (want < ((size_t)1 << 20)) want = (size_t)1 << 20;
        if (cudaHostAlloc(&p, want, cudaHostAllocPortable) != cudaSuccess) { set_cuda_error(cudaGetLastError(), "pinned staging"); p = nullptr; return nullptr; }
        cap = want;
        return (uint8_t*)p;
    }
};
thread_local PinnedArena t_pin;

int check_device(int device)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1 || device < 0 || device >= n) {
        cudaGetLastError();
        set_cuda_error(cudaErrorNoDevice, "no CUDA device (liborbb200 has no CPU fallback)");
        return ORBB_E_CUDA;
    }
    ORBB_CUDA(cudaSetDevice(device));
    return ORBB_OK;
}

}  // namespace

extern "C" int orbm_distance_pairs(const uint8_t* a_host, const uint8_t* b_host, int n, int32_t* out_host, int device)
{
    if (n < 0 || (n > 0 && (!a_host || !b_host || !out_host))) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    if (n == 0) return ORBB_OK;
    DevBuf a, b, o;
    if ((rc = a.alloc((size_t)n * 32)) || (rc = b.alloc((size_t)n * 32)) || (rc = o.alloc((size_t)n * 4))) return rc;
    ORBB_CUDA(cudaMemcpy(a.p, a_host, (size_t)n * 32, cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(b.p, b_host, (size_t)n * 32, cudaMemcpyHostToDevice));
    hamming_pairs_kernel<<<ceil_div(n, 256), 256>>>(a.as<uint4>(), b.as<uint4>(), n, o.as<int>());
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    ORBB_CUDA(cudaMemcpy(out_host, o.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

// ---- MapPoint::ComputeDistinctiveDescriptors, batched (MapPoint.cpp:302-380) --------------------------
extern "C" int orbm_distinctive_descriptors_device(const uint8_t* desc_dev, const int64_t* offsets_dev, int npoints,
                                                   int32_t* best_row_dev, int32_t* best_median_dev, void* stream)
{
    if (npoints < 0 || (npoints > 0 && (!desc_dev || !offsets_dev || !best_row_dev))) return ORBB_E_INVALID;
    if (npoints == 0) return ORBB_OK;
    if (((uintptr_t)desc_dev & 15) != 0) return ORBB_E_INVALID;
    distinctive_kernel<<<ceil_div(npoints, DD_WARPS), 32 * DD_WARPS, 0, (cudaStream_t)stream>>>(
        (const uint4*)desc_dev, (const long long*)offsets_dev, npoints, best_row_dev, best_median_dev);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_distinctive_descriptors(const uint8_t* desc_host, const int64_t* offsets_host, int npoints,
                                            int32_t* best_row_host, int32_t* best_median_host, int device)
{
    if (npoints < 0 || (npoints > 0 && (!offsets_host || !best_row_host))) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    if (npoints == 0) return ORBB_OK;
    for (int p = 0; p < npoints; p++) if (offsets_host[p + 1] < offsets_host[p]) return ORBB_E_INVALID;
    if (offsets_host[0] < 0) return ORBB_E_INVALID;
    const int64_t nrows = offsets_host[npoints];
    if (nrows > 0 && !desc_host) return ORBB_E_INVALID;
    DevBuf d, o, r, m;
    if ((rc = d.alloc((size_t)nrows * 32)) || (rc = o.alloc((size_t)(npoints + 1) * 8)) || (rc = r.alloc((size_t)npoints * 4)) ||
        (rc = m.alloc((size_t)npoints * 4))) return rc;
    if (nrows > 0) ORBB_CUDA(cudaMemcpy(d.p, desc_host, (size_t)nrows * 32, cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(o.p, offsets_host, (size_t)(npoints + 1) * 8, cudaMemcpyHostToDevice));
    rc = orbm_distinctive_descriptors_device(d.as<uint8_t>(), o.as<int64_t>(), npoints, r.as<int32_t>(), m.as<int32_t>(), nullptr);
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpy(best_row_host, r.p, (size_t)npoints * 4, cudaMemcpyDeviceToHost));
    if (best_median_host) ORBB_CUDA(cudaMemcpy(best_median_host, m.p, (size_t)npoints * 4, cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

// ---- Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cpp:462-536) ----------------------------------------------------
static int undistort_args(const float* K4, const float* dist, int ndist, UndistortArgs& a)
{
    if (!K4 || !dist || (ndist != 4 && ndist != 5 && ndist != 8 && ndist != 12) || K4[0] == 0.f || K4[1] == 0.f) return ORBB_E_INVALID;
    a.fx = K4[0]; a.fy = K4[1]; a.cx = K4[2]; a.cy = K4[3];
    a.ifx = 1. / a.fx; a.ify = 1. / a.fy;
    for (int i = 0; i < 12; i++) a.k[i] = i < ndist ? (double)dist[i] : 0.0;
    a.identity = dist[0] == 0.0;                            // Frame.cpp:465
    return ORBB_OK;
}

extern "C" int orbm_undistort_keypoints_device(const orbx_keypoint* kps_dev, int n, const float* K4, const float* dist, int ndist,
                                               orbx_keypoint* kps_un_dev, int device, void* stream)
{
    UndistortArgs a;
    if (n < 0 || (n > 0 && (!kps_dev || !kps_un_dev)) || undistort_args(K4, dist, ndist, a) != ORBB_OK) return ORBB_E_INVALID;
    const int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    if (n == 0) return ORBB_OK;
    undistort_kernel<<<ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(kps_dev, kps_un_dev, n, a);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_undistort_keypoints(const orbx_keypoint* kps_host, int n, const float* K4, const float* dist, int ndist,
                                        orbx_keypoint* kps_un_host, int device)
{
    UndistortArgs a;
    if (n < 0 || (n > 0 && (!kps_host || !kps_un_host)) || undistort_args(K4, dist, ndist, a) != ORBB_OK) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    if (n == 0) return ORBB_OK;
    DevBuf in, out;
    if ((rc = in.alloc((size_t)n * sizeof(orbx_keypoint))) || (rc = out.alloc((size_t)n * sizeof(orbx_keypoint)))) return rc;
    ORBB_CUDA(cudaMemcpy(in.p, kps_host, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice));
    undistort_kernel<<<ceil_div(n, 256), 256>>>(in.as<orbx_keypoint>(), out.as<orbx_keypoint>(), n, a);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    ORBB_CUDA(cudaMemcpy(kps_un_host, out.p, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

extern "C" int orbm_image_bounds(int cols, int rows, const float* K4, const float* dist, int ndist, float* bounds4, int device)
{
    UndistortArgs a;
    if (cols <= 0 || rows <= 0 || !bounds4 || undistort_args(K4, dist, ndist, a) != ORBB_OK) return ORBB_E_INVALID;
    if (a.identity) {                                        // Frame.cpp:527-533; no device work
        const int rc = check_device(device);
        if (rc != ORBB_OK) return rc;
        bounds4[0] = 0.0f; bounds4[1] = 0.0f; bounds4[2] = (float)cols; bounds4[3] = (float)rows;
        return ORBB_OK;
    }
    orbx_keypoint c[4], u[4];
    memset(c, 0, sizeof(c));
    c[1].x = (float)cols; c[2].y = (float)rows; c[3].x = (float)cols; c[3].y = (float)rows;       // :504-511
    const int rc = orbm_undistort_keypoints(c, 4, K4, dist, ndist, u, device);
    if (rc != ORBB_OK) return rc;
    bounds4[0] = std::min(u[0].x, u[2].x); bounds4[2] = std::max(u[1].x, u[3].x);                   // :518-521
    bounds4[1] = std::min(u[0].y, u[1].y); bounds4[3] = std::max(u[2].y, u[3].y);
    return ORBB_OK;
}

// ---- frame -------------------------------------------------------------------------------------------
struct orbm_frame {
    int device = 0, n = 0;
    FrameDev f{};
    // host copies for the order-dependent resolution loops that stay on the host (reloc / loop / fuse / initialisation);
    // a frame built from device arrays fetches them on first use
    mutable std::vector<orbx_keypoint> h_kps;
    mutable std::vector<uint8_t> h_desc;
    mutable bool host_valid = false;
    DevBuf kps, desc, cell_of, cell_start, cell_items, cursor;
    DevBuf uright_own; bool uright_valid = false;      // mvuRight computed by orbm_stereo_frame_create (ORBM_URIGHT_OF_FRAME)
    // workspace of the device resolve (orbm_search_by_projection_*_device); calls on one frame are serialised by the caller
    mutable DevBuf ws_counts, ws_offs, ws_idx, ws_dist, ws_choice, ws_taker, ws_uright, ws_mp;   // ws_mp: [mvpMapPoints | observations | 4 scalars] of the host-pointer searches
    // recorded after the copies and grid kernels of orbm_frame_create_device on the caller's stream: the host-pointer entry points
    // run on the legacy default stream, which does not order against cudaStreamNonBlocking streams, so they wait on it first
    cudaEvent_t ready = nullptr;
    // streams that may still hold queued work on this frame's buffers: drained by orbm_frame_destroy before the blocks go back to the cache
    mutable cudaStream_t s_create = nullptr, s_last = nullptr;
    ~orbm_frame() { if (ready) cudaEventDestroy(ready); }
};

// Host-form entry points: block until a frame built by orbm_frame_create_device is complete.
static int frame_wait_ready(const orbm_frame* fr)
{
    if (fr->ready) ORBB_CUDA(cudaEventSynchronize(fr->ready));
    return ORBB_OK;
}

static int frame_host_copies(const orbm_frame* fr)
{
    if (fr->host_valid) return ORBB_OK;
    int rcw = frame_wait_ready(fr);
    if (rcw != ORBB_OK) return rcw;
    fr->h_kps.resize(fr->n); fr->h_desc.resize((size_t)fr->n * 32);
    if (fr->n > 0) {
        ORBB_CUDA(cudaMemcpy(fr->h_kps.data(), fr->kps.p, (size_t)fr->n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost));
        ORBB_CUDA(cudaMemcpy(fr->h_desc.data(), fr->desc.p, (size_t)fr->n * 32, cudaMemcpyDeviceToHost));
    }
    fr->host_valid = true;
    return ORBB_OK;
}

// kind = cudaMemcpyHostToDevice: the classic entry point (synchronous); cudaMemcpyDeviceToDevice: everything is queued on `stream`.
static int frame_create_impl(orbm_frame** out, const orbx_keypoint* kps, const uint8_t* desc, int n,
                             float min_x, float min_y, float max_x, float max_y, int device, cudaMemcpyKind kind, cudaStream_t st)
{
    if (!out || n < 0 || (n > 0 && (!kps || !desc)) || !(max_x > min_x) || !(max_y > min_y)) return ORBB_E_INVALID;
    *out = nullptr;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    orbm_frame* fr = new (std::nothrow) orbm_frame();
    if (!fr) return ORBB_E_NOMEM;
    fr->device = device; fr->n = n; fr->s_create = st; fr->s_last = st;
    const bool from_host = kind == cudaMemcpyHostToDevice;
    if (from_host) {
        if (n > 0) { fr->h_kps.assign(kps, kps + n); fr->h_desc.assign(desc, desc + (size_t)n * 32); }
        fr->host_valid = true;
    }
    const int NC = GRID_COLS * GRID_ROWS;
    if ((rc = fr->kps.alloc((size_t)n * sizeof(orbx_keypoint))) || (rc = fr->desc.alloc((size_t)n * 32)) ||
        (rc = fr->cell_of.alloc((size_t)n * 4)) || (rc = fr->cell_start.alloc((size_t)(NC + 1) * 4)) ||
        (rc = fr->cell_items.alloc((size_t)n * 4)) || (rc = fr->cursor.alloc((size_t)NC * 4))) { delete fr; return rc; }
    FrameDev& f = fr->f;
    f.kps = fr->kps.as<orbx_keypoint>(); f.desc = fr->desc.as<uint4>(); f.n = n;
    f.minX = min_x; f.minY = min_y;
    // Frame.cpp:115-116 (mfGridElementWidthInv / HeightInv)
    f.invW = static_cast<float>(GRID_COLS) / static_cast<float>(max_x - min_x);
    f.invH = static_cast<float>(GRID_ROWS) / static_cast<float>(max_y - min_y);
    f.cell_of = fr->cell_of.as<int>(); f.cell_start = fr->cell_start.as<int>(); f.cell_items = fr->cell_items.as<int>();
    auto fail = [&](int code) { delete fr; return code; };
    if (n > 0) {
        const void *ksrc = kps, *dsrc = desc;
        if (from_host) {             // through this thread's pinned block: the copies are asynchronous and the call synchronises once
            uint8_t* pin = t_pin.get((size_t)n * (sizeof(orbx_keypoint) + 32));
            if (!pin) return fail(ORBB_E_NOMEM);
            memcpy(pin, kps, (size_t)n * sizeof(orbx_keypoint)); memcpy(pin + (size_t)n * sizeof(orbx_keypoint), desc, (size_t)n * 32);
            ksrc = pin; dsrc = pin + (size_t)n * sizeof(orbx_keypoint);
        }
        if (cudaMemcpyAsync(fr->kps.p, ksrc, (size_t)n * sizeof(orbx_keypoint), kind, st) != cudaSuccess) return fail(ORBB_E_CUDA);
        if (cudaMemcpyAsync(fr->desc.p, dsrc, (size_t)n * 32, kind, st) != cudaSuccess) return fail(ORBB_E_CUDA);
    }
    if (cudaMemsetAsync(fr->cursor.p, 0, (size_t)NC * 4, st) != cudaSuccess) return fail(ORBB_E_CUDA);
    if (n > 0) { grid_assign_kernel<<<ceil_div(n, 256), 256, 0, st>>>(f, fr->cursor.as<int>()); ORBB_LAUNCHED(); }
    grid_scan_kernel<<<1, 1024, 0, st>>>(fr->cursor.as<int>(), f.cell_start); ORBB_LAUNCHED();
    if (n > 0) {
        grid_fill_kernel<<<ceil_div(n, 256), 256, 0, st>>>(f, fr->cursor.as<int>()); ORBB_LAUNCHED();
        grid_sort_kernel<<<ceil_div(NC, 256), 256, 0, st>>>(f); ORBB_LAUNCHED();
    }
    cudaError_t e = from_host ? cudaStreamSynchronize(st) : cudaGetLastError();
    if (e == cudaSuccess && !from_host) {
        e = cudaEventCreateWithFlags(&fr->ready, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventRecord(fr->ready, st);
    }
    if (e != cudaSuccess) { set_cuda_error(e, "orbm_frame_create"); return fail(ORBB_E_CUDA); }
    *out = fr;
    return ORBB_OK;
}

extern "C" int orbm_frame_create(orbm_frame** out, const orbx_keypoint* kps, const uint8_t* desc, int n,
                                 float min_x, float min_y, float max_x, float max_y, int device)
{
    return frame_create_impl(out, kps, desc, n, min_x, min_y, max_x, max_y, device, cudaMemcpyHostToDevice, 0);
}

extern "C" int orbm_frame_create_device(orbm_frame** out, const orbx_keypoint* kps_dev, const uint8_t* desc_dev, int n,
                                        float min_x, float min_y, float max_x, float max_y, int device, void* stream)
{
    return frame_create_impl(out, kps_dev, desc_dev, n, min_x, min_y, max_x, max_y, device, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
}

extern "C" int orbm_frame_destroy(orbm_frame* f)
{
    if (!f) return ORBB_E_INVALID;
    cudaSetDevice(f->device);
    // asynchronous searches may still be queued (cudaFree used to wait for them implicitly; the block cache does not)
    cudaStreamSynchronize(f->s_create);
    if (f->s_last != f->s_create) cudaStreamSynchronize(f->s_last);
    cudaGetLastError();                 // a stream the caller already destroyed is not an error here
    delete f;
    return ORBB_OK;
}

extern "C" int orbm_window_search(const orbm_frame* fr, const orbm_query* q_host, int nq, int32_t* offs_host,
                                  int32_t* idx_host, int32_t* dist_host, int cap)
{
    if (!fr || nq < 0 || !offs_host || cap < 0 || (nq > 0 && !q_host) || (cap > 0 && (!idx_host || !dist_host))) return ORBB_E_INVALID;
    int rc = check_device(fr->device);
    if (rc != ORBB_OK || (rc = frame_wait_ready(fr)) != ORBB_OK) return rc;
    if (nq == 0) { offs_host[0] = 0; return ORBB_OK; }
    DevBuf q, counts, offs, idx, dist;
    if ((rc = q.alloc((size_t)nq * sizeof(orbm_query))) || (rc = counts.alloc((size_t)nq * 4)) || (rc = offs.alloc((size_t)(nq + 1) * 4)) ||
        (rc = idx.alloc((size_t)cap * 4)) || (rc = dist.alloc((size_t)cap * 4))) return rc;
    ORBB_CUDA(cudaMemcpy(q.p, q_host, (size_t)nq * sizeof(orbm_query), cudaMemcpyHostToDevice));
    const int wpb = 8;
    window_search_kernel<false><<<ceil_div(nq, wpb), wpb * 32>>>(fr->f, q.p, sizeof(orbm_query), nq, counts.as<int>(), nullptr, nullptr, nullptr, 0);
    ORBB_LAUNCHED();
    excl_scan_kernel<<<1, 1024>>>(counts.as<int>(), offs.as<int>(), nq);
    ORBB_LAUNCHED();
    window_search_kernel<true><<<ceil_div(nq, wpb), wpb * 32>>>(fr->f, q.p, sizeof(orbm_query), nq, nullptr, offs.as<int>(), idx.as<int>(), dist.as<int>(), cap);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    ORBB_CUDA(cudaMemcpy(offs_host, offs.p, (size_t)(nq + 1) * 4, cudaMemcpyDeviceToHost));
    const int total = offs_host[nq];
    const int m = total < cap ? total : cap;
    if (m > 0) {
        ORBB_CUDA(cudaMemcpy(idx_host, idx.p, (size_t)m * 4, cudaMemcpyDeviceToHost));
        ORBB_CUDA(cudaMemcpy(dist_host, dist.p, (size_t)m * 4, cudaMemcpyDeviceToHost));
    }
    return total > cap ? ORBB_E_CAPACITY : ORBB_OK;
}

// ---- full Search* functions: device candidates + host resolution -------------------------------------------
namespace {
constexpr int HISTO_LENGTH = 30;                 // ORBmatcher.cpp:39

// CSR candidates for a query list, growing the buffers until they fit.
int run_window_search(const orbm_frame* f, const std::vector<orbm_query>& q, std::vector<int>& offs, std::vector<int>& idx, std::vector<int>& dist)
{
    const int nq = (int)q.size();
    offs.assign(nq + 1, 0);
    int rc0 = check_device(f->device);
    if (rc0 != ORBB_OK || (rc0 = frame_host_copies(f)) != ORBB_OK) return rc0;    // the callers' resolve loops read h_kps
    int cap = 64 * nq + 1024;
    for (int attempt = 0; attempt < 2; attempt++) {
        idx.resize(cap); dist.resize(cap);
        const int rc = orbm_window_search(f, q.data(), nq, offs.data(), idx.data(), dist.data(), cap);
        if (rc == ORBB_OK) return ORBB_OK;
        if (rc != ORBB_E_CAPACITY) return rc;
        cap = offs[nq] + 16;
    }
    return ORBB_E_CAPACITY;
}

// ORBmatcher::ComputeThreeMaxima (ORBmatcher.cpp:1792-1833)
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
}

inline int rot_bin(float a1, float a2)
{
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * (HISTO_LENGTH / 360.0f));
    if (bin == HISTO_LENGTH) bin = 0;
    if (bin < 0 || bin > HISTO_LENGTH) bin = 0;        // angles outside [0, 360) are outside the reference's domain (it would index past rotHist): keep the C ABI memory-safe
    return bin;
}
}  // namespace

// ---- SearchByProjection(Frame&, vector<MapPoint*>) / (Frame&, const Frame&): candidates AND resolve on the device ------------
// (SURVEY 8(f)-1).  Everything is queued on `st`; nothing is read back here.
static int resolve_on_device(const orbm_frame* fr, int mode, const float* uright_dev, const orbm_proj_query* q_dev, int nq, float nnratio,
                             int check_ori, int32_t* frame_mp_dev, int32_t* frame_mp_obs_dev, int32_t* nmatches_dev, int32_t* status_dev,
                             int32_t* rounds_dev, int cand_capacity, cudaStream_t st, int max_dist = 100 /* TH_HIGH */, int32_t* total_out_dev = nullptr)
{
    const int cap = cand_capacity > 0 ? cand_capacity : 64 * nq + 1024;
    int rc;
    if (fr->ready) ORBB_CUDA(cudaStreamWaitEvent(st, fr->ready, 0));      // the frame may have been built on another stream
    if (fr->s_last != st && fr->s_last != fr->s_create) cudaStreamSynchronize(fr->s_last);       // a third stream: keep only two to drain at destroy
    fr->s_last = st;
    // ws_offs: [0, nq) first candidate of each query, [nq] total (host retry reads it), ws_counts: [0, nq) one past the last candidate
    if ((rc = fr->ws_counts.reserve((size_t)nq * 4)) ||
        (rc = fr->ws_offs.reserve((size_t)(nq + 1) * 4)) || (rc = fr->ws_idx.reserve((size_t)cap * 4)) ||
        (rc = fr->ws_dist.reserve((size_t)cap * 4)) || (rc = fr->ws_choice.reserve((size_t)nq * 4)) ||
        (rc = fr->ws_taker.reserve((size_t)fr->n * 4))) return rc;
    const int wpb = 8;
    const WindowFilter flt{frame_mp_dev, frame_mp_obs_dev, uright_dev};
    int* total_dev = total_out_dev ? total_out_dev : fr->ws_offs.as<int>() + nq;      // number of candidates found (> cap: overflow)
    ORBB_CUDA(cudaMemsetAsync(total_dev, 0, 4, st));
    window_search_onepass_kernel<true><<<ceil_div(nq, wpb), wpb * 32, 0, st>>>(fr->f, q_dev, sizeof(orbm_proj_query), nq, fr->ws_offs.as<int>(), fr->ws_counts.as<int>(),
                                                                             total_dev, fr->ws_idx.as<int>(), fr->ws_dist.as<int>(), cap, flt);
    ORBB_LAUNCHED();
    ResolveArgs a{};
    a.f = fr->f; a.q = q_dev; a.nq = nq; a.offs = fr->ws_offs.as<int>(); a.ends = fr->ws_counts.as<int>(); a.total = total_dev;
    a.idx = fr->ws_idx.as<int>(); a.dist = fr->ws_dist.as<int>(); a.cap = cap;
    a.uright = uright_dev; a.nnratio = nnratio; a.check_ori = check_ori; a.max_dist = max_dist; a.frame_mp = frame_mp_dev; a.frame_mp_obs = frame_mp_obs_dev;
    a.choice = fr->ws_choice.as<int>(); a.taker = fr->ws_taker.as<int>(); a.nmatches = nmatches_dev; a.status = status_dev; a.rounds = rounds_dev;
    const size_t smem = fr->n <= RESOLVE_SMEM_SLOTS ? (size_t)fr->n * 4 : 0;
    if (mode == 0) resolve_projection_kernel<0><<<1, 1024, smem, st>>>(a);
    else if (mode == 1) resolve_projection_kernel<1><<<1, 1024, smem, st>>>(a);
    else resolve_projection_kernel<2><<<1, 1024, smem, st>>>(a);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_search_by_projection_points_device(const orbm_frame* f, const float* mvuRight_dev, const orbm_proj_query* q_dev, int nq, float nnratio,
                                                       int32_t* frame_mp_dev, int32_t* frame_mp_obs_dev, int32_t* nmatches_dev, int32_t* status_dev,
                                                       int cand_capacity, void* stream)
{
    if (!f || nq <= 0 || !q_dev || !nmatches_dev || !status_dev || (f->n > 0 && (!frame_mp_dev || !frame_mp_obs_dev))) return ORBB_E_INVALID;
    const int rc = check_device(f->device);
    if (rc != ORBB_OK) return rc;
    return resolve_on_device(f, 0, mvuRight_dev, q_dev, nq, nnratio, 0, frame_mp_dev, frame_mp_obs_dev, nmatches_dev, status_dev, nullptr, cand_capacity,
                             (cudaStream_t)stream);
}

extern "C" int orbm_search_by_projection_frame_device(const orbm_frame* f, const float* mvuRight_dev, const orbm_proj_query* q_dev, int nq, int check_ori,
                                                      int32_t* frame_mp_dev, int32_t* frame_mp_obs_dev, int32_t* nmatches_dev, int32_t* status_dev,
                                                      int cand_capacity, void* stream)
{
    if (!f || nq <= 0 || !q_dev || !nmatches_dev || !status_dev || (f->n > 0 && (!frame_mp_dev || !frame_mp_obs_dev))) return ORBB_E_INVALID;
    const int rc = check_device(f->device);
    if (rc != ORBB_OK) return rc;
    return resolve_on_device(f, 1, mvuRight_dev, q_dev, nq, 0.f, check_ori, frame_mp_dev, frame_mp_obs_dev, nmatches_dev, status_dev, nullptr, cand_capacity,
                             (cudaStream_t)stream);
}

// Host-pointer forms: upload the queries and the Frame's match state, run the device path, read the state back.
static int search_projection_host(const orbm_frame* f, int mode, const float* uright, const orbm_proj_query* pq, int nq, float nnratio, int check_ori,
                                  int32_t* frame_mp, int32_t* frame_mp_obs, int* nmatches, int max_dist = 100 /* TH_HIGH */)
{
    if (!f || nq < 0 || !nmatches || (nq > 0 && !pq) || (f->n > 0 && (!frame_mp || (mode != 2 && !frame_mp_obs)))) return ORBB_E_INVALID;
    *nmatches = 0;
    int rc = check_device(f->device);
    if (rc != ORBB_OK) return rc;
    if (nq == 0) return ORBB_OK;
    const bool own_uright = uright == ORBM_URIGHT_OF_FRAME;            // mvuRight as orbm_stereo_frame_create left it on the device
    if (own_uright && !f->uright_valid && f->n > 0) return ORBB_E_STATE;
    const size_t nb = (size_t)f->n * 4, qb = (size_t)nq * sizeof(orbm_proj_query);
    const bool up_uright = uright && !own_uright && f->n > 0;
    // device state block [mvpMapPoints | observations | nmatches, status, rounds, candidates]: one copy in, one copy out
    DevBuf q;
    if ((rc = q.alloc(qb)) || (rc = f->ws_mp.reserve(2 * nb + 16)) || (up_uright && (rc = f->ws_uright.reserve(nb)))) return rc;
    // one pinned block per call: [queries | mvuRight | state block]; every copy is asynchronous on this thread's stream and the call
    // synchronises once per attempt
    uint8_t* pin = t_pin.get(qb + 3 * nb + 64);
    if (!pin) return ORBB_E_NOMEM;
    uint8_t *p_q = pin, *p_u = pin + qb, *p_mp = p_u + nb, *p_obs = p_mp + nb;
    int32_t* p_h = reinterpret_cast<int32_t*>(p_obs + nb);
    int32_t *d_mp = f->ws_mp.as<int32_t>(), *d_obs = d_mp + f->n, *sc = d_obs + f->n;
    cudaStream_t st = 0;                                               // this thread's default stream
    memcpy(p_q, pq, qb);
    ORBB_CUDA(cudaMemcpyAsync(q.p, p_q, qb, cudaMemcpyHostToDevice, st));
    if (up_uright) { memcpy(p_u, uright, nb); ORBB_CUDA(cudaMemcpyAsync(f->ws_uright.p, p_u, nb, cudaMemcpyHostToDevice, st)); }
    const float* ur_dev = own_uright ? (f->n > 0 ? (const float*)f->uright_own.p : nullptr) : (uright ? f->ws_uright.as<float>() : nullptr);
    int cap = 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        if (f->n > 0) {
            memcpy(p_mp, frame_mp, nb);
            if (frame_mp_obs) memcpy(p_obs, frame_mp_obs, nb);
            ORBB_CUDA(cudaMemcpyAsync(d_mp, p_mp, frame_mp_obs ? 2 * nb : nb, cudaMemcpyHostToDevice, st));
        }
        rc = resolve_on_device(f, mode, ur_dev, q.as<orbm_proj_query>(), nq, nnratio, check_ori, d_mp, frame_mp_obs ? d_obs : nullptr, sc, sc + 1, sc + 2, cap, st,
                               max_dist, sc + 3);
        if (rc != ORBB_OK) return rc;
        ORBB_CUDA(cudaMemcpyAsync(p_mp, d_mp, 2 * nb + 16, cudaMemcpyDeviceToHost, st));
        ORBB_CUDA(cudaStreamSynchronize(st));
        if (p_h[1] == ORBB_E_CAPACITY && attempt == 0) {            // more candidates than the default budget: size it exactly
            cap = p_h[3] + 16;
            continue;
        }
        if (p_h[1] != ORBB_OK) return p_h[1];
        *nmatches = p_h[0];
        if (f->n > 0) {
            memcpy(frame_mp, p_mp, nb);
            if (frame_mp_obs) memcpy(frame_mp_obs, p_obs, nb);
        }
        return ORBB_OK;
    }
    return ORBB_E_CAPACITY;
}

extern "C" int orbm_search_by_projection_points(const orbm_frame* f, const float* uright, const orbm_proj_query* pq, int nq,
                                                float nnratio, int32_t* frame_mp, int32_t* frame_mp_obs, int* nmatches)
{
    return search_projection_host(f, 0, uright, pq, nq, nnratio, 0, frame_mp, frame_mp_obs, nmatches);
}

extern "C" int orbm_search_by_projection_frame(const orbm_frame* f, const float* uright, const orbm_proj_query* pq, int nq,
                                               int check_ori, int32_t* frame_mp, int32_t* frame_mp_obs, int* nmatches)
{
    return search_projection_host(f, 1, uright, pq, nq, 0.f, check_ori, frame_mp, frame_mp_obs, nmatches);
}

// ---- relocalisation / loop-closing / fusion variants --------------------------
// They all project on the host (pose algebra, PredictScale) and share the device window search; only the candidate gates and
// the bookkeeping differ.

// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist) (ORBmatcher.cpp:1661-1789):
// a candidate is skipped when the Frame feature already has a MapPoint (:1727-1728), best distance <= ORBdist (:1742).
extern "C" int orbm_search_by_projection_reloc(const orbm_frame* f, const orbm_proj_query* pq, int nq, int max_dist, int check_ori,
                                               int32_t* frame_mp, int* nmatches)
{
    if (max_dist < 0) return ORBB_E_INVALID;
    // window search + resolve on the device (resolve_projection_kernel<2>: every occupied slot is skipped, any accepted match blocks its slot
    // for the later queries, best <= ORBdist, rotation histogram :1766-1786); a query without candidates never matches, whatever max_dist
    return search_projection_host(f, 2, nullptr, pq, nq, 0.f, check_ori, frame_mp, nullptr, nmatches, max_dist > 255 ? 255 : max_dist);
}

// The same with device arrays, queued on `stream` without host synchronisation (include/orbb200.h).
extern "C" int orbm_search_by_projection_reloc_device(const orbm_frame* f, const orbm_proj_query* q_dev, int nq, int max_dist, int check_ori,
                                                      int32_t* frame_mp_dev, int32_t* nmatches_dev, int32_t* status_dev, int cand_capacity, void* stream)
{
    if (!f || nq <= 0 || max_dist < 0 || !q_dev || !nmatches_dev || !status_dev || (f->n > 0 && !frame_mp_dev)) return ORBB_E_INVALID;
    const int rc = check_device(f->device);
    if (rc != ORBB_OK) return rc;
    return resolve_on_device(f, 2, nullptr, q_dev, nq, 0.f, check_ori, frame_mp_dev, nullptr, nmatches_dev, status_dev, nullptr, cand_capacity,
                             (cudaStream_t)stream, max_dist > 255 ? 255 : max_dist);
}

// SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, th)
// (ORBmatcher.cpp:352-472): KeyFrame::GetFeaturesInArea has no level filter, the level gate [p-1, p] follows (:441-444) - the
// same set as a window query with that level range; already matched features are skipped (:438-439); best <= TH_LOW (:457); no
// orientation check.  The relocalisation resolve with those parameters.
extern "C" int orbm_search_by_projection_loop(const orbm_frame* f, const orbm_proj_query* pq, int nq, int32_t* matched, int* nmatches)
{
    return search_projection_host(f, 2, nullptr, pq, nq, 0.f, 0, matched, nullptr, nmatches, TH_LOW);
}

// The per-point search of Fuse(KeyFrame*, const vector<MapPoint*>&, th) (ORBmatcher.cpp:962-1115; chi-square reprojection gate
// :1053-1078), Fuse(KeyFrame*, cv::Mat Scw, ...) (:1119-1251, no gate) and of each direction of SearchBySim3 (:1253-1500, no gate,
// TH_HIGH): first candidate with the smallest distance among those passing the level range of the query and, when
// inv_level_sigma2 is given, the gate; best_idx = -1 unless best distance <= max_dist.  Queries do not interact.
extern "C" int orbm_fuse_best(const orbm_frame* f, const orbm_proj_query* pq, int nq, const float* uright, const float* inv_level_sigma2,
                              int max_dist, int32_t* best_idx, int32_t* best_dist)
{
    if (!f || nq < 0 || (nq > 0 && (!pq || !best_idx))) return ORBB_E_INVALID;
    if (nq == 0) return ORBB_OK;
    int rc = check_device(f->device);
    if (rc != ORBB_OK || (rc = frame_wait_ready(f)) != ORBB_OK) return rc;
    // window search and the per-point selection on the device (fuse_best_kernel); only best_idx / best_dist come back
    int nlev = 0;
    if (inv_level_sigma2) {                               // the caller's table has one entry per pyramid level: upload as many as the frame's octaves need
        if ((rc = frame_host_copies(f)) != ORBB_OK) return rc;
        for (int k = 0; k < f->n; k++) nlev = std::max(nlev, f->h_kps[k].octave + 1);
        if (nlev > 64) return ORBB_E_INVALID;
    }
    DevBuf q, counts, offs, idx, dist, dur, dsig, dbi, dbd;
    const size_t nb = (size_t)f->n * 4;
    if ((rc = q.alloc((size_t)nq * sizeof(orbm_proj_query))) || (rc = counts.alloc((size_t)nq * 4)) || (rc = offs.alloc((size_t)(nq + 1) * 4)) ||
        (rc = dbi.alloc((size_t)nq * 4)) || (rc = dbd.alloc((size_t)nq * 4)) || (uright && inv_level_sigma2 && (rc = dur.alloc(nb))) ||
        (nlev > 0 && (rc = dsig.alloc((size_t)nlev * 4)))) return rc;
    ORBB_CUDA(cudaMemcpy(q.p, pq, (size_t)nq * sizeof(orbm_proj_query), cudaMemcpyHostToDevice));
    if (dur.p && f->n > 0) ORBB_CUDA(cudaMemcpy(dur.p, uright, nb, cudaMemcpyHostToDevice));
    if (dsig.p) ORBB_CUDA(cudaMemcpy(dsig.p, inv_level_sigma2, (size_t)nlev * 4, cudaMemcpyHostToDevice));
    const int wpb = 8;
    window_search_kernel<false><<<ceil_div(nq, wpb), wpb * 32>>>(f->f, q.p, sizeof(orbm_proj_query), nq, counts.as<int>(), nullptr, nullptr, nullptr, 0);
    ORBB_LAUNCHED();
    excl_scan_kernel<<<1, 1024>>>(counts.as<int>(), offs.as<int>(), nq);
    ORBB_LAUNCHED();
    int cap = 64 * nq + 1024;
    for (int attempt = 0; attempt < 2; attempt++) {
        if ((rc = idx.reserve((size_t)cap * 4)) || (rc = dist.reserve((size_t)cap * 4))) return rc;
        window_search_kernel<true><<<ceil_div(nq, wpb), wpb * 32>>>(f->f, q.p, sizeof(orbm_proj_query), nq, nullptr, offs.as<int>(), idx.as<int>(), dist.as<int>(), cap);
        ORBB_LAUNCHED();
        fuse_best_kernel<<<ceil_div(nq, 128), 128>>>(f->f, q.as<orbm_proj_query>(), nq, offs.as<int>(), idx.as<int>(), dist.as<int>(), cap,
                                                    dur.p ? dur.as<float>() : nullptr, dsig.p ? dsig.as<float>() : nullptr, max_dist, dbi.as<int>(), dbd.as<int>());
        ORBB_LAUNCHED();
        ORBB_CUDA(cudaGetLastError());
        int total = 0;
        ORBB_CUDA(cudaMemcpy(&total, offs.as<int>() + nq, 4, cudaMemcpyDeviceToHost));
        if (total <= cap) {
            ORBB_CUDA(cudaMemcpy(best_idx, dbi.p, (size_t)nq * 4, cudaMemcpyDeviceToHost));
            if (best_dist) ORBB_CUDA(cudaMemcpy(best_dist, dbd.p, (size_t)nq * 4, cudaMemcpyDeviceToHost));
            return ORBB_OK;
        }
        cap = total + 16;                                 // more candidates than the default budget: size it exactly
    }
    return ORBB_E_CAPACITY;
}

// ---- ORBmatcher::SearchByBoW ---------------------------------------------------------------------------
// Shared front half of both SearchByBoW overloads: walk the two FeatureVectors together (the lower_bound dance of :224-310 /
// :634-720 is a sorted-key intersection), make one join entry per A feature that carries a good MapPoint, and compute the
// all-pairs distances of every common vocabulary node on the device (bow_join_kernel).  Entries come out in (node, A feature)
// order, which is the order the reference's loops visit them.
static int bow_join(const uint8_t* desc_a, const uint8_t* has_a, int n_a, const int32_t* fa_nodes, const int32_t* fa_off, const int32_t* fa_feat, int nfa,
                    const uint8_t* desc_b, int n_b, const int32_t* fb_nodes, const int32_t* fb_off, const int32_t* fb_feat, int nfb,
                    int device, std::vector<int4>& entries, std::vector<uint16_t>& dist)
{
    if (n_a < 0 || n_b < 0 || nfa < 0 || nfb < 0) return ORBB_E_INVALID;
    if ((n_a > 0 && (!desc_a || !has_a)) || (n_b > 0 && !desc_b)) return ORBB_E_INVALID;
    if ((nfa > 0 && (!fa_nodes || !fa_off || !fa_feat)) || (nfb > 0 && (!fb_nodes || !fb_off || !fb_feat))) return ORBB_E_INVALID;
    for (int j = 0; j < nfa; j++) {
        if (fa_off[j + 1] < fa_off[j] || (j > 0 && fa_nodes[j] <= fa_nodes[j - 1])) return ORBB_E_INVALID;
        for (int t = fa_off[j]; t < fa_off[j + 1]; t++) if (fa_feat[t] < 0 || fa_feat[t] >= n_a) return ORBB_E_INVALID;
    }
    for (int j = 0; j < nfb; j++) {
        if (fb_off[j + 1] < fb_off[j] || (j > 0 && fb_nodes[j] <= fb_nodes[j - 1])) return ORBB_E_INVALID;
        for (int t = fb_off[j]; t < fb_off[j + 1]; t++) if (fb_feat[t] < 0 || fb_feat[t] >= n_b) return ORBB_E_INVALID;
    }
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    entries.clear();
    long long total = 0;
    for (int a = 0, b = 0; a < nfa && b < nfb;) {
        if (fa_nodes[a] < fb_nodes[b]) { a++; continue; }
        if (fa_nodes[a] > fb_nodes[b]) { b++; continue; }
        const int nB = fb_off[b + 1] - fb_off[b];
        for (int t = fa_off[a]; t < fa_off[a + 1]; t++) {
            const int iA = fa_feat[t];
            if (!has_a[iA]) continue;                                                   // :233-238 / :643-647
            if (total + nB > 0x7fffffffLL) return ORBB_E_CAPACITY;
            entries.push_back(make_int4(iA, fb_off[b], nB, (int)total));
            total += nB;
        }
        a++; b++;
    }
    dist.assign((size_t)total, 0);
    if (entries.empty() || total == 0) return ORBB_OK;
    DevBuf dA, dB, dF, dE, dD;
    const int nfeatB = fb_off[nfb];
    const size_t bA = (size_t)n_a * 32, bB = (size_t)n_b * 32, bF = (size_t)nfeatB * 4, bE = entries.size() * sizeof(int4), bD = (size_t)total * 2;
    if ((rc = dA.alloc(bA)) || (rc = dB.alloc(bB)) || (rc = dF.alloc(bF)) || (rc = dE.alloc(bE)) || (rc = dD.alloc(bD))) return rc;
    // one pinned block per call: [desc A | desc B | features of B | entries | distances]; asynchronous copies, one synchronisation
    uint8_t* pin = t_pin.get(bA + bB + bF + bE + bD + 64);
    if (!pin) return ORBB_E_NOMEM;
    uint8_t *pA = pin, *pB = pA + bA, *pF = pB + bB, *pE = pF + bF, *pD = pE + bE;
    cudaStream_t st = 0;
    memcpy(pA, desc_a, bA); memcpy(pB, desc_b, bB); memcpy(pF, fb_feat, bF); memcpy(pE, entries.data(), bE);
    ORBB_CUDA(cudaMemcpyAsync(dA.p, pA, bA, cudaMemcpyHostToDevice, st));
    ORBB_CUDA(cudaMemcpyAsync(dB.p, pB, bB, cudaMemcpyHostToDevice, st));
    ORBB_CUDA(cudaMemcpyAsync(dF.p, pF, bF, cudaMemcpyHostToDevice, st));
    ORBB_CUDA(cudaMemcpyAsync(dE.p, pE, bE, cudaMemcpyHostToDevice, st));
    bow_join_kernel<<<ceil_div((int)entries.size(), 8), 256, 0, st>>>(dA.as<uint4>(), dB.as<uint4>(), dF.as<int>(), dE.as<int4>(), (int)entries.size(),
                                                                      dD.as<uint16_t>());
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    ORBB_CUDA(cudaMemcpyAsync(pD, dD.p, bD, cudaMemcpyDeviceToHost, st));
    ORBB_CUDA(cudaStreamSynchronize(st));
    memcpy(dist.data(), pD, bD);
    return ORBB_OK;
}

static void bow_rot_filter(std::vector<int>* rotHist, int32_t* matches, int& nm)          // :313-345 / :722-740
{
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int b = 0; b < HISTO_LENGTH; b++)
        if (b != ind1 && b != ind2 && b != ind3)
            for (size_t j = 0; j < rotHist[b].size(); j++) { matches[rotHist[b][j]] = -1; nm--; }
}

// SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) (ORBmatcher.cpp:205-348).  The scan that skips already
// matched Frame features (:243-245) is order dependent and runs on the host exactly as written.
extern "C" int orbm_search_by_bow(const uint8_t* desc_kf, const float* angle_kf, const uint8_t* kf_has_mp, int n_kf,
                                  const int32_t* fv_kf_nodes, const int32_t* fv_kf_off, const int32_t* fv_kf_feat, int nfv_kf,
                                  const uint8_t* desc_f, const float* angle_f, int n_f,
                                  const int32_t* fv_f_nodes, const int32_t* fv_f_off, const int32_t* fv_f_feat, int nfv_f,
                                  float nnratio, int check_ori, int32_t* match_f_to_kf, int* nmatches, int device)
{
    if (!nmatches || (n_f > 0 && !match_f_to_kf)) return ORBB_E_INVALID;
    if (check_ori && ((n_kf > 0 && !angle_kf) || (n_f > 0 && !angle_f))) return ORBB_E_INVALID;
    std::vector<int4> entries; std::vector<uint16_t> dist;
    const int rc = bow_join(desc_kf, kf_has_mp, n_kf, fv_kf_nodes, fv_kf_off, fv_kf_feat, nfv_kf, desc_f, n_f, fv_f_nodes, fv_f_off, fv_f_feat, nfv_f,
                            device, entries, dist);
    if (rc != ORBB_OK) return rc;
    *nmatches = 0;
    for (int i = 0; i < n_f; i++) match_f_to_kf[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int nm = 0;
    for (const int4& en : entries) {
        int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
        for (int j = 0; j < en.z; j++) {
            const int realIdxF = fv_f_feat[en.y + j];
            if (match_f_to_kf[realIdxF] >= 0) continue;                                 // :243-245
            const int d = dist[(size_t)en.w + j];
            if (d < bestDist1) { bestDist2 = bestDist1; bestDist1 = d; bestIdxF = realIdxF; }
            else if (d < bestDist2) bestDist2 = d;
        }
        if (bestDist1 <= TH_LOW && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {   // :264-268
            match_f_to_kf[bestIdxF] = en.x;
            if (check_ori) rotHist[rot_bin(angle_kf[en.x], angle_f[bestIdxF])].push_back(bestIdxF);
            nm++;
        }
    }
    if (check_ori) bow_rot_filter(rotHist, match_f_to_kf, nm);
    *nmatches = nm;
    return ORBB_OK;
}

// SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) (ORBmatcher.cpp:604-743): both sides need a good
// MapPoint (:643-647, :660-666), KeyFrame-2 features are used once (vbMatched2, :662, :691), strict bestDist1 < TH_LOW (:686).
extern "C" int orbm_search_by_bow_kf(const uint8_t* desc1, const float* angle1, const uint8_t* has_mp1, int n1,
                                     const int32_t* fv1_nodes, const int32_t* fv1_off, const int32_t* fv1_feat, int nfv1,
                                     const uint8_t* desc2, const float* angle2, const uint8_t* has_mp2, int n2,
                                     const int32_t* fv2_nodes, const int32_t* fv2_off, const int32_t* fv2_feat, int nfv2,
                                     float nnratio, int check_ori, int32_t* match12, int* nmatches, int device)
{
    if (!nmatches || (n1 > 0 && !match12) || (n2 > 0 && !has_mp2)) return ORBB_E_INVALID;
    if (check_ori && ((n1 > 0 && !angle1) || (n2 > 0 && !angle2))) return ORBB_E_INVALID;
    std::vector<int4> entries; std::vector<uint16_t> dist;
    const int rc = bow_join(desc1, has_mp1, n1, fv1_nodes, fv1_off, fv1_feat, nfv1, desc2, n2, fv2_nodes, fv2_off, fv2_feat, nfv2, device, entries, dist);
    if (rc != ORBB_OK) return rc;
    *nmatches = 0;
    for (int i = 0; i < n1; i++) match12[i] = -1;
    std::vector<uint8_t> vbMatched2((size_t)n2, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nm = 0;
    for (const int4& en : entries) {
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        for (int j = 0; j < en.z; j++) {
            const int idx2 = fv2_feat[en.y + j];
            if (vbMatched2[idx2] || !has_mp2[idx2]) continue;                           // :660-666
            const int d = dist[(size_t)en.w + j];
            if (d < bestDist1) { bestDist2 = bestDist1; bestDist1 = d; bestIdx2 = idx2; }
            else if (d < bestDist2) bestDist2 = d;
        }
        if (bestDist1 < TH_LOW && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {    // :686-688
            match12[en.x] = bestIdx2;
            vbMatched2[bestIdx2] = 1;
            if (check_ori) rotHist[rot_bin(angle1[en.x], angle2[bestIdx2])].push_back(en.x);
            nm++;
        }
    }
    if (check_ori) bow_rot_filter(rotHist, match12, nm);
    *nmatches = nm;
    return ORBB_OK;
}

// ORBmatcher::CheckDistEpipolarLine (ORBmatcher.cpp:166-190): F12 row-major 3x3 floats.
static bool check_dist_epipolar_line(const orbx_keypoint& kp1, const orbx_keypoint& kp2, const float* F12, const float* level_sigma2_2)
{
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * level_sigma2_2[kp2.octave];
}

// ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t,size_t>>& vMatchedPairs,
// bOnlyStereo) (ORBmatcher.cpp:755-927): features WITHOUT a MapPoint on both sides (:793-796, :815-818), optional stereo-only,
// dist <= TH_LOW and <= the best so far (:829-830), the epipole proximity test for mono pairs (:834-840), the epipolar-line test
// (:842-846), vbMatched2, rotation histogram.  Distances of the common vocabulary nodes come from the device join.
extern "C" int orbm_search_for_triangulation(const uint8_t* desc1, const orbx_keypoint* kps_un1, const float* uright1, const uint8_t* has_mp1, int n1,
                                             const int32_t* fv1_nodes, const int32_t* fv1_off, const int32_t* fv1_feat, int nfv1,
                                             const uint8_t* desc2, const orbx_keypoint* kps_un2, const float* uright2, const uint8_t* has_mp2, int n2,
                                             const int32_t* fv2_nodes, const int32_t* fv2_off, const int32_t* fv2_feat, int nfv2,
                                             const float* F12, float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                             int only_stereo, int check_ori, int32_t* match12, int* nmatches, int device)
{
    if (!nmatches || !F12 || !scale_factors2 || !level_sigma2_2 || (n1 > 0 && (!match12 || !kps_un1 || !has_mp1)) || (n2 > 0 && (!kps_un2 || !has_mp2)))
        return ORBB_E_INVALID;
    if (only_stereo && ((n1 > 0 && !uright1) || (n2 > 0 && !uright2))) return ORBB_E_INVALID;
    // KeyFrame-1 features that take part: no MapPoint yet, and stereo when bOnlyStereo (:793-802)
    std::vector<uint8_t> a_ok((size_t)(n1 > 0 ? n1 : 0));
    for (int i = 0; i < n1; i++) a_ok[i] = (!has_mp1[i] && (!only_stereo || uright1[i] >= 0)) ? 1 : 0;
    std::vector<int4> entries; std::vector<uint16_t> dist;
    const int rc = bow_join(desc1, a_ok.data(), n1, fv1_nodes, fv1_off, fv1_feat, nfv1, desc2, n2, fv2_nodes, fv2_off, fv2_feat, nfv2, device, entries, dist);
    if (rc != ORBB_OK) return rc;
    *nmatches = 0;
    for (int i = 0; i < n1; i++) match12[i] = -1;
    std::vector<uint8_t> vbMatched2((size_t)n2, 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nm = 0;
    for (const int4& en : entries) {
        const int idx1 = en.x;
        const bool bStereo1 = uright1 && uright1[idx1] >= 0;
        const orbx_keypoint& kp1 = kps_un1[idx1];
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int j = 0; j < en.z; j++) {
            const int idx2 = fv2_feat[en.y + j];
            if (vbMatched2[idx2] || has_mp2[idx2]) continue;                            // :815-818
            const bool bStereo2 = uright2 && uright2[idx2] >= 0;
            if (only_stereo && !bStereo2) continue;
            const int d = dist[(size_t)en.w + j];
            if (d > TH_LOW || d > bestDist) continue;                                   // :829-830
            const orbx_keypoint& kp2 = kps_un2[idx2];
            if (!bStereo1 && !bStereo2) {                                               // :834-840
                const float distex = ex - kp2.x, distey = ey - kp2.y;
                if (distex * distex + distey * distey < 100 * scale_factors2[kp2.octave]) continue;
            }
            if (check_dist_epipolar_line(kp1, kp2, F12, level_sigma2_2)) { bestIdx2 = idx2; bestDist = d; }
        }
        if (bestIdx2 >= 0) {
            match12[idx1] = bestIdx2;
            vbMatched2[bestIdx2] = 1;
            nm++;
            if (check_ori) rotHist[rot_bin(kp1.angle, kps_un2[bestIdx2].angle)].push_back(idx1);
        }
    }
    if (check_ori) bow_rot_filter(rotHist, match12, nm);
    *nmatches = nm;
    return ORBB_OK;
}

extern "C" int orbm_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_xy, int32_t* m12,
                                              int window, float nnratio, int check_ori, int* nmatches)
{
    if (!f1 || !f2 || !nmatches || (f1->n > 0 && (!prev_xy || !m12))) return ORBB_E_INVALID;
    *nmatches = 0;
    const int n1 = f1->n, n2 = f2->n;
    for (int i = 0; i < n1; i++) m12[i] = -1;
    int rc0 = check_device(f1->device);
    if (rc0 != ORBB_OK || (rc0 = frame_host_copies(f1)) != ORBB_OK || (rc0 = frame_host_copies(f2)) != ORBB_OK) return rc0;
    // level-0 keypoints of F1 are the queries (:490-494)
    std::vector<orbm_query> q; std::vector<int> qi;
    for (int i = 0; i < n1; i++) {
        if (f1->h_kps[i].octave > 0) continue;
        orbm_query e;
        e.u = prev_xy[2 * i]; e.v = prev_xy[2 * i + 1]; e.r = (float)window; e.min_level = 0; e.max_level = 0;
        memcpy(e.desc, &f1->h_desc[(size_t)i * 32], 32);
        q.push_back(e); qi.push_back(i);
    }
    std::vector<int> offs, idx, dist;
    if (!q.empty()) { const int rc = run_window_search(f2, q, offs, idx, dist); if (rc != ORBB_OK) return rc; }
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(n2, 0x7fffffff), vnMatches21(n2, -1);
    int nm = 0;
    for (size_t qq = 0; qq < q.size(); qq++) {
        const int i1 = qi[qq];
        int bestDist = 0x7fffffff, bestDist2 = 0x7fffffff, bestIdx2 = -1;
        for (int c = offs[qq]; c < offs[qq + 1]; c++) {
            const int i2 = idx[c], d = dist[c];
            if (vMatchedDistance[i2] <= d) continue;                                    // :513-514
            if (d < bestDist) { bestDist2 = bestDist; bestDist = d; bestIdx2 = i2; }
            else if (d < bestDist2) bestDist2 = d;
        }
        if (bestDist <= TH_LOW && bestDist < (float)bestDist2 * nnratio) {              // :529-531
            if (vnMatches21[bestIdx2] >= 0) { m12[vnMatches21[bestIdx2]] = -1; nm--; }
            m12[i1] = bestIdx2; vnMatches21[bestIdx2] = i1; vMatchedDistance[bestIdx2] = bestDist;
            nm++;
            if (check_ori) rotHist[rot_bin(f1->h_kps[i1].angle, f2->h_kps[bestIdx2].angle)].push_back(i1);
        }
    }
    if (check_ori) {                                                                    // :558-580
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int b = 0; b < HISTO_LENGTH; b++) {
            if (b == ind1 || b == ind2 || b == ind3) continue;
            for (size_t j = 0; j < rotHist[b].size(); j++) {
                const int i1 = rotHist[b][j];
                if (m12[i1] >= 0) { m12[i1] = -1; nm--; }
            }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)                                                     // :584-587
        if (m12[i1] >= 0) { prev_xy[2 * i1] = f2->h_kps[m12[i1]].x; prev_xy[2 * i1 + 1] = f2->h_kps[m12[i1]].y; }
    *nmatches = nm;
    return ORBB_OK;
}

// ---- stereo ------------------------------------------------------------------------------------------
extern "C" int orbm_stereo_match(orbx_handle* left, orbx_handle* right, const orbx_keypoint* kL, const uint8_t* dL, int nL,
                                 const orbx_keypoint* kR, const uint8_t* dR, int nR, float mbf, float mb,
                                 float* uR_host, float* depth_host)
{
    if (!left || !right || nL < 0 || nR < 0 || (nL > 0 && (!kL || !dL || !uR_host || !depth_host)) || (nR > 0 && (!kR || !dR))) return ORBB_E_INVALID;
    if (nR >= 65536) return ORBB_E_CAPACITY;
    const LevelGeom *ghL, *gdL, *ghR, *gdR; int nlL, nlR, devL, devR; const uint8_t *s0L, *s0R, *pL, *pR; long long spL, spR;
    const float *scL, *iscL, *scR, *iscR; void *stL, *stR;
    int rc = orbx__pyramid_view(left, &ghL, &gdL, &nlL, &s0L, &spL, &pL, &devL, &scL, &iscL, &stL);
    if (rc != ORBB_OK) return rc;
    rc = orbx__pyramid_view(right, &ghR, &gdR, &nlR, &s0R, &spR, &pR, &devR, &scR, &iscR, &stR);
    if (rc != ORBB_OK) return rc;
    if (devL != devR || nlL != nlR || ghL[0].w != ghR[0].w || ghL[0].h != ghR[0].h) return ORBB_E_STATE;
    rc = check_device(devL);
    if (rc != ORBB_OK) return rc;
    if (nL == 0) return ORBB_OK;
    ORBB_CUDA(cudaStreamSynchronize((cudaStream_t)stL));
    ORBB_CUDA(cudaStreamSynchronize((cudaStream_t)stR));
    DevBuf bkL, bdL, bkR, bdR, bu, bd, bs, bi, bb;
    if ((rc = bkL.alloc((size_t)nL * sizeof(orbx_keypoint))) || (rc = bdL.alloc((size_t)nL * 32)) ||
        (rc = bkR.alloc((size_t)nR * sizeof(orbx_keypoint))) || (rc = bdR.alloc((size_t)nR * 32)) ||
        (rc = bu.alloc((size_t)nL * 4)) || (rc = bd.alloc((size_t)nL * 4)) || (rc = bs.alloc((size_t)nL * 4)) ||
        (rc = bi.alloc((size_t)nL * 4)) || (rc = bb.alloc((size_t)nL * 4))) return rc;
    // one pinned block per call: [kL | dL | kR | dR | uRight | depth]; asynchronous copies on this thread's stream, one synchronisation
    const size_t bk = sizeof(orbx_keypoint);
    uint8_t* pin = t_pin.get((size_t)(nL + nR) * (bk + 32) + (size_t)nL * 8);
    if (!pin) return ORBB_E_NOMEM;
    uint8_t *p_kL = pin, *p_dL = p_kL + (size_t)nL * bk, *p_kR = p_dL + (size_t)nL * 32, *p_dR = p_kR + (size_t)nR * bk, *p_u = p_dR + (size_t)nR * 32,
            *p_d = p_u + (size_t)nL * 4;
    cudaStream_t st = 0;
    memcpy(p_kL, kL, (size_t)nL * bk); memcpy(p_dL, dL, (size_t)nL * 32);
    ORBB_CUDA(cudaMemcpyAsync(bkL.p, p_kL, (size_t)nL * bk, cudaMemcpyHostToDevice, st));
    ORBB_CUDA(cudaMemcpyAsync(bdL.p, p_dL, (size_t)nL * 32, cudaMemcpyHostToDevice, st));
    if (nR > 0) {
        memcpy(p_kR, kR, (size_t)nR * bk); memcpy(p_dR, dR, (size_t)nR * 32);
        ORBB_CUDA(cudaMemcpyAsync(bkR.p, p_kR, (size_t)nR * bk, cudaMemcpyHostToDevice, st));
        ORBB_CUDA(cudaMemcpyAsync(bdR.p, p_dR, (size_t)nR * 32, cudaMemcpyHostToDevice, st));
    }
    StereoArgs a;
    memset(&a, 0, sizeof a);
    a.kL = bkL.as<orbx_keypoint>(); a.dL = bdL.as<uint4>(); a.nL = nL;
    a.kR = bkR.as<orbx_keypoint>(); a.dR = bdR.as<uint4>(); a.nR = nR;
    a.geom = gdL; a.nlevels = nlL; a.nRows = ghL[0].h;
    a.src0L = s0L; a.src0L_pitch = spL; a.pyrL = pL;
    a.src0R = s0R; a.src0R_pitch = spR; a.pyrR = pR;
    for (int i = 0; i < nlL; i++) { a.scale[i] = scL[i]; a.inv_scale[i] = iscL[i]; }
    // Frame.cpp:579-581: minZ = mb, minD = 0, maxD = mbf/minZ (mb == 0 at the reference's call => +inf)
    a.mbf = mbf; a.minD = 0.f; a.maxD = mbf / mb;
    a.uRight = bu.as<float>(); a.depth = bd.as<float>(); a.sad = bs.as<int>(); a.bestIdx = bi.as<int>(); a.bestDist = bb.as<int>();
    stereo_match_kernel<<<ceil_div(nL, 8), 256, 0, st>>>(a);
    ORBB_LAUNCHED();
    const float k21 = 1.5f * 1.4f;
    stereo_median_filter_kernel<<<1, 256, 0, st>>>(nL, a.sad, a.uRight, a.depth, k21);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    ORBB_CUDA(cudaMemcpyAsync(p_u, bu.p, (size_t)nL * 4, cudaMemcpyDeviceToHost, st));
    ORBB_CUDA(cudaMemcpyAsync(p_d, bd.p, (size_t)nL * 4, cudaMemcpyDeviceToHost, st));
    ORBB_CUDA(cudaStreamSynchronize(st));
    memcpy(uR_host, p_u, (size_t)nL * 4); memcpy(depth_host, p_d, (size_t)nL * 4);
    return ORBB_OK;
}

// Frame::Frame, stereo constructor (Frame.cpp:61-115) as ONE call: ExtractORB(0, left) || ExtractORB(1, right) as one two-frame batch of
// the handle, UndistortKeyPoints for rectified input (mDistCoef[0] == 0: mvKeysUn = mvKeys, Frame.cpp:467-472), ComputeStereoMatches
// and AssignFeaturesToGrid on the results where they lie in HBM.  Nothing is uploaded twice; the host gets every Frame member back
// (mvKeys, mDescriptors, mvKeysRight, mDescriptorsRight, mvuRight, mvDepth) and the device keeps the frame for the Search* calls.
extern "C" int orbm_stereo_frame_create(orbx_handle* h, const uint8_t* left_host, const uint8_t* right_host, int w, int ht, size_t stride,
                                        float mbf, float mb, float min_x, float min_y, float max_x, float max_y, orbm_frame** out,
                                        orbx_keypoint* kL_host, uint8_t* dL_host, int* nL_out, orbx_keypoint* kR_host, uint8_t* dR_host,
                                        int* nR_out, int capacity, float* uR_host, float* depth_host)
{
    if (!h || !left_host || !right_host || !out || !nL_out || !nR_out || capacity < 1 || !kL_host || !dL_host || !uR_host || !depth_host ||
        ((kR_host == nullptr) != (dR_host == nullptr))) return ORBB_E_INVALID;
    *out = nullptr; *nL_out = 0; *nR_out = 0;
    const uint8_t* imgs[2] = {left_host, right_host};
    int32_t cnt[2] = {0, 0};
    orbx_resident_view v{};
    static const bool trace = getenv("ORBB_TRACE_PAIR") != nullptr;      // debugging aid: host-side phase times on stderr every 100 calls
    static double acc[5] = {0, 0, 0, 0, 0}; static int ncalls = 0;
    auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    const double t0 = trace ? now() : 0.0;
    int rc = orbx__extract_resident(h, imgs, 2, w, ht, stride, cnt, &v);
    const double t1 = trace ? now() : 0.0;
    if (rc != ORBB_OK) { if (v.s_out) cudaStreamSynchronize((cudaStream_t)v.s_out); return rc; }
    cudaStream_t st = (cudaStream_t)v.stream, s_out = (cudaStream_t)v.s_out, s_aux = (cudaStream_t)v.s_aux;
    const int nL = cnt[0], nR = cnt[1];
    *nL_out = nL; *nR_out = nR;
    if (nL > capacity || (kR_host && nR > capacity) || nR >= 65536) { cudaStreamSynchronize(s_out); return ORBB_E_CAPACITY; }
    uint8_t* pin = t_pin.get((size_t)(nL > 0 ? nL : 1) * 8);
    if (!pin) { cudaStreamSynchronize(s_out); return ORBB_E_NOMEM; }
    // all three streams are idle here (the extraction was drained to read the counts): stereo matching runs on s_aux beside the grid on st
    DevBuf ur, bd, bs, bi, bb;
    if (nL > 0) {
        if ((rc = ur.alloc((size_t)nL * 4)) || (rc = bd.alloc((size_t)nL * 4)) || (rc = bs.alloc((size_t)nL * 4)) ||
            (rc = bi.alloc((size_t)nL * 4)) || (rc = bb.alloc((size_t)nL * 4))) { cudaStreamSynchronize(s_out); return rc; }
        StereoArgs a;
        memset(&a, 0, sizeof a);
        a.kL = v.kps_dev; a.dL = reinterpret_cast<const uint4*>(v.desc_dev); a.nL = nL;
        a.kR = v.kps_dev + v.devcap; a.dR = reinterpret_cast<const uint4*>(v.desc_dev + (size_t)v.devcap * 32); a.nR = nR;
        a.geom = v.geom_dev; a.nlevels = v.nlevels; a.nRows = v.geom_host[0].h;
        a.src0L = v.src0; a.src0L_pitch = v.src0_pitch; a.pyrL = v.pyr;
        a.src0R = v.src0 + v.src0_fstride; a.src0R_pitch = v.src0_pitch; a.pyrR = v.pyr + v.pyr_fstride;
        for (int i = 0; i < v.nlevels; i++) { a.scale[i] = v.scale[i]; a.inv_scale[i] = v.inv_scale[i]; }
        a.mbf = mbf; a.minD = 0.f; a.maxD = mbf / mb;                      // Frame.cpp:579-581
        a.uRight = ur.as<float>(); a.depth = bd.as<float>(); a.sad = bs.as<int>(); a.bestIdx = bi.as<int>(); a.bestDist = bb.as<int>();
        stereo_match_kernel<<<ceil_div(nL, 8), 256, 0, s_aux>>>(a);
        ORBB_LAUNCHED();
        stereo_median_filter_kernel<<<1, 256, 0, s_aux>>>(nL, a.sad, a.uRight, a.depth, 1.5f * 1.4f);
        ORBB_LAUNCHED();
        if (cudaMemcpyAsync(pin, a.uRight, (size_t)nL * 4, cudaMemcpyDeviceToHost, s_aux) != cudaSuccess ||
            cudaMemcpyAsync(pin + (size_t)nL * 4, a.depth, (size_t)nL * 4, cudaMemcpyDeviceToHost, s_aux) != cudaSuccess) {
            cudaStreamSynchronize(s_aux); cudaStreamSynchronize(s_out); return ORBB_E_CUDA;
        }
    }
    orbm_frame* fr = nullptr;
    rc = frame_create_impl(&fr, v.kps_dev, v.desc_dev, nL, min_x, min_y, max_x, max_y, v.device, cudaMemcpyDeviceToDevice, st);
    if (rc != ORBB_OK) { cudaStreamSynchronize(s_aux); cudaStreamSynchronize(s_out); return rc; }
    auto fail = [&](int code) { cudaStreamSynchronize(st); cudaStreamSynchronize(s_aux); cudaStreamSynchronize(s_out); orbm_frame_destroy(fr); return code; };
    const double t2 = trace ? now() : 0.0;
    cudaError_t e = cudaStreamSynchronize(s_out);              // keypoints and descriptors of both images are in the handle's pinned block
    if (e == cudaSuccess) e = cudaStreamSynchronize(s_aux);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_cuda_error(e, "orbm_stereo_frame_create"); return fail(ORBB_E_CUDA); }
    const double t3 = trace ? now() : 0.0;
    if (nL > 0) {                                              // mvuRight stays with the frame (ORBM_URIGHT_OF_FRAME)
        fr->uright_own.p = ur.p; fr->uright_own.cap = ur.cap; ur.p = nullptr; ur.cap = 0;
        fr->uright_valid = true;
    }
    if (nL > 0) {
        memcpy(kL_host, v.kps_pinned, (size_t)nL * sizeof(orbx_keypoint)); memcpy(dL_host, v.desc_pinned, (size_t)nL * 32);
        memcpy(uR_host, pin, (size_t)nL * 4); memcpy(depth_host, pin + (size_t)nL * 4, (size_t)nL * 4);
    }
    if (kR_host && nR > 0) {
        memcpy(kR_host, v.kps_pinned + v.devcap, (size_t)nR * sizeof(orbx_keypoint)); memcpy(dR_host, v.desc_pinned + (size_t)v.devcap * 32, (size_t)nR * 32);
    }
    if (trace) {
        const double t4 = now();
        acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2; acc[3] += t4 - t3;
        if (++ncalls % 100 == 0) {
            fprintf(stderr, "[pair trace] extract (stage, launch, wait for counts) %.1f us, stereo + grid queued %.1f us, drained %.1f us, results copied out %.1f us\n",
                    acc[0] / 100, acc[1] / 100, acc[2] / 100, acc[3] / 100);
            acc[0] = acc[1] = acc[2] = acc[3] = 0;
        }
    }
    *out = fr;
    return ORBB_OK;
}

// Frame::Frame, monocular constructor (Frame.cpp:171-224; the RGB-D one :117-169 differs by its depth lookup only) as ONE call: ExtractORB(0, imGray),
// UndistortKeyPoints (:462-496) and AssignFeaturesToGrid (:241-257) on the keypoints where the extraction left them.  K4 / dist may be NULL
// (or dist[0] == 0, :465): mvKeysUn = mvKeys.
extern "C" int orbm_mono_frame_create(orbx_handle* h, const uint8_t* img_host, int w, int ht, size_t stride, const float* K4, const float* dist, int ndist,
                                      float min_x, float min_y, float max_x, float max_y, orbm_frame** out, orbx_keypoint* kps_host,
                                      orbx_keypoint* kps_un_host, uint8_t* desc_host, int* n_out, int capacity)
{
    if (!h || !img_host || !out || !n_out || capacity < 1 || !kps_host || !kps_un_host || !desc_host) return ORBB_E_INVALID;
    *out = nullptr; *n_out = 0;
    UndistortArgs ua;
    const bool undist = K4 && dist && dist[0] != 0.0f;
    if (undist && undistort_args(K4, dist, ndist, ua) != ORBB_OK) return ORBB_E_INVALID;
    const uint8_t* imgs[1] = {img_host};
    int32_t cnt[1] = {0};
    orbx_resident_view v{};
    int rc = orbx__extract_resident(h, imgs, 1, w, ht, stride, cnt, &v);
    if (rc != ORBB_OK) { if (v.s_out) cudaStreamSynchronize((cudaStream_t)v.s_out); return rc; }
    cudaStream_t st = (cudaStream_t)v.stream, s_out = (cudaStream_t)v.s_out;
    const int n = cnt[0];
    *n_out = n;
    if (n > capacity) { cudaStreamSynchronize(s_out); return ORBB_E_CAPACITY; }
    const size_t bk = (size_t)n * sizeof(orbx_keypoint);
    uint8_t* pin = t_pin.get(bk + 16);
    if (!pin) { cudaStreamSynchronize(s_out); return ORBB_E_NOMEM; }
    DevBuf un;
    const orbx_keypoint* kun_dev = v.kps_dev;
    if (undist && n > 0) {
        if ((rc = un.alloc(bk))) { cudaStreamSynchronize(s_out); return rc; }
        undistort_kernel<<<ceil_div(n, 256), 256, 0, st>>>(v.kps_dev, un.as<orbx_keypoint>(), n, ua);
        ORBB_LAUNCHED();
        kun_dev = un.as<orbx_keypoint>();
        if (cudaMemcpyAsync(pin, un.p, bk, cudaMemcpyDeviceToHost, st) != cudaSuccess) { cudaStreamSynchronize(st); cudaStreamSynchronize(s_out); return ORBB_E_CUDA; }
    }
    orbm_frame* fr = nullptr;
    rc = frame_create_impl(&fr, kun_dev, v.desc_dev, n, min_x, min_y, max_x, max_y, v.device, cudaMemcpyDeviceToDevice, st);
    if (rc != ORBB_OK) { cudaStreamSynchronize(st); cudaStreamSynchronize(s_out); return rc; }
    cudaError_t e = cudaStreamSynchronize(s_out);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { set_cuda_error(e, "orbm_mono_frame_create"); orbm_frame_destroy(fr); return ORBB_E_CUDA; }
    if (n > 0) {
        memcpy(kps_host, v.kps_pinned, bk); memcpy(desc_host, v.desc_pinned, (size_t)n * 32);
        memcpy(kps_un_host, undist ? (const void*)pin : (const void*)v.kps_pinned, bk);
    }
    *out = fr;
    return ORBB_OK;
}

// ---- brute-force database ----------------------------------------------------------------------------
struct orbm_db {
    int device = 0;
    int64_t capacity = 0, rows = 0, row_offset = 0;
    uint4* d_rows = nullptr;
    // scratch for queries / partial results, grown on demand
    int nq_cap = 0, chunk_cap = 0;
    uint4* d_q = nullptr; unsigned *d_k1 = nullptr, *d_k2 = nullptr;
    int *d_d1 = nullptr, *d_d2 = nullptr; long long* d_i1 = nullptr;
    int* d_bound = nullptr;                    // [nq_cap] cross-chunk second-best bound of the CTA-pair kernel
    int variant = 0;
};

extern "C" int orbm_db_create(orbm_db** out, int64_t capacity_rows, int64_t row_offset, int device)
{
    if (!out || capacity_rows < 0 || row_offset < 0) return ORBB_E_INVALID;
    *out = nullptr;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    orbm_db* db = new (std::nothrow) orbm_db();
    if (!db) return ORBB_E_NOMEM;
    db->device = device; db->capacity = capacity_rows; db->row_offset = row_offset;
    const char* v = getenv("ORBB_BF_VARIANT");
    db->variant = v ? atoi(v) : -1;           // -1 (default): 4 when a call has more than 256 queries, else 3; 4: tcgen05 on a CTA pair (cta_group::2, 512 queries per pair), 3: tcgen05 / TMEM (warp-specialised, N = 256 MMAs), 2: legacy mma.sync (IMMA), 1: integer pipe with carry-save adders, 0: plain XOR + POPC
    if (db->variant < 0 || db->variant > 4) db->variant = -1;
    if (cudaMalloc((void**)&db->d_rows, (size_t)(capacity_rows > 0 ? capacity_rows : 1) * 32) != cudaSuccess) {
        set_cuda_error(cudaGetLastError(), "orbm_db_create: cudaMalloc");
        delete db;
        return ORBB_E_NOMEM;
    }
    *out = db;
    return ORBB_OK;
}

extern "C" int orbm_db_destroy(orbm_db* db)
{
    if (!db) return ORBB_E_INVALID;
    cudaSetDevice(db->device);
    cudaFree(db->d_rows); cudaFree(db->d_q); cudaFree(db->d_k1); cudaFree(db->d_k2); cudaFree(db->d_d1); cudaFree(db->d_d2); cudaFree(db->d_i1); cudaFree(db->d_bound);
    delete db;
    return ORBB_OK;
}

extern "C" int64_t orbm_db_rows(const orbm_db* db) { return db ? db->rows : -1; }

static int db_append(orbm_db* db, const uint8_t* rows, int64_t n, cudaMemcpyKind kind)
{
    if (!db || n < 0 || (n > 0 && !rows)) return ORBB_E_INVALID;
    if (db->rows + n > db->capacity) return ORBB_E_CAPACITY;
    ORBB_CUDA(cudaSetDevice(db->device));
    if (n > 0) ORBB_CUDA(cudaMemcpy((uint8_t*)db->d_rows + (size_t)db->rows * 32, rows, (size_t)n * 32, kind));
    db->rows += n;
    return ORBB_OK;
}
extern "C" int orbm_db_append(orbm_db* db, const uint8_t* rows_host, int64_t n) { return db_append(db, rows_host, n, cudaMemcpyHostToDevice); }
extern "C" int orbm_db_append_device(orbm_db* db, const uint8_t* rows_dev, int64_t n) { return db_append(db, rows_dev, n, cudaMemcpyDeviceToDevice); }

// Kernel of one search: ORBB_BF_VARIANT when given, else the CTA-pair kernel when its 512-query tile is more than half used (measured
// ahead of the single-CTA kernel from 300 queries x 100 k rows upwards) and the single-CTA tcgen05 kernel otherwise.
static int bf_variant(const orbm_db* db, int nq)
{
    if (db->variant >= 0) return db->variant;
    return (nq > 256 && db->rows >= 65536) ? 4 : 3;
}

static int db_plan(const orbm_db* db, int nq, long long* rows_per_chunk, int* nchunks)
{
    const int variant = bf_variant(db, nq);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, db->device);
    const int qtiles = ceil_div(nq, variant >= 4 ? B2_QTILE : variant >= 2 ? BM_QTILE : BF_QTILE);
    static const int waves4 = getenv("ORBB_BF_WAVES") ? atoi(getenv("ORBB_BF_WAVES")) : 1;
    if (variant >= 4 && waves4 == 1) {          // one wave of at most sms / 2 CTA pairs (whole query tiles): 2.72 ms against 2.80 ms in two waves (ORBB_BF_WAVES=2)
        int t1 = (sms / 2) / qtiles; if (t1 < 1) t1 = 1;
        long long rpc1 = (db->rows + t1 - 1) / t1; rpc1 = (rpc1 + 255) / 256 * 256;
        if (rpc1 < BF_ROWS) rpc1 = BF_ROWS;
        if (rpc1 <= (1ll << BF_IDX_BITS)) { *rows_per_chunk = rpc1; *nchunks = (int)((db->rows + rpc1 - 1) / rpc1); if (*nchunks < 1) *nchunks = 1; return ORBB_OK; }
    }
    int target = ceil_div(variant >= 4 ? sms : sms * (variant == 3 ? 2 : 4), qtiles);      // two waves (variant 4: of sms / 2 CTA pairs); integer-pipe kernels: ~4 CTAs per SM in flight
    long long rpc = (db->rows + target - 1) / target;
    rpc = (rpc + 255) / 256 * 256;                                  // a multiple of every kernel's stage size
    if (rpc < BF_ROWS) rpc = BF_ROWS;
    if (rpc > (1ll << BF_IDX_BITS)) rpc = 1ll << BF_IDX_BITS;
    *rows_per_chunk = rpc;
    *nchunks = (int)((db->rows + rpc - 1) / rpc);
    if (*nchunks < 1) *nchunks = 1;
    return ORBB_OK;
}

static int db_reserve(orbm_db* db, int nq, int nchunks)
{
    if (nq > db->nq_cap) {
        cudaFree(db->d_q); cudaFree(db->d_d1); cudaFree(db->d_d2); cudaFree(db->d_i1); cudaFree(db->d_bound);
        db->d_q = nullptr; db->d_d1 = db->d_d2 = nullptr; db->d_i1 = nullptr; db->d_bound = nullptr; db->nq_cap = 0;
        ORBB_CUDA(cudaMalloc((void**)&db->d_q, (size_t)nq * 32));
        ORBB_CUDA(cudaMalloc((void**)&db->d_d1, (size_t)nq * 4));
        ORBB_CUDA(cudaMalloc((void**)&db->d_d2, (size_t)nq * 4));
        ORBB_CUDA(cudaMalloc((void**)&db->d_i1, (size_t)nq * 8));
        ORBB_CUDA(cudaMalloc((void**)&db->d_bound, (size_t)nq * 4));
        db->nq_cap = nq;
        db->chunk_cap = 0;
    }
    if ((long long)nq * nchunks > (long long)db->chunk_cap) {
        cudaFree(db->d_k1); cudaFree(db->d_k2); db->d_k1 = db->d_k2 = nullptr; db->chunk_cap = 0;
        ORBB_CUDA(cudaMalloc((void**)&db->d_k1, (size_t)db->nq_cap * nchunks * 4));
        ORBB_CUDA(cudaMalloc((void**)&db->d_k2, (size_t)db->nq_cap * nchunks * 4));
        db->chunk_cap = db->nq_cap * nchunks;
    }
    return ORBB_OK;
}

// A search writes the database's per-chunk partial-key scratch (and may grow it): not const, and calls on one orbm_db are
// serialised by the caller (include/orbb200.h).
extern "C" int orbm_bf_top2_device(orbm_db* db, const uint8_t* q_dev, int nq, int32_t* d1, int64_t* i1, int32_t* d2, void* stream)
{
    if (!db || nq < 0 || (nq > 0 && (!q_dev || !d1 || !i1 || !d2))) return ORBB_E_INVALID;
    if (nq == 0) return ORBB_OK;
    ORBB_CUDA(cudaSetDevice(db->device));
    long long rpc; int nchunks;
    db_plan(db, nq, &rpc, &nchunks);
    int rc = db_reserve(db, nq, nchunks);
    if (rc != ORBB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    BfArgs a;
    a.db = db->d_rows; a.nrows = db->rows; a.q = (const uint4*)q_dev; a.nq = nq;
    a.rows_per_chunk = rpc; a.nchunks = nchunks; a.part_k1 = db->d_k1; a.part_k2 = db->d_k2; a.bound_d2 = nullptr;
    const int variant = bf_variant(db, nq);
    static const bool share = getenv("ORBB_BF_NO_BOUND") == nullptr;                 // experiment knob: chunks do not share their bounds
    if (share && variant >= 3) { a.bound_d2 = db->d_bound; ORBB_CUDA(cudaMemsetAsync(db->d_bound, 0x7f, (size_t)nq * 4, st)); }      // 0x7f7f7f7f: no bound yet
    dim3 grid(ceil_div(nq, variant >= 2 ? BM_QTILE : BF_QTILE), nchunks);
    if (variant == 4) {
        static std::atomic<bool> attr2_set[64];
        const int dv = db->device & 63;
        if (!attr2_set[dv].load(std::memory_order_acquire)) {
            ORBB_CUDA(cudaFuncSetAttribute(bf_top2_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B2_SMEM));
            attr2_set[dv].store(true, std::memory_order_release);
        }
        bf_top2_tc2_kernel<<<dim3(2 * ceil_div(nq, B2_QTILE), nchunks), B2_THREADS, B2_SMEM, st>>>(a);      // clusters of 2 along x
    }
    else if (variant == 3) {
        // more than 48 KB of dynamic shared memory is an opt-in per device
        static std::atomic<bool> attr_set[64];                   // setting it twice is harmless; the flag only saves the call
        const int dv = db->device & 63;
        if (!attr_set[dv].load(std::memory_order_acquire)) {
            ORBB_CUDA(cudaFuncSetAttribute(bf_top2_tcn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_SMEM));
            attr_set[dv].store(true, std::memory_order_release);
        }
        bf_top2_tcn_kernel<<<grid, BW_THREADS, BN_SMEM, st>>>(a);
    }
    else if (variant == 0) bf_top2_kernel<0><<<grid, BF_THREADS, 0, st>>>(a);
    else if (variant == 2) bf_top2_mma_kernel<<<grid, BM_THREADS, 0, st>>>(a);
    else bf_top2_kernel<1><<<grid, BF_THREADS, 0, st>>>(a);
    ORBB_LAUNCHED();
    bf_merge_chunks_kernel<<<ceil_div(nq, 256), 256, 0, st>>>(db->d_k1, db->d_k2, nchunks, nq, rpc, db->row_offset, d1, (long long*)i1, d2);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_bf_top2(orbm_db* db, const uint8_t* q_host, int nq, int32_t* d1, int64_t* i1, int32_t* d2)
{
    if (!db || nq < 0 || (nq > 0 && (!q_host || !d1 || !i1 || !d2))) return ORBB_E_INVALID;
    if (nq == 0) return ORBB_OK;
    ORBB_CUDA(cudaSetDevice(db->device));
    long long rpc; int nchunks;
    db_plan(db, nq, &rpc, &nchunks);
    int rc = db_reserve(db, nq, nchunks);
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpy(db->d_q, q_host, (size_t)nq * 32, cudaMemcpyHostToDevice));
    rc = orbm_bf_top2_device(db, (const uint8_t*)db->d_q, nq, db->d_d1, (int64_t*)db->d_i1, db->d_d2, nullptr);
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpy(d1, db->d_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost));
    ORBB_CUDA(cudaMemcpy(i1, db->d_i1, (size_t)nq * 8, cudaMemcpyDeviceToHost));
    ORBB_CUDA(cudaMemcpy(d2, db->d_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

// ---- brute-force top-2 without any allocation ---------------------------------------------------------
// Same scan + chunk merge as orbm_bf_top2_device, for callers that cannot call cudaMalloc (CUDA graph capture, serving loops): the
// per-(chunk, query) partial keys live in ONE file-scope device array that every device holding this library keeps until exit.
// Every call on a device shares it: calls that overlap on the GPU race (see include/orbb200.h).
constexpr int BF_NOALLOC_MAX_NQ = 4096;                   // queries per call
constexpr int BF_NOALLOC_MAX_KEYS = 1 << 19;              // nq * nchunks partial keys per array (2 MiB each, 4 MiB in total)
static __device__ __align__(256) unsigned g_bf_noalloc_scratch[2 * BF_NOALLOC_MAX_KEYS];

// The array's name is only a host-side stand-in: the device address comes from cudaGetSymbolAddress, once per device, with that
// device current.
static int bf_noalloc_scratch(int device, unsigned** out)
{
    static std::mutex mu;
    static unsigned* ptr[64] = {};
    if (device < 0 || device >= 64) return ORBB_E_INVALID;
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        ORBB_CUDA(cudaSetDevice(device));
        unsigned* p = nullptr;
        ORBB_CUDA(cudaGetSymbolAddress((void**)&p, g_bf_noalloc_scratch));
        // more than 48 KB of dynamic shared memory is an opt-in per device (not an allocation)
        ORBB_CUDA(cudaFuncSetAttribute(bf_top2_tcn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_SMEM));
        ptr[device] = p;
    }
    *out = ptr[device];
    return ORBB_OK;
}

extern "C" int orbm_bf_top2_noalloc(const uint8_t* rows_dev, int64_t nrows, int64_t row_offset, const uint8_t* q_dev, int nq,
                                    int32_t* d1, int64_t* i1, int32_t* d2, int device, void* stream)
{
    if (nrows < 0 || row_offset < 0 || nq < 0 || (nrows > 0 && !rows_dev) || (nq > 0 && (!q_dev || !d1 || !i1 || !d2))) return ORBB_E_INVALID;
    if (((uintptr_t)rows_dev | (uintptr_t)q_dev) & 15) return ORBB_E_INVALID;            // 16-byte row loads
    if (nq > BF_NOALLOC_MAX_NQ) {
        set_cuda_error(cudaErrorInvalidValue, "orbm_bf_top2_noalloc: nq exceeds ORBM_NOALLOC_MAX_QUERIES (4096)");
        return ORBB_E_CAPACITY;
    }
    if (nq == 0) return ORBB_OK;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    unsigned* scratch = nullptr;
    if ((rc = bf_noalloc_scratch(device, &scratch)) != ORBB_OK) return rc;
    // chunking as db_plan does for the tcgen05 kernel (two CTAs per SM over all query tiles), then capped by the scratch size
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int qtiles = ceil_div(nq, BM_QTILE);
    int target = ceil_div(sms * 2, qtiles);
    if ((long long)target * nq > BF_NOALLOC_MAX_KEYS) target = BF_NOALLOC_MAX_KEYS / nq;
    long long rpc = (nrows + target - 1) / target;
    rpc = (rpc + 255) / 256 * 256;
    if (rpc < BF_ROWS) rpc = BF_ROWS;
    if (rpc > (1ll << BF_IDX_BITS)) rpc = 1ll << BF_IDX_BITS;
    const long long nchunks_ll = nrows > 0 ? (nrows + rpc - 1) / rpc : 1;
    if (nchunks_ll * nq > BF_NOALLOC_MAX_KEYS) {
        set_cuda_error(cudaErrorInvalidValue, "orbm_bf_top2_noalloc: nrows x nq exceeds the static scratch (2^19 partial keys of <= 2^23 rows each)");
        return ORBB_E_CAPACITY;
    }
    const int nchunks = (int)nchunks_ll;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* k1 = scratch; unsigned* k2 = scratch + BF_NOALLOC_MAX_KEYS;
    // the array keeps the previous call's keys: start both halves of this call's region at "no row" (0xffffffff)
    ORBB_CUDA(cudaMemsetAsync(k1, 0xff, (size_t)nchunks * nq * 4, st));
    ORBB_CUDA(cudaMemsetAsync(k2, 0xff, (size_t)nchunks * nq * 4, st));
    BfArgs a;
    a.db = (const uint4*)rows_dev; a.nrows = nrows; a.q = (const uint4*)q_dev; a.nq = nq;
    a.rows_per_chunk = rpc; a.nchunks = nchunks; a.part_k1 = k1; a.part_k2 = k2; a.bound_d2 = nullptr;
    bf_top2_tcn_kernel<<<dim3(qtiles, nchunks), BW_THREADS, BN_SMEM, st>>>(a);
    ORBB_LAUNCHED();
    bf_merge_chunks_kernel<<<ceil_div(nq, 256), 256, 0, st>>>(k1, k2, nchunks, nq, rpc, row_offset, d1, (long long*)i1, d2);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_merge_top2_device(const int32_t* d1_all, const int64_t* i1_all, const int32_t* d2_all, int G, int nq,
                                      int32_t* d1, int64_t* i1, int32_t* d2, void* stream)
{
    if (G < 1 || nq < 0 || (nq > 0 && (!d1_all || !i1_all || !d2_all || !d1 || !i1 || !d2))) return ORBB_E_INVALID;
    if (nq == 0) return ORBB_OK;
    bf_merge_shards_kernel<<<ceil_div(nq, 256), 256, 0, (cudaStream_t)stream>>>(d1_all, (const long long*)i1_all, d2_all, G, nq, d1, (long long*)i1, d2);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbm_ratio_test(const int32_t* d1, const int32_t* d2, int nq, int th_low, float ratio, uint8_t* accepted)
{
    if (nq < 0 || (nq > 0 && (!d1 || !d2 || !accepted))) return ORBB_E_INVALID;
    // ORBmatcher.cpp:278-281: bestDist1 <= TH_LOW && (float)bestDist1 < mfNNratio*(float)bestDist2
    for (int i = 0; i < nq; i++) accepted[i] = (d1[i] <= th_low && static_cast<float>(d1[i]) < ratio * static_cast<float>(d2[i])) ? 1 : 0;
    return ORBB_OK;
}

// Measured POPC throughput (G popc/s) of a register-only loop: the roofline denominator for matching.
extern "C" int orbb200_measure_popc_peak(int device, double* gpopc_per_s)
{
    if (!gpopc_per_s) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    DevBuf o;
    const int blocks = sms * 8, threads = 256, iters = 4096;
    if ((rc = o.alloc((size_t)blocks * threads * 4))) return rc;
    cudaEvent_t e0, e1;
    ORBB_CUDA(cudaEventCreate(&e0)); ORBB_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        ORBB_CUDA(cudaEventRecord(e0));
        popc_peak_kernel<<<blocks, threads>>>(o.as<unsigned>(), iters, 12345u + rep);
        ORBB_LAUNCHED();
        ORBB_CUDA(cudaEventRecord(e1));
        ORBB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ORBB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double g = (double)blocks * threads * iters * 8 / (ms * 1e-3) / 1e9;
        if (rep > 0 && g > best) best = g;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *gpopc_per_s = best;
    return ORBB_OK;
}

// Measured throughput (1e12 int8 multiply-adds x 2 per second) of a register-only mma.sync.m16n8k32 loop: the tensor-pipe
// denominator of the brute-force top-2 kernel (one 256-bit comparison = 512 int8 operations).
extern "C" int orbb200_measure_imma_peak(int device, double* int8_tops)
{
    if (!int8_tops) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    DevBuf o;
    const int blocks = sms * 4, threads = 256, iters = 4096;
    if ((rc = o.alloc((size_t)blocks * threads * 4))) return rc;
    cudaEvent_t e0, e1;
    ORBB_CUDA(cudaEventCreate(&e0)); ORBB_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        ORBB_CUDA(cudaEventRecord(e0));
        imma_peak_kernel<<<blocks, threads>>>(o.as<int>(), iters, 12345u + rep);
        ORBB_LAUNCHED();
        ORBB_CUDA(cudaEventRecord(e1));
        ORBB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ORBB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tops = (double)blocks * (threads / 32) * iters * 8 * (2.0 * 16 * 8 * 32) / (ms * 1e-3) / 1e12;
        if (rep > 0 && tops > best) best = tops;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *int8_tops = best;
    return ORBB_OK;
}

// Measured tcgen05.mma kind::i8 rate (int8 TOP/s, 2 x multiply-adds) of the 128 x 256 x 32 shape the scan kernel issues, operands in
// shared memory, accumulators in TMEM, nothing else running: the denominator of the brute-force top-2 roofline.
extern "C" int orbb200_measure_tcgen05_i8_peak(int device, double* int8_tops)
{
    if (!int8_tops) return ORBB_E_INVALID;
    int rc = check_device(device);
    if (rc != ORBB_OK) return rc;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const bool pair = getenv("ORBB_PEAK_PAIR") != nullptr;             // experiment: the same probe with cta_group::2 on CTA pairs
    ORBB_CUDA(cudaFuncSetAttribute(tc_i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TCP_SMEM));
    ORBB_CUDA(cudaFuncSetAttribute(tc2_i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TCP_SMEM));
    cudaEvent_t e0, e1;
    ORBB_CUDA(cudaEventCreate(&e0)); ORBB_CUDA(cudaEventCreate(&e1));
    const int groups = 4096;                                  // x 16 MMAs of 2 * 128 * 256 * 32 operations per CTA
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        ORBB_CUDA(cudaEventRecord(e0));
        const int ncols = getenv("ORBB_PEAK_N") ? atoi(getenv("ORBB_PEAK_N")) : 256;
        if (pair) tc2_i8_peak_kernel<<<sms & ~1, 128, TCP_SMEM>>>(groups, ncols);      // each MMA is twice the work: same operations per CTA
        else tc_i8_peak_kernel<<<sms, 128, TCP_SMEM>>>(groups, nullptr);
        ORBB_LAUNCHED();
        ORBB_CUDA(cudaEventRecord(e1));
        ORBB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        ORBB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tops = (double)sms * groups * 16 * (2.0 * 128 * (pair && getenv("ORBB_PEAK_N") ? atoi(getenv("ORBB_PEAK_N")) : 256) * 32) / (ms * 1e-3) / 1e12;
        if (rep > 0 && tops > best) best = tops;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    ORBB_CUDA(cudaGetLastError());
    *int8_tops = best;
    return ORBB_OK;
}

#ifdef RESOLVE_TIMING
extern "C" int orbb200__resolve_timing(unsigned long long* out) { return cudaMemcpyFromSymbol(out, g_resolve_t, sizeof(unsigned long long) * 8) == cudaSuccess ? 0 : -2; }
#endif
#ifdef B2_TIMING
extern "C" int orbb200__b2_gtiming(long long* out) { return cudaMemcpyFromSymbol(out, g_b2_g, sizeof(long long) * 2 * 32 * 2 * 8) == cudaSuccess ? 0 : -2; }
extern "C" int orbb200__b2_timing(long long* out) { return cudaMemcpyFromSymbol(out, g_b2_t, sizeof(long long) * 32 * 2 * 8) == cudaSuccess ? 0 : -2; }
#endif
