// extract.cu — host side of the extractor half of liborbb200: handle, geometry/tables, launches, C ABI.
// Replaces ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111, src/ORBextractor.cpp:412-472, :1053-1148).
#include "extract_kernels.cuh"
#include <math.h>
#include <stdlib.h>

namespace orbb {

std::atomic<uint64_t> g_launches{0};
static std::mutex g_err_mu;
static char g_err[512] = "";
namespace {
struct ScratchCache {
    static constexpr int NDEV = 64, NCLASS = 27;          // classes 2^8 .. 2^26 bytes
    std::mutex mu;
    std::vector<void*> free_[NDEV][NCLASS];
    size_t cached[NDEV] = {};
};
ScratchCache& scratch_cache() { static ScratchCache* c = new ScratchCache(); return *c; }      // never destroyed: outlives every handle
int scratch_class(size_t bytes) { int c = 8; while (((size_t)1 << c) < bytes) c++; return c; }
}  // namespace

void* scratch_alloc(size_t bytes, size_t* granted)
{
    if (bytes < 256) bytes = 256;
    int dev = 0;
    void* p = nullptr;
    if (cudaGetDevice(&dev) == cudaSuccess && dev < ScratchCache::NDEV && bytes <= ((size_t)1 << 26)) {
        const int c = scratch_class(bytes);
        ScratchCache& sc = scratch_cache();
        {
            std::lock_guard<std::mutex> lk(sc.mu);
            std::vector<void*>& v = sc.free_[dev][c];
            if (!v.empty()) { p = v.back(); v.pop_back(); sc.cached[dev] -= (size_t)1 << c; }
        }
        *granted = (size_t)1 << c;
        if (p) return p;
        bytes = *granted;
    } else {
        *granted = bytes;
    }
    const cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) { set_cuda_error(e, "cudaMalloc (scratch)"); cudaGetLastError(); return nullptr; }
    return p;
}

void scratch_free(void* p, size_t granted)
{
    if (!p) return;
    int dev = 0;
    const bool pow2 = granted >= 256 && (granted & (granted - 1)) == 0 && granted <= ((size_t)1 << 26);
    if (pow2 && cudaGetDevice(&dev) == cudaSuccess && dev < ScratchCache::NDEV) {
        ScratchCache& sc = scratch_cache();
        std::lock_guard<std::mutex> lk(sc.mu);
        if (sc.cached[dev] + granted <= ((size_t)1 << 30)) {
            sc.free_[dev][scratch_class(granted)].push_back(p);
            sc.cached[dev] += granted;
            return;
        }
    }
    cudaFree(p);
}

void scratch_trim(void)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev >= ScratchCache::NDEV) return;
    ScratchCache& sc = scratch_cache();
    std::lock_guard<std::mutex> lk(sc.mu);
    for (auto& v : sc.free_[dev]) { for (void* p : v) cudaFree(p); v.clear(); }
    sc.cached[dev] = 0;
}

void set_cuda_error(cudaError_t e, const char* what)
{
    std::lock_guard<std::mutex> lk(g_err_mu);
    snprintf(g_err, sizeof g_err, "%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
}

static const int8_t h_pattern[1024] = {
#include "orb_pattern.inc"
};

static inline int cv_round_f(float v) { return (int)nearbyintf(v); }
static inline int cv_round_d(double v) { return (int)nearbyint(v); }

}  // namespace orbb

using namespace orbb;

struct orbx_handle {
    orbx_config cfg;
    int nlevels;
    cudaStream_t stream = nullptr;
    cudaStream_t s_in = nullptr, s_out = nullptr;      // copy streams of the host-buffer batch pipeline
    cudaStream_t s_k[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // extra kernel streams: chunks rotate over stream + these (latency-bound tails overlap the next chunk)
    int n_kstreams = 4;
    // per kernel stream: a side stream on which the octree of level L runs while the main stream works on levels L+1..
    cudaStream_t s_side[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t s_side2[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // odd levels' octrees
    cudaEvent_t ev_side2[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_side[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::vector<cudaEvent_t> ev_lvl[8];
    bool oct_overlap = true;
    std::vector<cudaEvent_t> ev_in, ev_done;          // per chunk: frames landed in HBM / results ready
    // ORBextractor tables (ORBextractor.cpp:417-448, :456-471)
    float scale[ORBB_MAX_LEVELS], inv_scale[ORBB_MAX_LEVELS], sigma2[ORBB_MAX_LEVELS], inv_sigma2[ORBB_MAX_LEVELS];
    int per_level[ORBB_MAX_LEVELS], umax[16];
    int q[7];
    int l3_ow[ORBB_MAX_LEVELS];   // owned words per tile row of each level: 30 (120-pixel tiles) or 32 (128-pixel tiles, where that saves a tile column)
    int l3_warps = 4;          // warps per level-kernel CTA (tile = 120 x 8*warps pixels); ORBB_L3_WARPS = 2 | 4 | 8
    bool l3_warps_forced = false;
    int sm_count = 148;
    float factorPI;
    // geometry for the current image shape
    int cur_w = 0, cur_h = 0;
    LevelGeom geom[ORBB_MAX_LEVELS];
    long long node_off[ORBB_MAX_LEVELS], cell_off[ORBB_MAX_LEVELS];
    size_t pyr_fstride = 0, rec_total = 0, node_total = 0, cell_total = 0, out_total = 0, tab_total = 0, tile_total = 0;
    // capacities (sized for max_width x max_height at creation)
    size_t cap_pyr = 0, cap_rec = 0, cap_node = 0, cap_cell = 0, cap_out = 0, cap_tab = 0, cap_tile = 0;
    int dev_capacity = 0;         // per-frame capacity of the internal final output buffers
    // device buffers
    uint8_t *d_pyr = nullptr, *d_blur = nullptr, *d_cellflag = nullptr, *d_desc = nullptr;
    uint32_t *d_rec = nullptr, *d_pts = nullptr;
    uint4* d_nodes = nullptr;
    int *d_iscr = nullptr, *d_rec_count = nullptr, *d_out_count = nullptr, *d_status = nullptr, *d_counts = nullptr, *d_tiles = nullptr;
    int4* d_out = nullptr;
    int2* d_tab = nullptr;
    LevelGeom* d_geom = nullptr;
    long long *d_node_off = nullptr, *d_cell_off = nullptr;
    orbx_keypoint* d_kps = nullptr;
    // pinned staging
    uint8_t *h_img = nullptr, *h_desc = nullptr;
    uint8_t* d_color = nullptr;          // colour / raw frames of orbx_extract_batch_color / _rectified (allocated on first use)
    int2* d_map = nullptr;               // rectification map in cv::remap's fixed point {cvRound(mapx*32), cvRound(mapy*32)}
    int map_w = 0, map_h = 0, map_sw = 0, map_sh = 0;
    orbx_keypoint* h_kps = nullptr;
    int *h_counts = nullptr, *h_status = nullptr;
    // optional per-kernel timing (bench.py roofline): events around each kernel class
    bool profiling = false;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double prof_ms[3] = {0, 0, 0};     // level kernels, octree, orientation+descriptor
    int prof_calls = 0;
    // last extraction
    int last_frames = 0;
    const uint8_t* last_src = nullptr; long long last_src_pitch = 0, last_src_fstride = 0;
    const orbx_keypoint* last_kps_dev = nullptr; int last_capacity = 0;
    // orbx_graph_capture: the launch sequence of one device-resident extraction as an executable CUDA graph
    cudaGraphExec_t graph_exec = nullptr;
    int graph_frames = 0, graph_kernels = 0;
    cudaEvent_t ev_fork = nullptr, ev_join[7] = {};           // chunked device-resident batches
    cudaEvent_t ev_fan[ORBB_MAX_LEVELS] = {};                  // latency mode: one octree stream per level, joined before the descriptors
};

static void level_sizes(const orbx_handle* h, int w, int ht, int* lw, int* lh)
{
    for (int l = 0; l < h->nlevels; l++) {
        lw[l] = cv_round_f((float)w * h->inv_scale[l]);     // ORBextractor.cpp:1128
        lh[l] = cv_round_f((float)ht * h->inv_scale[l]);
    }
}

// cv::resize INTER_LINEAR coefficient tables for src -> dst along one axis: tab[d] = {source index, a0 | a1 << 16}.
// `padded` >= dst entries are written (the tail replicates the last real entry: destination words are stored whole).
// own_first[t] = first destination UNIT (unit = `unit` destination entries; x: a 4-pixel word, y: a row) whose first
// entry has source index >= t * tile - bias; own_first[ntiles] = number of units.
static void build_axis_table(int src, int dst, int padded, int2* tab, int tile, int bias, int unit, int* own_first, int ntiles)
{
    const double inv_scale = (double)dst / src;
    const double sc = 1.0 / inv_scale;
    for (int d = 0; d < dst; d++) {
        float f = (float)((d + 0.5) * sc - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= src - 1) { s = src - 1; f = 0.f; }
        const int a0 = cv_round_f((1.f - f) * 2048.f), a1 = cv_round_f(f * 2048.f);
        tab[d] = make_int2(s, (a0 & 0xffff) | (a1 << 16));
    }
    for (int d = dst; d < padded; d++) tab[d] = tab[dst - 1];
    const int nunits = ceil_div(dst, unit);
    int t = 0;
    for (int u = 0; u < nunits; u++) {
        const int s = tab[u * unit].x;
        while (t < ntiles && s >= t * tile - bias) own_first[t++] = u;
    }
    own_first[0] = 0;
    while (t <= ntiles) own_first[t++] = nunits;
    own_first[ntiles] = nunits;
}

// Computes the per-shape geometry; returns false if the shape cannot be handled by this handle's buffers.
static int setup_geometry(orbx_handle* h, int w, int ht, bool allocate_only)
{
    int lw[ORBB_MAX_LEVELS], lh[ORBB_MAX_LEVELS];
    level_sizes(h, w, ht, lw, lh);
    size_t off = 0, rec = 0, node = 0, cell = 0, out = 0, tab = 0, tile = 0;
    for (int l = 0; l < h->nlevels; l++) {
        LevelGeom& g = h->geom[l];
        memset(&g, 0, sizeof g);
        if (lw[l] < 1 || lh[l] < 1) return ORBB_E_BAD_SHAPE;
        g.w = lw[l]; g.h = lh[l];
        g.pitch = (int)align_up((size_t)g.w, 128);
        g.off = off; off += (size_t)g.pitch * g.h;
        // cell grid (ORBextractor.cpp:781-795)
        const int maxBX = g.w - 16, maxBY = g.h - 16;
        const float width = (float)(maxBX - 16), height = (float)(maxBY - 16);
        int nCols = width > 0 ? (int)(width / 30.f) : 0, nRows = height > 0 ? (int)(height / 30.f) : 0;
        if (nCols < 1 || nRows < 1) { nCols = nRows = 0; }     // level smaller than one cell: no keypoints (reference: UB)
        g.nCols = nCols; g.nRows = nRows;
        g.wCell = nCols ? (int)ceilf(width / nCols) : 1;
        g.hCell = nRows ? (int)ceilf(height / nRows) : 1;
        g.magicW = (unsigned)((0x100000000ull + g.wCell - 1) / g.wCell); g.magicH = (unsigned)((0x100000000ull + g.hCell - 1) / g.hCell);
        // octree roots (ORBextractor.cpp:545-547)
        int nIni = nCols ? (int)roundf((float)(maxBX - 16) / (float)(maxBY - 16)) : 1;
        if (nIni < 1) nIni = 1;
        if (nIni > 256) return ORBB_E_BAD_SHAPE;
        g.nIni = nIni;
        g.hX = (float)(maxBX - 16) / nIni;
        g.quota = h->per_level[l];
        g.rec_cap = (int)((size_t)g.w * g.h / 10 + 1024);
        g.rec_off = rec; rec += g.rec_cap;
        g.node_cap = 6 * g.quota + 8 * nIni + 64;
        g.out_cap = g.node_cap;
        g.out_off = out; out += g.out_cap;
        h->node_off[l] = (long long)node; node += g.node_cap;
        h->cell_off[l] = (long long)cell; cell += align_up((size_t)nCols * nRows + 16, 16);
        g.scale = h->scale[l];
        g.size = (float)(int)(31 * h->scale[l]);              // ORBextractor.cpp:847
        {   // 128-pixel tiles where they pay (640 -> 5 instead of 6 columns), else the cheaper 120-pixel tiles; ORBB_L3_OW = 30 | 32 forces one
            static const int ow_env = getenv("ORBB_L3_OW") ? atoi(getenv("ORBB_L3_OW")) : 0;
            // measured cost model (512-frame batches): a full 128-pixel tile costs 1.10 of a full 120-pixel one; a partly filled tile
            // 0.55 + 0.45 x its fill (the candidate-dependent phases scale with the pixels it really holds)
            auto row_cost = [](int w, int tw) { const int full = w / tw, rem = w % tw; return full + (rem ? 0.55 + 0.45 * rem / tw : 0.0); };
            int ow = 1.10 * row_cost(g.w, 128) < 0.98 * row_cost(g.w, 120) ? 32 : 30;
            if (ow_env == 30 || ow_env == 32) ow = ow_env;
            if (h->l3_warps != 4) ow = 30;
            h->l3_ow[l] = ow;
        }
        if (l + 1 < h->nlevels) {
            g.xt_off = (int)tab; tab += align_up((size_t)lw[l + 1], 4);
            g.yt_off = (int)tab; tab += align_up((size_t)lh[l + 1], 4);      // keeps every xt 16-byte aligned (int4 loads)
            g.dxs_off = (int)tile; tile += ceil_div(g.w, 4 * 30) + 1;       // (room for the narrower tile's count)
            g.dys_off = (int)tile; tile += ceil_div(g.h, L3_RG) + 1;
        }
    }
    if (allocate_only) {
        // other shapes (<= max) can have more octree roots per level (aspect ratio): reserve room for 16 roots/level
        const size_t root_margin = (size_t)h->nlevels * 8 * 16;
        node += root_margin; out += root_margin;
        h->cap_pyr = off; h->cap_rec = rec; h->cap_node = node; h->cap_cell = cell; h->cap_out = out;
        h->cap_tab = tab + 64; h->cap_tile = tile + 64;
        return ORBB_OK;
    }
    if (off > h->cap_pyr || rec > h->cap_rec || node > h->cap_node || cell > h->cap_cell || out > h->cap_out ||
        tab > h->cap_tab || tile > h->cap_tile)
        return ORBB_E_BAD_SHAPE;
    h->pyr_fstride = off; h->rec_total = rec; h->node_total = node; h->cell_total = cell; h->out_total = out;
    h->tab_total = tab; h->tile_total = tile;
    return ORBB_OK;
}

static int upload_geometry(orbx_handle* h)
{
    std::vector<int2> tab(h->tab_total + 1);
    std::vector<int> tiles(h->tile_total + 1);
    for (int l = 0; l + 1 < h->nlevels; l++) {
        const LevelGeom& g = h->geom[l];
        const LevelGeom& n = h->geom[l + 1];
        // x: a tile column owns the destination WORDS whose first source column is in [x0 - 2, x0 + 118)
        const int twpx = 4 * h->l3_ow[l];
        build_axis_table(g.w, n.w, (int)align_up((size_t)n.w, 4), tab.data() + g.xt_off, twpx, 2, 4, tiles.data() + g.dxs_off, ceil_div(g.w, twpx));
        // y: an 8-row group owns the destination rows whose (upper) source row lies in it
        build_axis_table(g.h, n.h, n.h, tab.data() + g.yt_off, L3_RG, 0, 1, tiles.data() + g.dys_off, ceil_div(g.h, L3_RG));
    }
    // plain synchronous copies: shape changes are rare (once per camera), and the stream must not race them
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    ORBB_CUDA(cudaMemcpy(h->d_tab, tab.data(), h->tab_total * sizeof(int2), cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(h->d_tiles, tiles.data(), h->tile_total * sizeof(int), cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(h->d_geom, h->geom, sizeof(LevelGeom) * h->nlevels, cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(h->d_node_off, h->node_off, sizeof(long long) * h->nlevels, cudaMemcpyHostToDevice));
    ORBB_CUDA(cudaMemcpy(h->d_cell_off, h->cell_off, sizeof(long long) * h->nlevels, cudaMemcpyHostToDevice));
    return ORBB_OK;
}

// Per-shape geometry of the handle, rebuilt when the frame shape changes.  setup_geometry writes h->geom[] in place, so a rejected
// shape (ORBB_E_BAD_SHAPE) leaves it partly overwritten: the current shape is invalidated first and only set again on success, so
// the next call with the previous, valid shape rebuilds and re-uploads instead of running on the rejected shape's tables.
static int ensure_geometry(orbx_handle* h, int w, int ht)
{
    if (w == h->cur_w && ht == h->cur_h) return ORBB_OK;
    h->cur_w = h->cur_h = 0;
    int rc = setup_geometry(h, w, ht, false);
    if (rc != ORBB_OK) return rc;
    rc = upload_geometry(h);
    if (rc != ORBB_OK) return rc;
    h->cur_w = w; h->cur_h = ht;
    return ORBB_OK;
}

static std::mutex g_const_mu;
static bool g_const_done[64] = {false};

static int init_constants(orbx_handle* h)
{
    std::lock_guard<std::mutex> lk(g_const_mu);
    const int dev = h->cfg.device;
    if (dev < 64 && g_const_done[dev]) return ORBB_OK;
    // cv::fastAtan2 polynomial, degrees
    const float sc = (float)(180.0 / 3.14159265358979323846);
    const float at[4] = {0.9997878412794807f * sc, -0.3258083974640975f * sc, 0.1555786518463281f * sc, -0.04432655554792128f * sc};
    ORBB_CUDA(cudaMemcpyToSymbol(c_pattern, h_pattern, sizeof h_pattern));
    ORBB_CUDA(cudaMemcpyToSymbol(c_umax, h->umax, sizeof(int) * 16));
    ORBB_CUDA(cudaMemcpyToSymbol(c_atan, at, sizeof at));
    if (dev < 64) g_const_done[dev] = true;
    return ORBB_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda).
typedef CUresult (*tmap_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tmap_encode_fn get_tmap_encode()
{
    static tmap_encode_fn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (tmap_encode_fn)p;
        cudaGetLastError();
    });
    return fn;
}

// Level image as a 3-D u8 tensor (x, y, frame) with a 144 B x box_rows x 1 box.  Returns false when the layout cannot be
// described (base or strides not 16-byte aligned): the kernel then stages its tile with ordinary loads.
static bool encode_level_tmap(CUtensorMap* tm, const uint8_t* base, int w, int ht, int nframes, size_t pitch, size_t fstride, int box_rows, int box_bytes = 144)
{
    static const bool off = getenv("ORBB_NO_TMA") != nullptr;
    tmap_encode_fn enc = get_tmap_encode();
    if (off || !enc || ((uintptr_t)base & 15) || (pitch & 15) || (nframes > 1 && (fstride & 15)) || pitch == 0) return false;
    if (fstride == 0 || (fstride & 15)) fstride = pitch * (size_t)ht;           // single frame: any valid stride
    if (fstride & 15) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)ht, (cuuint64_t)(nframes > 0 ? nframes : 1)};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)fstride};
    const cuuint32_t box[3] = {(cuuint32_t)box_bytes, (cuuint32_t)box_rows, 1u};   // tile + halo, start rounded down to 16 B: 144 B (120-pixel tiles) or 160 B (128)
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <typename T> static int dev_alloc(T** p, size_t n)
{
    ORBB_CUDA(cudaMalloc((void**)p, (n ? n : 1) * sizeof(T)));
    return ORBB_OK;
}
template <typename T> static int host_alloc(T** p, size_t n)
{
    ORBB_CUDA(cudaMallocHost((void**)p, (n ? n : 1) * sizeof(T)));
    return ORBB_OK;
}

extern "C" int orbx_create(const orbx_config* cfg, orbx_handle** out)
{
    if (!cfg || !out) return ORBB_E_INVALID;
    *out = nullptr;
    if (cfg->nlevels < 1 || cfg->nlevels > ORBB_MAX_LEVELS || cfg->nfeatures < 0 || !(cfg->scale_factor > 1.0f) ||
        cfg->min_th_fast < 0 || cfg->ini_th_fast < cfg->min_th_fast || cfg->ini_th_fast > 254 || cfg->max_batch < 1)
        return ORBB_E_INVALID;
    if (cfg->max_width < 1 || cfg->max_height < 1 || cfg->max_width > 4096 || cfg->max_height > 4096) return ORBB_E_BAD_SHAPE;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || cfg->device < 0 || cfg->device >= ndev) {
        set_cuda_error(cudaErrorNoDevice, "orbx_create: no CUDA device (this library has no CPU fallback)");
        return ORBB_E_CUDA;
    }
    ORBB_CUDA(cudaSetDevice(cfg->device));
    orbx_handle* h = new (std::nothrow) orbx_handle();
    if (!h) return ORBB_E_NOMEM;
    h->cfg = *cfg;
    const int nl = h->nlevels = cfg->nlevels;
    // --- ORBextractor constructor tables (ORBextractor.cpp:417-448); scaleFactor member is a double holding the float ---
    const double scaleFactor = (double)cfg->scale_factor;
    h->scale[0] = 1.0f; h->sigma2[0] = 1.0f;
    for (int i = 1; i < nl; i++) {
        h->scale[i] = (float)(h->scale[i - 1] * scaleFactor);
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    for (int i = 0; i < nl; i++) { h->inv_scale[i] = 1.0f / h->scale[i]; h->inv_sigma2[i] = 1.0f / h->sigma2[i]; }
    {
        const float factor = (float)(1.0f / scaleFactor);
        float nDesired = cfg->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
        int sum = 0;
        for (int l = 0; l < nl - 1; l++) {
            h->per_level[l] = cv_round_f(nDesired);
            sum += h->per_level[l];
            nDesired *= factor;
        }
        h->per_level[nl - 1] = cfg->nfeatures - sum > 0 ? cfg->nfeatures - sum : 0;
    }
    {   // umax (ORBextractor.cpp:456-471)
        int v, v0;
        const int vmax = (int)floor(15 * sqrt(2.f) / 2 + 1), vmin = (int)ceil(15 * sqrt(2.f) / 2);
        for (v = 0; v <= vmax; ++v) h->umax[v] = cv_round_d(sqrt(225.0 - v * v));
        for (v = 15, v0 = 0; v >= vmin; --v) {
            while (h->umax[v0] == h->umax[v0 + 1]) ++v0;
            h->umax[v] = v0;
            ++v0;
        }
    }
    {
        static const int Q[7] = {18, 34, 48, 56, 48, 34, 18}, QL[7] = {18, 34, 49, 55, 49, 34, 18};
        for (int i = 0; i < 7; i++) h->q[i] = cfg->blur_legacy ? QL[i] : Q[i];
    }
    h->factorPI = (float)(3.14159265358979323846 / 180.f);
    { const char* e = getenv("ORBB_L3_WARPS"); if (e && (atoi(e) == 8 || atoi(e) == 2 || atoi(e) == 4)) { h->l3_warps = atoi(e); h->l3_warps_forced = true; } }

    int rc = setup_geometry(h, cfg->max_width, cfg->max_height, true);
    if (rc != ORBB_OK) { delete h; return rc; }
    // non-max-size shapes can need slightly different totals (rounding): keep a margin
    h->cap_pyr += 64 * 1024; h->cap_rec += 16 * 1024; h->cap_cell += 4096;
    h->dev_capacity = (int)h->cap_out;
    const size_t B = (size_t)cfg->max_batch;
    rc = ORBB_OK;
    auto ok = [&](int r) { if (rc == ORBB_OK) rc = r; };
    ORBB_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    ORBB_CUDA(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
    ORBB_CUDA(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 7; i++) ORBB_CUDA(cudaStreamCreateWithFlags(&h->s_k[i], cudaStreamNonBlocking));
    // ORBB_OCT_PRIO=1: octree streams at the highest priority (experiment: the octrees then finish earlier but take the SM slots from the
    // level kernels - C4 step 7.06 vs 7.03 ms, C1 2.45 vs 2.37 ms - so plain streams stay the default)
    int prio_lo = 0, prio_hi = 0;
    if (const char* e = getenv("ORBB_OCT_PRIO")) { if (atoi(e) == 1) cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi); }
    for (int i = 0; i < 8; i++) {
        ORBB_CUDA(cudaStreamCreateWithPriority(&h->s_side[i], cudaStreamNonBlocking, prio_hi));
        ORBB_CUDA(cudaEventCreateWithFlags(&h->ev_side[i], cudaEventDisableTiming));
        ORBB_CUDA(cudaStreamCreateWithPriority(&h->s_side2[i], cudaStreamNonBlocking, prio_hi));
        ORBB_CUDA(cudaEventCreateWithFlags(&h->ev_side2[i], cudaEventDisableTiming));
    }
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, cfg->device);
    if (getenv("ORBB_NO_OCT_OVERLAP")) h->oct_overlap = false;
    if (const char* e = getenv("ORBB_KSTREAMS")) { const int v = atoi(e); if (v >= 1 && v <= 8) h->n_kstreams = v; }
    ok(dev_alloc(&h->d_pyr, B * h->cap_pyr)); ok(dev_alloc(&h->d_blur, B * h->cap_pyr));
    ok(dev_alloc(&h->d_rec, B * h->cap_rec)); ok(dev_alloc(&h->d_pts, 2 * B * h->cap_rec));
    ok(dev_alloc(&h->d_nodes, B * h->cap_node)); ok(dev_alloc(&h->d_iscr, B * h->cap_node * ISCR_PER_NODE));
    ok(dev_alloc(&h->d_cellflag, B * h->cap_cell));
    ok(dev_alloc(&h->d_out, B * h->cap_out));
    ok(dev_alloc(&h->d_rec_count, B * ORBB_MAX_LEVELS)); ok(dev_alloc(&h->d_out_count, B * ORBB_MAX_LEVELS));
    ok(dev_alloc(&h->d_status, B)); ok(dev_alloc(&h->d_counts, B));
    ok(dev_alloc(&h->d_tab, h->cap_tab)); ok(dev_alloc(&h->d_tiles, h->cap_tile));
    ok(dev_alloc(&h->d_geom, (size_t)ORBB_MAX_LEVELS));
    ok(dev_alloc(&h->d_node_off, (size_t)ORBB_MAX_LEVELS)); ok(dev_alloc(&h->d_cell_off, (size_t)ORBB_MAX_LEVELS));
    ok(dev_alloc(&h->d_kps, B * h->dev_capacity)); ok(dev_alloc(&h->d_desc, B * h->dev_capacity * 32));
    ok(host_alloc(&h->h_img, B * (size_t)cfg->max_width * cfg->max_height));
    ok(host_alloc(&h->h_kps, B * h->dev_capacity)); ok(host_alloc(&h->h_desc, B * h->dev_capacity * 32));
    ok(host_alloc(&h->h_counts, B)); ok(host_alloc(&h->h_status, B));
    if (rc == ORBB_OK) rc = init_constants(h);
    if (rc != ORBB_OK) { orbx_destroy(h); return rc; }
    *out = h;
    return ORBB_OK;
}

extern "C" int orbx_destroy(orbx_handle* h)
{
    if (!h) return ORBB_E_INVALID;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_color); cudaFree(h->d_map); cudaFree(h->d_pyr); cudaFree(h->d_blur); cudaFree(h->d_rec); cudaFree(h->d_pts); cudaFree(h->d_nodes);
    cudaFree(h->d_iscr); cudaFree(h->d_cellflag); cudaFree(h->d_out); cudaFree(h->d_rec_count); cudaFree(h->d_out_count);
    cudaFree(h->d_status); cudaFree(h->d_counts); cudaFree(h->d_tab); cudaFree(h->d_tiles); cudaFree(h->d_geom);
    cudaFree(h->d_node_off); cudaFree(h->d_cell_off); cudaFree(h->d_kps); cudaFree(h->d_desc);
    cudaFreeHost(h->h_img); cudaFreeHost(h->h_kps); cudaFreeHost(h->h_desc); cudaFreeHost(h->h_counts); cudaFreeHost(h->h_status);
    for (int i = 0; i < 4; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (cudaEvent_t e : h->ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : h->ev_done) cudaEventDestroy(e);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    for (int i = 0; i < 7; i++) if (h->s_k[i]) cudaStreamDestroy(h->s_k[i]);
    for (int i = 0; i < 8; i++) {
        if (h->s_side[i]) cudaStreamDestroy(h->s_side[i]);
        if (h->s_side2[i]) cudaStreamDestroy(h->s_side2[i]);
        if (h->ev_side2[i]) cudaEventDestroy(h->ev_side2[i]);
        if (h->ev_side[i]) cudaEventDestroy(h->ev_side[i]);
        for (cudaEvent_t e : h->ev_lvl[i]) cudaEventDestroy(e);
    }
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    for (cudaEvent_t e : h->ev_join) if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : h->ev_fan) if (e) cudaEventDestroy(e);
    if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return ORBB_OK;
}

extern "C" int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                               int32_t* per_level, int32_t* umax)
{
    if (!h) return ORBB_E_INVALID;
    for (int i = 0; i < h->nlevels; i++) {
        if (scale) scale[i] = h->scale[i];
        if (inv_scale) inv_scale[i] = h->inv_scale[i];
        if (sigma2) sigma2[i] = h->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = h->inv_sigma2[i];
        if (per_level) per_level[i] = h->per_level[i];
    }
    if (umax) for (int i = 0; i < 16; i++) umax[i] = h->umax[i];
    return ORBB_OK;
}

// Enqueue the extraction of frames [f0, f0 + nframes) of a device-resident batch on the handle's stream.  imgs_dev / kps_dev /
// desc_dev / counts_dev address frame 0 of the batch; every per-frame scratch buffer of the handle is indexed the same way, so
// chunks of one batch can be enqueued independently (orbx_extract_batch pipelines them against the host copies).
static int enqueue_extract(orbx_handle* h, const uint8_t* imgs_dev0, int f0, int nframes, int w, int ht, size_t stride, size_t fstride,
                           orbx_keypoint* kps_dev0, uint8_t* desc_dev0, int32_t* counts_dev0, int capacity, int slot = 0, bool side_octree = true)
{
    { const int rc = ensure_geometry(h, w, ht); if (rc != ORBB_OK) return rc; }
    const int nl = h->nlevels;
    cudaStream_t st = slot ? h->s_k[slot - 1] : h->stream;
    cudaStream_t side = h->s_side[slot], side2 = h->s_side2[slot];
    // the octree of a level only needs that level's FAST records: it runs beside the next levels' kernels (not while profiling:
    // the per-kernel-class event times are taken with the classes back to back).  Even and odd levels use two side streams, so
    // a long octree (level 0 of a large frame, few frames per launch) does not hold up the octrees of the levels after it.
    // Latency mode (one frame, a stereo pair: at most two level-0 CTAs per SM): every level's octree gets a stream of its own, so that
    // only the LAST level's (small) octree is left on the critical path behind the resize chain (ORBB_OCT_FAN = 0 switches it off)
    static const bool fan_on = getenv("ORBB_OCT_FAN") ? atoi(getenv("ORBB_OCT_FAN")) != 0 : true;
    const bool fan = fan_on && slot == 0 && h->oct_overlap && !h->profiling && nl <= 16 &&
                     (long long)ceil_div(h->geom[0].w, 4 * h->l3_ow[0]) * ceil_div(h->geom[0].h, h->l3_warps * L3_RG) * nframes <= 2ll * h->sm_count;
    const bool overlap = (h->oct_overlap && side_octree && !h->profiling) || fan;
    if (fan) for (int l = 0; l < nl; l++) if (!h->ev_fan[l]) ORBB_CUDA(cudaEventCreateWithFlags(&h->ev_fan[l], cudaEventDisableTiming));
    while ((int)h->ev_lvl[slot].size() < nl) {
        cudaEvent_t e;
        ORBB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_lvl[slot].push_back(e);
    }
    const size_t F0 = (size_t)f0;
    const uint8_t* imgs_dev = imgs_dev0 + F0 * fstride;
    orbx_keypoint* kps_dev = kps_dev0 + F0 * capacity; uint8_t* desc_dev = desc_dev0 + F0 * capacity * 32; int32_t* counts_dev = counts_dev0 + F0;
    uint8_t* const d_pyr = h->d_pyr + F0 * h->pyr_fstride; uint8_t* const d_blur = h->d_blur + F0 * h->pyr_fstride;
    uint32_t* const d_rec = h->d_rec + F0 * h->rec_total; uint32_t* const d_pts = h->d_pts + F0 * 2 * h->rec_total;
    uint4* const d_nodes = h->d_nodes + F0 * h->node_total; int* const d_iscr = h->d_iscr + F0 * h->node_total * ISCR_PER_NODE;
    uint8_t* const d_cellflag = h->d_cellflag + F0 * h->cell_total; int4* const d_out = h->d_out + F0 * h->out_total;
    int* const d_rec_count = h->d_rec_count + F0 * nl; int* const d_out_count = h->d_out_count + F0 * nl; int* const d_status = h->d_status + F0;
    ORBB_CUDA(cudaMemsetAsync(d_rec_count, 0, sizeof(int) * (size_t)nframes * nl, st));
    ORBB_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int) * (size_t)nframes, st));
    OctArgs oa;
    memset(&oa, 0, sizeof oa);
    oa.geom = h->d_geom; oa.nlevels = nl; oa.iniTh = h->cfg.ini_th_fast;
    oa.rec = d_rec; oa.rec_fstride = (long long)h->rec_total; oa.rec_count = d_rec_count;
    oa.pts = d_pts; oa.pts_fstride = 2 * (long long)h->rec_total;
    oa.nodes = d_nodes; oa.nodes_fstride = (long long)h->node_total;
    oa.iscr = d_iscr; oa.iscr_fstride = (long long)h->node_total * ISCR_PER_NODE;
    oa.cellflag = d_cellflag; oa.cellflag_fstride = (long long)h->cell_total;
    oa.node_off = h->d_node_off; oa.cell_off = h->d_cell_off;
    oa.out = d_out; oa.out_fstride = (long long)h->out_total;
    oa.out_count = d_out_count; oa.status = d_status;
    static const bool tl_on = getenv("ORBB_TIMELINE") != nullptr;     // debugging aid: device timeline of one call on stderr
    static cudaEvent_t tl[2 + 2 * ORBB_MAX_LEVELS + 1];
    if (tl_on && !tl[0]) for (auto& e : tl) ORBB_CUDA(cudaEventCreate(&e));
    if (tl_on) ORBB_CUDA(cudaEventRecord(tl[0], st));
    if (h->profiling) ORBB_CUDA(cudaEventRecord(h->ev[0], st));
    for (int l = 0; l < nl; l++) {
        const LevelGeom& g = h->geom[l];
        Level3Args a;
        memset(&a, 0, sizeof a);
        if (l == 0) { a.src = imgs_dev; a.src_pitch = (long long)stride; a.src_fstride = (long long)fstride; }
        else { a.src = d_pyr + g.off; a.src_pitch = g.pitch; a.src_fstride = (long long)h->pyr_fstride; }
        a.w = g.w; a.h = g.h;
        a.status = d_status;
        const int nw = h->l3_warps;          // (half-height tiles, ORBB_L3_WARPS=2, do not shorten a single frame's chain: 209 vs 207 us; neither does building the
                                             //  pyramid first with a light resize kernel and running the levels side by side on eight streams: 242 vs 179 us)
        const int ow = h->l3_ow[l];
        a.use_tma = encode_level_tmap(&a.tmap, a.src, g.w, g.h, nframes, (size_t)a.src_pitch, (size_t)a.src_fstride, 8 * nw + 8, 4 * l3_sp(ow)) ? 1 : 0;
        a.blur = d_blur + g.off; a.blur_pitch = g.pitch; a.blur_fstride = (long long)h->pyr_fstride;
        const bool has_next = l + 1 < nl;
        // the fused resize needs the 4 source columns of a destination word inside the 4-pixel halo: scale <= 1.25
        const bool fused_resize = has_next && (double)g.w <= 1.25 * h->geom[l + 1].w && (double)g.h <= 1.25 * h->geom[l + 1].h;
        if (has_next) {
            a.xt = h->d_tab + g.xt_off; a.yt = h->d_tab + g.yt_off;
            a.xws = h->d_tiles + g.dxs_off; a.yrs = h->d_tiles + g.dys_off;
        }
        if (fused_resize) {
            const LevelGeom& n = h->geom[l + 1];
            a.dst = d_pyr + n.off; a.dst_pitch = n.pitch; a.dst_fstride = (long long)h->pyr_fstride;
        }
        a.fast_on = g.nCols > 0; a.wCell = g.wCell; a.hCell = g.hCell;
        // a corner is recorded when its arc score m exceeds minThFAST; cv::FAST's 3x3 NMS keeps a pixel only if its score m - 1 is strictly above
        // its neighbours', and a non-corner neighbour scores 0: at minThFAST = 0 the corners with m = 1 (score 0) never survive it
        a.minTh = h->cfg.min_th_fast > 1 ? h->cfg.min_th_fast : 1;
        a.magicW = g.magicW; a.magicH = g.magicH;
        a.rec = d_rec + g.rec_off; a.rec_cap = g.rec_cap; a.rec_fstride = (long long)h->rec_total;
        a.rec_count = d_rec_count + l; a.rec_count_stride = nl;
        for (int i = 0; i < 7; i++) a.q[i] = h->q[i];
        l3_blur_taps(a.q, a.kq);
        const bool legacy = h->cfg.blur_legacy != 0;
        {   // quick-reject threshold t' = 2^k - 1 <= minTh (k >= 1); minTh == 0 disables the reject
            int k = 0;
            while ((2 << k) - 1 <= h->cfg.min_th_fast && k < 7) k++;
            if (k >= 1) { a.qmask = ((0xffu << k) & 0xffu) * 0x01010101u; a.qadd = (0x80u - (1u << (k - 1))) * 0x01010101u; }
            else { a.qmask = 0; a.qadd = 0x80808080u; }           // every valid pixel passes
        }
        {
            dim3 grid(ceil_div(g.w, 4 * ow), ceil_div(g.h, nw * L3_RG), nframes);
            // at most two CTAs per SM in the whole launch (one frame, a stereo pair): the level is bound by the serial chain of a tile, not by
            // throughput - twice the warps per tile, FAST reject beside blur + resize (ORBB_L3_SPLIT = 0 | 1 forces the choice)
            static const int split_env = getenv("ORBB_L3_SPLIT") ? atoi(getenv("ORBB_L3_SPLIT")) : -1;
            const bool split = nw == 4 && (split_env >= 0 ? split_env != 0 : (long long)grid.x * grid.y * grid.z <= 2ll * h->sm_count);
            if (split) {
                // levels >= 1 follow their source level by programmatic dependent launch: launch latency and prologue leave the chain
                static const bool pdl_on = getenv("ORBB_L3_PDL") ? atoi(getenv("ORBB_L3_PDL")) != 0 : true;
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = grid; cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = st;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at; cfg.numAttrs = (pdl_on && l > 0) ? 1 : 0;
                if (ow == 32) {
                    if (legacy) ORBB_CUDA(cudaLaunchKernelEx(&cfg, level3_kernel<4, true, true, 32>, a));
                    else ORBB_CUDA(cudaLaunchKernelEx(&cfg, level3_kernel<4, false, true, 32>, a));
                } else if (legacy) ORBB_CUDA(cudaLaunchKernelEx(&cfg, level3_kernel<4, true, true>, a));
                else ORBB_CUDA(cudaLaunchKernelEx(&cfg, level3_kernel<4, false, true>, a));
            }
            else if (ow == 32) { if (legacy) level3_kernel<4, true, false, 32><<<grid, 128, 0, st>>>(a); else level3_kernel<4, false, false, 32><<<grid, 128, 0, st>>>(a); }
            else if (nw == 8) { if (legacy) level3_kernel<8, true><<<grid, 256, 0, st>>>(a); else level3_kernel<8, false><<<grid, 256, 0, st>>>(a); }
            else if (nw == 2) { if (legacy) level3_kernel<2, true><<<grid, 64, 0, st>>>(a); else level3_kernel<2, false><<<grid, 64, 0, st>>>(a); }
            else if (legacy) level3_kernel<4, true><<<grid, 128, 0, st>>>(a);
            else level3_kernel<4, false><<<grid, 128, 0, st>>>(a);
        }
        ORBB_LAUNCHED();
        if (has_next && !fused_resize) {
            const LevelGeom& n = h->geom[l + 1];
            dim3 grid(ceil_div(n.w, 128), n.h, nframes);
            resize_plain_kernel<<<grid, 128, 0, st>>>(a.src, a.src_pitch, a.src_fstride, g.w, g.h, d_pyr + n.off, n.pitch,
                                                       (long long)h->pyr_fstride, n.w, n.h, a.xt, a.yt);
            ORBB_LAUNCHED();
        }
        if (overlap) {
            cudaStream_t sl = fan ? (l < 8 ? h->s_side[l] : h->s_side2[l - 8]) : ((l & 1) ? side2 : side);
            ORBB_CUDA(cudaEventRecord(h->ev_lvl[slot][l], st));
            ORBB_CUDA(cudaStreamWaitEvent(sl, h->ev_lvl[slot][l], 0));
            oa.level0 = l;
            octree_kernel<<<dim3(1, nframes), NTO, 0, sl>>>(oa);
            ORBB_LAUNCHED();
            if (tl_on) ORBB_CUDA(cudaEventRecord(tl[2 + ORBB_MAX_LEVELS + l], sl));
        }
        if (tl_on) ORBB_CUDA(cudaEventRecord(tl[2 + l], st));
    }
    if (h->profiling) ORBB_CUDA(cudaEventRecord(h->ev[1], st));
    if (!overlap) {
        octree_kernel<<<dim3(nl, nframes), NTO, 0, st>>>(oa);
        ORBB_LAUNCHED();
    } else if (fan) {
        for (int l = 0; l < nl; l++) {
            ORBB_CUDA(cudaEventRecord(h->ev_fan[l], l < 8 ? h->s_side[l] : h->s_side2[l - 8]));
            ORBB_CUDA(cudaStreamWaitEvent(st, h->ev_fan[l], 0));
        }
    } else {
        ORBB_CUDA(cudaEventRecord(h->ev_side[slot], side));
        ORBB_CUDA(cudaStreamWaitEvent(st, h->ev_side[slot], 0));
        if (nl > 1) {
            ORBB_CUDA(cudaEventRecord(h->ev_side2[slot], side2));
            ORBB_CUDA(cudaStreamWaitEvent(st, h->ev_side2[slot], 0));
        }
    }
    if (h->profiling) ORBB_CUDA(cudaEventRecord(h->ev[2], st));
    {
        DescArgs a;
        memset(&a, 0, sizeof a);
        a.geom = h->d_geom; a.nlevels = nl;
        a.src0 = imgs_dev; a.src0_pitch = (long long)stride; a.src0_fstride = (long long)fstride;
        a.pyr = d_pyr; a.pyr_fstride = (long long)h->pyr_fstride;
        a.blur = d_blur; a.blur_fstride = (long long)h->pyr_fstride;
        a.lvl = d_out; a.lvl_fstride = (long long)h->out_total; a.lvl_count = d_out_count;
        a.kps = kps_dev; a.desc = desc_dev; a.counts = counts_dev; a.capacity = capacity;
        a.factorPI = h->factorPI;
        const int maxk = capacity < (int)h->out_total ? capacity : (int)h->out_total;
        // a warp walks its keypoints one after the other (~1.5 us each): with few frames in the launch, fewer keypoints per warp
        // shorten that serial chain (one 1000-keypoint frame: 49 -> 12 us) at no cost, the SMs being mostly idle
        a.kpw = 32;
        while (a.kpw > 4 && (long long)ceil_div(maxk > 0 ? maxk : 1, a.kpw) * nframes < 16ll * h->sm_count) a.kpw >>= 1;
        orient_desc_kernel<<<dim3(ceil_div(ceil_div(maxk > 0 ? maxk : 1, a.kpw), OD_WARPS), nframes), 32 * OD_WARPS, 0, st>>>(a);
        ORBB_LAUNCHED();
    }
    if (tl_on) {
        ORBB_CUDA(cudaEventRecord(tl[1], st));
        ORBB_CUDA(cudaEventSynchronize(tl[1]));
        float ms = 0;
        fprintf(stderr, "[timeline] frames %d:", nframes);
        for (int l = 0; l < nl; l++) { cudaEventElapsedTime(&ms, tl[0], tl[2 + l]); fprintf(stderr, " L%d %.3f", l, ms); }
        if (overlap) for (int l = 0; l < nl; l++) { cudaEventElapsedTime(&ms, tl[0], tl[2 + ORBB_MAX_LEVELS + l]); fprintf(stderr, " O%d %.3f", l, ms); }
        cudaEventElapsedTime(&ms, tl[0], tl[1]); fprintf(stderr, " end %.3f\n", ms);
    }
    if (h->profiling) {
        ORBB_CUDA(cudaEventRecord(h->ev[3], st));
        ORBB_CUDA(cudaEventSynchronize(h->ev[3]));
        for (int i = 0; i < 3; i++) { float ms = 0; ORBB_CUDA(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1])); h->prof_ms[i] += ms; }
        h->prof_calls++;
    }
    ORBB_CUDA(cudaGetLastError());
    h->last_frames = f0 + nframes;
    h->last_src = imgs_dev0; h->last_src_pitch = (long long)stride; h->last_src_fstride = (long long)fstride;
    h->last_kps_dev = kps_dev0; h->last_capacity = capacity;
    return ORBB_OK;
}

extern "C" int orbx_extract_batch_device(orbx_handle* h, const uint8_t* imgs_dev, int nframes, int w, int ht,
                                         size_t stride, size_t fstride, orbx_keypoint* kps_dev, uint8_t* desc_dev,
                                         int32_t* counts_dev, int capacity)
{
    if (!h || !imgs_dev || !kps_dev || !desc_dev || !counts_dev || capacity < 1) return ORBB_E_INVALID;
    if (nframes < 1 || nframes > h->cfg.max_batch) return ORBB_E_INVALID;
    if (w < 1 || ht < 1 || w > h->cfg.max_width || ht > h->cfg.max_height || stride < (size_t)w) return ORBB_E_BAD_SHAPE;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    // A batch of 128 frames or more is cut into two chunks of frames on two kernel streams (forked from and joined back into the handle's
    // stream): the latency-bound tail of one chunk (small levels, octree, descriptors) overlaps the big level kernels of the other, as in
    // the host-buffer pipeline.  Per-frame scratch is indexed by frame, so chunks are independent.  512 C1 frames: 2.366 -> 2.308 ms;
    // three chunks the same, four with side-stream octrees 2.455; 64 4K frames are better left whole (6.93 vs 7.20 ms).  ORBB_DEV_CHUNKS=n.
    static const int want = getenv("ORBB_DEV_CHUNKS") ? atoi(getenv("ORBB_DEV_CHUNKS")) : 2;
    int nch = want;
    if (nch > h->n_kstreams) nch = h->n_kstreams;
    if (nch > 1 && (nframes < 64 * nch || h->profiling)) nch = 1;
    if (nch <= 1) return enqueue_extract(h, imgs_dev, 0, nframes, w, ht, stride, fstride, kps_dev, desc_dev, counts_dev, capacity);
    { const int rc = ensure_geometry(h, w, ht); if (rc != ORBB_OK) return rc; }
    if (!h->ev_fork) {
        ORBB_CUDA(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
        for (int i = 0; i < 7; i++) ORBB_CUDA(cudaEventCreateWithFlags(&h->ev_join[i], cudaEventDisableTiming));
    }
    ORBB_CUDA(cudaEventRecord(h->ev_fork, h->stream));
    const int per = (nframes + nch - 1) / nch;
    for (int c = 0; c < nch; c++) {
        const int f0 = c * per, nf = std::min(per, nframes - f0);
        if (nf <= 0) break;
        if (c > 0) ORBB_CUDA(cudaStreamWaitEvent(h->s_k[c - 1], h->ev_fork, 0));
        const int rc = enqueue_extract(h, imgs_dev, f0, nf, w, ht, stride, fstride, kps_dev, desc_dev, counts_dev, capacity, c, false);
        if (rc != ORBB_OK) return rc;
        if (c > 0) {
            ORBB_CUDA(cudaEventRecord(h->ev_join[c - 1], h->s_k[c - 1]));
            ORBB_CUDA(cudaStreamWaitEvent(h->stream, h->ev_join[c - 1], 0));
        }
    }
    h->last_frames = nframes;
    return ORBB_OK;
}

// The launch sequence of orbx_extract_batch_device (2 memsets, the level kernels, the octrees on their side streams, the
// orientation + descriptor kernel) captured once as a CUDA graph: a tracking thread that extracts one frame at a time from fixed
// device buffers replays it with a single launch instead of ~25 runtime calls (Frame::ExtractORB, Frame.cpp:259-265, is on the
// latency path of every frame).  The graph bakes in the pointers, the shape and the capacity given here.
extern "C" int orbx_graph_capture(orbx_handle* h, const uint8_t* imgs_dev, int nframes, int w, int ht, size_t stride, size_t fstride,
                                  orbx_keypoint* kps_dev, uint8_t* desc_dev, int32_t* counts_dev, int capacity)
{
    if (!h || !imgs_dev || !kps_dev || !desc_dev || !counts_dev || capacity < 1) return ORBB_E_INVALID;
    if (nframes < 1 || nframes > h->cfg.max_batch) return ORBB_E_INVALID;
    if (w < 1 || ht < 1 || w > h->cfg.max_width || ht > h->cfg.max_height || stride < (size_t)w) return ORBB_E_BAD_SHAPE;
    if (h->profiling) return ORBB_E_STATE;                      // the profiling path synchronises: not capturable
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    { const int rc = ensure_geometry(h, w, ht); if (rc != ORBB_OK) return rc; }     // table uploads happen outside the capture
    while ((int)h->ev_lvl[0].size() < h->nlevels) {
        cudaEvent_t e;
        ORBB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->ev_lvl[0].push_back(e);
    }
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    if (h->graph_exec) { cudaGraphExecDestroy(h->graph_exec); h->graph_exec = nullptr; h->graph_frames = 0; }
    ORBB_CUDA(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
    const uint64_t launched = g_launches.load();
    const int rc = enqueue_extract(h, imgs_dev, 0, nframes, w, ht, stride, fstride, kps_dev, desc_dev, counts_dev, capacity);
    h->graph_kernels = (int)(g_launches.load() - launched);
    g_launches.store(launched);                                 // captured, not launched
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(h->stream, &graph);
    if (rc != ORBB_OK) { if (graph) cudaGraphDestroy(graph); cudaGetLastError(); return rc; }
    ORBB_CUDA(ce);
    const cudaError_t ci = cudaGraphInstantiate(&h->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    ORBB_CUDA(ci);
    h->graph_frames = nframes;
    return ORBB_OK;
}

extern "C" int orbx_graph_launch(orbx_handle* h)
{
    if (!h) return ORBB_E_INVALID;
    if (!h->graph_exec) return ORBB_E_STATE;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    ORBB_CUDA(cudaGraphLaunch(h->graph_exec, h->stream));
    g_launches.fetch_add((uint64_t)h->graph_kernels, std::memory_order_relaxed);          // the kernels one replay launches
    h->last_frames = h->graph_frames;
    return ORBB_OK;
}

extern "C" int orbx_sync(orbx_handle* h)
{
    if (!h) return ORBB_E_INVALID;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    if (h->last_frames > 0) {
        ORBB_CUDA(cudaMemcpy(h->h_status, h->d_status, sizeof(int) * h->last_frames, cudaMemcpyDeviceToHost));
        for (int f = 0; f < h->last_frames; f++) if (h->h_status[f]) return ORBB_E_CAPACITY;
    }
    return ORBB_OK;
}

extern "C" int orbx_set_profiling(orbx_handle* h, int on)
{
    if (!h) return ORBB_E_INVALID;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    if (on && !h->ev[0]) for (int i = 0; i < 4; i++) ORBB_CUDA(cudaEventCreate(&h->ev[i]));
    h->profiling = on != 0;
    h->prof_ms[0] = h->prof_ms[1] = h->prof_ms[2] = 0; h->prof_calls = 0;
    return ORBB_OK;
}
extern "C" int orbx_get_profile(orbx_handle* h, double* ms3, int* calls)
{
    if (!h || !ms3) return ORBB_E_INVALID;
    for (int i = 0; i < 3; i++) ms3[i] = h->prof_ms[i];
    if (calls) *calls = h->prof_calls;
    return ORBB_OK;
}

extern "C" void* orbx_stream(orbx_handle* h) { return h ? (void*)h->stream : nullptr; }

// Frames per pipeline chunk of orbx_extract_batch: large enough to fill the GPU (64 VGA frames = 46 k CTAs on level 0), small
// enough that the H2D copy of chunk c+1 and the D2H copy of chunk c-1 overlap the kernels of chunk c.
static int pipeline_chunk(int w, int ht)
{
    const long long px = (long long)w * ht;
    int c = (int)((64ll * 640 * 480) / (px > 0 ? px : 1));
    if (c < 1) c = 1;
    if (c > 64) c = 64;
    if (const char* e = getenv("ORBB_CHUNK")) { const int v = atoi(e); if (v > 0) c = v; }
    return c;
}

// Host-buffer batch pipeline.  channels == 1: grey frames (the reference's operator() input).  channels == 3 | 4: colour frames,
// converted on the device like Tracking::GrabImage* does with cv::cvtColor before the extractor runs (Tracking.cpp:202-227).
static int extract_batch_impl(orbx_handle* h, const uint8_t* imgs_host, int nframes, int w, int ht, size_t stride,
                              size_t fstride, int channels, int rgb_order, int raw_w, int raw_h, orbx_keypoint* kps_host, uint8_t* desc_host,
                              int32_t* counts_host, int capacity)
{
    if (!h || !counts_host || capacity < 0 || (capacity > 0 && (!kps_host || !desc_host))) return ORBB_E_INVALID;
    if (channels != 1 && channels != 3 && channels != 4) return ORBB_E_INVALID;
    const size_t CH_ = (size_t)channels;
    const bool remap = raw_w > 0;           // raw camera frames of raw_w x raw_h, rectified on the device into w x ht (orbx_set_rectification)
    if (remap && (channels != 1 || raw_h < 1 || !h || !h->d_map || h->map_w != w || h->map_h != ht || h->map_sw != raw_w || h->map_sh != raw_h)) return ORBB_E_STATE;
    const int in_w = remap ? raw_w : w, in_h = remap ? raw_h : ht;     // what crosses PCIe
    const bool staged = channels > 1 || remap;
    if (nframes < 1 || nframes > h->cfg.max_batch) return ORBB_E_INVALID;
    if (w == 0 || ht == 0 || !imgs_host) {              // empty image: silent return (ORBextractor.cpp:1056-1057)
        for (int f = 0; f < nframes; f++) counts_host[f] = 0;
        return ORBB_OK;
    }
    if (w < 1 || ht < 1 || w > h->cfg.max_width || ht > h->cfg.max_height || stride < (size_t)in_w * CH_) return ORBB_E_BAD_SHAPE;
    if (remap && (size_t)in_w * in_h > 4 * (size_t)h->cfg.max_width * h->cfg.max_height) return ORBB_E_BAD_SHAPE;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    cudaStream_t st = h->stream;
    if (staged && !h->d_color) {      // colour landing buffer, allocated on first use
        const int rc = dev_alloc(&h->d_color, (size_t)h->cfg.max_batch * h->cfg.max_width * h->cfg.max_height * 4);
        if (rc != ORBB_OK) return rc;
    }
    // H2D straight into the level-0 slots of the pyramid (pitch = geom[0].pitch)
    { const int rc = ensure_geometry(h, w, ht); if (rc != ORBB_OK) return rc; }
    const LevelGeom& g0 = h->geom[0];
    cudaPointerAttributes attr;
    const bool pinned = cudaPointerGetAttributes(&attr, imgs_host) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const int devcap = capacity < h->dev_capacity ? (capacity > 0 ? capacity : 1) : h->dev_capacity;
    // results: straight into the caller's buffers when they are pinned and share the device layout, else via staging
    bool direct = capacity == devcap && capacity > 0;
    if (direct) {
        cudaPointerAttributes a1, a2;
        direct = cudaPointerGetAttributes(&a1, kps_host) == cudaSuccess && a1.type == cudaMemoryTypeHost &&
                 cudaPointerGetAttributes(&a2, desc_host) == cudaSuccess && a2.type == cudaMemoryTypeHost;
        cudaGetLastError();
    }
    orbx_keypoint* kdst = direct ? kps_host : h->h_kps;
    uint8_t* ddst = direct ? desc_host : h->h_desc;

    // ---- three-stage pipeline over chunks of frames: H2D (s_in) -> kernels (stream) -> D2H (s_out) ----
    ORBB_CUDA(cudaStreamSynchronize(st));       // an earlier asynchronous device-API call may still be reading the level-0 slots
    const int CH = pipeline_chunk(w, ht);
    // chunk schedule: full chunks in the middle, shorter ones at both ends (the first kernels start sooner, and less work is
    // left when the last input copy lands)
    std::vector<int> cstart;
    {
        static const int ramp = getenv("ORBB_RAMP") ? atoi(getenv("ORBB_RAMP")) : 1;
        std::vector<int> head, tail;
        int left = nframes;
        if (ramp && nframes >= 4 * CH && CH >= 8) {
            head = {CH / 4, CH / 2}; tail = {CH / 2, CH / 4};
            if (ramp >= 2) { head = {CH / 4, CH / 4, CH / 2}; tail = {CH / 2, CH / 4, CH / 4}; }
            for (int v : head) left -= v;
            for (int v : tail) left -= v;
        }
        int f = 0;
        for (int v : head) { cstart.push_back(f); f += v; }
        const int nmid = ceil_div(left, CH);
        for (int i = 0; i < nmid; i++) { cstart.push_back(f); f += (left - i * CH < CH ? left - i * CH : CH); }
        for (int v : tail) { cstart.push_back(f); f += v; }
        cstart.push_back(nframes);
    }
    const int nchunks = (int)cstart.size() - 1;
    static const bool trace = getenv("ORBB_TRACE") != nullptr;      // debugging aid: per-chunk timeline on stderr
    static cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
    if (trace && !ev_t0) { cudaEventCreate(&ev_t0); cudaEventCreate(&ev_t1); }
    while ((int)h->ev_in.size() < nchunks) {
        cudaEvent_t e1, e2;
        ORBB_CUDA(cudaEventCreateWithFlags(&e1, trace ? cudaEventDefault : cudaEventDisableTiming));
        ORBB_CUDA(cudaEventCreateWithFlags(&e2, trace ? cudaEventDefault : cudaEventDisableTiming));
        h->ev_in.push_back(e1); h->ev_done.push_back(e2);
    }
    const bool prof = h->profiling;      // per-kernel-class event timing needs the whole batch in one enqueue
    // Level-0 layout on the device: rows packed at pitch w when that keeps the TMA path (16-byte pitch); then a chunk of
    // packed host frames travels as ONE copy whose "rows" are whole frames (the copy engine runs at PCIe speed on large rows;
    // one 2-D copy per frame with 752-byte rows reaches about half of it).
    const bool tight = (w % 16) == 0 && (size_t)w * ht <= (size_t)g0.pitch * g0.h;
    const size_t dpitch0 = tight ? (size_t)w : (size_t)g0.pitch;
    const bool packed = tight && (!pinned || (stride == (size_t)w && (nframes == 1 || fstride >= (size_t)w * ht)));
    // one chunk (a single frame, a few frames): copies and kernels on ONE stream - no event hops between the three stages
    const bool one = nchunks == 1 && !prof && !trace;
    cudaStream_t sin_ = one ? st : h->s_in, sout_ = one ? st : h->s_out;
    if (trace) cudaEventRecord(ev_t0, sin_);
    for (int c = 0; c < nchunks; c++) {
        const int f0 = cstart[c], nf = cstart[c + 1] - f0;
        if (staged) {          // colour / raw frames land packed in d_color (pageable sources are staged by the driver)
            const size_t fbytes = (size_t)in_w * in_h * CH_;
            if (stride == (size_t)in_w * CH_ && (nframes == 1 || fstride >= fbytes)) {
                ORBB_CUDA(cudaMemcpy2DAsync(h->d_color + (size_t)f0 * fbytes, fbytes, imgs_host + (size_t)f0 * fstride,
                                            nf > 1 ? fstride : fbytes, fbytes, nf, cudaMemcpyHostToDevice, sin_));
            } else {
                for (int f = f0; f < f0 + nf; f++)
                    ORBB_CUDA(cudaMemcpy2DAsync(h->d_color + (size_t)f * fbytes, (size_t)in_w * CH_, imgs_host + (size_t)f * fstride, stride,
                                                (size_t)in_w * CH_, in_h, cudaMemcpyHostToDevice, sin_));
            }
            if (!one) ORBB_CUDA(cudaEventRecord(h->ev_in[c], sin_));
            continue;
        }
        if (!pinned) {         // stage through pinned memory so the copy is truly asynchronous
            for (int f = f0; f < f0 + nf; f++) {
                const uint8_t* src = imgs_host + (size_t)f * fstride;
                uint8_t* stg = h->h_img + (size_t)f * w * ht;
                if (stride == (size_t)w) memcpy(stg, src, (size_t)w * ht);
                else for (int y = 0; y < ht; y++) memcpy(stg + (size_t)y * w, src + (size_t)y * stride, w);
            }
        }
        const uint8_t* cbase = pinned ? imgs_host : h->h_img;
        const size_t cstride = pinned ? stride : (size_t)w, cfstride = pinned ? fstride : (size_t)w * ht;
        if (packed) {
            ORBB_CUDA(cudaMemcpy2DAsync(h->d_pyr + (size_t)f0 * h->pyr_fstride + g0.off, h->pyr_fstride, cbase + (size_t)f0 * cfstride,
                                        nf > 1 ? cfstride : (size_t)w * ht, (size_t)w * ht, nf, cudaMemcpyHostToDevice, sin_));
        } else {
            for (int f = f0; f < f0 + nf; f++)
                ORBB_CUDA(cudaMemcpy2DAsync(h->d_pyr + (size_t)f * h->pyr_fstride + g0.off, dpitch0, cbase + (size_t)f * cfstride, cstride, w, ht,
                                            cudaMemcpyHostToDevice, sin_));
        }
        if (!one) ORBB_CUDA(cudaEventRecord(h->ev_in[c], sin_));
    }
    // every input copy is queued before the first kernel: the copy engine never waits for the host to finish launching a chunk
    for (int c = 0; c < nchunks && !prof; c++) {
        const int f0 = cstart[c], nf = cstart[c + 1] - f0;
        const int ks = c % h->n_kstreams;
        cudaStream_t kst = ks ? h->s_k[ks - 1] : st;
        if (!one) ORBB_CUDA(cudaStreamWaitEvent(kst, h->ev_in[c], 0));
        if (remap) {
            const size_t fbytes = (size_t)in_w * in_h;
            launch_remap(h->d_color + (size_t)f0 * fbytes, (long long)in_w, (long long)fbytes, in_w, in_h, h->d_map, w, ht,
                         h->d_pyr + (size_t)f0 * h->pyr_fstride + g0.off, (long long)dpitch0, (long long)h->pyr_fstride, nf, kst);
            ORBB_LAUNCHED();
        } else if (channels > 1) {
            const size_t fbytes = (size_t)w * ht * CH_;
            launch_cvt_gray(h->d_color + (size_t)f0 * fbytes, (long long)((size_t)w * CH_), (long long)fbytes, w, ht, channels, rgb_order,
                            h->d_pyr + (size_t)f0 * h->pyr_fstride + g0.off, (long long)dpitch0, (long long)h->pyr_fstride, nf, kst);
            ORBB_LAUNCHED();
        }
        const int rc = enqueue_extract(h, h->d_pyr + g0.off, f0, nf, w, ht, dpitch0, h->pyr_fstride, h->d_kps, h->d_desc, h->d_counts, devcap, ks, false);   // chunks already overlap each other: side octrees only add launches
        if (rc != ORBB_OK) return rc;
        if (!one) {
            ORBB_CUDA(cudaEventRecord(h->ev_done[c], kst));
            ORBB_CUDA(cudaStreamWaitEvent(sout_, h->ev_done[c], 0));
        }
        ORBB_CUDA(cudaMemcpyAsync(h->h_counts + f0, h->d_counts + f0, sizeof(int) * nf, cudaMemcpyDeviceToHost, sout_));
        ORBB_CUDA(cudaMemcpyAsync(h->h_status + f0, h->d_status + f0, sizeof(int) * nf, cudaMemcpyDeviceToHost, sout_));
        ORBB_CUDA(cudaMemcpyAsync(kdst + (size_t)f0 * devcap, h->d_kps + (size_t)f0 * devcap, sizeof(orbx_keypoint) * (size_t)nf * devcap, cudaMemcpyDeviceToHost, sout_));
        ORBB_CUDA(cudaMemcpyAsync(ddst + (size_t)f0 * devcap * 32, h->d_desc + (size_t)f0 * devcap * 32, (size_t)32 * nf * devcap, cudaMemcpyDeviceToHost, sout_));
    }
    if (prof) {
        ORBB_CUDA(cudaStreamWaitEvent(st, h->ev_in[nchunks - 1], 0));
        if (remap) {
            launch_remap(h->d_color, (long long)in_w, (long long)((size_t)in_w * in_h), in_w, in_h, h->d_map, w, ht, h->d_pyr + g0.off,
                         (long long)dpitch0, (long long)h->pyr_fstride, nframes, st);
            ORBB_LAUNCHED();
        } else if (channels > 1) {
            const size_t fbytes = (size_t)w * ht * CH_;
            launch_cvt_gray(h->d_color, (long long)((size_t)w * CH_), (long long)fbytes, w, ht, channels, rgb_order, h->d_pyr + g0.off,
                            (long long)dpitch0, (long long)h->pyr_fstride, nframes, st);
            ORBB_LAUNCHED();
        }
        const int rc = enqueue_extract(h, h->d_pyr + g0.off, 0, nframes, w, ht, dpitch0, h->pyr_fstride, h->d_kps, h->d_desc, h->d_counts, devcap);
        if (rc != ORBB_OK) return rc;
        ORBB_CUDA(cudaMemcpyAsync(h->h_counts, h->d_counts, sizeof(int) * nframes, cudaMemcpyDeviceToHost, st));
        ORBB_CUDA(cudaMemcpyAsync(h->h_status, h->d_status, sizeof(int) * nframes, cudaMemcpyDeviceToHost, st));
        ORBB_CUDA(cudaMemcpyAsync(kdst, h->d_kps, sizeof(orbx_keypoint) * (size_t)nframes * devcap, cudaMemcpyDeviceToHost, st));
        ORBB_CUDA(cudaMemcpyAsync(ddst, h->d_desc, (size_t)32 * nframes * devcap, cudaMemcpyDeviceToHost, st));
        ORBB_CUDA(cudaStreamSynchronize(st));
    }
    if (trace) cudaEventRecord(ev_t1, h->s_out);
    if (!one) {
        ORBB_CUDA(cudaStreamSynchronize(h->s_out));
        for (int i = 0; i < 7; i++) ORBB_CUDA(cudaStreamSynchronize(h->s_k[i]));
    }
    if (trace && !prof) {
        float t = 0.f, u = 0.f;
        for (int c = 0; c < nchunks; c++) {
            cudaEventElapsedTime(&t, ev_t0, h->ev_in[c]); cudaEventElapsedTime(&u, ev_t0, h->ev_done[c]);
            fprintf(stderr, "[orbb trace] chunk %2d frames %3d..%3d  h2d done %.3f ms  kernels done %.3f ms\n", c, cstart[c], cstart[c + 1], t, u);
        }
        cudaEventElapsedTime(&t, ev_t0, ev_t1);
        fprintf(stderr, "[orbb trace] results on host %.3f ms\n", t);
    }
    ORBB_CUDA(cudaStreamSynchronize(st));
    int ret = ORBB_OK;
    for (int f = 0; f < nframes; f++) {
        const int n = h->h_counts[f];
        counts_host[f] = n;
        if (h->h_status[f]) ret = ORBB_E_CAPACITY;
        const int m = n < capacity ? n : capacity;
        if (n > capacity) ret = ORBB_E_CAPACITY;
        if (m > 0 && !direct) {
            memcpy(kps_host + (size_t)f * capacity, h->h_kps + (size_t)f * devcap, sizeof(orbx_keypoint) * m);
            memcpy(desc_host + (size_t)f * capacity * 32, h->h_desc + (size_t)f * devcap * 32, (size_t)32 * m);
        }
    }
    return ret;
}

extern "C" int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs_host, int nframes, int w, int ht, size_t stride,
                                  size_t fstride, orbx_keypoint* kps_host, uint8_t* desc_host, int32_t* counts_host, int capacity)
{
    return extract_batch_impl(h, imgs_host, nframes, w, ht, stride, fstride, 1, 0, 0, 0, kps_host, desc_host, counts_host, capacity);
}

extern "C" int orbx_extract_batch_color(orbx_handle* h, const uint8_t* imgs_host, int nframes, int w, int ht, size_t stride,
                                        size_t fstride, int channels, int rgb_order, orbx_keypoint* kps_host, uint8_t* desc_host,
                                        int32_t* counts_host, int capacity)
{
    if (channels != 3 && channels != 4) return ORBB_E_INVALID;
    return extract_batch_impl(h, imgs_host, nframes, w, ht, stride, fstride, channels, rgb_order ? 1 : 0, 0, 0, kps_host, desc_host, counts_host, capacity);
}

// cv::remap's conversion of float maps to fixed point (imgproc/imgwarp.cpp, INTER_BITS = 5): cvRound(map * 32), round half to even.
static void fixed_point_map(const float* mapx, const float* mapy, size_t n, std::vector<int2>& out)
{
    out.resize(n);
    for (size_t i = 0; i < n; i++) out[i] = make_int2((int)nearbyintf(mapx[i] * 32.f), (int)nearbyintf(mapy[i] * 32.f));
}

extern "C" int orbx_set_rectification(orbx_handle* h, const float* mapx_host, const float* mapy_host, int dst_w, int dst_h, int src_w, int src_h)
{
    if (!h) return ORBB_E_INVALID;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    if (!mapx_host && !mapy_host) {             // clear
        cudaFree(h->d_map); h->d_map = nullptr; h->map_w = h->map_h = h->map_sw = h->map_sh = 0;
        return ORBB_OK;
    }
    if (!mapx_host || !mapy_host) return ORBB_E_INVALID;
    if (dst_w < 1 || dst_h < 1 || src_w < 1 || src_h < 1 || dst_w > h->cfg.max_width || dst_h > h->cfg.max_height) return ORBB_E_BAD_SHAPE;
    std::vector<int2> m;
    fixed_point_map(mapx_host, mapy_host, (size_t)dst_w * dst_h, m);
    cudaFree(h->d_map); h->d_map = nullptr;
    const int rc = dev_alloc(&h->d_map, m.size());
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpy(h->d_map, m.data(), m.size() * sizeof(int2), cudaMemcpyHostToDevice));
    h->map_w = dst_w; h->map_h = dst_h; h->map_sw = src_w; h->map_sh = src_h;
    return ORBB_OK;
}

extern "C" int orbx_extract_batch_rectified(orbx_handle* h, const uint8_t* raw_host, int nframes, size_t stride, size_t fstride,
                                            orbx_keypoint* kps_host, uint8_t* desc_host, int32_t* counts_host, int capacity)
{
    if (!h) return ORBB_E_INVALID;
    if (!h->d_map) return ORBB_E_STATE;
    return extract_batch_impl(h, raw_host, nframes, h->map_w, h->map_h, stride, fstride, 1, 0, h->map_sw, h->map_sh, kps_host, desc_host, counts_host, capacity);
}

extern "C" int orbx_remap_device(const uint8_t* src_dev, int nframes, int src_w, int src_h, size_t src_stride, size_t src_fstride,
                                 const float* mapx_host, const float* mapy_host, int dst_w, int dst_h,
                                 uint8_t* dst_dev, size_t dst_stride, size_t dst_fstride, void* stream)
{
    if (!src_dev || !dst_dev || !mapx_host || !mapy_host || nframes < 1) return ORBB_E_INVALID;
    if (src_w < 1 || src_h < 1 || dst_w < 1 || dst_h < 1 || src_stride < (size_t)src_w || dst_stride < (size_t)dst_w) return ORBB_E_BAD_SHAPE;
    std::vector<int2> m;
    fixed_point_map(mapx_host, mapy_host, (size_t)dst_w * dst_h, m);
    int2* d_map = nullptr;
    ORBB_CUDA(cudaMalloc((void**)&d_map, m.size() * sizeof(int2)));
    cudaError_t e = cudaMemcpyAsync(d_map, m.data(), m.size() * sizeof(int2), cudaMemcpyHostToDevice, (cudaStream_t)stream);
    if (e == cudaSuccess) {
        launch_remap(src_dev, (long long)src_stride, (long long)src_fstride, src_w, src_h, d_map, dst_w, dst_h, dst_dev, (long long)dst_stride,
                     (long long)dst_fstride, nframes, (cudaStream_t)stream);
        ORBB_LAUNCHED();
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);      // the temporary map must outlive the kernel
    cudaFree(d_map);
    if (e != cudaSuccess) { set_cuda_error(e, "orbx_remap_device"); return ORBB_E_CUDA; }
    return ORBB_OK;
}

extern "C" int orbx_cvt_gray_device(const uint8_t* src_dev, int nframes, int w, int ht, size_t src_stride, size_t src_fstride,
                                    int channels, int rgb_order, uint8_t* dst_dev, size_t dst_stride, size_t dst_fstride, void* stream)
{
    if (!src_dev || !dst_dev || nframes < 1 || (channels != 3 && channels != 4)) return ORBB_E_INVALID;
    if (w < 1 || ht < 1 || src_stride < (size_t)w * channels || dst_stride < (size_t)w) return ORBB_E_BAD_SHAPE;
    launch_cvt_gray(src_dev, (long long)src_stride, (long long)src_fstride, w, ht, channels, rgb_order ? 1 : 0, dst_dev, (long long)dst_stride,
                    (long long)dst_fstride, nframes, (cudaStream_t)stream);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbx_extract(orbx_handle* h, const uint8_t* img_host, int w, int ht, size_t stride,
                            orbx_keypoint* kps_host, uint8_t* desc_host, int capacity, int* n_out)
{
    if (!n_out) return ORBB_E_INVALID;
    int32_t n = 0;
    const int rc = orbx_extract_batch(h, img_host, 1, w, ht, stride, 0, kps_host, desc_host, &n, capacity);
    *n_out = n;
    return rc;
}

extern "C" int orbx_pyramid_level(orbx_handle* h, int frame, int level, uint8_t* dst_host, size_t dst_stride, int* w, int* ht)
{
    if (!h || level < 0 || level >= h->nlevels || frame < 0) return ORBB_E_INVALID;
    if (h->last_frames < 1 || frame >= h->last_frames) return ORBB_E_STATE;
    const LevelGeom& g = h->geom[level];
    if (w) *w = g.w;
    if (ht) *ht = g.h;
    if (!dst_host) return ORBB_OK;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    if (dst_stride == 0) dst_stride = g.w;
    if (level == 0)
        ORBB_CUDA(cudaMemcpy2D(dst_host, dst_stride, h->last_src + (size_t)frame * h->last_src_fstride, h->last_src_pitch, g.w, g.h, cudaMemcpyDeviceToHost));
    else
        ORBB_CUDA(cudaMemcpy2D(dst_host, dst_stride, h->d_pyr + (size_t)frame * h->pyr_fstride + g.off, g.pitch, g.w, g.h, cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

extern "C" int orbx_debug_stage(orbx_handle* h, int frame, int level, int stage, void* dst_host, size_t dst_bytes, size_t* bytes_out)
{
    if (!h || level < 0 || level >= h->nlevels || frame < 0) return ORBB_E_INVALID;
    if (h->last_frames < 1 || frame >= h->last_frames) return ORBB_E_STATE;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    ORBB_CUDA(cudaStreamSynchronize(h->stream));
    const LevelGeom& g = h->geom[level];
    const int nl = h->nlevels;
    if (stage == ORBX_STAGE_BLUR) {
        const size_t need = (size_t)g.w * g.h;
        if (bytes_out) *bytes_out = need;
        if (!dst_host) return ORBB_OK;
        if (dst_bytes < need) return ORBB_E_CAPACITY;
        ORBB_CUDA(cudaMemcpy2D(dst_host, g.w, h->d_blur + (size_t)frame * h->pyr_fstride + g.off, g.pitch, g.w, g.h, cudaMemcpyDeviceToHost));
        return ORBB_OK;
    }
    if (stage == ORBX_STAGE_FASTKPS) {
        // the records after the per-cell threshold decision = the octree's point buffer is permuted in place, so
        // re-derive the decision here from the raw records (same rule as octree_kernel).
        int nraw = 0;
        ORBB_CUDA(cudaMemcpy(&nraw, h->d_rec_count + (size_t)frame * nl + level, sizeof(int), cudaMemcpyDeviceToHost));
        if (nraw > g.rec_cap) nraw = g.rec_cap;
        std::vector<uint32_t> rec(nraw > 0 ? nraw : 1);
        if (nraw > 0)
            ORBB_CUDA(cudaMemcpy(rec.data(), h->d_rec + (size_t)frame * h->rec_total + g.rec_off, sizeof(uint32_t) * nraw, cudaMemcpyDeviceToHost));
        std::vector<uint8_t> flag((size_t)(g.nCols * g.nRows) + 1, 0);
        auto cell = [&](uint32_t r) { const int x = (r >> 8) & 0xfff, y = r >> 20; return ((y - 19) / g.hCell) * g.nCols + (x - 19) / g.wCell; };
        for (int i = 0; i < nraw; i++) if ((int)(rec[i] & 0xff) > h->cfg.ini_th_fast) flag[cell(rec[i])] = 1;
        std::vector<int32_t> outv;
        for (int i = 0; i < nraw; i++) {
            const int m = rec[i] & 0xff;
            if (m > h->cfg.ini_th_fast || !flag[cell(rec[i])]) {
                outv.push_back((int)((rec[i] >> 8) & 0xfff) - 16); outv.push_back((int)(rec[i] >> 20) - 16); outv.push_back(m - 1);
            }
        }
        const size_t need = outv.size() * sizeof(int32_t);
        if (bytes_out) *bytes_out = need;
        if (!dst_host) return ORBB_OK;
        if (dst_bytes < need) return ORBB_E_CAPACITY;
        if (need) memcpy(dst_host, outv.data(), need);
        return ORBB_OK;
    }
    if (stage == ORBX_STAGE_LEVELKPS) {
        std::vector<int> cnt(nl);
        ORBB_CUDA(cudaMemcpy(cnt.data(), h->d_out_count + (size_t)frame * nl, sizeof(int) * nl, cudaMemcpyDeviceToHost));
        const int n = cnt[level];
        const size_t need = (size_t)n * sizeof(orbx_keypoint);
        if (bytes_out) *bytes_out = need;
        if (!dst_host) return ORBB_OK;
        if (dst_bytes < need) return ORBB_E_CAPACITY;
        if (!n) return ORBB_OK;
        std::vector<int4> lv(n);
        ORBB_CUDA(cudaMemcpy(lv.data(), h->d_out + (size_t)frame * h->out_total + g.out_off, sizeof(int4) * n, cudaMemcpyDeviceToHost));
        int before = 0;
        for (int l = 0; l < level; l++) before += cnt[l];
        std::vector<orbx_keypoint> fin(n);
        if (before + n > h->last_capacity) return ORBB_E_CAPACITY;
        ORBB_CUDA(cudaMemcpy(fin.data(), h->last_kps_dev + (size_t)frame * h->last_capacity + before, sizeof(orbx_keypoint) * n, cudaMemcpyDeviceToHost));
        orbx_keypoint* o = (orbx_keypoint*)dst_host;
        for (int i = 0; i < n; i++) { o[i] = fin[i]; o[i].x = (float)lv[i].x; o[i].y = (float)lv[i].y; }
        return ORBB_OK;
    }
    return ORBB_E_INVALID;
}

extern "C" const char* orbb200_strerror(int code)
{
    switch (code) {
        case ORBB_OK: return "ok";
        case ORBB_E_INVALID: return "invalid argument";
        case ORBB_E_CUDA: return "CUDA error (see orbb200_last_cuda_error)";
        case ORBB_E_CAPACITY: return "output or internal buffer capacity exceeded";
        case ORBB_E_BAD_SHAPE: return "unsupported image shape for this handle";
        case ORBB_E_NOMEM: return "out of memory";
        case ORBB_E_STATE: return "call order violated";
        default: return "unknown error";
    }
}
extern "C" const char* orbb200_last_cuda_error(void) { return g_err; }
extern "C" int orbb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
extern "C" uint64_t orbb200_launch_count(void) { return g_launches.load(); }

// Internal (not part of include/orbb200.h): lets match.cu read a handle's pyramid in place for ComputeStereoMatches.
extern "C" int orbx__pyramid_view(orbx_handle* h, const LevelGeom** geom_host, const LevelGeom** geom_dev, int* nlevels,
                                  const uint8_t** src0, long long* src0_pitch, const uint8_t** pyr, int* device,
                                  const float** scale, const float** inv_scale, void** stream)
{
    if (!h) return ORBB_E_INVALID;
    if (h->last_frames < 1) return ORBB_E_STATE;
    *geom_host = h->geom; *geom_dev = h->d_geom; *nlevels = h->nlevels;
    *src0 = h->last_src; *src0_pitch = h->last_src_pitch; *pyr = h->d_pyr; *device = h->cfg.device;
    *scale = h->scale; *inv_scale = h->inv_scale; *stream = (void*)h->stream;
    return ORBB_OK;
}

// Internal (not part of include/orbb200.h): the extraction half of orbm_stereo_frame_create (match.cu).  `nframes` separate host
// images of one shape are extracted as ONE batch of the handle; on return the per-frame counts are on the host, the keypoints and
// descriptors are still on the device (view->kps_dev / desc_dev, `devcap` rows per frame) and a copy of them into the handle's
// pinned buffers is in flight on view->s_out (exactly counts[f] rows per frame).  Only view->stream was synchronised.
extern "C" int orbx__extract_resident(orbx_handle* h, const uint8_t* const* imgs_host, int nframes, int w, int ht, size_t stride,
                                      int32_t* counts_host, orbx_resident_view* v)
{
    if (!h || !imgs_host || !counts_host || !v || nframes < 1 || nframes > h->cfg.max_batch) return ORBB_E_INVALID;
    for (int f = 0; f < nframes; f++) if (!imgs_host[f]) return ORBB_E_INVALID;
    if (w < 1 || ht < 1 || w > h->cfg.max_width || ht > h->cfg.max_height || stride < (size_t)w) return ORBB_E_BAD_SHAPE;
    ORBB_CUDA(cudaSetDevice(h->cfg.device));
    { const int rc = ensure_geometry(h, w, ht); if (rc != ORBB_OK) return rc; }
    cudaStream_t st = h->stream;
    static const bool trace = getenv("ORBB_TRACE_PAIR") != nullptr;
    static double acc[3] = {0, 0, 0}; static int ncalls = 0;
    auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    const double t0 = trace ? now() : 0.0;
    ORBB_CUDA(cudaStreamSynchronize(st));       // an earlier asynchronous call may still be reading the level-0 slots
    const LevelGeom& g0 = h->geom[0];
    const bool tight = (w % 16) == 0 && (size_t)w * ht <= (size_t)g0.pitch * g0.h;
    const size_t dpitch0 = tight ? (size_t)w : (size_t)g0.pitch;
    for (int f = 0; f < nframes; f++) {         // pageable images go through the pinned staging block, so that every copy is asynchronous
        cudaPointerAttributes attr;
        const bool pinned = cudaPointerGetAttributes(&attr, imgs_host[f]) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        const uint8_t* src = imgs_host[f]; size_t spitch = stride;
        if (!pinned) {
            uint8_t* stg = h->h_img + (size_t)f * w * ht;
            if (stride == (size_t)w) memcpy(stg, imgs_host[f], (size_t)w * ht);
            else for (int y = 0; y < ht; y++) memcpy(stg + (size_t)y * w, imgs_host[f] + (size_t)y * stride, w);
            src = stg; spitch = (size_t)w;
        }
        ORBB_CUDA(cudaMemcpy2DAsync(h->d_pyr + (size_t)f * h->pyr_fstride + g0.off, dpitch0, src, spitch, w, ht, cudaMemcpyHostToDevice, st));
    }
    const int devcap = h->dev_capacity;
    const double t1 = trace ? now() : 0.0;
    const int rc = enqueue_extract(h, h->d_pyr + g0.off, 0, nframes, w, ht, dpitch0, h->pyr_fstride, h->d_kps, h->d_desc, h->d_counts, devcap, 0, false);   // (small launches fan their octrees out anyway)
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpyAsync(h->h_counts, h->d_counts, sizeof(int) * nframes, cudaMemcpyDeviceToHost, st));
    ORBB_CUDA(cudaMemcpyAsync(h->h_status, h->d_status, sizeof(int) * nframes, cudaMemcpyDeviceToHost, st));
    const double t2 = trace ? now() : 0.0;
    ORBB_CUDA(cudaStreamSynchronize(st));
    if (trace) {
        const double t3 = now();
        acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2;
        if (++ncalls % 100 == 0) {
            fprintf(stderr, "[pair trace] images staged and queued %.1f us, extraction launched %.1f us, waited %.1f us\n", acc[0] / 100, acc[1] / 100, acc[2] / 100);
            acc[0] = acc[1] = acc[2] = 0;
        }
    }
    int ret = ORBB_OK;
    for (int f = 0; f < nframes; f++) {
        const int n = h->h_counts[f] < devcap ? h->h_counts[f] : devcap;
        counts_host[f] = n;
        if (h->h_status[f] || h->h_counts[f] > devcap) ret = ORBB_E_CAPACITY;
        if (n > 0) {
            ORBB_CUDA(cudaMemcpyAsync(h->h_kps + (size_t)f * devcap, h->d_kps + (size_t)f * devcap, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyDeviceToHost, h->s_out));
            ORBB_CUDA(cudaMemcpyAsync(h->h_desc + (size_t)f * devcap * 32, h->d_desc + (size_t)f * devcap * 32, (size_t)32 * n, cudaMemcpyDeviceToHost, h->s_out));
        }
    }
    v->kps_dev = h->d_kps; v->desc_dev = h->d_desc; v->kps_pinned = h->h_kps; v->desc_pinned = h->h_desc; v->devcap = devcap;
    v->geom_host = h->geom; v->geom_dev = h->d_geom; v->nlevels = h->nlevels;
    v->src0 = h->last_src; v->src0_pitch = h->last_src_pitch; v->src0_fstride = h->last_src_fstride;
    v->pyr = h->d_pyr; v->pyr_fstride = (long long)h->pyr_fstride;
    v->scale = h->scale; v->inv_scale = h->inv_scale; v->device = h->cfg.device; v->stream = (void*)st; v->s_out = (void*)h->s_out; v->s_aux = (void*)h->s_in;
    return ret;
}
