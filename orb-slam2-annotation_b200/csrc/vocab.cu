// vocab.cu — DBoW2 vocabulary-tree descent of Frame::ComputeBoW on the device (SURVEY 8(f)-2, C ABI section "vocabulary").
// Replaces TemplatedVocabulary<FORB::TDescriptor, FORB>::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1151-1218 and
// :1242-1285) as called by Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cpp:452-459, src/KeyFrame.cpp:74-85).
//
// Layout: the tree is flattened once into a child CSR — the children of a node sit next to each other, descriptors in that
// order (64 B apart would waste sectors: 32 B rows, a 10-child node is 320 contiguous bytes).  One warp walks one feature down
// the tree: lane j takes child j (k <= 32 in one pass), the best child is a packed (distance << 16 | child) minimum, which is
// also the reference's "first child with the smallest distance" (strict <, :1266-1270).  L dependent steps per feature, 60
// Hamming distances for the ORB-SLAM2 vocabulary (k = 10, L = 6).
#include "common.cuh"
#include <map>
#include <vector>
#include <math.h>

using namespace orbb;

namespace {

struct VocabDev {
    const uint4* desc;          // [2 * (n_nodes - 1)] child-CSR order
    const int* child_start;     // [n_nodes]
    const int* child_count;     // [n_nodes]
    const int* child_id;        // [n_nodes - 1] node id of CSR position
    const int* word_id;         // [n_nodes]  (-1 for inner nodes)
    const double* weight;       // [n_nodes]
};

constexpr int VT_WARPS = 8;

__global__ void __launch_bounds__(32 * VT_WARPS) vocab_transform_kernel(const VocabDev v, const uint4* __restrict__ feats, int n, int nid_level,
                                                                        int* __restrict__ word_id, double* __restrict__ weight, int* __restrict__ node_id)
{
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * VT_WARPS + (threadIdx.x >> 5);
    if (f >= n) return;
    const uint4 a0 = feats[2 * f], a1 = feats[2 * f + 1];
    int final_id = 0, level = 0, nid = 0;
    for (;;) {                                              // do { ... } while (!isLeaf)  (:1256-1281); the root is never a leaf here
        const int cnt = v.child_count[final_id], start = v.child_start[final_id];
        ++level;
        unsigned key = 0xffffffffu;
        for (int j = lane; j < cnt; j += 32) {
            const int d = hamming256(a0, a1, v.desc[2 * (start + j)], v.desc[2 * (start + j) + 1]);
            key = min(key, ((unsigned)d << 16) | (unsigned)j);
        }
        key = __reduce_min_sync(0xffffffffu, key);
        final_id = v.child_id[start + (int)(key & 0xffffu)];
        if (level == nid_level) nid = final_id;
        if (v.child_count[final_id] == 0) break;
    }
    if (lane == 0) {
        word_id[f] = v.word_id[final_id];
        weight[f] = v.weight[final_id];
        if (node_id) node_id[f] = nid;
    }
}

struct DevBuf {                       // blocks come from the library's scratch cache (common.cuh): the per-frame transform calls do not cudaMalloc
    void* p = nullptr;
    size_t cap = 0;
    ~DevBuf() { if (p) scratch_free(p, cap); }
    int alloc(size_t bytes) { p = scratch_alloc(bytes ? bytes : 16, &cap); if (!p) { cap = 0; return ORBB_E_NOMEM; } return ORBB_OK; }
    template <typename T> T* as() { return (T*)p; }
};

}  // namespace

struct orbv_vocab {
    int device = 0, k = 0, L = 0, n_nodes = 0, n_words = 0, weighting = 0, norm = 1;
    DevBuf desc, child_start, child_count, child_id, word_id, weight;
    VocabDev dev{};
};

extern "C" int orbv_vocab_create(orbv_vocab** out, int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_leaf,
                                 const uint8_t* desc, const double* weight, int weighting, int norm, int device)
{
    if (!out || !parent || !is_leaf || !desc || !weight) return ORBB_E_INVALID;
    if (k < 1 || k > 65535 || L < 1 || n_nodes < 2 || weighting < 0 || weighting > 3 || norm < 0 || norm > 2) return ORBB_E_INVALID;
    // child CSR in loadFromTextFile order: node nid (>= 1) is appended to its parent's children (:1412-1415)
    std::vector<int> count(n_nodes, 0), start(n_nodes + 1, 0), cursor(n_nodes), child_id(n_nodes - 1), word(n_nodes, -1);
    for (int i = 1; i < n_nodes; i++) {
        if (parent[i] < 0 || parent[i] >= i) return ORBB_E_INVALID;
        count[parent[i]]++;
    }
    for (int i = 0; i < n_nodes; i++) { start[i + 1] = start[i] + count[i]; cursor[i] = start[i]; }
    std::vector<uint8_t> cdesc((size_t)(n_nodes - 1) * 32);
    for (int i = 1; i < n_nodes; i++) {
        const int pos = cursor[parent[i]]++;
        child_id[pos] = i;
        memcpy(&cdesc[(size_t)pos * 32], desc + (size_t)i * 32, 32);
    }
    int nwords = 0;
    for (int i = 1; i < n_nodes; i++) {
        // a word is a node flagged as leaf in the file (:1432-1439); the descent stops at childless nodes (:1281): they must agree
        if ((is_leaf[i] != 0) != (count[i] == 0)) return ORBB_E_INVALID;
        if (is_leaf[i]) word[i] = nwords++;
    }
    if (count[0] == 0 || nwords == 0) return ORBB_E_INVALID;
    for (int i = 0; i < n_nodes; i++) if (count[i] > 65535) return ORBB_E_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1 || device < 0 || device >= ndev) {
        cudaGetLastError();
        set_cuda_error(cudaErrorNoDevice, "no CUDA device (liborbb200 has no CPU fallback)");
        return ORBB_E_CUDA;
    }
    ORBB_CUDA(cudaSetDevice(device));
    orbv_vocab* v = new orbv_vocab();
    v->device = device; v->k = k; v->L = L; v->n_nodes = n_nodes; v->n_words = nwords; v->weighting = weighting; v->norm = norm;
    int rc;
    if ((rc = v->desc.alloc(cdesc.size())) || (rc = v->child_start.alloc((size_t)n_nodes * 4)) || (rc = v->child_count.alloc((size_t)n_nodes * 4)) ||
        (rc = v->child_id.alloc((size_t)(n_nodes - 1) * 4)) || (rc = v->word_id.alloc((size_t)n_nodes * 4)) || (rc = v->weight.alloc((size_t)n_nodes * 8))) {
        delete v;
        return rc;
    }
    bool ok = cudaMemcpy(v->desc.p, cdesc.data(), cdesc.size(), cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(v->child_start.p, start.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(v->child_count.p, count.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(v->child_id.p, child_id.data(), (size_t)(n_nodes - 1) * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(v->word_id.p, word.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    ok = ok && cudaMemcpy(v->weight.p, weight, (size_t)n_nodes * 8, cudaMemcpyHostToDevice) == cudaSuccess;
    if (!ok) { set_cuda_error(cudaGetLastError(), "orbv_vocab_create upload"); delete v; return ORBB_E_CUDA; }
    v->dev.desc = v->desc.as<uint4>(); v->dev.child_start = v->child_start.as<int>(); v->dev.child_count = v->child_count.as<int>();
    v->dev.child_id = v->child_id.as<int>(); v->dev.word_id = v->word_id.as<int>(); v->dev.weight = v->weight.as<double>();
    *out = v;
    return ORBB_OK;
}

extern "C" int orbv_vocab_destroy(orbv_vocab* v)
{
    if (!v) return ORBB_OK;
    cudaSetDevice(v->device);
    cudaDeviceSynchronize();            // orbv_transform_device calls may still be queued (cudaFree used to wait for them; the block cache does not)
    delete v;
    return ORBB_OK;
}

extern "C" int orbv_vocab_words(const orbv_vocab* v) { return v ? v->n_words : 0; }

extern "C" int orbv_transform_device(const orbv_vocab* v, const uint8_t* desc_dev, int n, int levelsup, int32_t* word_id_dev,
                                     double* weight_dev, int32_t* node_id_dev, void* stream)
{
    if (!v || n < 0 || (n > 0 && (!desc_dev || !word_id_dev || !weight_dev))) return ORBB_E_INVALID;
    if (n == 0) return ORBB_OK;
    if (((uintptr_t)desc_dev & 15) != 0) return ORBB_E_INVALID;
    vocab_transform_kernel<<<ceil_div(n, VT_WARPS), 32 * VT_WARPS, 0, (cudaStream_t)stream>>>(v->dev, (const uint4*)desc_dev, n, v->L - levelsup,
                                                                                             word_id_dev, weight_dev, node_id_dev);
    ORBB_LAUNCHED();
    ORBB_CUDA(cudaGetLastError());
    return ORBB_OK;
}

extern "C" int orbv_transform(const orbv_vocab* v, const uint8_t* desc_host, int n, int levelsup, int32_t* word_id, double* weight, int32_t* node_id)
{
    if (!v || n < 0 || (n > 0 && (!desc_host || !word_id || !weight))) return ORBB_E_INVALID;
    if (n == 0) return ORBB_OK;
    ORBB_CUDA(cudaSetDevice(v->device));
    DevBuf d, w, x, nd;
    int rc;
    if ((rc = d.alloc((size_t)n * 32)) || (rc = w.alloc((size_t)n * 4)) || (rc = x.alloc((size_t)n * 8)) || (rc = nd.alloc((size_t)n * 4))) return rc;
    ORBB_CUDA(cudaMemcpy(d.p, desc_host, (size_t)n * 32, cudaMemcpyHostToDevice));
    rc = orbv_transform_device(v, d.as<uint8_t>(), n, levelsup, w.as<int32_t>(), x.as<double>(), nd.as<int32_t>(), nullptr);
    if (rc != ORBB_OK) return rc;
    ORBB_CUDA(cudaMemcpy(word_id, w.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    ORBB_CUDA(cudaMemcpy(weight, x.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
    if (node_id) ORBB_CUDA(cudaMemcpy(node_id, nd.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    return ORBB_OK;
}

// transform(features, BowVector, FeatureVector, levelsup): the descent runs on the device; the two ordered maps are assembled on
// the host in feature order, because BowVector values are doubles accumulated and normalised in map order (:1181-1217).
extern "C" int orbv_compute_bow(const orbv_vocab* v, const uint8_t* desc_host, int n, int levelsup,
                                int32_t* bow_ids, double* bow_vals, int32_t* n_bow,
                                int32_t* fv_nodes, int32_t* fv_offsets, int32_t* fv_features, int32_t* n_fv)
{
    if (!v || n < 0 || !n_bow || !n_fv || (n > 0 && (!desc_host || !bow_ids || !bow_vals || !fv_nodes || !fv_offsets || !fv_features))) return ORBB_E_INVALID;
    *n_bow = 0; *n_fv = 0;
    if (fv_offsets) fv_offsets[0] = 0;
    if (n == 0) return ORBB_OK;
    std::vector<int32_t> wid(n), nid(n);
    std::vector<double> w(n);
    const int rc = orbv_transform(v, desc_host, n, levelsup, wid.data(), w.data(), nid.data());
    if (rc != ORBB_OK) return rc;
    std::map<int32_t, double> bow;
    std::map<int32_t, std::vector<int32_t> > fv;
    const bool tf = v->weighting == 0 || v->weighting == 1;         // TF_IDF, TF accumulate; IDF, BINARY keep the first weight
    for (int i = 0; i < n; i++) {
        if (!(w[i] > 0)) continue;                                  // stopped word (:1181)
        auto it = bow.find(wid[i]);
        if (it == bow.end()) bow.emplace(wid[i], w[i]);
        else if (tf) it->second += w[i];
        fv[nid[i]].push_back(i);
    }
    const bool must = v->norm != 0;
    if (tf && !bow.empty() && !must) { const double nd = (double)bow.size(); for (auto& e : bow) e.second /= nd; }
    if (must) {
        double nrm = 0.0;
        if (v->norm == 1) { for (auto& e : bow) nrm += fabs(e.second); }
        else { for (auto& e : bow) nrm += e.second * e.second; nrm = sqrt(nrm); }
        if (nrm > 0.0) for (auto& e : bow) e.second /= nrm;
    }
    int i = 0;
    for (auto& e : bow) { bow_ids[i] = e.first; bow_vals[i] = e.second; i++; }
    *n_bow = i;
    int j = 0, pos = 0;
    for (auto& e : fv) {
        fv_nodes[j] = e.first;
        for (int32_t f : e.second) fv_features[pos++] = f;
        fv_offsets[++j] = pos;
    }
    *n_fv = j;
    return ORBB_OK;
}
