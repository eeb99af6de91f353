// oracle/dbow_oracle.hpp — TEST INFRASTRUCTURE (CPU oracle), never linked into the product.
//
// CPU restatement of the DBoW2 vocabulary-tree descent ORB-SLAM2 runs per frame (Frame::ComputeBoW,
// ORB-SLAM2/src/Frame.cpp:452-459).  Reference files followed (third-party code vendored in the reference tree):
//   ORB-SLAM2/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h  :312-336 (Node), :1151-1218 (transform -> BowVector + FeatureVector),
//                                                            :1242-1285 (transform of one feature), :1362-1448 (node order of
//                                                            loadFromTextFile: id = line number, children in file order)
//   ORB-SLAM2/Thirdparty/DBoW2/DBoW2/BowVector.cpp           addWeight, addIfNotExist, normalize
//   ORB-SLAM2/Thirdparty/DBoW2/DBoW2/FeatureVector.cpp       addFeature
//   ORB-SLAM2/Thirdparty/DBoW2/DBoW2/FORB.cpp :82-102        distance (twin of ORBmatcher::DescriptorDistance)
// PINNED: the reference ships no vocabulary file (Vocabulary/ holds only bin_vocabulary.cpp) and no tests, so the vendored DBoW2 itself
// was compiled unmodified (oracle/_ref/liborb_refmatch.so) and run on a seeded vocabulary written in its loadFromTextFile format:
// tests/golden/ref_match.npz cases "vocab", "bow", "triang_*" (tests/test_ref_match_golden.py); a second, independent numpy statement
// in tests/test_vocab.py covers more tree shapes.
#pragma once
#include <cstdint>
#include <cmath>
#include <map>
#include <vector>
#include "orb_match_oracle.hpp"

namespace orbo {

enum { DBOW_TF_IDF = 0, DBOW_TF = 1, DBOW_IDF = 2, DBOW_BINARY = 3 };      // WeightingType (BowVector.h)
enum { DBOW_NORM_NONE = 0, DBOW_NORM_L1 = 1, DBOW_NORM_L2 = 2 };           // what the scoring object asks for (ScoringObject.h)

struct VocNode {                      // TemplatedVocabulary.h:312-336
    int parent = 0;
    std::vector<int> children;
    uint8_t descriptor[32];
    double weight = 0;
    int word_id = -1;
    bool isLeaf() const { return children.empty(); }
};

struct Vocabulary {
    int k = 0, L = 0, weighting = DBOW_TF_IDF, norm = DBOW_NORM_L1;
    std::vector<VocNode> nodes;       // nodes[0] = root
    int nwords = 0;

    // loadFromTextFile order (:1400-1443): node i (i >= 1) names its parent, children are appended in that order, leaves get
    // consecutive word ids.
    void build(int k_, int L_, int n, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight)
    {
        k = k_; L = L_;
        nodes.assign(n, VocNode());
        nwords = 0;
        for (int nid = 1; nid < n; nid++) {
            nodes[nid].parent = parent[nid];
            nodes[parent[nid]].children.push_back(nid);
            for (int b = 0; b < 32; b++) nodes[nid].descriptor[b] = desc[(size_t)nid * 32 + b];
            nodes[nid].weight = weight[nid];
            if (is_leaf[nid]) nodes[nid].word_id = nwords++;
        }
    }

    // transform(feature, word_id, weight, nid, levelsup)  (:1242-1285)
    void transform(const uint8_t* feature, int& word_id, double& weight, int* nid, int levelsup) const
    {
        const int nid_level = L - levelsup;
        if (nid) *nid = 0;                        // (the reference sets it only when nid_level <= 0; see include/orbb200.h)
        int final_id = 0, current_level = 0;
        do {
            ++current_level;
            const std::vector<int>& ch = nodes[final_id].children;
            final_id = ch[0];
            double best_d = DescriptorDistance(feature, nodes[final_id].descriptor);
            for (size_t i = 1; i < ch.size(); i++) {
                const int id = ch[i];
                const double d = DescriptorDistance(feature, nodes[id].descriptor);
                if (d < best_d) { best_d = d; final_id = id; }
            }
            if (nid != nullptr && current_level == nid_level) *nid = final_id;
        } while (!nodes[final_id].isLeaf());
        word_id = nodes[final_id].word_id;
        weight = nodes[final_id].weight;
    }

    // transform(features, BowVector, FeatureVector, levelsup)  (:1151-1218)
    void transform_all(const uint8_t* feats, int n, int levelsup, std::map<unsigned, double>& v,
                       std::map<unsigned, std::vector<unsigned> >& fv) const
    {
        v.clear(); fv.clear();
        if (nwords == 0) return;
        const bool must = norm != DBOW_NORM_NONE;
        if (weighting == DBOW_TF || weighting == DBOW_TF_IDF) {
            for (int i = 0; i < n; i++) {
                int id, nid; double w;
                transform(feats + (size_t)i * 32, id, w, &nid, levelsup);
                if (w > 0) {
                    std::map<unsigned, double>::iterator vit = v.lower_bound(id);          // BowVector::addWeight
                    if (vit != v.end() && !(v.key_comp()(id, vit->first))) vit->second += w;
                    else v.insert(vit, std::make_pair((unsigned)id, w));
                    fv[nid].push_back(i);                                                  // FeatureVector::addFeature
                }
            }
            if (!v.empty() && !must) {
                const double nd = v.size();
                for (std::map<unsigned, double>::iterator vit = v.begin(); vit != v.end(); vit++) vit->second /= nd;
            }
        } else {
            for (int i = 0; i < n; i++) {
                int id, nid; double w;
                transform(feats + (size_t)i * 32, id, w, &nid, levelsup);
                if (w > 0) {
                    std::map<unsigned, double>::iterator vit = v.lower_bound(id);          // BowVector::addIfNotExist
                    if (vit == v.end() || (v.key_comp()(id, vit->first))) v.insert(vit, std::make_pair((unsigned)id, w));
                    fv[nid].push_back(i);
                }
            }
        }
        if (must) {                                                                        // BowVector::normalize
            double nrm = 0.0;
            if (norm == DBOW_NORM_L1) { for (auto& e : v) nrm += fabs(e.second); }
            else { for (auto& e : v) nrm += e.second * e.second; nrm = sqrt(nrm); }
            if (nrm > 0.0) for (auto& e : v) e.second /= nrm;
        }
    }
};

}  // namespace orbo
