// bow_ref_wrap.cpp — TEST INFRASTRUCTURE (oracle side), not product code.
//
// extern "C" façade over the reference's own, unmodified DBoW2 (ORB_SLAM3/Thirdparty/DBoW2/DBoW2/
// TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp) compiled against the minicv
// shim: the vocabulary is loaded with the reference's loadFromTextFile (the loader ORB-SLAM3 uses, System.cc:118) and
// features go through TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)
// (TemplatedVocabulary.h:1126-1196) exactly as Frame::ComputeBoW does (src/Frame.cc:738-745).
// Built only where /root/reference exists (oracle/Makefile -> oracle/_ref/libbowref.so).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "DBoW2/FORB.h"
#include "DBoW2/TemplatedVocabulary.h"

namespace {
typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> OrbVocabulary;  // ORBVocabulary.h:27-28
}

extern "C" {

void* bowref_load_text(const char* path) {
    OrbVocabulary* v = new OrbVocabulary();
    if (!v->loadFromTextFile(path)) {
        delete v;
        return nullptr;
    }
    return v;
}

void bowref_destroy(void* h) { delete static_cast<OrbVocabulary*>(h); }

int bowref_words(void* h) { return (int)static_cast<OrbVocabulary*>(h)->size(); }

// desc: n x 32 bytes.  Outputs in std::map iteration order (ascending ids): BowVector as (word_ids, word_vals)[*n_words];
// FeatureVector as CSR: node_ids[*n_nodes], node_begin[*n_nodes + 1], feat_idx[node_begin[*n_nodes]].
// Every output array needs room for n entries (node_begin: n + 1).
void bowref_transform(void* h, const uint8_t* desc, int n, int levelsup, int32_t* word_ids, double* word_vals, int32_t* n_words,
                      int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx, int32_t* n_nodes) {
    OrbVocabulary* v = static_cast<OrbVocabulary*>(h);
    // Converter::toDescriptorVector (src/Converter.cc:30-39): one 1x32 row header per descriptor
    cv::Mat all(n > 0 ? n : 1, 32, CV_8U);
    if (n > 0) std::memcpy(all.data, desc, (size_t)n * 32);
    std::vector<cv::Mat> feats;
    feats.reserve(n);
    for (int i = 0; i < n; ++i) feats.push_back(all.row(i));
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    v->transform(feats, bv, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++k) {
        word_ids[k] = (int32_t)it->first;
        word_vals[k] = it->second;
    }
    *n_words = k;
    int m = 0, pos = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++m) {
        node_ids[m] = (int32_t)it->first;
        node_begin[m] = pos;
        for (size_t j = 0; j < it->second.size(); ++j) feat_idx[pos++] = (int32_t)it->second[j];
    }
    node_begin[m] = pos;
    *n_nodes = m;
}

// L1 score between two BowVectors (ScoringObject.cpp L1Scoring::score), the score ORB-SLAM3's KeyFrameDatabase uses
double bowref_score(void* h, const int32_t* ids1, const double* vals1, int n1, const int32_t* ids2, const double* vals2, int n2) {
    OrbVocabulary* v = static_cast<OrbVocabulary*>(h);
    DBoW2::BowVector a, b;
    for (int i = 0; i < n1; ++i) a.addWeight((DBoW2::WordId)ids1[i], vals1[i]);
    for (int i = 0; i < n2; ++i) b.addWeight((DBoW2::WordId)ids2[i], vals2[i]);
    return v->score(a, b);
}

}  // extern "C"
