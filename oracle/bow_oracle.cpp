// bow_oracle.cpp — TEST INFRASTRUCTURE (oracle side), not product code.  Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline leg may call into this file.
//
// CPU restatement of Frame::ComputeBoW (ORB_SLAM3/src/Frame.cc:738-745) =
// DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, BowVector&, FeatureVector&, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1126-1196) for the ORB vocabulary's settings (TF_IDF weighting, L1_NORM
// scoring: the "10 6 0 0" header of ORBvoc.txt), on a vocabulary given as flat arrays.
//
// Pinned by oracle/_ref/libbowref.so — the reference's own DBoW2 sources compiled unchanged — in
// tests/test_oracle_bow.py (equal on every case, doubles compared bit for bit) and by tests/golden/bow.npz generated from it.
//
// Flat vocabulary: node 0 is the root; parent[i] < i for i >= 1 (loadFromTextFile numbers nodes in file order and
// appends each to its parent's children, TemplatedVocabulary.h:1373-1388, so children are visited in increasing id
// order); a node without children is a leaf = a word; word ids are assigned to leaves in node order (:1405-1411).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <vector>

namespace {

// FORB::distance (Thirdparty/DBoW2/DBoW2/FORB.cpp:81-100): the SWAR popcount over 8 x 32-bit words
int forbDistance(const uint8_t* a, const uint8_t* b) {
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32);
    std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
struct Voc {
    int L = 0;
    std::vector<std::vector<int>> children;
    std::vector<int> wordOf;
    std::vector<uint8_t> desc;
    std::vector<double> weight;
};

Voc* buildVoc(int n_nodes, const int32_t* parent, const uint8_t* node_desc, const double* weight, int L) {
    Voc* v = new Voc();
    v->L = L;
    v->children.resize(n_nodes);
    for (int i = 1; i < n_nodes; ++i) {
        if (parent[i] < 0 || parent[i] >= i) {
            delete v;
            return nullptr;
        }
        v->children[parent[i]].push_back(i);
    }
    v->wordOf.assign(n_nodes, -1);
    int nw = 0;
    for (int i = 1; i < n_nodes; ++i)
        if (v->children[i].empty()) v->wordOf[i] = nw++;
    v->desc.assign(node_desc, node_desc + (size_t)n_nodes * 32);
    v->weight.assign(weight, weight + n_nodes);
    return v;
}

int transformVoc(const Voc& V, const uint8_t* desc, int n, int levelsup, int32_t* word_ids, double* word_vals, int32_t* n_words,
                 int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx, int32_t* n_nodes_out) {
    const std::vector<std::vector<int>>& children = V.children;
    const uint8_t* node_desc = V.desc.data();
    std::map<unsigned, double> bow;                     // BowVector  (BowVector.h:52-54)
    std::map<unsigned, std::vector<unsigned>> fvec;     // FeatureVector (FeatureVector.h:24-25)
    if (n > 0 && children[0].empty()) return -1;
    for (int f = 0; f < n; ++f) {
        // transform(feature, word_id, weight, nid, levelsup)  (TemplatedVocabulary.h:1217-1263)
        const uint8_t* feat = desc + (size_t)f * 32;
        const int nidLevel = V.L - levelsup;
        unsigned nid = 0;  // "if(nid_level <= 0 && nid != NULL) *nid = 0" — and 0 where the reference leaves it unset
        int cur = 0, level = 0;
        do {
            ++level;
            const std::vector<int>& ch = children[cur];
            cur = ch[0];
            double best = forbDistance(feat, node_desc + (size_t)cur * 32);
            for (size_t j = 1; j < ch.size(); ++j) {
                const double d = forbDistance(feat, node_desc + (size_t)ch[j] * 32);
                if (d < best) {
                    best = d;
                    cur = ch[j];
                }
            }
            if (level == nidLevel) nid = (unsigned)cur;
        } while (!children[cur].empty());
        const double w = V.weight[cur];
        if (w > 0) {  // "not stopped" (:1157)
            bow[(unsigned)V.wordOf[cur]] += w;  // BowVector::addWeight (BowVector.cpp:31-43): insert or +=
            fvec[nid].push_back((unsigned)f);   // FeatureVector::addFeature (FeatureVector.cpp:33-47)
        }
    }
    // L1 normalisation (L1Scoring::mustNormalize -> BowVector::normalize, BowVector.cpp:62-84)
    double norm = 0.0;
    for (auto& kv : bow) norm += std::fabs(kv.second);
    if (norm > 0.0)
        for (auto& kv : bow) kv.second /= norm;

    int k = 0;
    for (auto& kv : bow) {
        word_ids[k] = (int32_t)kv.first;
        word_vals[k] = kv.second;
        ++k;
    }
    *n_words = k;
    int m = 0, pos = 0;
    for (auto& kv : fvec) {
        node_ids[m] = (int32_t)kv.first;
        node_begin[m] = pos;
        for (unsigned i : kv.second) feat_idx[pos++] = (int32_t)i;
        ++m;
    }
    node_begin[m] = pos;
    *n_nodes_out = m;
    return 0;
}

}  // namespace

extern "C" {

// A vocabulary kept between calls (bench.py's cpu_baseline leg); NULL for a malformed tree.
void* orbo_bow_create(int n_nodes, const int32_t* parent, const uint8_t* node_desc, const double* weight, int L) {
    return buildVoc(n_nodes, parent, node_desc, weight, L);
}
void orbo_bow_destroy(void* h) { delete static_cast<Voc*>(h); }

// Outputs as in oracle/bow_ref_wrap.cpp: BowVector (word_ids, word_vals)[*n_words] ascending; FeatureVector CSR
// node_ids[*n_nodes], node_begin[*n_nodes + 1], feat_idx[...].  Returns 0, or -1 for a malformed tree.
int orbo_bow_transform_h(void* h, const uint8_t* desc, int n, int levelsup, int32_t* word_ids, double* word_vals,
                         int32_t* n_words, int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx, int32_t* n_nodes_out) {
    return transformVoc(*static_cast<Voc*>(h), desc, n, levelsup, word_ids, word_vals, n_words, node_ids, node_begin, feat_idx,
                        n_nodes_out);
}

int orbo_bow_transform(int n_nodes, const int32_t* parent, const uint8_t* node_desc, const double* weight, int L,
                       const uint8_t* desc, int n, int levelsup, int32_t* word_ids, double* word_vals, int32_t* n_words,
                       int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx, int32_t* n_nodes_out) {
    Voc* v = buildVoc(n_nodes, parent, node_desc, weight, L);
    if (!v) return -1;
    const int rc = transformVoc(*v, desc, n, levelsup, word_ids, word_vals, n_words, node_ids, node_begin, feat_idx, n_nodes_out);
    delete v;
    return rc;
}

}  // extern "C"
