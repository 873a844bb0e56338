// orb_oracle.cpp — TEST INFRASTRUCTURE (the "port" oracle), not product code.
//
// CPU restatement of the reference's ORB frontend hot path, written independently of the
// reference's data structures (flat arrays, round-synchronous quadtree, own port of
// libstdc++'s introsort) so that it is a second derivation of the same results:
//   * pinned against the reference's own ORBextractor.cc compiled unchanged on the minicv shim
//     (oracle/_ref/liborbref_*.so) by tests/test_oracle_vs_ref.py, and against committed
//     end-to-end fixtures generated from it (tests/golden/e2e_*.npz, tools/gen_golden.py);
//   * the OpenCV primitives underneath (oracle/minicv) are pinned to cv2 4.13.0 goldens.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may call into this library.  Citations are relative to /root/reference/ORB_SLAM3/.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include <atomic>
#include <thread>

#include "minicv.hpp"
#include "orb_pattern.inc"  // kOrbPattern[256][4] = (x0,y0,x1,y1) of bit_pattern_31_ (src/ORBextractor.cc:149-407)

namespace {

// Minimal fork-join helper for the multi-threaded CPU baseline (dynamic chunks of 1).
template <typename F> void parallelFor(int n, int threads, F&& body) {
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads <= 1 || n <= 1) { for (int i = 0; i < n; ++i) body(i, 0); return; }
    if (threads > n) threads = n;
    std::atomic<int> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t]() { for (int i; (i = next.fetch_add(1)) < n;) body(i, t); });
    for (auto& th : pool) th.join();
}

constexpr int kEdge = 19;       // EDGE_THRESHOLD   src/ORBextractor.cc:73
constexpr int kHalfPatch = 15;  // HALF_PATCH_SIZE  src/ORBextractor.cc:72
constexpr int kPatch = 31;      // PATCH_SIZE       src/ORBextractor.cc:71

struct Kp {  // cv::KeyPoint layout (28 bytes)
    float x, y, size, angle, response;
    int octave, class_id;
};

struct Cand {  // FAST candidate in cell-offset coordinates (origin = (16,16) of the level)
    int x, y, resp;
};

// ---------------------------------------------------------------------------------------------
// libstdc++ std::sort (GCC 13 bits/stl_algo.h:1792-1950, bits/stl_heap.h) restated on an array of
// (key, payload) with comp(a,b) = a.key < b.key.  The reference's DistributeOctTree sorts
// (point count, node) pairs with compareNodes = lexicographic (count, UL.x) (src/ORBextractor.cc:538-553,700);
// ties are common and the unstable result is part of parity, so the exact move
// sequence of introsort is reproduced here and checked against std::sort in tests.
// ---------------------------------------------------------------------------------------------
struct SortElem {
    int64_t key;
    int payload;
};
inline bool lessKey(const SortElem& a, const SortElem& b) { return a.key < b.key; }

void adjustHeap(SortElem* f, long hole, long len, SortElem value) {
    const long top = hole;
    long child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (lessKey(f[child], f[child - 1])) child--;
        f[hole] = f[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        f[hole] = f[child - 1];
        hole = child - 1;
    }
    long parent = (hole - 1) / 2;  // __push_heap
    while (hole > top && lessKey(f[parent], value)) {
        f[hole] = f[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    f[hole] = value;
}

void heapSortRange(SortElem* f, long len) {  // __partial_sort(first, last, last)
    if (len >= 2) {                          // __make_heap
        long parent = (len - 2) / 2;
        while (true) {
            SortElem v = f[parent];
            adjustHeap(f, parent, len, v);
            if (parent == 0) break;
            parent--;
        }
    }
    long last = len;  // __sort_heap
    while (last > 1) {
        --last;
        SortElem v = f[last];
        f[last] = f[0];
        adjustHeap(f, 0, last, v);
    }
}

void unguardedLinearInsert(SortElem* last) {
    SortElem v = *last;
    SortElem* next = last - 1;
    while (lessKey(v, *next)) {
        *last = *next;
        last = next;
        --next;
    }
    *last = v;
}

void insertionSort(SortElem* first, SortElem* last) {
    if (first == last) return;
    for (SortElem* i = first + 1; i != last; ++i) {
        if (lessKey(*i, *first)) {
            SortElem v = *i;
            std::memmove(first + 1, first, (size_t)(i - first) * sizeof(SortElem));
            *first = v;
        } else {
            unguardedLinearInsert(i);
        }
    }
}

void introsortLoop(SortElem* first, SortElem* last, int depth) {
    while (last - first > 16) {
        if (depth == 0) {
            heapSortRange(first, last - first);
            return;
        }
        --depth;
        SortElem* mid = first + (last - first) / 2;
        SortElem *a = first + 1, *b = mid, *c = last - 1;  // __move_median_to_first(first, a, b, c)
        if (lessKey(*a, *b)) {
            if (lessKey(*b, *c)) std::swap(*first, *b);
            else if (lessKey(*a, *c)) std::swap(*first, *c);
            else std::swap(*first, *a);
        } else if (lessKey(*a, *c)) std::swap(*first, *a);
        else if (lessKey(*b, *c)) std::swap(*first, *c);
        else std::swap(*first, *b);
        SortElem *lo = first + 1, *hi = last;  // __unguarded_partition(first+1, last, pivot=first)
        while (true) {
            while (lessKey(*lo, *first)) ++lo;
            --hi;
            while (lessKey(*first, *hi)) --hi;
            if (!(lo < hi)) break;
            std::swap(*lo, *hi);
            ++lo;
        }
        introsortLoop(lo, last, depth);
        last = lo;
    }
}

void stdSortPort(SortElem* first, long n) {
    if (n == 0) return;
    int lg = 0;
    for (long t = n; t > 1; t >>= 1) ++lg;
    introsortLoop(first, first + n, 2 * lg);
    if (n > 16) {
        insertionSort(first, first + 16);
        for (SortElem* i = first + 16; i != first + n; ++i) unguardedLinearInsert(i);
    } else {
        insertionSort(first, first + n);
    }
}

// ---------------------------------------------------------------------------------------------
// DistributeOctTree (src/ORBextractor.cc:555-779) + DivideNode (:480-536), round-synchronous form
// (SURVEY.md Appendix B/D.3).  The node list is an ordered vector rebuilt every round.
// ---------------------------------------------------------------------------------------------
struct QNode {
    int ulx, urx, uly, bry;
    std::vector<int> pts;  // candidate indices, original order preserved
    bool noMore;
};

void splitNode(const QNode& p, const std::vector<Cand>& c, QNode out[4]) {
    const int halfX = (int)std::ceil((float)(p.urx - p.ulx) / 2);  // :482
    const int halfY = (int)std::ceil((float)(p.bry - p.uly) / 2);  // :483
    const int mx = p.ulx + halfX, my = p.uly + halfY;
    out[0] = QNode{p.ulx, mx, p.uly, my, {}, false};  // n1 UL quadrant
    out[1] = QNode{mx, p.urx, p.uly, my, {}, false};  // n2 UR
    out[2] = QNode{p.ulx, mx, my, p.bry, {}, false};  // n3 BL
    out[3] = QNode{mx, p.urx, my, p.bry, {}, false};  // n4 BR
    for (int i : p.pts) {                             // :513-527 (float coords are integer valued)
        int q = (c[i].x < mx ? 0 : 1) + (c[i].y < my ? 0 : 2);
        out[q].pts.push_back(i);
    }
    for (int q = 0; q < 4; ++q) out[q].noMore = out[q].pts.size() == 1;  // :529-536
}

std::vector<int> distributeQuadTree(const std::vector<Cand>& c, int width, int height, int N) {
    std::vector<int> result;
    if (c.empty()) return result;
    const int nIni = (int)std::round((float)width / (float)height);  // :559
    const float hX = (float)width / (float)nIni;                      // :561
    std::vector<QNode> nodes;                                         // arena
    std::vector<int> order;                                           // the std::list, front..back
    if (nIni <= 0) return result;  // the reference would divide by zero; no config reaches this
    for (int i = 0; i < nIni; ++i)
        nodes.push_back(QNode{(int)(hX * (float)i), (int)(hX * (float)(i + 1)), 0, height, {}, false});  // :568-580
    for (int i = 0; i < (int)c.size(); ++i) nodes[(size_t)((float)c[i].x / hX)].pts.push_back(i);  // :583-587
    for (int i = 0; i < nIni; ++i) {                                                               // :589-602
        if (nodes[i].pts.empty()) continue;
        nodes[i].noMore = nodes[i].pts.size() == 1;
        order.push_back(i);
    }
    std::vector<SortElem> expandable;  // vSizeAndPointerToNode
    bool finish = false;
    while (!finish) {
        const int prevSize = (int)order.size();
        // ---- one breadth pass (:618-681): every splittable node is replaced by its children, which are
        // pushed to the FRONT one by one; untouched (noMore) nodes keep their relative order behind them.
        std::vector<int> created, kept;
        expandable.clear();
        for (int id : order) {
            if (nodes[id].noMore) { kept.push_back(id); continue; }
            QNode ch[4];
            splitNode(nodes[id], c, ch);
            for (int q = 0; q < 4; ++q) {
                if (ch[q].pts.empty()) continue;
                int nid = (int)nodes.size();
                if (ch[q].pts.size() > 1)
                    expandable.push_back(SortElem{((int64_t)ch[q].pts.size() << 32) | (uint32_t)ch[q].ulx, nid});
                nodes.push_back(std::move(ch[q]));
                created.push_back(nid);
            }
        }
        order.assign(created.rbegin(), created.rend());
        order.insert(order.end(), kept.begin(), kept.end());
        const int nToExpand = (int)expandable.size();
        if ((int)order.size() >= N || (int)order.size() == prevSize) {  // :685
            finish = true;
        } else if ((int)order.size() + nToExpand * 3 > N) {             // :689
            while (!finish) {                                           // :692-753
                const int prev2 = (int)order.size();
                std::vector<SortElem> prev = expandable;
                expandable.clear();
                stdSortPort(prev.data(), (long)prev.size());  // std::sort(..., compareNodes), :700
                std::vector<int> created2;
                std::vector<char> erased(nodes.size() + 4 * prev.size() + 4, 0);
                int size = prev2;
                for (int j = (int)prev.size() - 1; j >= 0; --j) {
                    QNode ch[4];
                    splitNode(nodes[prev[j].payload], c, ch);
                    for (int q = 0; q < 4; ++q) {
                        if (ch[q].pts.empty()) continue;
                        int nid = (int)nodes.size();
                        if (ch[q].pts.size() > 1)
                            expandable.push_back(SortElem{((int64_t)ch[q].pts.size() << 32) | (uint32_t)ch[q].ulx, nid});
                        nodes.push_back(std::move(ch[q]));
                        created2.push_back(nid);
                        ++size;
                    }
                    erased[prev[j].payload] = 1;
                    --size;
                    if (size >= N) break;  // :746-747
                }
                std::vector<int> next(created2.rbegin(), created2.rend());
                for (int id : order)
                    if (!erased[id]) next.push_back(id);
                order.swap(next);
                if ((int)order.size() >= N || (int)order.size() == prev2) finish = true;  // :750-751
            }
        }
    }
    // :757-776 best response per node, first wins ties, in list order
    for (int id : order) {
        const std::vector<int>& p = nodes[id].pts;
        int best = p[0];
        for (size_t k = 1; k < p.size(); ++k)
            if (c[p[k]].resp > c[best].resp) best = p[k];
        result.push_back(best);
    }
    return result;
}

// ---------------------------------------------------------------------------------------------
// The extractor
// ---------------------------------------------------------------------------------------------
struct Oracle {
    int nfeatures, nlevels, iniTh, minTh, fpMode;  // fpMode 0 = STRICT, 1 = NATIVE (fma form)
    double scaleFactor;
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> nPerLevel;
    int umax[kHalfPatch + 1];
    // last call
    std::vector<std::vector<uchar>> padded;  // (w+38)*(h+38)
    std::vector<int> lw, lh;
    std::vector<std::vector<Cand>> cands;     // per level, reference order
    std::vector<std::vector<Kp>> levelKps;    // per level after quadtree + orientation (level coordinates)
};

// ORBextractor::ORBextractor (src/ORBextractor.cc:409-469)
void initTables(Oracle& o) {
    const int nl = o.nlevels;
    o.sf.assign(nl, 1.f); o.sig2.assign(nl, 1.f); o.isf.assign(nl, 1.f); o.isig2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        o.sf[i] = (float)(o.sf[i - 1] * o.scaleFactor);  // float * double member, rounded to float (:420)
        o.sig2[i] = o.sf[i] * o.sf[i];
    }
    for (int i = 0; i < nl; ++i) { o.isf[i] = 1.0f / o.sf[i]; o.isig2[i] = 1.0f / o.sig2[i]; }
    o.nPerLevel.assign(nl, 0);
    float factor = (float)(1.0f / o.scaleFactor);  // :435
    float want = (float)(o.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl)));  // :436
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        o.nPerLevel[l] = cv::cvRound(want);
        sum += o.nPerLevel[l];
        want *= factor;
    }
    o.nPerLevel[nl - 1] = std::max(o.nfeatures - sum, 0);
    // umax (:453-468)
    int vmax = cv::cvFloor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    int vmin = cv::cvCeil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) o.umax[v] = cv::cvRound(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (o.umax[v0] == o.umax[v0 + 1]) ++v0;
        o.umax[v] = v0;
        ++v0;
    }
}

// ComputePyramid (src/ORBextractor.cc:1170-1195)
void buildPyramid(Oracle& o, const uchar* img, int w, int h, int stride) {
    const int nl = o.nlevels;
    o.padded.assign(nl, {}); o.lw.assign(nl, 0); o.lh.assign(nl, 0);
    for (int l = 0; l < nl; ++l) {
        const float s = o.isf[l];
        const int lw = cv::cvRound((float)w * s), lh = cv::cvRound((float)h * s);  // from the ORIGINAL size (:1175)
        o.lw[l] = lw; o.lh[l] = lh;
        o.padded[l].assign((size_t)(lw + 2 * kEdge) * (lh + 2 * kEdge), 0);
        cv::Mat whole(lh + 2 * kEdge, lw + 2 * kEdge, CV_8UC1, o.padded[l].data());
        cv::Mat inner = whole(cv::Rect(kEdge, kEdge, lw, lh));
        if (l == 0) {
            cv::Mat src(h, w, CV_8UC1, const_cast<uchar*>(img), (size_t)stride);
            cv::copyMakeBorder(src, whole, kEdge, kEdge, kEdge, kEdge, cv::BORDER_REFLECT_101);
        } else {
            cv::Mat prevWhole(o.lh[l - 1] + 2 * kEdge, o.lw[l - 1] + 2 * kEdge, CV_8UC1, o.padded[l - 1].data());
            cv::Mat prev = prevWhole(cv::Rect(kEdge, kEdge, o.lw[l - 1], o.lh[l - 1]));
            cv::resize(prev, inner, cv::Size(lw, lh), 0, 0, cv::INTER_LINEAR);
            cv::copyMakeBorder(inner, whole, kEdge, kEdge, kEdge, kEdge, cv::BORDER_REFLECT_101 + cv::BORDER_ISOLATED);
        }
    }
}

// IC_Angle (src/ORBextractor.cc:76-103)
float intensityCentroidAngle(const Oracle& o, int l, int px, int py) {
    const int step = o.lw[l] + 2 * kEdge;
    const uchar* c = o.padded[l].data() + (size_t)(py + kEdge) * step + (px + kEdge);
    int m01 = 0, m10 = 0;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = o.umax[v];
        for (int u = -d; u <= d; ++u) {
            int plus = c[u + v * step], minus = c[u - v * step];
            vsum += plus - minus;
            m10 += u * (plus + minus);
        }
        m01 += v * vsum;
    }
    return cv::fastAtan2((float)m01, (float)m10);
}

// computeOrbDescriptor (src/ORBextractor.cc:107-146).  cos/sin resolve to the float overloads → libm
// cosf/sinf.  fpMode selects how `x*b + y*a` / `x*a - y*b` are rounded (F8).
void rbrief(const uchar* blurred, int step, int px, int py, float angleDeg, int fpMode, uchar* desc) {
    const float factorPI = (float)(CV_PI / 180.f);
    const float angle = angleDeg * factorPI;
    const float a = std::cos(angle), b = std::sin(angle);
    const uchar* center = blurred + (size_t)py * step + px;
    auto sample = [&](int x, int y) -> int {
        const float fx = (float)x, fy = (float)y;
        int r, cidx;
        if (fpMode == 0) {
            volatile float p0 = fx * b, p1 = fy * a, p2 = fx * a, p3 = fy * b;  // each product rounded
            r = cv::cvRound(p0 + p1);
            cidx = cv::cvRound(p2 - p3);
        } else {
            volatile float p1 = fy * a, p3 = fy * b;
            r = cv::cvRound(fmaf(fx, b, p1));    // vfmadd231ss
            cidx = cv::cvRound(fmaf(fx, a, -p3));  // vfmsub132ss
        }
        return center[r * step + cidx];
    };
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            const signed char* p = kOrbPattern[i * 8 + k];
            int t0 = sample(p[0], p[1]), t1 = sample(p[2], p[3]);
            val |= (t0 < t1) << k;
        }
        desc[i] = (uchar)val;
    }
}

}  // namespace

extern "C" {

void* orbo_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int fpMode) {
    Oracle* o = new Oracle();
    o->nfeatures = nfeatures; o->scaleFactor = scaleFactor; o->nlevels = nlevels;
    o->iniTh = iniTh; o->minTh = minTh; o->fpMode = fpMode;
    initTables(*o);
    return o;
}
void orbo_destroy(void* h) { delete static_cast<Oracle*>(h); }

int orbo_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* nfeat, int* umax16) {
    Oracle& o = *static_cast<Oracle*>(h);
    for (int i = 0; i < o.nlevels; ++i) {
        scale[i] = o.sf[i]; inv_scale[i] = o.isf[i]; sigma2[i] = o.sig2[i]; inv_sigma2[i] = o.isig2[i];
        nfeat[i] = o.nPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax16[i] = o.umax[i];
    return o.nlevels;
}

// operator() (src/ORBextractor.cc:1086-1168).  Returns monoIndex (-1 on empty image).
int orbo_extract(void* h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, void* kps_out,
                 uint8_t* desc_out, int cap, int* n_out) {
    Oracle& o = *static_cast<Oracle*>(h);
    if (n_out) *n_out = 0;
    if (!img || w <= 0 || hgt <= 0) return -1;  // :1090-1091
    buildPyramid(o, img, w, hgt, stride);
    const int nl = o.nlevels;
    o.cands.assign(nl, {});
    o.levelKps.assign(nl, {});
    // ComputeKeyPointsOctTree (:781-896)
    for (int l = 0; l < nl; ++l) {
        const int lw = o.lw[l], lh = o.lh[l], step = lw + 2 * kEdge;
        const int minB = kEdge - 3, maxBX = lw - kEdge + 3, maxBY = lh - kEdge + 3;
        const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
        const int nCols = (int)(width / 35.f), nRows = (int)(height / 35.f);
        std::vector<Cand>& cand = o.cands[l];
        if (nCols > 0 && nRows > 0) {
            const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
            cv::Mat whole(lh + 2 * kEdge, lw + 2 * kEdge, CV_8UC1, o.padded[l].data());
            cv::Mat level = whole(cv::Rect(kEdge, kEdge, lw, lh));
            (void)step;
            for (int i = 0; i < nRows; ++i) {
                const int iniY = minB + i * hCell;
                int maxY = iniY + hCell + 6;
                if (iniY >= maxBY - 3) continue;
                if (maxY > maxBY) maxY = maxBY;
                for (int j = 0; j < nCols; ++j) {
                    const int iniX = minB + j * wCell;
                    int maxX = iniX + wCell + 6;
                    if (iniX >= maxBX - 6) continue;
                    if (maxX > maxBX) maxX = maxBX;
                    cv::Mat cell = level.rowRange(iniY, maxY).colRange(iniX, maxX);
                    std::vector<cv::KeyPoint> k;
                    cv::FAST(cell, k, o.iniTh, true);
                    if (k.empty()) cv::FAST(cell, k, o.minTh, true);
                    for (const cv::KeyPoint& p : k)
                        cand.push_back(Cand{(int)p.pt.x + j * wCell, (int)p.pt.y + i * hCell, (int)p.response});
                }
            }
        }
        std::vector<int> keep = distributeQuadTree(cand, maxBX - minB, maxBY - minB, o.nPerLevel[l]);
        const int scaledPatch = (int)(kPatch * o.sf[l]);  // :879
        for (int idx : keep) {
            Kp k;
            k.x = (float)(cand[idx].x + minB); k.y = (float)(cand[idx].y + minB);
            k.size = (float)scaledPatch; k.angle = -1.f; k.response = (float)cand[idx].resp;
            k.octave = l; k.class_id = -1;
            o.levelKps[l].push_back(k);
        }
    }
    for (int l = 0; l < nl; ++l)  // computeOrientation (:894-895)
        for (Kp& k : o.levelKps[l]) k.angle = intensityCentroidAngle(o, l, (int)k.x, (int)k.y);

    int total = 0;
    for (int l = 0; l < nl; ++l) total += (int)o.levelKps[l].size();
    if (n_out) *n_out = total;
    if (total > cap) return -2;
    Kp* out = static_cast<Kp*>(kps_out);
    int mono = 0, stereo = total - 1;
    for (int l = 0; l < nl; ++l) {
        if (o.levelKps[l].empty()) continue;
        const int lw = o.lw[l], lh = o.lh[l];
        cv::Mat whole(lh + 2 * kEdge, lw + 2 * kEdge, CV_8UC1, o.padded[l].data());
        cv::Mat work = whole(cv::Rect(kEdge, kEdge, lw, lh)).clone();  // :1132
        cv::GaussianBlur(work, work, cv::Size(7, 7), 2, 2, cv::BORDER_REFLECT_101);
        const float scale = o.sf[l];
        for (const Kp& k0 : o.levelKps[l]) {
            uchar d[32];
            rbrief(work.data, (int)work.step, (int)k0.x, (int)k0.y, k0.angle, o.fpMode, d);
            Kp k = k0;
            if (l != 0) { k.x *= scale; k.y *= scale; }  // :1149-1151
            int slot = (k.x >= lap0 && k.x <= lap1) ? stereo-- : mono++;  // :1153-1162
            out[slot] = k;
            std::memcpy(desc_out + (size_t)slot * 32, d, 32);
        }
    }
    return mono;
}

int orbo_pyramid(void* h, int level, uint8_t* out, int* w, int* hgt) {
    Oracle& o = *static_cast<Oracle*>(h);
    if (level < 0 || level >= (int)o.padded.size() || o.padded[level].empty()) return -1;
    if (w) *w = o.lw[level];
    if (hgt) *hgt = o.lh[level];
    if (out) std::memcpy(out, o.padded[level].data(), o.padded[level].size());
    return 0;
}

// FAST candidates of the last call, reference order, cell-offset coordinates; rows (x, y, response).
int orbo_candidates(void* h, int level, int32_t* out, int cap) {
    Oracle& o = *static_cast<Oracle*>(h);
    if (level < 0 || level >= (int)o.cands.size()) return -1;
    const std::vector<Cand>& c = o.cands[level];
    for (int i = 0; i < (int)c.size() && i < cap; ++i) { out[3 * i] = c[i].x; out[3 * i + 1] = c[i].y; out[3 * i + 2] = c[i].resp; }
    return (int)c.size();
}

// Per-level keypoints of the last call after the quadtree and orientation (level coordinates, list order).
int orbo_level_keypoints(void* h, int level, void* kps, int cap) {
    Oracle& o = *static_cast<Oracle*>(h);
    if (level < 0 || level >= (int)o.levelKps.size()) return -1;
    const std::vector<Kp>& k = o.levelKps[level];
    if (kps) std::memcpy(kps, k.data(), sizeof(Kp) * std::min<size_t>(k.size(), (size_t)cap));
    return (int)k.size();
}

// Stand-alone quadtree on arbitrary candidates (x, y, response triplets) → indices of the retained points.
int orbo_quadtree(const int32_t* cand, int n, int width, int height, int N, int32_t* keep, int cap) {
    std::vector<Cand> c(n);
    for (int i = 0; i < n; ++i) c[i] = Cand{cand[3 * i], cand[3 * i + 1], cand[3 * i + 2]};
    std::vector<int> r = distributeQuadTree(c, width, height, N);
    for (int i = 0; i < (int)r.size() && i < cap; ++i) keep[i] = r[i];
    return (int)r.size();
}

// Sort check: permutation produced by the introsort port and by the host libstdc++ std::sort.
void orbo_sort_check(const int64_t* keys, int n, int32_t* perm_port, int32_t* perm_std) {
    std::vector<SortElem> a(n), b(n);
    for (int i = 0; i < n; ++i) a[i] = b[i] = SortElem{keys[i], i};
    stdSortPort(a.data(), n);
    std::sort(b.begin(), b.end(), lessKey);
    for (int i = 0; i < n; ++i) { perm_port[i] = a[i].payload; perm_std[i] = b[i].payload; }
}

void orbo_rbrief(const uint8_t* blurred, int step, int px, int py, float angle, int fpMode, uint8_t* desc) {
    rbrief(blurred, step, px, py, angle, fpMode, desc);
}

void orbo_sincosf(const float* x, float* s, float* c, int n) {
    for (int i = 0; i < n; ++i) { s[i] = std::sin(x[i]); c[i] = std::cos(x[i]); }
}

// ---------------------------------------------------------------------------------------------
// Matching: DescriptorDistance (src/ORBmatcher.cc:2058-2074) inside the strict-< best/second loop
// every Search* uses (e.g. src/ORBmatcher.cc:83-120, :273-294), candidates in ascending index.
// Result per query: (idx1,d1,idx2,d2); idx = -1 / d = 256 when fewer than 1 / 2 candidates exist.
// This equals cv::BFMatcher(NORM_HAMMING).knnMatch(k=2) order (distance asc, trainIdx asc; src/Frame.cc:1144).
// ---------------------------------------------------------------------------------------------
int orbo_hamming(const uint8_t* a, const uint8_t* b) {
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32); std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        unsigned v = (unsigned)(pa[i] ^ pb[i]);
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

void orbo_match_top2(const uint8_t* q, int nq, const uint8_t* db, long nd, int32_t* idx1, int32_t* d1, int32_t* idx2,
                     int32_t* d2, int threads) {
    parallelFor(nq, threads, [&](int i, int) {
        int b1 = 256, b2 = 256, i1 = -1, i2 = -1;
        for (long j = 0; j < nd; ++j) {
            int d = orbo_hamming(q + (size_t)i * 32, db + (size_t)j * 32);
            if (d < b1) { b2 = b1; i2 = i1; b1 = d; i1 = (int)j; }
            else if (d < b2) { b2 = d; i2 = (int)j; }
        }
        idx1[i] = i1; d1[i] = b1; idx2[i] = i2; d2[i] = b2;
    });
}

// Frame::ComputeStereoMatches Hamming stage (src/Frame.cc:811-894): right keypoints bucketed by row
// band ±2·scale[octave]; per left keypoint the best right candidate with |octave diff| <= 1 and
// uR in [uL - maxD, uL], strict < from bestDist = TH_HIGH (100), ascending iR.
// kpsL/kpsR are cv::KeyPoint arrays (28 B).  best_idx = 0 / best_dist = 100 when nothing beats TH_HIGH.
void orbo_match_stereo_banded(const void* kpsL_, const uint8_t* descL, int nL, const void* kpsR_,
                              const uint8_t* descR, int nR, const float* scaleFactors, int nRows, float maxD,
                              int32_t* best_idx, int32_t* best_dist) {
    const Kp* kl = static_cast<const Kp*>(kpsL_);
    const Kp* kr = static_cast<const Kp*>(kpsR_);
    std::vector<std::vector<int>> rows(nRows);
    for (int iR = 0; iR < nR; ++iR) {
        const float r = 2.0f * scaleFactors[kr[iR].octave];
        const int maxr = (int)std::ceil(kr[iR].y + r), minr = (int)std::floor(kr[iR].y - r);
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < nRows) rows[yi].push_back(iR);
    }
    for (int iL = 0; iL < nL; ++iL) {
        best_idx[iL] = 0; best_dist[iL] = 100;
        const int row = (int)kl[iL].y;
        if (row < 0 || row >= nRows) continue;
        const float uL = kl[iL].x, minU = uL - maxD, maxU = uL;
        if (maxU < 0) continue;
        int bd = 100, bi = 0;
        for (int iR : rows[row]) {
            if (kr[iR].octave < kl[iL].octave - 1 || kr[iR].octave > kl[iL].octave + 1) continue;
            const float uR = kr[iR].x;
            if (uR >= minU && uR <= maxU) {
                int d = orbo_hamming(descL + (size_t)iL * 32, descR + (size_t)iR * 32);
                if (d < bd) { bd = d; bi = iR; }
            }
        }
        best_idx[iL] = bi; best_dist[iL] = bd;
    }
}

// Frame-parallel extraction for the CPU baseline: one extractor instance per thread over disjoint
// frames (the reference's own stereo pattern, src/Frame.cc:122-125).  Returns total keypoints.
long orbo_extract_many(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, int fpMode,
                       const uint8_t* imgs, int nframes, int w, int hgt, int lap0, int lap1, int threads) {
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    const int cap = nfeatures + 8 * nlevels + 64;
    std::vector<void*> ex(threads, nullptr);  // one extractor instance per thread
    std::vector<std::vector<Kp>> k(threads, std::vector<Kp>(cap));
    std::vector<std::vector<uint8_t>> d(threads, std::vector<uint8_t>((size_t)cap * 32));
    std::vector<long> sub(threads, 0);
    for (int t = 0; t < threads; ++t) ex[t] = orbo_create(nfeatures, scaleFactor, nlevels, iniTh, minTh, fpMode);
    parallelFor(nframes, threads, [&](int f, int t) {
        int n = 0;
        orbo_extract(ex[t], imgs + (size_t)f * w * hgt, w, hgt, w, lap0, lap1, k[t].data(), d[t].data(), cap, &n);
        sub[t] += n;
    });
    long total = 0;
    for (int t = 0; t < threads; ++t) { total += sub[t]; orbo_destroy(ex[t]); }
    return total;
}

}  // extern "C"
