// refslice_wrap.cpp — TEST INFRASTRUCTURE (oracle side), not product code.
//
// The reference's OWN function bodies, cut out of /root/reference/ORB_SLAM3/src/{Frame,ORBmatcher}.cc by oracle/Makefile
// (line ranges below, each checked against its expected first line), compiled unchanged inside the class shells of
// shim.h, behind an extern "C" façade:
//   ORBmatcher.cc:35-37      TH_HIGH / TH_LOW / HISTO_LENGTH
//   ORBmatcher.cc:2058-2074  ORBmatcher::DescriptorDistance
//   ORBmatcher.cc:223-425    ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)
//   ORBmatcher.cc:2012-2053  ORBmatcher::ComputeThreeMaxima
//   ORBmatcher.cc:648-763    ORBmatcher::SearchForInitialization
//   ORBmatcher.cc:765-905    ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)
//   ORBmatcher.cc:1676-1887  ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono)  (TrackWithMotionModel's matcher)
//   Pinhole.cpp:43-49        Pinhole::project(const Eigen::Vector3f&)
//   ORBmatcher.cc:43-221     ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints) +
//                            RadiusByViewingCos  (Tracking::SearchLocalPoints' matcher)
//   Frame.cc:385-416         Frame::AssignFeaturesToGrid
//   Frame.cc:657-723         Frame::GetFeaturesInArea
//   Frame.cc:725-735         Frame::PosInGrid
//   Frame.cc:811-981         Frame::ComputeStereoMatches
//   Frame.cc:747-780         Frame::UndistortKeyPoints   (cv::undistortPoints = minicv's, pinned to cv2 goldens)
//   Frame.cc:782-809         Frame::ComputeImageBounds
//   ORBmatcher.cc:123-128, 327-330, 523-524, 636-637, 702-705, 848-851, 1312-1313, 1437-1438, 1569-1570, 1649-1650,
//   1770-1771 and Frame.cc:1151   the ACCEPT CONDITIONS on (best, second) of every matcher: the `if`
//                            lines as written, up to the opening brace of the accepted branch (AcceptShell below)
// They pin oracle/match_oracle.cpp (tests/test_oracle_refslices.py).  Built only where /root/reference exists.
#include <cstdint>
#include <cstring>

#include "shim.h"

namespace ORB_SLAM3 {
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv, Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
#include "ORBmatcher_consts.inc"
#include "ORBmatcher_DescriptorDistance.inc"
#include "Frame_AssignFeaturesToGrid.inc"
#include "Frame_GetFeaturesInArea.inc"
#include "Frame_PosInGrid.inc"
#include "Frame_ComputeStereoMatches.inc"
#include "Frame_UndistortKeyPoints.inc"
#include "Frame_ComputeImageBounds.inc"
#include "ORBmatcher_SearchByBoW.inc"
#include "ORBmatcher_ComputeThreeMaxima.inc"
#include "ORBmatcher_SearchForInitialization.inc"
#include "ORBmatcher_SearchByBoW_KFKF.inc"
#include "Pinhole_project.inc"
#include "ORBmatcher_SearchByProjection_FF.inc"
#include "ORBmatcher_SearchByProjection_FM.inc"
}  // namespace ORB_SLAM3

using ORB_SLAM3::Frame;
using ORB_SLAM3::ORBmatcher;

namespace cv {
struct DMatch {  // the one member Frame.cc:1151 reads
    float distance;
};
}  // namespace cv

namespace {
// The accept conditions of the reference's matchers, each compiled from its own source lines.  A slice ends at the opening
// brace of the accepted branch, so the shell only adds `return true;` and the closing braces; the member / local names are
// the ones the reference uses at that site.
struct AcceptShell {
    static constexpr int TH_HIGH = 100, TH_LOW = 50;  // checked against the sliced ORBmatcher.cc:35-37 in refslice_accept
    float mfNNratio;
    bool projMap(int bestDist, int bestDist2, int bestLevel, int bestLevel2) const {  // ORBmatcher.cc:123-128
        for (int once = 0; once < 1; ++once) {
#include "accept_proj_map_123.inc"
                return true; } }
        }
        return false;
    }
    bool bowKfF(int bestDist1, int bestDist2) const {  // :327-330
#include "accept_bow_kf_f_327.inc"
            return true; } }
        return false;
    }
    bool lowRatio523(int bestDist, float ratioHamming) const {  // :523-524
#include "accept_low_ratio_523.inc"
            return true; }
        return false;
    }
    bool lowRatio636(int bestDist, float ratioHamming) const {  // :636-637
#include "accept_low_ratio_636.inc"
            return true; }
        return false;
    }
    bool init(int bestDist, int bestDist2) const {  // :702-705
#include "accept_init_702.inc"
            return true; } }
        return false;
    }
    bool bowKfKf(int bestDist1, int bestDist2) const {  // :848-851
#include "accept_bow_kf_kf_848.inc"
            return true; } }
        return false;
    }
    bool bestLow1312(int bestDist) const {
#include "accept_best_low_1312.inc"
            return true; }
        return false;
    }
    bool bestLow1437(int bestDist) const {
#include "accept_best_low_1437.inc"
            return true; }
        return false;
    }
    bool bestHigh1569(int bestDist) const {
#include "accept_best_high_1569.inc"
            return true; }
        return false;
    }
    bool bestHigh1649(int bestDist) const {
#include "accept_best_high_1649.inc"
            return true; }
        return false;
    }
    bool bestHigh1770(int bestDist) const {
#include "accept_best_high_1770.inc"
            return true; }
        return false;
    }
    bool knnRatio(int d1, int d2, int results) const {  // Frame.cc:1151, `it` walks the knnMatch result
        std::vector<std::vector<cv::DMatch>> matches(1);
        if (results >= 1) matches[0].push_back(cv::DMatch{(float)d1});
        if (results >= 2) matches[0].push_back(cv::DMatch{(float)d2});
        std::vector<std::vector<cv::DMatch>>::iterator it = matches.begin();
#include "accept_knn_ratio_1151.inc"
            return true; }
        return false;
    }
};
}  // namespace

namespace {
static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
std::vector<cv::KeyPoint> keypoints(const void* p, int n) {
    std::vector<cv::KeyPoint> v(n);
    if (n) std::memcpy(v.data(), p, (size_t)n * sizeof(cv::KeyPoint));
    return v;
}
cv::Mat descriptors(const uint8_t* p, int n) {
    cv::Mat m(n > 0 ? n : 1, 32, CV_8U);
    if (n) std::memcpy(m.data, p, (size_t)n * 32);
    return m;
}
}  // namespace

extern "C" {

// site = the reference line of the condition; ratio = mfNNratio (or ratioHamming at 523 / 636); results = entries of the
// knnMatch row (1151 only).  -1 = unknown site.
int refslice_accept(int site, float ratio, int d1, int d2, int level1, int level2, int results) {
    static_assert(AcceptShell::TH_HIGH == 100 && AcceptShell::TH_LOW == 50, "");
    if (ORBmatcher::TH_HIGH != AcceptShell::TH_HIGH || ORBmatcher::TH_LOW != AcceptShell::TH_LOW) return -2;
    AcceptShell a{ratio};
    switch (site) {
        case 123: return a.projMap(d1, d2, level1, level2);
        case 327: return a.bowKfF(d1, d2);
        case 523: return a.lowRatio523(d1, ratio);
        case 636: return a.lowRatio636(d1, ratio);
        case 702: return a.init(d1, d2);
        case 848: return a.bowKfKf(d1, d2);
        case 1312: return a.bestLow1312(d1);
        case 1437: return a.bestLow1437(d1);
        case 1569: return a.bestHigh1569(d1);
        case 1649: return a.bestHigh1649(d1);
        case 1770: return a.bestHigh1770(d1);
        case 1151: return a.knnRatio(d1, d2, results);
    }
    return -1;
}

// the same over every (d1, d2) in [0, 256]^2: out[d1 * 257 + d2]
void refslice_accept_grid(int site, float ratio, int level1, int level2, int results, uint8_t* out) {
    for (int d1 = 0; d1 <= 256; ++d1)
        for (int d2 = 0; d2 <= 256; ++d2) out[d1 * 257 + d2] = (uint8_t)refslice_accept(site, ratio, d1, d2, level1, level2, results);
}

int refslice_th_low() { return ORBmatcher::TH_LOW; }
int refslice_th_high() { return ORBmatcher::TH_HIGH; }

int refslice_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    alignas(4) uint8_t aa[32], bb[32];
    std::memcpy(aa, a, 32);
    std::memcpy(bb, b, 32);
    return ORBmatcher::DescriptorDistance(cv::Mat(1, 32, CV_8U, aa), cv::Mat(1, 32, CV_8U, bb));
}

// Frame::ComputeStereoMatches on keypoints / descriptors / padded pyramid planes handed in by the caller.
// planes_l / planes_r: nlevels pointers to (h_l + 38) x (w_l + 38) padded planes (mvImagePyramid[l] is the ROI at (19, 19),
// ORBextractor.cc:1176-1178).  u_right / depth: nl floats.
void refslice_compute_stereo_matches(const void* kps_l, const uint8_t* desc_l, int nl, const void* kps_r, const uint8_t* desc_r,
                                     int nr, const float* scale, const float* inv_scale, int nlevels, const uint8_t* const* planes_l,
                                     const uint8_t* const* planes_r, const int* w, const int* h, float bf, float b, float* u_right,
                                     float* depth) {
    ORB_SLAM3::ORBextractor L, R;
    for (int l = 0; l < nlevels; ++l) {
        const int pw = w[l] + 38, ph = h[l] + 38;
        cv::Mat pl(ph, pw, CV_8U, const_cast<uint8_t*>(planes_l[l]), (size_t)pw), pr(ph, pw, CV_8U, const_cast<uint8_t*>(planes_r[l]), (size_t)pw);
        L.mvImagePyramid.push_back(pl(cv::Rect(19, 19, w[l], h[l])));
        R.mvImagePyramid.push_back(pr(cv::Rect(19, 19, w[l], h[l])));
    }
    Frame F;
    F.mpORBextractorLeft = &L; F.mpORBextractorRight = &R;
    F.N = nl;
    F.mvKeys = keypoints(kps_l, nl); F.mvKeysRight = keypoints(kps_r, nr);
    F.mDescriptors = descriptors(desc_l, nl); F.mDescriptorsRight = descriptors(desc_r, nr);
    F.mvScaleFactors.assign(scale, scale + nlevels); F.mvInvScaleFactors.assign(inv_scale, inv_scale + nlevels);
    F.mbf = bf; F.mb = b;
    F.ComputeStereoMatches();
    std::memcpy(u_right, F.mvuRight.data(), (size_t)nl * 4);
    std::memcpy(depth, F.mvDepth.data(), (size_t)nl * 4);
}

namespace {
void pinholeFrame(Frame& F, ORB_SLAM3::Pinhole& cam, const float* K, const float* dist, int ndist, const float* P) {
    cam.K = cv::Mat(3, 3, CV_32F, const_cast<float*>(K)).clone();
    F.mpCamera = &cam;
    F.mK = cv::Mat(3, 3, CV_32F, const_cast<float*>(P)).clone();
    F.mDistCoef = cv::Mat(ndist, 1, CV_32F, const_cast<float*>(dist)).clone();  // Tracking.cc builds it as ndist x 1 CV_32F
}
}  // namespace

// Frame::UndistortKeyPoints: mvKeys = kps -> mvKeysUn = out (n 28-byte cv::KeyPoint records).  K, P: 3 x 3 row-major.
void refslice_undistort_keypoints(const void* kps, int n, const float* K, const float* dist, int ndist, const float* P, void* out) {
    Frame F;
    ORB_SLAM3::Pinhole cam;
    pinholeFrame(F, cam, K, dist, ndist, P);
    F.N = n;
    F.mvKeys = keypoints(kps, n);
    F.UndistortKeyPoints();
    if (n) std::memcpy(out, F.mvKeysUn.data(), (size_t)n * sizeof(cv::KeyPoint));
}

// Frame::ComputeImageBounds(imLeft) for an image of cols x rows -> bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)
void refslice_image_bounds(int cols, int rows, const float* K, const float* dist, int ndist, const float* P, float* bounds) {
    Frame F;
    ORB_SLAM3::Pinhole cam;
    pinholeFrame(F, cam, K, dist, ndist, P);
    cv::Mat im;
    im.cols = cols; im.rows = rows;
    F.ComputeImageBounds(im);
    bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY;
}

// Monocular frame (Nleft == -1): AssignFeaturesToGrid over mvKeysUn = kps_t, then for every query GetFeaturesInArea and the
// strict-< best/second loop of ORBmatcher::SearchByProjection (ORBmatcher.cc:75-117) with the reference's DescriptorDistance.
// queries: nq x (u, v, r, min_level, max_level) as 3 floats + 2 int32 (20 bytes).  cand_count (may be NULL): candidates per query.
void refslice_grid_top2(const void* kps_t, const uint8_t* desc_t, int nt, float min_x, float max_x, float min_y, float max_y,
                        const uint8_t* queries, const uint8_t* desc_q, int nq, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                        int32_t* cand_count) {
    Frame F;
    F.N = nt;
    F.Nleft = -1;
    F.mvKeysUn = keypoints(kps_t, nt);
    F.mDescriptors = descriptors(desc_t, nt);
    Frame::mnMinX = min_x; Frame::mnMaxX = max_x; Frame::mnMinY = min_y; Frame::mnMaxY = max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);   // Frame.cc:251
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);  // Frame.cc:252
    F.AssignFeaturesToGrid();
    cv::Mat Q = descriptors(desc_q, nq);
    for (int q = 0; q < nq; ++q) {
        float u, v, r;
        int32_t lo, hi;
        std::memcpy(&u, queries + (size_t)q * 20, 4); std::memcpy(&v, queries + (size_t)q * 20 + 4, 4);
        std::memcpy(&r, queries + (size_t)q * 20 + 8, 4); std::memcpy(&lo, queries + (size_t)q * 20 + 12, 4);
        std::memcpy(&hi, queries + (size_t)q * 20 + 16, 4);
        const std::vector<size_t> cand = F.GetFeaturesInArea(u, v, r, lo, hi);
        int best = 256, best2 = 256, bi = -1, bi2 = -1;
        const cv::Mat dq = Q.row(q);
        for (size_t j = 0; j < cand.size(); ++j) {
            const int dist = ORBmatcher::DescriptorDistance(dq, F.mDescriptors.row((int)cand[j]));
            if (dist < best) { best2 = best; bi2 = bi; best = dist; bi = (int)cand[j]; }
            else if (dist < best2) { best2 = dist; bi2 = (int)cand[j]; }
        }
        idx1[q] = bi; d1[q] = best; idx2[q] = bi2; d2[q] = best2;
        if (cand_count) cand_count[q] = (int32_t)cand.size();
    }
}

// ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches) for a monocular keyframe / frame (Nleft == -1, mpCamera2 == NULL).
// valid_kf[i] != 0: keyframe feature i has a (good) MapPoint; 2 = its MapPoint isBad().  The two FeatureVectors arrive as
// the CSR arrays of the BoW transform.  match_f[j] = keyframe feature whose MapPoint was given to frame feature j, or -1.
int refslice_search_by_bow(const uint8_t* desc_kf, int n_kf, const uint8_t* valid_kf, const float* angle_kf, const int32_t* kf_node_ids,
                           const int32_t* kf_node_begin, const int32_t* kf_feat_idx, int kf_nodes, const uint8_t* desc_f, int n_f,
                           const float* angle_f, const int32_t* f_node_ids, const int32_t* f_node_begin, const int32_t* f_feat_idx,
                           int f_nodes, float nnratio, int check_orientation, int32_t* match_f) {
    ORB_SLAM3::KeyFrame KF;
    Frame F;
    std::vector<ORB_SLAM3::MapPoint> points(n_kf);
    KF.mvpMapPoints.assign(n_kf, nullptr);
    KF.mvKeysUn.resize(n_kf);
    for (int i = 0; i < n_kf; ++i) {
        points[i].index = i;
        points[i].mbBad = valid_kf[i] == 2;
        if (valid_kf[i]) KF.mvpMapPoints[i] = &points[i];
        KF.mvKeysUn[i].angle = angle_kf[i];
    }
    KF.mDescriptors = descriptors(desc_kf, n_kf);
    for (int j = 0; j < kf_nodes; ++j)
        for (int p = kf_node_begin[j]; p < kf_node_begin[j + 1]; ++p) KF.mFeatVec.addFeature((DBoW2::NodeId)kf_node_ids[j], (unsigned)kf_feat_idx[p]);
    F.N = n_f;
    F.Nleft = -1;
    F.mvKeys.resize(n_f);
    for (int i = 0; i < n_f; ++i) F.mvKeys[i].angle = angle_f[i];
    F.mDescriptors = descriptors(desc_f, n_f);
    for (int j = 0; j < f_nodes; ++j)
        for (int p = f_node_begin[j]; p < f_node_begin[j + 1]; ++p) F.mFeatVec.addFeature((DBoW2::NodeId)f_node_ids[j], (unsigned)f_feat_idx[p]);
    ORBmatcher matcher(nnratio, check_orientation != 0);
    std::vector<ORB_SLAM3::MapPoint*> out;
    const int n = matcher.SearchByBoW(&KF, F, out);
    for (int j = 0; j < n_f; ++j) match_f[j] = out[j] ? out[j]->index : -1;
    return n;
}

// ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) on two monocular frames: F2's grid is
// built by the reference's AssignFeaturesToGrid over kps2 (= mvKeysUn) inside `bounds`.  prev_matched: n1 x 2 floats, in / out.
int refslice_search_for_initialization(const void* kps1, const uint8_t* desc1, int n1, const void* kps2, const uint8_t* desc2, int n2,
                                       float min_x, float max_x, float min_y, float max_y, float* prev_matched, int window_size,
                                       float nnratio, int check_orientation, int32_t* matches12) {
    Frame F1, F2;
    F1.N = n1; F1.Nleft = -1;
    F1.mvKeysUn = keypoints(kps1, n1);
    F1.mDescriptors = descriptors(desc1, n1);
    F2.N = n2; F2.Nleft = -1;
    F2.mvKeysUn = keypoints(kps2, n2);
    F2.mDescriptors = descriptors(desc2, n2);
    Frame::mnMinX = min_x; Frame::mnMaxX = max_x; Frame::mnMinY = min_y; Frame::mnMaxY = max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);   // Frame.cc:251
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);  // Frame.cc:252
    F2.AssignFeaturesToGrid();
    std::vector<cv::Point2f> prev(n1);
    for (int i = 0; i < n1; ++i) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchForInitialization(F1, F2, prev, m12, window_size);
    for (int i = 0; i < n1; ++i) {
        matches12[i] = m12[i];
        prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y;
    }
    return n;
}

// ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) for two monocular keyframes (NLeft == -1).  valid1 / valid2: 0 = no MapPoint,
// 1 = good MapPoint, 2 = isBad().  match12[i] = feature of keyframe 2 whose MapPoint was given to feature i of keyframe 1, or -1.
int refslice_search_by_bow_kf(const uint8_t* desc1, int n1, const uint8_t* valid1, const float* angle1, const int32_t* node_ids1,
                              const int32_t* node_begin1, const int32_t* feat_idx1, int nodes1, const uint8_t* desc2, int n2,
                              const uint8_t* valid2, const float* angle2, const int32_t* node_ids2, const int32_t* node_begin2,
                              const int32_t* feat_idx2, int nodes2, float nnratio, int check_orientation, int32_t* match12) {
    ORB_SLAM3::KeyFrame KF1, KF2;
    std::vector<ORB_SLAM3::MapPoint> p1(n1), p2(n2);
    auto fill = [](ORB_SLAM3::KeyFrame& KF, std::vector<ORB_SLAM3::MapPoint>& pts, const uint8_t* desc, int n, const uint8_t* valid,
                   const float* angle, const int32_t* ids, const int32_t* begin, const int32_t* feat, int nodes) {
        KF.mvpMapPoints.assign(n, nullptr);
        KF.mvKeysUn.resize(n);
        for (int i = 0; i < n; ++i) {
            pts[i].index = i;
            pts[i].mbBad = valid[i] == 2;
            if (valid[i]) KF.mvpMapPoints[i] = &pts[i];
            KF.mvKeysUn[i].angle = angle[i];
        }
        KF.mDescriptors = descriptors(desc, n);
        for (int j = 0; j < nodes; ++j)
            for (int p = begin[j]; p < begin[j + 1]; ++p) KF.mFeatVec.addFeature((DBoW2::NodeId)ids[j], (unsigned)feat[p]);
    };
    fill(KF1, p1, desc1, n1, valid1, angle1, node_ids1, node_begin1, feat_idx1, nodes1);
    fill(KF2, p2, desc2, n2, valid2, angle2, node_ids2, node_begin2, feat_idx2, nodes2);
    ORBmatcher matcher(nnratio, check_orientation != 0);
    std::vector<ORB_SLAM3::MapPoint*> out;
    const int n = matcher.SearchByBoW(&KF1, &KF2, out);
    for (int i = 0; i < n1; ++i) match12[i] = out[i] ? out[i]->index : -1;
    return n;
}

// ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) for Nleft == -1 frames (monocular / rectified stereo / RGB-D).
// last_flags[i]: bit 0 = LastFrame.mvpMapPoints[i] != NULL and !mvbOutlier[i], bit 1 = that map point's Observations() > 0;
// world: n_last x 3; mp_desc: the map points' descriptors; pose = (qx, qy, qz, qw, tx, ty, tz); cur_taken[i2] != 0: the current
// frame's feature already holds a map point with Observations() > 0; cur_u_right may be NULL (all -1).  match_cur[i2] = index of the
// last-frame feature whose map point CurrentFrame.mvpMapPoints[i2] ends up holding, or -1.  Returns nmatches.
int refslice_search_by_projection_ff(const void* kps_last, const uint8_t* last_flags, const float* world, const uint8_t* mp_desc, int n_last,
                                     const void* kps_cur, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken,
                                     int n_cur, const float* pose_cw, const float* pose_lw, const float* cam, float min_x, float max_x,
                                     float min_y, float max_y, const float* scale_factors, int nlevels, float mbf, float mb, float th,
                                     int b_mono, float nnratio, int check_orientation, int32_t* match_cur) {
    Frame Cur, Last;
    auto pose = [](const float* p) {
        Sophus::SE3f T;
        T.qx = p[0]; T.qy = p[1]; T.qz = p[2]; T.qw = p[3]; T.t[0] = p[4]; T.t[1] = p[5]; T.t[2] = p[6];
        return T;
    };
    ORB_SLAM3::Pinhole camera;
    camera.mvParameters = {cam[0], cam[1], cam[2], cam[3]};
    Cur.N = n_cur; Cur.Nleft = -1;
    Cur.mvKeysUn = keypoints(kps_cur, n_cur);
    Cur.mvKeys = Cur.mvKeysUn;
    Cur.mDescriptors = descriptors(desc_cur, n_cur);
    Cur.mvuRight.assign(n_cur, -1.f);
    if (cur_u_right) Cur.mvuRight.assign(cur_u_right, cur_u_right + n_cur);
    Cur.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    Cur.mbf = mbf; Cur.mb = mb;
    Cur.mTcw = pose(pose_cw);
    Cur.mpCamera = &camera;
    Frame::mnMinX = min_x; Frame::mnMaxX = max_x; Frame::mnMinY = min_y; Frame::mnMaxY = max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);   // Frame.cc:251
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);  // Frame.cc:252
    Cur.AssignFeaturesToGrid();
    std::vector<ORB_SLAM3::MapPoint> takenPts(n_cur);
    Cur.mvpMapPoints.assign(n_cur, nullptr);
    for (int i = 0; i < n_cur; ++i)
        if (cur_taken && cur_taken[i]) {
            takenPts[i].nObs = 1; takenPts[i].index = -2;
            Cur.mvpMapPoints[i] = &takenPts[i];
        }
    Last.N = n_last; Last.Nleft = -1;
    Last.mvKeysUn = keypoints(kps_last, n_last);
    Last.mvKeys = Last.mvKeysUn;
    Last.mTcw = pose(pose_lw);
    std::vector<ORB_SLAM3::MapPoint> pts(n_last);
    Last.mvpMapPoints.assign(n_last, nullptr);
    Last.mvbOutlier.assign(n_last, false);
    for (int i = 0; i < n_last; ++i) {
        pts[i].index = i;
        pts[i].nObs = (last_flags[i] & 2) ? 3 : 0;
        pts[i].mWorldPos = Eigen::Vector3f(world[3 * i], world[3 * i + 1], world[3 * i + 2]);
        pts[i].mDescriptor = descriptors(mp_desc + (size_t)i * 32, 1);
        if (last_flags[i] & 1) Last.mvpMapPoints[i] = &pts[i];
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    const int n = matcher.SearchByProjection(Cur, Last, th, b_mono != 0);
    for (int i = 0; i < n_cur; ++i) match_cur[i] = (Cur.mvpMapPoints[i] && Cur.mvpMapPoints[i]->index >= 0) ? Cur.mvpMapPoints[i]->index : -1;
    return n;
}

// ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) for an Nleft == -1 frame (Tracking::SearchLocalPoints).
// Per map point: track = (mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos, mTrackDepth), level = mnTrackScaleLevel,
// flags bit 0 = mbTrackInView, bit 1 = Observations() > 0, bit 2 = isBad().  cur_taken / cur_u_right as in
// refslice_search_by_projection_ff.  match_cur[i] = index of the map point F.mvpMapPoints[i] ends up holding, or -1.
int refslice_search_by_projection_map(const float* track, const int32_t* level, const uint8_t* flags, const uint8_t* mp_desc, int n_mp,
                                      const void* kps_cur, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken,
                                      int n_cur, float min_x, float max_x, float min_y, float max_y, const float* scale_factors, int nlevels,
                                      float th, int far_points, float th_far_points, float nnratio, int32_t* match_cur) {
    Frame F;
    F.N = n_cur; F.Nleft = -1;
    F.mvKeysUn = keypoints(kps_cur, n_cur);
    F.mvKeys = F.mvKeysUn;
    F.mDescriptors = descriptors(desc_cur, n_cur);
    F.mvuRight.assign(n_cur, -1.f);
    if (cur_u_right) F.mvuRight.assign(cur_u_right, cur_u_right + n_cur);
    F.mvScaleFactors.assign(scale_factors, scale_factors + nlevels);
    Frame::mnMinX = min_x; Frame::mnMaxX = max_x; Frame::mnMinY = min_y; Frame::mnMaxY = max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);   // Frame.cc:251
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);  // Frame.cc:252
    F.AssignFeaturesToGrid();
    std::vector<ORB_SLAM3::MapPoint> takenPts(n_cur), pts(n_mp);
    F.mvpMapPoints.assign(n_cur, nullptr);
    for (int i = 0; i < n_cur; ++i)
        if (cur_taken && cur_taken[i]) {
            takenPts[i].nObs = 1; takenPts[i].index = -2;
            F.mvpMapPoints[i] = &takenPts[i];
        }
    std::vector<ORB_SLAM3::MapPoint*> vp(n_mp);
    for (int i = 0; i < n_mp; ++i) {
        ORB_SLAM3::MapPoint& p = pts[i];
        p.index = i;
        p.mTrackProjX = track[5 * i]; p.mTrackProjY = track[5 * i + 1]; p.mTrackProjXR = track[5 * i + 2];
        p.mTrackViewCos = track[5 * i + 3]; p.mTrackDepth = track[5 * i + 4];
        p.mnTrackScaleLevel = level[i];
        p.mbTrackInView = (flags[i] & 1) != 0; p.mbTrackInViewR = false;
        p.nObs = (flags[i] & 2) ? 3 : 0;
        p.mbBad = (flags[i] & 4) != 0;
        p.mDescriptor = descriptors(mp_desc + (size_t)i * 32, 1);
        vp[i] = &p;
    }
    ORBmatcher matcher(nnratio, true);
    const int n = matcher.SearchByProjection(F, vp, th, far_points != 0, th_far_points);
    for (int i = 0; i < n_cur; ++i) match_cur[i] = (F.mvpMapPoints[i] && F.mvpMapPoints[i]->index >= 0) ? F.mvpMapPoints[i]->index : -1;
    return n;
}

}  // extern "C"
