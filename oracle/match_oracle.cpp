// match_oracle.cpp — CPU restatement (TEST INFRASTRUCTURE, never on the product path) of the matcher-side rows that
// sit next to the extractor: the accept rules of ORBmatcher, the guided search built on Frame::GetFeaturesInArea, and
// the whole of Frame::ComputeStereoMatches.  The reference's ORBmatcher.cc / Frame.cc cannot be compiled here (they
// pull Eigen, Sophus, DBoW2, g2o, KeyFrame, MapPoint), so these functions follow the cited lines statement by
// statement.  PARITY UNPINNED for these rows: the reference holds no tests or golden vectors for them.
// Citations are file:line under /root/reference/ORB_SLAM3/.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <utility>
#include <vector>

extern "C" int orbo_hamming(const uint8_t* a, const uint8_t* b);

namespace {
struct Kp {  // cv::KeyPoint, 28 bytes
    float x, y, size, angle, response;
    int32_t octave, class_id;
};
struct GridQuery {
    float u, v, r;
    int32_t minLevel, maxLevel;
    float ur;
};
constexpr int TH_HIGH = 100, TH_LOW = 50;         // src/ORBmatcher.cc:35-36
constexpr int GRID_ROWS = 48, GRID_COLS = 64;     // include/Frame.h:44-45
}  // namespace

extern "C" {

// rules: see include/orbx.h (orbx_accept_rule); each line is the reference's own condition
void orbo_accept_top2(int rule, float nnratio, const int32_t* d1, const int32_t* d2, const int64_t* idx2,
                      const int32_t* level1, const int32_t* level2, int n, uint8_t* accept) {
    for (int i = 0; i < n; ++i) {
        const int bestDist = d1[i], bestDist2 = d2[i];
        bool ok = false;
        switch (rule) {
            case 0: {  // src/ORBmatcher.cc:123-128
                const int bestLevel = level1 ? level1[i] : 0, bestLevel2 = level2 ? level2[i] : 0;
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) break;
                    if (bestLevel != bestLevel2 || bestDist <= nnratio * bestDist2) ok = true;
                }
                break;
            }
            case 1:  // :327-329
                if (bestDist <= TH_LOW) ok = static_cast<float>(bestDist) < nnratio * static_cast<float>(bestDist2);
                break;
            case 2:  // :702-705
                if (bestDist <= TH_LOW) ok = bestDist < (float)bestDist2 * nnratio;
                break;
            case 3:  // :848-850
                if (bestDist < TH_LOW) ok = static_cast<float>(bestDist) < nnratio * static_cast<float>(bestDist2);
                break;
            case 4: ok = bestDist <= TH_HIGH; break;  // :1770, :1569, :1649
            case 5: ok = bestDist <= TH_LOW; break;   // :1312, :1437
            case 6: {                                 // src/Frame.cc:1151 (DMatch::distance is float, 0.7 a double)
                const float f1 = (float)bestDist, f2 = (float)bestDist2;
                ok = (!idx2 || idx2[i] >= 0) && f1 < f2 * 0.7;
                break;
            }
            case 7: ok = bestDist <= TH_LOW * nnratio; break;  // :523, :636 (nnratio carries ratioHamming)
        }
        accept[i] = ok ? 1 : 0;
    }
}

// Frame::AssignFeaturesToGrid (src/Frame.cc:385-416), Frame::PosInGrid (:725-736), Frame::GetFeaturesInArea
// (:657-723) and the best/second loop of ORBmatcher::SearchByProjection (src/ORBmatcher.cc:75-117; the frame-to-frame
// and initialization variants :1740-1765, :669-695 visit the same list with the same strict-< updates).
void orbo_search_grid_top2(const void* kpsT_, const uint8_t* descT, int nt, float mnMinX, float mnMaxX, float mnMinY,
                           float mnMaxY, const uint8_t* skipT, const float* uRightT, const void* queries_,
                           const uint8_t* descQ, int nq, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                           int32_t* lvl1, int32_t* lvl2) {
    const Kp* kt = static_cast<const Kp*>(kpsT_);
    const GridQuery* qs = static_cast<const GridQuery*>(queries_);
    const float mfGridElementWidthInv = static_cast<float>(GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);  // :251
    const float mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    std::vector<std::vector<int>> grid((size_t)GRID_COLS * GRID_ROWS);
    for (int i = 0; i < nt; ++i) {
        const int posX = (int)std::round((kt[i].x - mnMinX) * mfGridElementWidthInv);
        const int posY = (int)std::round((kt[i].y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid[(size_t)posX * GRID_ROWS + posY].push_back(i);
    }
    for (int q = 0; q < nq; ++q) {
        const float x = qs[q].u, y = qs[q].v, r = qs[q].r;
        const int minLevel = qs[q].minLevel, maxLevel = qs[q].maxLevel;
        int bestDist = 256, bestDist2 = 256, bestIdx = -1, bestIdx2 = -1, bestLevel = -1, bestLevel2 = -1;
        auto done = [&]() {
            idx1[q] = bestIdx; d1[q] = bestDist; idx2[q] = bestIdx2; d2[q] = bestDist2;
            if (lvl1) lvl1[q] = bestLevel;
            if (lvl2) lvl2[q] = bestLevel2;
        };
        const float factorX = r, factorY = r;
        const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
        if (nMinCellX >= GRID_COLS) { done(); continue; }
        const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
        if (nMaxCellX < 0) { done(); continue; }
        const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
        if (nMinCellY >= GRID_ROWS) { done(); continue; }
        const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
        if (nMaxCellY < 0) { done(); continue; }
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                for (int idx : grid[(size_t)ix * GRID_ROWS + iy]) {
                    const Kp& kpUn = kt[idx];
                    if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.x - x, disty = kpUn.y - y;
                    if (!(std::fabs(distx) < factorX && std::fabs(disty) < factorY)) continue;
                    // ---- the matcher's loop body ----
                    if (skipT && skipT[idx]) continue;  // src/ORBmatcher.cc:87-89
                    if (uRightT && uRightT[idx] > 0) {  // :91-96
                        const float er = std::fabs(qs[q].ur - uRightT[idx]);
                        if (er > r) continue;
                    }
                    const int dist = orbo_hamming(descQ + (size_t)q * 32, descT + (size_t)idx * 32);
                    if (dist < bestDist) {
                        bestDist2 = bestDist; bestIdx2 = bestIdx; bestLevel2 = bestLevel;
                        bestDist = dist; bestIdx = idx; bestLevel = kpUn.octave;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist; bestIdx2 = idx; bestLevel2 = kpUn.octave;
                    }
                }
        done();
    }
}

// Frame::ComputeStereoMatches (src/Frame.cc:811-980).  pyrL/pyrR[level] point at pixel (0,0) of mvImagePyramid[level]
// of the left/right extractor (a ROI at (19,19) of the padded plane), strideL/strideR[level] are their row pitches and
// cols[level] the level widths.  Compiled with -ffp-contract=off (STRICT).
void orbo_compute_stereo_matches(const void* kpsL_, const uint8_t* descL, int N, const void* kpsR_, const uint8_t* descR,
                                 int Nr, const float* mvScaleFactors, const float* mvInvScaleFactors, int nlevels,
                                 const uint8_t* const* pyrL, const int* strideL, const uint8_t* const* pyrR,
                                 const int* strideR, const int* cols, int nRows, float mbf, float mb, float* mvuRight,
                                 float* mvDepth) {
    (void)nlevels;
    const Kp* mvKeys = static_cast<const Kp*>(kpsL_);
    const Kp* mvKeysRight = static_cast<const Kp*>(kpsR_);
    for (int i = 0; i < N; ++i) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; }  // :813-814
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    std::vector<std::vector<size_t>> vRowIndices(nRows, std::vector<size_t>());
    for (int iR = 0; iR < Nr; iR++) {
        const Kp& kp = mvKeysRight[iR];
        const float& kpY = kp.y;
        const float r = 2.0f * mvScaleFactors[mvKeysRight[iR].octave];
        const int maxr = ceil(kpY + r);
        const int minr = floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // (the reference indexes unchecked)
    }
    const float minZ = mb;
    const float minD = 0;
    const float maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {
        const Kp& kpL = mvKeys[iL];
        const int& levelL = kpL.octave;
        const float& vL = kpL.y;
        const float& uL = kpL.x;
        if (!(vL >= 0 && vL < nRows)) continue;
        const std::vector<size_t>& vCandidates = vRowIndices[(size_t)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD;
        const float maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const size_t iR = vCandidates[iC];
            const Kp& kpR = mvKeysRight[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float& uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = orbo_hamming(descL + (size_t)iL * 32, descR + iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {  // :896
            const float uR0 = mvKeysRight[bestIdxR].x;
            const float scaleFactor = mvInvScaleFactors[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            const uint8_t* IL = pyrL[kpL.octave] + (ptrdiff_t)(int)(scaledvL - w) * strideL[kpL.octave] + (int)(scaleduL - w);
            int bestDist = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            std::vector<float> vDists;
            vDists.resize(2 * L + 1);
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= cols[kpL.octave]) continue;
            for (int incR = -L; incR <= +L; incR++) {
                const uint8_t* IR = pyrR[kpL.octave] + (ptrdiff_t)(int)(scaledvL - w) * strideR[kpL.octave] +
                                    (int)(scaleduR0 + incR - w);
                int l1 = 0;  // cv::norm(IL, IR, cv::NORM_L1) over the 11 x 11 window
                for (int yy = 0; yy < 2 * w + 1; ++yy)
                    for (int xx = 0; xx < 2 * w + 1; ++xx)
                        l1 += std::abs((int)IL[(ptrdiff_t)yy * strideL[kpL.octave] + xx] - (int)IR[(ptrdiff_t)yy * strideR[kpL.octave] + xx]);
                float dist = (float)l1;
                if (dist < bestDist) { bestDist = dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = mvScaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestDist, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;  // (the reference would index an empty vector here)
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        mvuRight[vDistIdx[i].second] = -1;
        mvDepth[vDistIdx[i].second] = -1;
    }
}


// ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) (ORBmatcher.cc:223-425) for a monocular keyframe / frame
// (F.Nleft == -1, mpCamera2 == NULL): features are only compared inside the same vocabulary node (the two FeatureVectors,
// walked in ascending node id, :243-246, :388-400); keyframe features are visited in their node's order, frame features that
// already received a MapPoint are skipped (:279), strict-< best / second (:287-296), accept best1 <= TH_LOW and
// (float)best1 < mfNNratio * (float)best2 (:323-327), then the rotation histogram with factor 1/HISTO_LENGTH (:234, :345-352)
// and ComputeThreeMaxima (:2012-2053) remove the matches outside the three dominant bins (:403-420).
// valid_kf: 0 = no MapPoint, 1 = good MapPoint, 2 = MapPoint isBad().  match_f[j] = keyframe feature index or -1.
// Pinned by oracle/_ref/librefslices.so (the reference's own body), tests/test_oracle_refslices.py.
int orbo_search_by_bow(const uint8_t* desc_kf, int n_kf, const uint8_t* valid_kf, const float* angle_kf, const int32_t* kf_node_ids,
                       const int32_t* kf_node_begin, const int32_t* kf_feat_idx, int kf_nodes, const uint8_t* desc_f, int n_f,
                       const float* angle_f, const int32_t* f_node_ids, const int32_t* f_node_begin, const int32_t* f_feat_idx,
                       int f_nodes, float mfNNratio, int mbCheckOrientation, int32_t* match_f) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    (void)n_kf;
    for (int j = 0; j < n_f; ++j) match_f[j] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[30];
    const float factor = 1.0f / HISTO_LENGTH;
    int a = 0, b = 0;
    while (a < kf_nodes && b < f_nodes) {
        if (kf_node_ids[a] == f_node_ids[b]) {
            for (int iKF = kf_node_begin[a]; iKF < kf_node_begin[a + 1]; iKF++) {
                const int realIdxKF = kf_feat_idx[iKF];
                if (valid_kf[realIdxKF] != 1) continue;  // !pMP, or pMP->isBad()
                const uint8_t* dKF = desc_kf + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (int iF = f_node_begin[b]; iF < f_node_begin[b + 1]; iF++) {
                    const int realIdxF = f_feat_idx[iF];
                    if (match_f[realIdxF] >= 0) continue;
                    const int dist = orbo_hamming(dKF, desc_f + (size_t)realIdxF * 32);
                    if (dist < bestDist1) {
                        bestDist2 = bestDist1;
                        bestDist1 = dist;
                        bestIdxF = realIdxF;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist;
                    }
                }
                if (bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        match_f[bestIdxF] = realIdxKF;
                        if (mbCheckOrientation) {
                            float rot = angle_kf[realIdxKF] - angle_f[bestIdxF];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                }
            }
            a++;
            b++;
        } else if (kf_node_ids[a] < f_node_ids[b]) {
            while (a < kf_nodes && kf_node_ids[a] < f_node_ids[b]) a++;  // lower_bound
        } else {
            while (b < f_nodes && f_node_ids[b] < kf_node_ids[a]) b++;
        }
    }
    if (mbCheckOrientation) {
        // ComputeThreeMaxima (ORBmatcher.cc:2012-2053)
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) {
                max3 = max2; max2 = max1; max1 = s;
                ind3 = ind2; ind2 = ind1; ind1 = i;
            } else if (s > max2) {
                max3 = max2; max2 = s;
                ind3 = ind2; ind2 = i;
            } else if (s > max3) {
                max3 = s;
                ind3 = i;
            }
        }
        if (max2 < 0.1f * (float)max1) {
            ind2 = -1;
            ind3 = -1;
        } else if (max3 < 0.1f * (float)max1) {
            ind3 = -1;
        }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                match_f[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// ---- Frame::UndistortKeyPoints (src/Frame.cc:747-780) and Frame::ComputeImageBounds (:782-809), pinhole camera ------------
// cv::undistortPoints is OpenCV's (un-vendored): minicv's restatement, pinned to cv2 4.13.0 by tests/golden/prim_undistort.npz.
// K = Pinhole::toK(), P = mK: 3 x 3 float row-major; dist = mDistCoef (ndist floats).  PINNED to the reference's own bodies
// (oracle/_ref/librefslices.so) by tests/test_oracle_refslices.py.
void minicv_undistort_points(const float* pts, int n, const float* K, const float* dist, int ndist, const float* P, float* out);

void orbo_undistort_keypoints(const void* kps_in, int n, const float* K, const float* dist, int ndist, const float* P, void* kps_out) {
    const Kp* in = (const Kp*)kps_in;
    Kp* out = (Kp*)kps_out;
    if (dist[0] == 0.0) {  // :749-753 mvKeysUn = mvKeys
        if (n) std::memcpy(out, in, (size_t)n * sizeof(Kp));
        return;
    }
    std::vector<float> mat((size_t)n * 2), un((size_t)n * 2);
    for (int i = 0; i < n; i++) {  // :758-762
        mat[2 * i] = in[i].x;
        mat[2 * i + 1] = in[i].y;
    }
    minicv_undistort_points(mat.data(), n, K, dist, ndist, P, un.data());  // :765-767
    for (int i = 0; i < n; i++) {  // :771-778 every other field is carried over
        Kp kp = in[i];
        kp.x = un[2 * i];
        kp.y = un[2 * i + 1];
        out[i] = kp;
    }
}

// bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)
void orbo_image_bounds(int cols, int rows, const float* K, const float* dist, int ndist, const float* P, float* bounds) {
    if (dist[0] != 0.0) {
        float mat[8] = {0.0f, 0.0f, (float)cols, 0.0f, 0.0f, (float)rows, (float)cols, (float)rows}, un[8];  // :786-790
        minicv_undistort_points(mat, 4, K, dist, ndist, P, un);  // :793-796
        bounds[0] = std::min(un[0], un[4]);  // :799 min(mat(0,0), mat(2,0))
        bounds[1] = std::max(un[2], un[6]);  // :800 max(mat(1,0), mat(3,0))
        bounds[2] = std::min(un[1], un[3]);  // :801 min(mat(0,1), mat(1,1))
        bounds[3] = std::max(un[5], un[7]);  // :802 max(mat(2,1), mat(3,1))
    } else {  // :804-809
        bounds[0] = 0.0f;
        bounds[1] = (float)cols;
        bounds[2] = 0.0f;
        bounds[3] = (float)rows;
    }
}


// ---- ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:648-763), monocular frames --------------------------------
// F2's grid = Frame::AssignFeaturesToGrid over kps2 (mvKeysUn) inside (mnMinX, mnMaxX, mnMinY, mnMaxY); the window query is
// Frame::GetFeaturesInArea(x, y, windowSize, level1, level1) (Frame.cc:657-723).  prev_matched: n1 x 2 floats, in / out
// (:756-759).  matches12: n1 ints (vnMatches12).  Returns nmatches.  PINNED to the reference's own body by
// tests/test_oracle_refslices.py.
static int g_init_withdrawn = 0;  // how often :705-709 fired in the last call (lets the tests prove the branch is exercised)
int orbo_search_for_initialization_withdrawn() { return g_init_withdrawn; }
int orbo_search_for_initialization(const void* kps1_, const uint8_t* desc1, int n1, const void* kps2_, const uint8_t* desc2, int n2,
                                   float mnMinX, float mnMaxX, float mnMinY, float mnMaxY, float* prev_matched, int windowSize,
                                   float mfNNratio, int mbCheckOrientation, int32_t* vnMatches12) {
    const Kp* k1 = static_cast<const Kp*>(kps1_);
    const Kp* k2 = static_cast<const Kp*>(kps2_);
    constexpr int HISTO_LENGTH = 30;  // :37
    const float mfGridElementWidthInv = static_cast<float>(GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);  // Frame.cc:251
    const float mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    std::vector<std::vector<int>> grid((size_t)GRID_COLS * GRID_ROWS);
    for (int i = 0; i < n2; ++i) {  // Frame.cc:385-416, :725-735
        const int posX = (int)std::round((k2[i].x - mnMinX) * mfGridElementWidthInv);
        const int posY = (int)std::round((k2[i].y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid[(size_t)posX * GRID_ROWS + posY].push_back(i);
    }
    int nmatches = 0;
    g_init_withdrawn = 0;
    for (int i = 0; i < n1; ++i) vnMatches12[i] = -1;  // :651
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;  // :656
    std::vector<int> vMatchedDistance(n2, INT_MAX), vnMatches21(n2, -1);
    for (int i1 = 0; i1 < n1; i1++) {
        const Kp kp1 = k1[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;  // :665-666
        // GetFeaturesInArea(vbPrevMatched[i1].x, .y, windowSize, level1, level1)
        const float x = prev_matched[2 * i1], y = prev_matched[2 * i1 + 1], r = (float)windowSize;
        const int minLevel = level1, maxLevel = level1;
        std::vector<int> vIndices2;
        {
            const float factorX = r, factorY = r;
            const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
            const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                        for (int idx : grid[(size_t)ix * GRID_ROWS + iy]) {
                            const Kp& kpUn = k2[idx];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - x, disty = kpUn.y - y;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices2.push_back(idx);
                        }
            }
        }
        if (vIndices2.empty()) continue;  // :670-671
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = orbo_hamming(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;  // :687-688
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestIdx2 = i2;
            } else if (dist < bestDist2) {
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) {  // :705-709 an earlier, worse match onto the same point is withdrawn
                    vnMatches12[vnMatches21[bestIdx2]] = -1;
                    nmatches--;
                    g_init_withdrawn++;
                }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (mbCheckOrientation) {
                    float rot = k1[i1].angle - k2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (mbCheckOrientation) {
        // ComputeThreeMaxima (ORBmatcher.cc:2012-2053)
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) {
                max3 = max2; max2 = max1; max1 = s;
                ind3 = ind2; ind2 = ind1; ind1 = i;
            } else if (s > max2) {
                max3 = max2; max2 = s;
                ind3 = ind2; ind2 = i;
            } else if (s > max3) {
                max3 = s;
                ind3 = i;
            }
        }
        if (max2 < 0.1f * (float)max1) {
            ind2 = -1;
            ind3 = -1;
        } else if (max3 < 0.1f * (float)max1) {
            ind3 = -1;
        }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                const int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) {  // :745-749 (it may already have been withdrawn)
                    vnMatches12[idx1] = -1;
                    nmatches--;
                }
            }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)  // :756-759
        if (vnMatches12[i1] >= 0) {
            prev_matched[2 * i1] = k2[vnMatches12[i1]].x;
            prev_matched[2 * i1 + 1] = k2[vnMatches12[i1]].y;
        }
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) (src/ORBmatcher.cc:1676-1887) ----
// TrackWithMotionModel's matcher (Tracking.cc:2876, :2906, :2914), frames with Nleft == -1 (monocular, rectified stereo, RGB-D).
// Restated statement by statement; the Sophus / Eigen operations are written out in the order the vendored headers write them
// (Thirdparty/Sophus/sophus/so3.hpp:358-368, se3.hpp:321-325 and inverse(); Eigen's cross product), every operation rounded once
// (this file is compiled with -ffp-contract=off): the declared STRICT arithmetic, the same as oracle/refslice/shim.h.
// Arguments as refslice_search_by_projection_ff (oracle/refslice/refslice_wrap.cpp).  PINNED to the reference's own body by
// tests/test_oracle_refslices.py.
namespace {
struct V3 { float x, y, z; };
inline V3 crossq(float ax, float ay, float az, V3 b) { return {ay * b.z - az * b.y, az * b.x - ax * b.z, ax * b.y - ay * b.x}; }
inline V3 rotq(const float* q, V3 p) {  // so3.hpp:366-368
    V3 uv = crossq(q[0], q[1], q[2], p);
    uv.x += uv.x; uv.y += uv.y; uv.z += uv.z;
    const V3 c = crossq(q[0], q[1], q[2], uv);
    return {(p.x + q[3] * uv.x) + c.x, (p.y + q[3] * uv.y) + c.y, (p.z + q[3] * uv.z) + c.z};
}
inline V3 se3mul(const float* T, V3 p) {  // se3.hpp:324
    const V3 r = rotq(T, p);
    return {r.x + T[4], r.y + T[5], r.z + T[6]};
}
}  // namespace

int orbo_search_by_projection_ff(const void* kps_last_, const uint8_t* last_flags, const float* world, const uint8_t* mp_desc, int n_last,
                                 const void* kps_cur_, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken, int n_cur,
                                 const float* pose_cw, const float* pose_lw, const float* cam, float mnMinX, float mnMaxX, float mnMinY,
                                 float mnMaxY, const float* mvScaleFactors, int nlevels, float mbf, float mb, float th, int bMono,
                                 float mfNNratio, int mbCheckOrientation, int32_t* match_cur) {
    (void)mfNNratio; (void)nlevels;
    const Kp* kl = static_cast<const Kp*>(kps_last_);
    const Kp* kc = static_cast<const Kp*>(kps_cur_);
    constexpr int HISTO_LENGTH = 30;
    const float mfGridElementWidthInv = static_cast<float>(GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);  // Frame.cc:251
    const float mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    std::vector<std::vector<int>> grid((size_t)GRID_COLS * GRID_ROWS);
    for (int i = 0; i < n_cur; ++i) {  // Frame.cc:385-416, :725-735
        const int posX = (int)std::round((kc[i].x - mnMinX) * mfGridElementWidthInv);
        const int posY = (int)std::round((kc[i].y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid[(size_t)posX * GRID_ROWS + posY].push_back(i);
    }
    // CurrentFrame.mvpMapPoints: -1 = NULL, -2 = a map point of the caller with Observations() > 0, i >= 0 = LastFrame's map point i
    for (int i = 0; i < n_cur; ++i) match_cur[i] = (cur_taken && cur_taken[i]) ? -2 : -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    // :1686-1693
    float invq[7] = {-pose_cw[0], -pose_cw[1], -pose_cw[2], pose_cw[3], 0, 0, 0};
    const V3 twc = rotq(invq, {pose_cw[4] * -1.f, pose_cw[5] * -1.f, pose_cw[6] * -1.f});  // Tcw.inverse().translation()
    const V3 tlc = se3mul(pose_lw, twc);
    const bool bForward = tlc.z > mb && !bMono;
    const bool bBackward = -tlc.z > mb && !bMono;
    for (int i = 0; i < n_last; i++) {
        if (!(last_flags[i] & 1)) continue;  // pMP && !mvbOutlier[i]
        const V3 x3Dc = se3mul(pose_cw, {world[3 * i], world[3 * i + 1], world[3 * i + 2]});
        const float invzc = 1.0 / x3Dc.z;  // double division, rounded to float (:1709)
        if (invzc < 0) continue;
        const float u = cam[0] * x3Dc.x / x3Dc.z + cam[2], v = cam[1] * x3Dc.y / x3Dc.z + cam[3];  // Pinhole.cpp:45-46
        if (u < mnMinX || u > mnMaxX) continue;
        if (v < mnMinY || v > mnMaxY) continue;
        const int nLastOctave = kl[i].octave;
        const float radius = th * mvScaleFactors[nLastOctave];
        int minLevel, maxLevel;
        if (bForward) { minLevel = nLastOctave; maxLevel = -1; }
        else if (bBackward) { minLevel = 0; maxLevel = nLastOctave; }
        else { minLevel = nLastOctave - 1; maxLevel = nLastOctave + 1; }
        std::vector<int> vIndices2;
        {
            const float x = u, y = v, factorX = radius, factorY = radius;
            const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
            const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                        for (int idx : grid[(size_t)ix * GRID_ROWS + iy]) {
                            const Kp& kpUn = kc[idx];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - x, disty = kpUn.y - y;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices2.push_back(idx);
                        }
            }
        }
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            // :1745-1747 a feature that already holds a map point with observations is skipped
            if (match_cur[i2] == -2 || (match_cur[i2] >= 0 && (last_flags[match_cur[i2]] & 2))) continue;
            if (cur_u_right && cur_u_right[i2] > 0) {  // :1749-1755 (Nleft == -1)
                const float ur = u - mbf * invzc;
                const float er = std::fabs(ur - cur_u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = orbo_hamming(mp_desc + (size_t)i * 32, desc_cur + (size_t)i2 * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx2 = i2;
            }
        }
        if (bestDist <= TH_HIGH) {
            match_cur[bestIdx2] = i;
            nmatches++;
            if (mbCheckOrientation) {
                float rot = kl[i].angle - kc[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (mbCheckOrientation) {
        // ComputeThreeMaxima (ORBmatcher.cc:2012-2053)
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) {
                max3 = max2; max2 = max1; max1 = s;
                ind3 = ind2; ind2 = ind1; ind1 = i;
            } else if (s > max2) {
                max3 = max2; max2 = s;
                ind3 = ind2; ind2 = i;
            } else if (s > max3) {
                max3 = s;
                ind3 = i;
            }
        }
        if (max2 < 0.1f * (float)max1) {
            ind2 = -1;
            ind3 = -1;
        } else if (max3 < 0.1f * (float)max1) {
            ind3 = -1;
        }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {  // :1877-1881: unconditional, entry by entry
                match_cur[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    for (int i = 0; i < n_cur; ++i)
        if (match_cur[i] == -2) match_cur[i] = -1;  // the caller's own map points are not reported
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, th, bFarPoints, thFarPoints) -------------------
// (src/ORBmatcher.cc:43-210 + RadiusByViewingCos :215-221), Tracking::SearchLocalPoints' matcher (Tracking.cc:3410, :3433), frames with
// Nleft == -1.  Arguments as refslice_search_by_projection_map.  PINNED to the reference's own body by tests/test_oracle_refslices.py.
int orbo_search_by_projection_map(const float* track, const int32_t* level, const uint8_t* flags, const uint8_t* mp_desc, int n_mp,
                                  const void* kps_cur_, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken, int n_cur,
                                  float mnMinX, float mnMaxX, float mnMinY, float mnMaxY, const float* mvScaleFactors, int nlevels, float th,
                                  int bFarPoints, float thFarPoints, float mfNNratio, int32_t* match_cur) {
    (void)nlevels;
    const Kp* kc = static_cast<const Kp*>(kps_cur_);
    const float mfGridElementWidthInv = static_cast<float>(GRID_COLS) / static_cast<float>(mnMaxX - mnMinX);  // Frame.cc:251
    const float mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / static_cast<float>(mnMaxY - mnMinY);
    std::vector<std::vector<int>> grid((size_t)GRID_COLS * GRID_ROWS);
    for (int i = 0; i < n_cur; ++i) {  // Frame.cc:385-416, :725-735
        const int posX = (int)std::round((kc[i].x - mnMinX) * mfGridElementWidthInv);
        const int posY = (int)std::round((kc[i].y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        grid[(size_t)posX * GRID_ROWS + posY].push_back(i);
    }
    // F.mvpMapPoints: -1 = NULL, -2 = a map point of the caller with Observations() > 0, i >= 0 = vpMapPoints[i]
    for (int i = 0; i < n_cur; ++i) match_cur[i] = (cur_taken && cur_taken[i]) ? -2 : -1;
    int nmatches = 0;
    const bool bFactor = th != 1.0;  // :47
    for (int iMP = 0; iMP < n_mp; iMP++) {
        if (!(flags[iMP] & 1)) continue;                                  // :52-53 (mbTrackInViewR is false for Nleft == -1 frames)
        if (bFarPoints && track[5 * iMP + 4] > thFarPoints) continue;     // :55-56
        if (flags[iMP] & 4) continue;                                     // :58-59 isBad()
        const int nPredictedLevel = level[iMP];
        float r = track[5 * iMP + 3] > 0.998 ? 2.5f : 4.0f;               // RadiusByViewingCos: a float compared with a double literal
        if (bFactor) r *= th;
        const float x = track[5 * iMP], y = track[5 * iMP + 1];
        const float factorX = r * mvScaleFactors[nPredictedLevel], factorY = factorX;
        const int minLevel = nPredictedLevel - 1, maxLevel = nPredictedLevel;
        std::vector<int> vIndices;
        {
            const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
            const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
                        for (int idx : grid[(size_t)ix * GRID_ROWS + iy]) {
                            const Kp& kpUn = kc[idx];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - x, disty = kpUn.y - y;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices.push_back(idx);
                        }
            }
        }
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (match_cur[idx] == -2 || (match_cur[idx] >= 0 && (flags[match_cur[idx]] & 2))) continue;  // :84-86
            if (cur_u_right && cur_u_right[idx] > 0) {                                                     // :88-93
                const float er = std::fabs(track[5 * iMP + 2] - cur_u_right[idx]);
                if (er > r * mvScaleFactors[nPredictedLevel]) continue;
            }
            const int dist = orbo_hamming(mp_desc + (size_t)iMP * 32, desc_cur + (size_t)idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestLevel2 = bestLevel;
                bestLevel = kc[idx].octave;
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = kc[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {  // :123-128
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
                match_cur[bestIdx] = iMP;
                nmatches++;
            }
        }
    }
    for (int i = 0; i < n_cur; ++i)
        if (match_cur[i] == -2) match_cur[i] = -1;
    return nmatches;
}

// ---- ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) (src/ORBmatcher.cc:765-905), monocular keyframes ----
// The keyframe-to-keyframe twin of orbo_search_by_bow (loop closing / merging): candidates of keyframe 2 need a good MapPoint
// and must not be matched yet (vbMatched2, :826-830), the accept rule is STRICT in TH_LOW (bestDist1 < TH_LOW, :848), the
// result is indexed by keyframe 1's features.  valid: 0 = no MapPoint, 1 = good, 2 = isBad().  match12[i] = feature of
// keyframe 2 or -1.  PINNED to the reference's own body by tests/test_oracle_refslices.py.
int orbo_search_by_bow_kf(const uint8_t* desc1, int n1, const uint8_t* valid1, const float* angle1, const int32_t* node_ids1,
                          const int32_t* node_begin1, const int32_t* feat_idx1, int nodes1, const uint8_t* desc2, int n2,
                          const uint8_t* valid2, const float* angle2, const int32_t* node_ids2, const int32_t* node_begin2,
                          const int32_t* feat_idx2, int nodes2, float mfNNratio, int mbCheckOrientation, int32_t* match12) {
    const int HISTO_LENGTH = 30;
    for (int i = 0; i < n1; ++i) match12[i] = -1;  // :777
    std::vector<char> vbMatched2((size_t)std::max(n2, 1), 0);
    std::vector<int> rotHist[30];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    int a = 0, b = 0;
    while (a < nodes1 && b < nodes2) {
        if (node_ids1[a] == node_ids2[b]) {
            for (int i1 = node_begin1[a]; i1 < node_begin1[a + 1]; i1++) {
                const int idx1 = feat_idx1[i1];
                if (valid1[idx1] != 1) continue;  // :803-807 !pMP1 or isBad()
                const uint8_t* d1 = desc1 + (size_t)idx1 * 32;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int i2 = node_begin2[b]; i2 < node_begin2[b + 1]; i2++) {
                    const int idx2 = feat_idx2[i2];
                    if (vbMatched2[idx2] || valid2[idx2] == 0) continue;  // :826-827
                    if (valid2[idx2] == 2) continue;                      // :829-830 isBad()
                    const int dist = orbo_hamming(d1, desc2 + (size_t)idx2 * 32);
                    if (dist < bestDist1) {
                        bestDist2 = bestDist1;
                        bestDist1 = dist;
                        bestIdx2 = idx2;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist;
                    }
                }
                if (bestDist1 < TH_LOW) {  // :848 strict
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        match12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = 1;
                        if (mbCheckOrientation) {
                            float rot = angle1[idx1] - angle2[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(idx1);
                        }
                        nmatches++;
                    }
                }
            }
            a++;
            b++;
        } else if (node_ids1[a] < node_ids2[b]) {
            while (a < nodes1 && node_ids1[a] < node_ids2[b]) a++;  // lower_bound
        } else {
            while (b < nodes2 && node_ids2[b] < node_ids1[a]) b++;
        }
    }
    if (mbCheckOrientation) {
        // ComputeThreeMaxima (ORBmatcher.cc:2012-2053)
        int ind1 = -1, ind2 = -1, ind3 = -1, max1 = 0, max2 = 0, max3 = 0;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = (int)rotHist[i].size();
            if (s > max1) {
                max3 = max2; max2 = max1; max1 = s;
                ind3 = ind2; ind2 = ind1; ind1 = i;
            } else if (s > max2) {
                max3 = max2; max2 = s;
                ind3 = ind2; ind2 = i;
            } else if (s > max3) {
                max3 = s;
                ind3 = i;
            }
        }
        if (max2 < 0.1f * (float)max1) {
            ind2 = -1;
            ind3 = -1;
        } else if (max3 < 0.1f * (float)max1) {
            ind3 = -1;
        }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                match12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

}  // extern "C"
