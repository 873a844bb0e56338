// refslice/shim.h — TEST INFRASTRUCTURE (oracle side), not product code.
//
// Declarations that let SLICES of the reference's own src/Frame.cc and src/ORBmatcher.cc compile on their own: the
// whole translation units need Eigen, Sophus, DBoW2, g2o, KeyFrame, MapPoint ..., but the functions on
// this path only touch a handful of members.  oracle/Makefile cuts the function bodies out of the reference files
// where they lie (sed by line range, checked against the expected first line) into oracle/_ref/slices/*.inc and
// compiles them, unchanged, inside the class shells below.  Member names and types follow include/Frame.h,
// include/ORBmatcher.h and include/ORBextractor.h; nothing here restates an algorithm.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstddef>
#include <utility>
#include <vector>

#include <opencv2/core/core.hpp>

#include "DBoW2/FeatureVector.h"  // the reference's own (Thirdparty/DBoW2), on the boost::serialization stub of minicv

using namespace std;  // the reference's headers do this (ORBmatcher.cc:30, and Frame.cc relies on it through its includes)

#define FRAME_GRID_ROWS 48  // include/Frame.h:44-45
#define FRAME_GRID_COLS 64

// ---- Eigen / Sophus stand-ins for ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono) (ORBmatcher.cc:1676-1887).
// Eigen and Sophus' dependencies are not installed, so the few operations that body uses are written out here in single-precision
// operations in the order the vendored Sophus writes them (Thirdparty/Sophus/sophus/so3.hpp:358-368: uv = q.vec x p; uv += uv;
// p + w * uv + q.vec x uv;  se3.hpp:321-325: + translation;  se3.hpp inverse(): (q*, q* (-t))) with Eigen's cross product
// (a1 b2 - a2 b1, a2 b0 - a0 b2, a0 b1 - a1 b0).  Compiled with -ffp-contract=off like every slice: the declared STRICT arithmetic
// (a Release build with -march=native may contract some of these into FMAs).
namespace Eigen {
struct Vector3f {
    float v[3] = {0, 0, 0};
    Vector3f() {}
    Vector3f(float a, float b, float c) : v{a, b, c} {}
    float operator()(int i) const { return v[i]; }
    float& operator()(int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
    float& operator[](int i) { return v[i]; }
};
struct Vector2f {
    float v[2] = {0, 0};
    float operator()(int i) const { return v[i]; }
    float& operator()(int i) { return v[i]; }
    float operator[](int i) const { return v[i]; }
    float& operator[](int i) { return v[i]; }
};
}  // namespace Eigen
namespace Sophus {
struct SE3f {
    float qx = 0, qy = 0, qz = 0, qw = 1;  // unit quaternion of so3()
    float t[3] = {0, 0, 0};                // translation()
    static Eigen::Vector3f cross(float ax, float ay, float az, const Eigen::Vector3f& b) {
        return Eigen::Vector3f(ay * b.v[2] - az * b.v[1], az * b.v[0] - ax * b.v[2], ax * b.v[1] - ay * b.v[0]);
    }
    Eigen::Vector3f rotate(const Eigen::Vector3f& p) const {  // so3.hpp:366-368
        Eigen::Vector3f uv = cross(qx, qy, qz, p);
        uv.v[0] += uv.v[0]; uv.v[1] += uv.v[1]; uv.v[2] += uv.v[2];
        const Eigen::Vector3f c = cross(qx, qy, qz, uv);
        return Eigen::Vector3f((p.v[0] + qw * uv.v[0]) + c.v[0], (p.v[1] + qw * uv.v[1]) + c.v[1], (p.v[2] + qw * uv.v[2]) + c.v[2]);
    }
    Eigen::Vector3f operator*(const Eigen::Vector3f& p) const {  // se3.hpp:324
        const Eigen::Vector3f r = rotate(p);
        return Eigen::Vector3f(r.v[0] + t[0], r.v[1] + t[1], r.v[2] + t[2]);
    }
    SE3f inverse() const {  // se3.hpp: SE3(invR, invR * (translation() * Scalar(-1)))
        SE3f o;
        o.qx = -qx; o.qy = -qy; o.qz = -qz; o.qw = qw;
        const Eigen::Vector3f r = o.rotate(Eigen::Vector3f(t[0] * -1.f, t[1] * -1.f, t[2] * -1.f));
        o.t[0] = r.v[0]; o.t[1] = r.v[1]; o.t[2] = r.v[2];
        return o;
    }
    Eigen::Vector3f translation() const { return Eigen::Vector3f(t[0], t[1], t[2]); }
};
}  // namespace Sophus

namespace ORB_SLAM3 {

struct ORBextractor {  // include/ORBextractor.h:83
    std::vector<cv::Mat> mvImagePyramid;
};

class GeometricCamera {  // include/CameraModels/GeometricCamera.h: the static_cast target below and project()
public:
    virtual ~GeometricCamera() {}
    virtual Eigen::Vector2f project(const Eigen::Vector3f& v3D) { (void)v3D; return Eigen::Vector2f(); }
};
class Pinhole : public GeometricCamera {  // include/CameraModels/Pinhole.h:81 toK()
public:
    cv::Mat toK() { return K.clone(); }
    cv::Mat K;  // (shim) 3 x 3 CV_32F, what Pinhole::toK builds from mvParameters (src/CameraModels/Pinhole.cpp:134-138)
    std::vector<float> mvParameters;  // fx, fy, cx, cy
    Eigen::Vector2f project(const Eigen::Vector3f& v3D) override;  // body = src/CameraModels/Pinhole.cpp:43-49, sliced
};
class Frame;

class MapPoint {  // include/MapPoint.h: isBad(), GetWorldPos(), GetDescriptor(), Observations() are used on this path
public:
    bool isBad() { return mbBad; }
    Eigen::Vector3f GetWorldPos() { return mWorldPos; }
    cv::Mat GetDescriptor() { return mDescriptor.clone(); }
    int Observations() { return nObs; }
    bool mbBad = false;
    int index = -1;  // (shim) the keyframe feature this map point was created for
    // tracking variables set by Frame::isInFrustum (include/MapPoint.h:141-152), read by SearchByProjection(Frame&, vector<MapPoint*>&)
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 0, mTrackProjXR = 0, mTrackProjYR = 0, mTrackViewCos = 0, mTrackViewCosR = 0;
    bool mbTrackInView = false, mbTrackInViewR = false;
    int mnTrackScaleLevel = 0, mnTrackScaleLevelR = -1;
    Eigen::Vector3f mWorldPos;
    cv::Mat mDescriptor;
    int nObs = 0;
};

class KeyFrame {  // include/KeyFrame.h
public:
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mDescriptors;
    GeometricCamera* mpCamera2 = nullptr;
    std::vector<cv::KeyPoint> mvKeysUn, mvKeys, mvKeysRight;
    int NLeft = -1;
    std::vector<MapPoint*> mvpMapPoints;
};

class ORBmatcher {  // include/ORBmatcher.h:38-103
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10);
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3, const bool bFarPoints = false,
                           const float thFarPoints = 50.0f);
    float RadiusByViewingCos(const float& viewCos);
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);
    float mfNNratio;
    bool mbCheckOrientation;
};

class Frame {  // include/Frame.h
public:
    void ComputeStereoMatches();
    void AssignFeaturesToGrid();
    vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                     const int maxLevel = -1, const bool bRight = false) const;
    bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY);
    void UndistortKeyPoints();
    void ComputeImageBounds(const cv::Mat& imLeft);
    cv::Mat mK, mDistCoef;
    GeometricCamera* mpCamera = nullptr;

    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
    float mbf = 0, mb = 0;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight, mvKeysUn;
    std::vector<float> mvuRight, mvDepth;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    int Nleft = -1, Nright = -1;
    std::vector<std::size_t> mGridRight[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    DBoW2::FeatureVector mFeatVec;
    GeometricCamera* mpCamera2 = nullptr;
    // SearchByProjection(Frame&, const Frame&, ...)
    Sophus::SE3f GetPose() const { return mTcw; }
    Sophus::SE3f GetRelativePoseTrl() { return mTrl; }
    Sophus::SE3f mTcw, mTrl;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
};

}  // namespace ORB_SLAM3
