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

namespace ORB_SLAM3 {

struct ORBextractor {  // include/ORBextractor.h:83
    std::vector<cv::Mat> mvImagePyramid;
};

class GeometricCamera {};  // include/CameraModels/GeometricCamera.h: only the static_cast target below is used
class Pinhole : public GeometricCamera {  // include/CameraModels/Pinhole.h:81 toK()
public:
    cv::Mat toK() { return K.clone(); }
    cv::Mat K;  // (shim) 3 x 3 CV_32F, what Pinhole::toK builds from mvParameters (src/CameraModels/Pinhole.cpp:134-138)
};
class Frame;

class MapPoint {  // include/MapPoint.h: only isBad() is used on this path
public:
    bool isBad() { return mbBad; }
    bool mbBad = false;
    int index = -1;  // (shim) the keyframe feature this map point was created for
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
};

}  // namespace ORB_SLAM3
