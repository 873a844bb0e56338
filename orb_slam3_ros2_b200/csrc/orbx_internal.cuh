// orbx_internal.cuh — shared declarations of the sm_100a ORB frontend (not part of the C ABI).
//
// Data layout in HBM (one "batch" = F equally sized frames, everything indexed [frame][level]):
//   pyr    F x frameBytes     padded pyramid planes, plane l = (h_l+38) rows of pitch_l bytes (pitch multiple
//                             of 128 B so TMA boxes and 16-B vector rows are legal); the level itself is the
//                             ROI at (19,19)  — this IS mvImagePyramid (ORBextractor.cc:1176-1178)
//   blur   F x blurBytes      GaussianBlur'ed levels (no border), pitch multiple of 128 B
//   cand   F x candWords      per FAST cell a fixed slot of cellCap_l packed candidates
//                             (resp<<24 | y<<12 | x, cell-offset coordinates) + cellCnt F x nCells
//   lkp    F x lkpCap         keypoints surviving the quadtree, per level slot of (N_l+4), list order
//   out    F x cap            final cv::KeyPoint records + F x cap x 32 descriptor bytes, in operator()'s order
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>

namespace orbx {

constexpr int kMaxLevels = 16;
constexpr int kEdge = 19;       // EDGE_THRESHOLD  (ORBextractor.cc:73)
constexpr int kHalfPatch = 15;  // HALF_PATCH_SIZE (ORBextractor.cc:72)
constexpr int kPatch = 31;      // PATCH_SIZE      (ORBextractor.cc:71)
constexpr int kMinBorder = kEdge - 3;  // 16 (ORBextractor.cc:789)
constexpr float kCellW = 35.f;         // ORBextractor.cc:785

struct LevelPlan {
    int w, h;            // level size
    int pitch;           // padded plane pitch (bytes)
    int bpitch;          // blurred plane pitch
    unsigned planeOff;   // byte offset of the padded plane inside a frame's pyramid block
    unsigned blurOff;    // byte offset of the blurred plane inside a frame's blur block
    int nCols, nRows, wCell, hCell;  // FAST grid (ORBextractor.cc:797-803)
    int cellCap;         // upper bound of NMS survivors per cell = ceil(wCell/2)*ceil(hCell/2)
    int cellBase;        // first cell of this level in the frame's cell arrays
    unsigned candOff;    // word offset of this level's candidate slots inside a frame's cand block
    int maxCand;         // nCols*nRows*cellCap (capacity of the contiguous candidate list)
    unsigned gathOff;    // word offset of the level's gathered candidate list in a frame's gather block
    int nFeat;           // mnFeaturesPerLevel[l]
    int lkpBase;         // first slot of this level in the frame's level-keypoint array
    int lkpCap;          // max(nFeat + 4, 4 * nIni + 1) (F5: the quadtree returns N..N+3, or the 4 * nIni children of the roots)
    int tabX, tabY;      // offsets into the resize coefficient table (int16 triplets ofs,a0,a1)
    int gtabOff, gtabStrip;  // word offsets of the level's column-group entries / strip entries in the resize group table (0 = none)
    int blurTile0, blurTilesX;  // first 128x32 blur tile of the level in the launch's tile list, tiles per row
    int stripCells, stripSegs, stripPitch;  // FAST strips (k_fast_strip.cu): cells per strip, strips per cell row, tile pitch (bytes)
    float scale;         // mvScaleFactor[l]
    float patchSize;     // (float)(int)(31*scale) (ORBextractor.cc:879)
};

struct FramePlan {
    int nlevels;
    int W, H;
    int nCellsTotal;
    int lkpTotal;       // sum of lkpCap
    int outCap;         // keypoint capacity per frame in the final arrays
    unsigned frameBytes, blurBytes, candWords, gathWords;
    int iniTh, minTh;
    int fpMode;
    int blurTiles;      // total 128x32 tiles over all levels
    int stripTiles;     // FAST strips per frame over all levels
    LevelPlan lv[kMaxLevels];
};

// Keypoint kept by the quadtree, level coordinates (already offset by minBorder).
struct LevelKp {
    short x, y;
    short resp;
    short level;
};

struct Buffers {
    uint8_t* img;        // staged input frames (F x H x W compact) when the caller passes host memory
    uint8_t* pyr;
    uint8_t* blur;
    uint32_t* cand;
    int* cellCnt;        // F x nCellsTotal
    uint32_t* gath;      // gathered candidates F x gathWords
    uint16_t* nid;       // node id per gathered candidate, F x gathWords
    LevelKp* lkp;        // F x lkpTotal
    int* lkpCnt;         // F x nlevels
    LevelKp* sorted;     // F x outCap : the kept keypoints in operator()'s output order (slotKernel)
    void* outKps;        // F x outCap x 28 B
    uint8_t* outDesc;    // F x outCap x 32 B
    int* outN;           // F
    int* outMono;        // F
    int16_t* tab;        // resize coefficient tables
    const uint32_t* pyrGroupTab;  // per level: packed windows / selectors / coefficients per 4 padded columns (buildPyrGroupTable)
    int* errFlag;        // device-side invariant violations (capacity overflow)
    uint32_t* fastTab;   // FAST cell table of the plan (buildFastCellTable)
    int* workCtr2;       // work counter of the early level-0 FAST launch
    int* workCtr;        // two work counters of the persistent FAST launches (distinct per concurrently running sub-batch)
    const void* fastMaps;  // host copy of the per-level TMA tensor maps over the whole pyramid allocation (buildFastMaps);
                           // nullptr = stage FAST tiles with ordinary loads (ORBX_FAST_TMA=0)
    const void* blurMaps;  // the same for the blur's 144 x 38 input tiles (buildBlurMaps); nullptr = ordinary loads
    const void* pyrMaps;   // source-level maps of the resize kernels' staged rows (buildPyrMaps); nullptr = ordinary loads
    uint32_t* blurTab;   // blur tile table of the plan (buildBlurTileTable)
    const void* stripMaps;  // per-level tensor maps of the FAST strips (buildStripMaps); nullptr = per-cell FAST kernels only
    uint32_t* stripTab;     // strip table of the plan (buildFastStripTable)
    unsigned long long* fbList;  // [2][fbCap] cells the strip kernel hands to the per-cell kernel (narrow / wide class)
    int* fbCount;           // [2]
    int fbCap;              // frames x cells per frame
    int frame0;          // first frame of these buffers inside the allocation (z coordinate base of the tensor maps)
};

// kernel launchers (each in its own .cu); all asynchronous on `s`
// levels: -1 = whole pyramid; 0 = only level 0 (copy + border); 1 = only levels >= 1 (level 0 must be complete on the stream)
// firstLevel / lastLevel: only the levels in [firstLevel, lastLevel] (the chain below firstLevel must be complete on the stream)
void launchPyramid(const FramePlan& P, const Buffers& B, const uint8_t* dImgs, int stride, size_t frameStride, int F,
                   cudaStream_t s, int levels = -1, int firstLevel = 0, int lastLevel = kMaxLevels);
bool fastLevel0Early(const FramePlan& P);
void launchFast(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int which = -1);
// strip-wise FAST (k_fast_strip.cu): the whole stage = strips at iniThFAST + the per-cell kernel on what they leave behind
void planFastStrips(FramePlan& P);
void buildFastStripTable(const FramePlan& P, std::vector<uint32_t>& tab);
int buildStripMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out);
// part: 0 = everything on s (clear the lists, strips of all levels, fallback); the split form runs the strips of the levels
// <= splitLevel beside the rest of the resize chain: 1 = clear the lists + strips of the levels <= splitLevel, 2 = strips of the
// levels > splitLevel, 3 = the fallback kernels (after 1 and 2 have completed)
void launchFastStrips(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int part = 0, int splitLevel = 3);
void launchFastFallback(const FramePlan& P, const Buffers& B, int F, const unsigned long long* fbList, const int* fbCount, int cap,
                        cudaStream_t s);
bool fastWideLevel(const LevelPlan& L);
bool fastHasBothClasses(const FramePlan& P);
void buildFastCellTable(const FramePlan& P, std::vector<uint32_t>& tab);
size_t fastMapsBytes();
size_t levelMapsBytes();
void buildBlurTileTable(const FramePlan& P, std::vector<uint32_t>& tab);
void buildPyrGroupTable(const FramePlan& P, const std::vector<int16_t>& tab, std::vector<uint32_t>& out, int* offsets, int* stripOfs);
int buildPyrMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out);
int buildBlurMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out);
int buildFastMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out);
void launchQuadtree(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int lBase = 0, int nLv = -1, const int* gate = nullptr,
                    int gateMode = 0, int gateLevels = 0);
void launchBlur(const FramePlan& P, const Buffers& B, int F, cudaStream_t s);
void launchDescribe(const FramePlan& P, const Buffers& B, int F, int lap0, int lap1, cudaStream_t s);
size_t quadtreeSmemBytes(const FramePlan& P);

void launchSynth(uint8_t* dOut, int k0, int F, int w, int h, uint64_t seed, int right, cudaStream_t s);

int matchNumChunks(int nq, int64_t nd, int sms);
void launchMatchTop2(const uint8_t* dq, int nq, const uint8_t* ddb, int64_t nd, int64_t base, uint64_t* dKeys1,
                     uint64_t* dKeys2, cudaStream_t s);
void launchUnpackTop2(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, int64_t* idx1, int32_t* d1,
                      int64_t* idx2, int32_t* d2, cudaStream_t s);
void launchUnpackTop2Strided(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, size_t setStride, int64_t* idx1,
                             int32_t* d1, int64_t* idx2, int32_t* d2, cudaStream_t s);
void launchReduceKeys(const uint64_t* k1, const uint64_t* k2, int nsets, int nq, uint64_t* outK1, uint64_t* outK2,
                      cudaStream_t s);
void launchMatchConsecutive(const uint8_t* dDesc, const int* dN, int F, int cap, uint64_t* keys1, uint64_t* keys2,
                            int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2, cudaStream_t s);
// Batched ComputeStereoMatches: pair f of F reads keypoints / descriptors [f][cap], counts nl[f] / nr[f] and pyramid blocks
// pyr + f * pyrStride; every per-keypoint output is [f][cap].  nl == nullptr = one pair (the scalar arguments apply).
struct StereoBatch {
    const int* nl;
    const int* nr;
    int cap, frames;
    size_t pyrStrideL, pyrStrideR;
};
void launchStereoBanded(const void* kl, const uint8_t* dl, int nl, const void* kr, const uint8_t* dr, int nr,
                        const float* sf, int nRows, float maxD, int32_t* bi, int32_t* bd, cudaStream_t s,
                        const StereoBatch* batch = nullptr, void* rowScratch = nullptr, int bandMax = 0);
// scratch of the row-list form of the search (bandMax = stereoBandMax(largest scale factor))
size_t stereoRowScratchBytes(int pairs, int cap, int nRows, int bandMax);
inline int stereoBandMax(float sfMax) { return (int)(4.0f * sfMax) + 4; }  // rows of [floor(v - 2 sf), ceil(v + 2 sf)] <= 4 sf + 3
void launchAccept(int rule, float nnratio, const int32_t* d1, const int32_t* d2, const int64_t* idx2, const int32_t* l1,
                  const int32_t* l2, int n, uint8_t* accept, cudaStream_t s);
int gridSortCapacity();
void launchGridSearch(const void* kpsT, const uint8_t* descT, int nt, float minX, float maxX, float minY, float maxY,
                      const uint8_t* skipT, const float* uRightT, const void* queries, const uint8_t* descQ, int nq,
                      uint32_t* sortedScratch, int* cellStart, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2,
                      int32_t* lvl1, int32_t* lvl2, cudaStream_t s);
void launchStereoRefine(const FramePlan& P, const float* sf, const float* isf, const uint8_t* pyrL, const uint8_t* pyrR,
                        const void* kl, const void* kr, int nl, const int* bestIdx, const int* bestDist, float mbf, float maxD,
                        float* uRight, float* depth, int* sad, cudaStream_t s, const StereoBatch* batch = nullptr);
void launchSearchInit(const void* kps1, const uint8_t* desc1, int n1, const void* kps2, const uint8_t* desc2, int n2, const int* nPerFrame,
                      int frames, int cap, float minX, float maxX, float minY, float maxY, int windowSize, float nnratio, int checkOri,
                      int prevFromKps, uint32_t* sorted, int* cellStart, float* prev, int* m12, int* scratch3, int* nMatches,
                      cudaStream_t s, uint32_t* candList = nullptr, int* candCount = nullptr);
constexpr int kSearchInitListCap = 64;  // = kInitListCap of k_guided.cu
// grids of `frames` keypoint sets [frame][cap] (Frame::AssignFeaturesToGrid as one sort per frame): sorted frames x cap words,
// cellStart frames x 3073 ints
void launchGridSortBatch(const void* kps, const int* nPerFrame, int frames, int cap, float minX, float maxX, float minY, float maxY,
                         uint32_t* sorted, int* cellStart, cudaStream_t s);
// ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (k_project.cu)
struct ProjCam {
    float fx, fy, cx, cy, minX, maxX, minY, maxY, mbf, mb, th;
    int mono, checkOri, nlevels;
    float sf[kMaxLevels];
};
size_t projScratchWordsPerLast();
size_t mapScratchWordsPerPoint();
void launchSearchByProjectionMap(const float* track, const int* level, const uint8_t* flags, const uint8_t* mpDesc, const int* nMp,
                                 const void* kpsCur, const uint8_t* descCur, const float* uRightCur, const uint8_t* takenCur, const int* nCur,
                                 int pairs, int capMp, int capCur, const ProjCam& C, float nnratio, int farPoints, float thFar,
                                 uint32_t* sorted, int* cellStart, uint32_t* scratch, int* matchCur, int* nMatches, cudaStream_t s);
void launchSearchByProjection(const void* kpsLast, const uint8_t* lastFlags, const float* world, const uint8_t* mpDesc, const int* nLast,
                              const void* kpsCur, const uint8_t* descCur, const float* uRightCur, const uint8_t* takenCur, const int* nCur,
                              const float* poses, int pairs, int capLast, int capCur, const ProjCam& C, uint32_t* sorted, int* cellStart,
                              uint32_t* scratch, int* matchCur, int* nMatches, cudaStream_t s);
// Frame::UndistortKeyPoints: the pinhole model in doubles, as cv::undistortPoints converts it (k = k1 k2 p1 p2 k3 k4 k5 k6 s1..s4)
struct UndistortCam {
    double ifx, ify, cx, cy;
    double k[12];
    double rr[9];  // P (* R = I), row-major
};
void launchUndistort(const void* kpsIn, void* kpsOut, const int* nPerFrame, int nSingle, int frames, int cap, const UndistortCam& C,
                     cudaStream_t s);
void launchStage(const uint8_t* dSrc, int sw, int sh, int sstride, int ch, int rgb, uint8_t* dDst, int dw, int dh,
                 const int16_t* tabX, const int16_t* tabY, cudaStream_t s);
// flat DBoW2 vocabulary on the device (k_bow.cu): children of node c are CSR entries [childBegin[c], childBegin[c+1]) in the
// reference's visiting order, their descriptors stored contiguously in that order
struct BowVocDev {
    int nNodes, L;
    const int* childBegin;     // nNodes + 1
    const int* childId;        // nNodes - 1
    const uint8_t* childDesc;  // (nNodes - 1) x 32, 16-B aligned
    const double* weight;      // nNodes
    const int* wordOf;         // nNodes, -1 for inner nodes
};
int bowSortCapacity();
void launchBowTransform(const BowVocDev& V, const uint8_t* dDesc, const int* dN, int cap, int F, int levelsup, int* leaf,
                        int* nid, int32_t* wordIds, double* wordVals, int32_t* nWords, int32_t* nodeIds, int32_t* nodeBegin,
                        int32_t* featIdx, int32_t* nNodes, cudaStream_t s);
void launchBowSearch(const uint8_t* descKF, const uint8_t* validKF, const float* angleKF, const int* kfNodeIds, const int* kfNodeBegin,
                     const int* kfFeatIdx, int kfNodes, const uint8_t* descF, int nF, const float* angleF, const int* fNodeIds,
                     const int* fNodeBegin, const int* fFeatIdx, int fNodes, float nnratio, int checkOrientation, int* matchF, int* binF,
                     int* nMatches, cudaStream_t s, const uint8_t* validF = nullptr, int strictLow = 0);
cudaError_t uploadDescribeTables();
double benchPopcRate(cudaStream_t s);

void countLaunch(int n = 1);

}  // namespace orbx
