// orbx_api.cu — host side of the C ABI (include/orbx.h): extractor handle, per-resolution plan, device buffers,
// stage orchestration on one CUDA stream, and the matching entry points.  No CPU fallback anywhere: every compute
// entry point needs a CUDA device and fails with ORBX_ERR_CUDA otherwise.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/orbx.h"
#include "orbx_internal.cuh"

namespace orbx {

static std::atomic<unsigned long long> gLaunches{0};
void countLaunch(int n) { gLaunches += (unsigned long long)n; }

static thread_local std::string tlsError;
static int fail(int code, const std::string& msg) {
    tlsError = msg;
    return code;
}
#define ORBX_CUDA(expr)                                                                                   \
    do {                                                                                                  \
        cudaError_t e_ = (expr);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            cudaGetLastError();                                                                           \
            return fail(ORBX_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));               \
        }                                                                                                 \
    } while (0)

static inline int cvRoundF(float v) { return (int)lrintf(v); }  // cvRound: round half to even
static inline unsigned alignUp(unsigned v, unsigned a) { return (v + a - 1) / a * a; }

// ORBX_GUARD=1 (debugging; compute-sanitizer is closed on this pool): every device buffer gets a 256-byte guard band behind it,
// filled with 0xA5; orbx_debug_check_guards verifies the bands after a run (an out-of-bounds WRITE of a kernel lands there)
static bool guardOn() {
    static const bool on = [] { const char* e = std::getenv("ORBX_GUARD"); return e && e[0] == '1'; }();
    return on;
}
constexpr size_t kGuardBytes = 256;

struct DevMem {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t need) {
        if (need <= bytes) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaMalloc(&p, need + (guardOn() ? kGuardBytes : 0));
        if (e == cudaSuccess) bytes = need;
        if (e == cudaSuccess && guardOn()) e = cudaMemset(static_cast<uint8_t*>(p) + need, 0xA5, kGuardBytes);
        return e;
    }
    // 0 = intact (or guards off / nothing allocated), 1 = overwritten, -1 = CUDA error
    int guardIntact() const {
        if (!guardOn() || !p) return 0;
        uint8_t host[kGuardBytes];
        if (cudaMemcpy(host, static_cast<const uint8_t*>(p) + bytes, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
        for (size_t i = 0; i < kGuardBytes; ++i)
            if (host[i] != 0xA5) return 1;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
};

}  // namespace orbx

using namespace orbx;

struct orbx_extractor {
    orbx_params prm;
    double scaleFactorD;  // the reference stores scaleFactor as a double member (ORBextractor.h:96)
    int device = 0;
    std::vector<float> sf, isf, sig2, isig2;
    std::vector<int> nPerLevel;
    cudaStream_t stream = nullptr;
    cudaStream_t aux = nullptr;  // blur runs here, beside FAST + quadtree on `stream` (both only need the pyramid)
    cudaEvent_t ev[12] = {};     // 0..5 stage boundaries on `stream`; 6 pyramid done -> aux, 7/8 blur begin/end on aux, 9 wide FAST done on aux, 10 level 0 of the pyramid done, 11 (unused)
    // sub-batch mode (orbx_set_subbatches): part p of a batch runs the whole pipeline on subS[p] (+ subX[p] for the blur)
    int nSub = 1;
    bool lastSplit = false;
    cudaStream_t subS[8] = {}, subX[8] = {};
    cudaEvent_t subFork = nullptr, subBlur0[8] = {}, subBlur1[8] = {}, subDone[8] = {}, subPyr[8] = {}, subWide[8] = {};
    // plan + buffers of the current (W, H, capacity F)
    FramePlan plan{};
    bool havePlan = false;
    int capFrames = 0;
    int lastFrames = 0;
    size_t tabEntries = 0;
    Buffers buf{};
    DevMem mImg, mPyr, mBlur, mCand, mCellCnt, mGath, mNid, mLkp, mLkpCnt, mSlot, mOutK, mOutD, mOutN, mOutM, mTab, mErr;
    DevMem mKeys1, mKeys2, mMatch;
    DevMem mFastTab;
    // TMA tensor maps of the FAST tile staging (one per level, over the whole mPyr allocation); rebuilt with the buffers
    alignas(64) unsigned char fastMaps[kMaxLevels * 128];
    alignas(64) unsigned char blurMaps[kMaxLevels * 128];
    alignas(64) unsigned char pyrMaps[kMaxLevels * 128];
    DevMem mBlurTab;
    alignas(64) unsigned char stripMaps[kMaxLevels * 128];
    DevMem mStripTab, mFb, mGroupTab;
    // orbx_extract_frame: page-locked staging (input frame | counts + flags | keypoints | descriptors | padded pyramid block)
    void* hostStage = nullptr;
    size_t hostStageBytes = 0;
    bool useStrips = true;  // FAST over strips of the cell grid (k_fast_strip.cu); ORBX_FAST_STRIP=0 = per-cell kernels only
    bool useTma = true;
    DevMem mKpsUn;           // orbx_undistort_batch: mvKeysUn of the last batch, [frame][cap]
    bool kpsUnValid = false;  // mKpsUn belongs to the CURRENT batch (cleared by every extraction)
    DevMem mRaw, mStageTab;  // orbx_extract_color: raw colour frame and the staging resize tables
    int stagedW = 0, stagedH = 0;
    std::vector<uint8_t> hStage;  // pinned staging would be the caller's; this is only for get_* taps
};

namespace orbx {

// ORBextractor::ORBextractor (ORB_SLAM3/src/ORBextractor.cc:409-445)
static void initTables(orbx_extractor* h) {
    const int nl = h->prm.nlevels;
    h->sf.assign(nl, 1.f); h->sig2.assign(nl, 1.f); h->isf.assign(nl, 1.f); h->isig2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        h->sf[i] = (float)(h->sf[i - 1] * h->scaleFactorD);  // :420
        h->sig2[i] = h->sf[i] * h->sf[i];
    }
    for (int i = 0; i < nl; ++i) { h->isf[i] = 1.0f / h->sf[i]; h->isig2[i] = 1.0f / h->sig2[i]; }
    h->nPerLevel.assign(nl, 0);
    const float factor = (float)(1.0f / h->scaleFactorD);  // :435
    float want = (float)(h->prm.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl)));  // :436
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        h->nPerLevel[l] = cvRoundF(want);
        sum += h->nPerLevel[l];
        want *= factor;
    }
    h->nPerLevel[nl - 1] = std::max(h->prm.nfeatures - sum, 0);  // :445
}

// cv::resize INTER_LINEAR coefficient generation (OpenCV imgproc/resize.cpp)
static void resizeTable(int srcN, int dstN, bool xAxis, std::vector<int16_t>& out) {
    const double scale = 1.0 / ((double)dstN / srcN);
    for (int d = 0; d < dstN; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= s;
        // the x axis clamps the offset and resets the fraction; the y axis keeps its fraction and the kernel clips
        // the two row indices instead (OpenCV resize.cpp does exactly this; the roundings differ)
        if (xAxis) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= srcN - 1) { f = 0; s = srcN - 1; }
        }
        auto sat = [](int v) { return (int16_t)(v < -32768 ? -32768 : (v > 32767 ? 32767 : v)); };
        out.push_back((int16_t)s);
        out.push_back(sat(cvRoundF((1.f - f) * 2048.f)));
        out.push_back(sat(cvRoundF(f * 2048.f)));
        out.push_back(0);
    }
}

// per-level keypoint capacity: DistributeOctTree returns at most max(N + 3, 4 * nIni) nodes — phase 2 stops at the first split
// that reaches N (+3 at most), but the first breadth pass splits every root whatever N is (ORBextractor.cc:610-681)
static int levelKpCap(int nFeat, int lw, int lh) {
    const float width = (float)(lw - 2 * kMinBorder), height = (float)(lh - 2 * kMinBorder);
    const int nIni = (width > 0 && height > 0) ? (int)std::round(width / height) : 0;  // :559
    return std::max(nFeat + 4, 4 * nIni + 1);
}

static int buildPlan(orbx_extractor* h, int W, int H, std::vector<int16_t>& tab) {
    FramePlan& P = h->plan;
    std::memset(&P, 0, sizeof(P));
    const int nl = h->prm.nlevels;
    P.nlevels = nl; P.W = W; P.H = H;
    P.iniTh = std::min(std::max(h->prm.ini_th_fast, 0), 255);
    P.minTh = std::min(std::max(h->prm.min_th_fast, 0), 255);
    P.fpMode = h->prm.fp_mode;
    unsigned pyrOff = 0, blurOff = 0, candOff = 0, gathOff = 0;
    int cellBase = 0, lkpBase = 0, blurTiles = 0;
    for (int l = 0; l < nl; ++l) {
        LevelPlan& L = P.lv[l];
        L.w = cvRoundF((float)W * h->isf[l]);  // ORBextractor.cc:1175 (from the ORIGINAL size)
        L.h = cvRoundF((float)H * h->isf[l]);
        if (L.w < 1 || L.h < 1) return fail(ORBX_ERR_ARG, "image too small for the number of pyramid levels");
        L.pitch = (int)alignUp(L.w + 2 * kEdge, 128);
        L.bpitch = (int)alignUp(L.w, 128);
        L.planeOff = pyrOff; pyrOff += (unsigned)L.pitch * (L.h + 2 * kEdge);
        L.blurOff = blurOff; blurOff += (unsigned)L.bpitch * L.h;
        L.blurTile0 = blurTiles; L.blurTilesX = (L.w + 127) / 128; blurTiles += L.blurTilesX * ((L.h + 31) / 32);
        // FAST grid (ORBextractor.cc:789-803)
        const float width = (float)(L.w - 2 * kMinBorder), height = (float)(L.h - 2 * kMinBorder);
        L.nCols = width > 0 ? (int)(width / kCellW) : 0;
        L.nRows = height > 0 ? (int)(height / kCellW) : 0;
        if (L.nCols <= 0 || L.nRows <= 0) { L.nCols = L.nRows = 0; L.wCell = L.hCell = 0; }
        else { L.wCell = (int)std::ceil(width / L.nCols); L.hCell = (int)std::ceil(height / L.nRows); }
        L.cellCap = ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);
        L.cellBase = cellBase; cellBase += L.nCols * L.nRows;
        L.candOff = candOff; L.maxCand = L.nCols * L.nRows * L.cellCap; candOff += (unsigned)L.maxCand;
        L.gathOff = gathOff; gathOff += (unsigned)L.maxCand;
        if (L.maxCand >= (1 << 24)) return fail(ORBX_ERR_ARG, "level too large (candidate index exceeds 24 bits)");
        L.nFeat = h->nPerLevel[l];
        L.lkpBase = lkpBase; L.lkpCap = levelKpCap(L.nFeat, L.w, L.h); lkpBase += L.lkpCap;
        L.scale = h->sf[l];
        L.patchSize = (float)(int)(kPatch * h->sf[l]);  // :879
        if (l > 0) {
            L.tabX = (int)(tab.size() / 4); resizeTable(P.lv[l - 1].w, L.w, true, tab);
            L.tabY = (int)(tab.size() / 4); resizeTable(P.lv[l - 1].h, L.h, false, tab);
        }
    }
    if (W > 4000 || H > 4000) return fail(ORBX_ERR_ARG, "images larger than 4000 px are not supported (12-bit packed coordinates)");
    P.frameBytes = alignUp(pyrOff + 256, 256);
    P.blurBytes = alignUp(blurOff + 256, 256);
    P.candWords = candOff; P.gathWords = alignUp(gathOff, 8);
    P.nCellsTotal = cellBase;
    P.blurTiles = blurTiles;
    P.lkpTotal = lkpBase;
    P.outCap = lkpBase;
    planFastStrips(P);
    if (quadtreeSmemBytes(P) > 220 * 1024) return fail(ORBX_ERR_ARG, "nfeatures too large for the quadtree kernel's shared memory");
    return ORBX_OK;
}

static int ensureCapacity(orbx_extractor* h, int W, int H, int F) {
    if (W < 1 || H < 1 || F < 1) return fail(ORBX_ERR_ARG, "bad frame size / count");
    if (!h->havePlan || h->plan.W != W || h->plan.H != H) {
        std::vector<int16_t> tab;
        const int rc = buildPlan(h, W, H, tab);
        if (rc != ORBX_OK) return rc;
        ORBX_CUDA(h->mTab.reserve(std::max<size_t>(tab.size() * sizeof(int16_t), 16)));
        ORBX_CUDA(cudaMemcpyAsync(h->mTab.p, tab.data(), tab.size() * sizeof(int16_t), cudaMemcpyHostToDevice, h->stream));
        std::vector<uint32_t> gtab;
        {
            int go[kMaxLevels], gs[kMaxLevels];
            buildPyrGroupTable(h->plan, tab, gtab, go, gs);
            for (int l = 0; l < kMaxLevels; ++l) { h->plan.lv[l].gtabOff = go[l]; h->plan.lv[l].gtabStrip = gs[l]; }
        }
        ORBX_CUDA(h->mGroupTab.reserve(std::max<size_t>(gtab.size() * 4, 16)));
        ORBX_CUDA(cudaMemcpyAsync(h->mGroupTab.p, gtab.data(), gtab.size() * 4, cudaMemcpyHostToDevice, h->stream));
        std::vector<uint32_t> cells;
        buildFastCellTable(h->plan, cells);
        ORBX_CUDA(h->mFastTab.reserve(std::max<size_t>(cells.size() * 4, 16)));
        if (!cells.empty())
            ORBX_CUDA(cudaMemcpyAsync(h->mFastTab.p, cells.data(), cells.size() * 4, cudaMemcpyHostToDevice, h->stream));
        std::vector<uint32_t> btiles;
        buildBlurTileTable(h->plan, btiles);
        if ((int)btiles.size() != h->plan.blurTiles) return fail(ORBX_ERR_CUDA, "blur tile table does not match the plan");
        ORBX_CUDA(h->mBlurTab.reserve(std::max<size_t>(btiles.size() * 4, 16)));
        ORBX_CUDA(cudaMemcpyAsync(h->mBlurTab.p, btiles.data(), btiles.size() * 4, cudaMemcpyHostToDevice, h->stream));
        std::vector<uint32_t> strips;
        buildFastStripTable(h->plan, strips);
        if (strips.size() != (size_t)h->plan.stripTiles * 16) return fail(ORBX_ERR_CUDA, "FAST strip table does not match the plan");
        ORBX_CUDA(h->mStripTab.reserve(std::max<size_t>(strips.size() * 4, 16)));
        if (!strips.empty())
            ORBX_CUDA(cudaMemcpyAsync(h->mStripTab.p, strips.data(), strips.size() * 4, cudaMemcpyHostToDevice, h->stream));
        ORBX_CUDA(cudaStreamSynchronize(h->stream));
        h->havePlan = true;
        h->capFrames = 0;
    }
    const FramePlan& P = h->plan;
    if (F > h->capFrames) {
        const size_t f = (size_t)F;
        ORBX_CUDA(h->mPyr.reserve(f * P.frameBytes + 4096));
        ORBX_CUDA(h->mBlur.reserve(f * P.blurBytes + 4096));
        ORBX_CUDA(h->mCand.reserve(std::max<size_t>(f * P.candWords * 4, 16)));
        ORBX_CUDA(h->mCellCnt.reserve(std::max<size_t>(f * P.nCellsTotal * 4, 16)));
        ORBX_CUDA(h->mGath.reserve(std::max<size_t>(f * P.gathWords * 4, 16)));
        ORBX_CUDA(h->mNid.reserve(std::max<size_t>(f * P.gathWords * 2, 16)));
        ORBX_CUDA(h->mLkp.reserve(f * P.lkpTotal * sizeof(LevelKp)));
        ORBX_CUDA(h->mLkpCnt.reserve(f * P.nlevels * 4));
        ORBX_CUDA(h->mSlot.reserve(f * P.outCap * sizeof(LevelKp)));
        ORBX_CUDA(h->mOutK.reserve(f * P.outCap * 28));
        ORBX_CUDA(h->mOutD.reserve(f * P.outCap * 32));
        ORBX_CUDA(h->mOutN.reserve(f * 4));
        ORBX_CUDA(h->mOutM.reserve(f * 4));
        ORBX_CUDA(h->mFb.reserve(std::max<size_t>(2 * f * P.nCellsTotal * 8, 16)));
        ORBX_CUDA(h->mErr.reserve(256));  // [0] error flags, [4 + 2p ..] work counters of persistent kernels, sub-batch p; [32 + p] early level-0 FAST; [40 + 2p ..] fallback list counts
        ORBX_CUDA(cudaMemsetAsync(h->mErr.p, 0, 256, h->stream));
        h->capFrames = F;
        if (h->useTma) {
            if (fastMapsBytes() > sizeof(h->fastMaps)) return fail(ORBX_ERR_CUDA, "tensor map table larger than expected");
            const int rc = buildFastMaps(P, (const uint8_t*)h->mPyr.p, F, h->fastMaps);
            if (rc != 0) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string(rc) + ")");
            if (levelMapsBytes() > sizeof(h->blurMaps)) return fail(ORBX_ERR_CUDA, "tensor map table larger than expected");
            const int rb = buildBlurMaps(P, (const uint8_t*)h->mPyr.p, F, h->blurMaps);
            if (rb != 0) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled failed for the blur tiles (" + std::to_string(rb) + ")");
            const int rp = buildPyrMaps(P, (const uint8_t*)h->mPyr.p, F, h->pyrMaps);
            if (rp != 0) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled failed for the resize rows (" + std::to_string(rp) + ")");
            if (h->useStrips) {
                const int rs = buildStripMaps(P, (const uint8_t*)h->mPyr.p, F, h->stripMaps);
                if (rs != 0) return fail(ORBX_ERR_CUDA, "cuTensorMapEncodeTiled failed for the FAST strips (" + std::to_string(rs) + ")");
            }
        }
    }
    Buffers& B = h->buf;
    B.img = (uint8_t*)h->mImg.p; B.pyr = (uint8_t*)h->mPyr.p; B.blur = (uint8_t*)h->mBlur.p;
    B.cand = (uint32_t*)h->mCand.p; B.cellCnt = (int*)h->mCellCnt.p; B.gath = (uint32_t*)h->mGath.p;
    B.nid = (uint16_t*)h->mNid.p; B.lkp = (LevelKp*)h->mLkp.p; B.lkpCnt = (int*)h->mLkpCnt.p; B.sorted = (LevelKp*)h->mSlot.p;
    B.outKps = h->mOutK.p; B.outDesc = (uint8_t*)h->mOutD.p; B.outN = (int*)h->mOutN.p; B.outMono = (int*)h->mOutM.p;
    B.pyrGroupTab = (const uint32_t*)h->mGroupTab.p;
    B.tab = (int16_t*)h->mTab.p; B.fastTab = (uint32_t*)h->mFastTab.p; B.errFlag = (int*)h->mErr.p; B.workCtr = B.errFlag + 4; B.workCtr2 = B.errFlag + 32;
    B.fastMaps = h->useTma ? h->fastMaps : nullptr; B.frame0 = 0;
    B.blurMaps = h->useTma ? h->blurMaps : nullptr; B.pyrMaps = h->useTma ? h->pyrMaps : nullptr; B.blurTab = (uint32_t*)h->mBlurTab.p;
    B.stripMaps = (h->useTma && h->useStrips) ? h->stripMaps : nullptr; B.stripTab = (uint32_t*)h->mStripTab.p;
    B.fbList = (unsigned long long*)h->mFb.p; B.fbCount = B.errFlag + 40; B.fbCap = h->capFrames * P.nCellsTotal;
    return ORBX_OK;
}

// Buffers of frames [f0, ...) of the batch, with the work counters of sub-batch `part`
static Buffers offsetBuffers(const Buffers& B, const FramePlan& P, int f0, int part, int nFrames) {
    Buffers O = B;
    const size_t f = (size_t)f0;
    O.pyr += f * P.frameBytes; O.blur += f * P.blurBytes; O.cand += f * P.candWords; O.cellCnt += f * P.nCellsTotal;
    O.gath += f * P.gathWords; O.nid += f * P.gathWords; O.lkp += f * P.lkpTotal; O.lkpCnt += f * P.nlevels;
    O.sorted += f * P.outCap; O.outKps = (uint8_t*)O.outKps + f * P.outCap * 28; O.outDesc += f * P.outCap * 32;
    O.outN += f; O.outMono += f;
    O.workCtr = B.errFlag + 4 + 2 * part;
    O.workCtr2 = B.errFlag + 32 + part;
    O.fbList = B.fbList + 2 * f * P.nCellsTotal; O.fbCount = B.errFlag + 40 + 2 * part; O.fbCap = nFrames * P.nCellsTotal;
    O.frame0 = B.frame0 + f0;
    return O;
}

// ORBX_FAST_FORK=0 keeps the two FAST cell classes back to back on one stream (A/B measurements)
static bool forkWideFast() {
    static const bool on = [] { const char* e = std::getenv("ORBX_FAST_FORK"); return !e || std::atoi(e) != 0; }();
    return on;
}

// ORBX_FAST_EARLY=1 starts level 0's FAST beside the resize chain (A/B measurements; measured no gain: pyramid + FAST
// 0.396 + 0.506 ms apart, 0.543 + 0.371 ms overlapped, profiles/r01_v15_e2e_sweep.txt) - off by default
static bool earlyLevel0Fast() {
    static const bool on = [] { const char* e = std::getenv("ORBX_FAST_EARLY"); return e && std::atoi(e) != 0; }();
    return on;
}

// Sub-batch mode: the batch is cut into nSub contiguous parts, each running the whole pipeline on its own stream pair.
// Frames are independent, so the parts only share the SMs: one part's latency-bound stages (quadtree, describe) run
// beside another part's issue-bound ones (FAST, blur, resize) instead of each stage owning the whole GPU in turn.
static int runPipelineSplit(orbx_extractor* h, const uint8_t* dImgs, int stride, size_t frameStride, int F, int lap0, int lap1) {
    const FramePlan& P = h->plan;
    cudaStream_t s = h->stream;
    const int n = h->nSub;
    const bool forkWide = forkWideFast() && fastHasBothClasses(P);
    cudaEventRecord(h->ev[0], s);
    cudaEventRecord(h->subFork, s);
    for (int p = 0; p < n; ++p) {
        const int f0 = (int)((long long)F * p / n), f1 = (int)((long long)F * (p + 1) / n);
        if (f1 <= f0) continue;
        const Buffers B = offsetBuffers(h->buf, P, f0, p, f1 - f0);
        const uint8_t* img = dImgs + (size_t)f0 * frameStride;
        cudaStream_t a = h->subS[p], x = h->subX[p];
        cudaStreamWaitEvent(a, h->subFork, 0);
        launchPyramid(P, B, img, stride, frameStride, f1 - f0, a);
        if (B.stripMaps) {
            launchFastStrips(P, B, f1 - f0, a);
        } else if (forkWide) {  // the few wide cells (small levels) run on the blur's stream beside the narrow class
            cudaEventRecord(h->subPyr[p], a);
            cudaStreamWaitEvent(x, h->subPyr[p], 0);
            launchFast(P, B, f1 - f0, x, 1);
            cudaEventRecord(h->subWide[p], x);
            launchFast(P, B, f1 - f0, a, 0);
            cudaStreamWaitEvent(a, h->subWide[p], 0);
        } else {
            launchFast(P, B, f1 - f0, a);
        }
        cudaEventRecord(h->subBlur0[p], a);
        cudaStreamWaitEvent(x, h->subBlur0[p], 0);
        launchBlur(P, B, f1 - f0, x);
        cudaEventRecord(h->subBlur1[p], x);
        launchQuadtree(P, B, f1 - f0, a);
        cudaStreamWaitEvent(a, h->subBlur1[p], 0);
        launchDescribe(P, B, f1 - f0, lap0, lap1, a);
        cudaEventRecord(h->subDone[p], a);
        cudaStreamWaitEvent(s, h->subDone[p], 0);
    }
    cudaEventRecord(h->ev[5], s);
    h->lastFrames = F;
    h->lastSplit = true;
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// the whole operator() pipeline for F device-resident frames, asynchronous on the handle's stream
static int runPipeline(orbx_extractor* h, const uint8_t* dImgs, int stride, size_t frameStride, int F, int lap0, int lap1) {
    h->kpsUnValid = false;
    if (h->nSub > 1 && F >= 8 * h->nSub) return runPipelineSplit(h, dImgs, stride, frameStride, F, lap0, lap1);
    h->lastSplit = false;
    const FramePlan& P = h->plan;
    const Buffers& B = h->buf;
    cudaStream_t s = h->stream;
    cudaStream_t x = h->aux;
    cudaEventRecord(h->ev[0], s);
    const bool forkWide = forkWideFast() && fastHasBothClasses(P);
    // ORBX_PIPE (A/B measurements): 0 = pyramid -> FAST -> (quadtree || blur) -> describe; 1 = the blur starts as soon as the pyramid is
    // complete, beside FAST (it only needs the pyramid); 2 = additionally the strips of the levels 0..3 start when level 3 is written and
    // run beside the resize of the small levels (launch- and latency-bound kernels of 20-40 us).  Measured (frames/s of the whole step):
    // 0 156.1 k, 1 152.6 k (FAST + blur together 0.64 ms, quadtree alone 0.19), 2 152.8 k: the stages share the same integer pipes, so
    // running them side by side moves time around without saving any; 3 = see below; default 0.
    static const int pipeMode = [] { const char* e = std::getenv("ORBX_PIPE"); return e ? std::atoi(e) : 0; }();
    bool blurLaunched = false, quadSplit = false;
    constexpr int kEarlyLevels = 2;  // pipeMode 3: the levels whose quadtree starts before FAST has finished the others
    int* gate = B.errFlag + 56;
    if (B.stripMaps && pipeMode == 3 && P.nlevels > kEarlyLevels) {
        // The quadtree is a chain of short rounds per (frame, level): latency-bound, 30 % of the issue slots, and the two big levels set its
        // length.  Their FAST strips run first; their quadtree then starts on the auxiliary stream beside the strips of the small levels
        // (gate = snapshot of the strips' overflow counts: if a strip of the big levels overflowed, their cells wait for the per-cell
        // kernel and the late pass takes those levels as well); the blur follows it on that stream.
        // Measured (parity-green incl. the overflow cases): 170.4-170.5 k against 171.3-172.1 k frames/s, one handle's batch 1.41 against
        // 1.36 ms - beside the strips the quadtree costs FAST what it saves itself; not the default.
        launchFastStrips(P, B, F, s, 4, kEarlyLevels - 1);
        launchPyramid(P, B, dImgs, stride, frameStride, F, s);
        cudaEventRecord(h->ev[1], s);
        launchFastStrips(P, B, F, s, 1, kEarlyLevels - 1);
        cudaMemcpyAsync(gate, B.fbCount, 2 * sizeof(int), cudaMemcpyDeviceToDevice, s);
        cudaEventRecord(h->ev[10], s);
        cudaStreamWaitEvent(x, h->ev[10], 0);
        launchQuadtree(P, B, F, x, 0, kEarlyLevels, gate, 1, kEarlyLevels);
        cudaEventRecord(h->ev[7], x);
        launchBlur(P, B, F, x);
        cudaEventRecord(h->ev[8], x);
        blurLaunched = true;
        quadSplit = true;
        launchFastStrips(P, B, F, s, 2, kEarlyLevels - 1);
        launchFastStrips(P, B, F, s, 3, kEarlyLevels - 1);
    } else if (B.stripMaps && pipeMode == 2 && P.nlevels > 4) {
        launchFastStrips(P, B, F, s, 4, 3);   // list counts cleared before either part can append
        launchPyramid(P, B, dImgs, stride, frameStride, F, s, -1, 0, 3);
        cudaEventRecord(h->ev[10], s);
        cudaStreamWaitEvent(x, h->ev[10], 0);
        launchFastStrips(P, B, F, x, 1, 3);   // levels 0..3 on the auxiliary stream
        cudaEventRecord(h->ev[9], x);
        launchPyramid(P, B, dImgs, stride, frameStride, F, s, 1, 4, kMaxLevels);
        cudaEventRecord(h->ev[1], s);
        cudaStreamWaitEvent(x, h->ev[1], 0);
        cudaEventRecord(h->ev[7], x);
        launchBlur(P, B, F, x);
        cudaEventRecord(h->ev[8], x);
        blurLaunched = true;
        launchFastStrips(P, B, F, s, 2, 3);   // levels 4.. on the main stream
        cudaStreamWaitEvent(s, h->ev[9], 0);
        launchFastStrips(P, B, F, s, 3, 3);   // what the strips left behind
    } else if (B.stripMaps) {
        launchPyramid(P, B, dImgs, stride, frameStride, F, s);
        cudaEventRecord(h->ev[1], s);
        if (pipeMode == 1) {
            cudaStreamWaitEvent(x, h->ev[1], 0);
            cudaEventRecord(h->ev[7], x);
            launchBlur(P, B, F, x);
            cudaEventRecord(h->ev[8], x);
            blurLaunched = true;
        }
        launchFastStrips(P, B, F, s);
    } else if (earlyLevel0Fast() && fastLevel0Early(P)) {
        // level 0 holds a third of the cells and needs no resize: its FAST starts on the auxiliary stream as soon as the level-0
        // plane is written and runs (on 2 CTAs per SM) beside the resize chain of levels 1..7, whose kernels are barrier- and
        // latency-bound (60 % issue) and leave it the issue slots
        launchPyramid(P, B, dImgs, stride, frameStride, F, s, 0);
        cudaEventRecord(h->ev[10], s);
        cudaStreamWaitEvent(x, h->ev[10], 0);
        launchFast(P, B, F, x, 2);
        launchPyramid(P, B, dImgs, stride, frameStride, F, s, 1);
        cudaEventRecord(h->ev[1], s);
        if (fastHasBothClasses(P)) {
            cudaStreamWaitEvent(x, h->ev[1], 0);
            launchFast(P, B, F, x, 1);
        }
        cudaEventRecord(h->ev[9], x);
        launchFast(P, B, F, s, 3);
        cudaStreamWaitEvent(s, h->ev[9], 0);
    } else {
        launchPyramid(P, B, dImgs, stride, frameStride, F, s);
        cudaEventRecord(h->ev[1], s);
        if (forkWide) {
            // the wide class (few, expensive cells of the small levels: low occupancy, latency-bound) starts first on the
            // auxiliary stream and the narrow class fills the rest of every SM beside it
            cudaStreamWaitEvent(x, h->ev[1], 0);
            launchFast(P, B, F, x, 1);
            cudaEventRecord(h->ev[9], x);
            launchFast(P, B, F, s, 0);
            cudaStreamWaitEvent(s, h->ev[9], 0);
        } else {
            launchFast(P, B, F, s);
        }
    }
    cudaEventRecord(h->ev[2], s);
    // fork: the quadtree is latency-bound (one CTA per (frame, level), serial rounds with barriers, ~30 % issue rate), the
    // blur is issue-bound and independent of it: run them side by side so the blur fills the quadtree's bubbles
    if (!blurLaunched) {
        cudaStreamWaitEvent(x, h->ev[2], 0);
        cudaEventRecord(h->ev[7], x);
        launchBlur(P, B, F, x);
        cudaEventRecord(h->ev[8], x);
    }
    ORBX_CUDA(cudaGetLastError());  // a failed launch of the pyramid / FAST / blur stages is reported as such, not carried further
    if (quadSplit) launchQuadtree(P, B, F, s, 0, -1, gate, 2, kEarlyLevels);
    else launchQuadtree(P, B, F, s);
    cudaEventRecord(h->ev[3], s);
    cudaStreamWaitEvent(s, h->ev[8], 0);  // join
    cudaEventRecord(h->ev[4], s);
    launchDescribe(P, B, F, lap0, lap1, s);
    cudaEventRecord(h->ev[5], s);
    h->lastFrames = F;
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

static int checkDeviceFlags(orbx_extractor* h) {
    int flag = 0;
    if (!h->mErr.p) {  // nothing has run on this handle yet
        ORBX_CUDA(cudaStreamSynchronize(h->stream));
        return ORBX_OK;
    }
    ORBX_CUDA(cudaMemcpyAsync(&flag, h->mErr.p, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (flag) {
        cudaMemsetAsync(h->mErr.p, 0, 16, h->stream);
        char msg[96];
        snprintf(msg, sizeof msg, "device-side capacity invariant violated (flags 0x%x)", flag);
        return fail(ORBX_ERR_CAPACITY, msg);
    }
    return ORBX_OK;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (dev >= 0 && dev != prev) cudaSetDevice(dev);
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace orbx
extern "C" {

const char* orbx_last_error(void) { return tlsError.c_str(); }
int orbx_version(void) { return 100; }
unsigned long long orbx_launch_count(void) { return gLaunches.load(); }

int orbx_create(const orbx_params* params, orbx_handle* out) {
    if (!params || !out) return fail(ORBX_ERR_ARG, "null argument");
    if (params->nlevels < 1 || params->nlevels > kMaxLevels) return fail(ORBX_ERR_ARG, "nlevels must be in [1,16]");
    if (params->nfeatures < 1 || !(params->scale_factor > 1.0f)) return fail(ORBX_ERR_ARG, "bad nfeatures / scale_factor");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    int dev = params->device;
    if (dev < 0) ORBX_CUDA(cudaGetDevice(&dev));
    if (dev >= ndev) return fail(ORBX_ERR_ARG, "device ordinal out of range");
    DeviceGuard guard(dev);
    orbx_extractor* h = new orbx_extractor();
    h->prm = *params;
    h->device = dev;
    h->scaleFactorD = (double)params->scale_factor;
    initTables(h);
    {
        const char* t = std::getenv("ORBX_FAST_TMA");  // 0 = stage FAST tiles with ordinary loads (A/B measurements only)
        h->useTma = !(t && t[0] == '0');
        const char* st = std::getenv("ORBX_FAST_STRIP");  // 0 = the per-cell FAST kernels only (A/B measurements)
        h->useStrips = !(st && st[0] == '0');
    }
    // ORBX_PRIO (A/B measurements): 1 = the main stream (pyramid, FAST, quadtree, describe) above the auxiliary one (blur), so the
    // quadtree's long serial CTAs are placed first and the blur fills in around them; 2 = the other way round; default 0 = equal
    static const int prio = [] { const char* v = std::getenv("ORBX_PRIO"); return v ? std::atoi(v) : 0; }();
    int prLeast = 0, prGreatest = 0;
    cudaDeviceGetStreamPriorityRange(&prLeast, &prGreatest);
    cudaError_t e = cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, prio == 1 ? prGreatest : prLeast);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&h->aux, cudaStreamNonBlocking, prio == 2 ? prGreatest : prLeast);
    for (int i = 0; i < 12 && e == cudaSuccess; ++i) e = cudaEventCreate(&h->ev[i]);
    if (e != cudaSuccess) {
        orbx_destroy(h);  // releases whatever streams / events were created
        return fail(ORBX_ERR_CUDA, std::string("stream/event creation: ") + cudaGetErrorString(e));
    }
    e = uploadDescribeTables();
    if (e != cudaSuccess) {
        orbx_destroy(h);
        return fail(ORBX_ERR_CUDA, std::string("uploading the rBRIEF pattern / disc tables: ") + cudaGetErrorString(e));
    }
    *out = h;
    return ORBX_OK;
}

int orbx_capacity(orbx_handle h, int w, int hgt, int* cap) {
    if (!h || !cap) return fail(ORBX_ERR_ARG, "null argument");
    if (w < 1 || hgt < 1) return fail(ORBX_ERR_ARG, "bad frame size");
    int sum = 0;
    for (int l = 0; l < h->prm.nlevels; ++l)
        sum += levelKpCap(h->nPerLevel[l], cvRoundF((float)w * h->isf[l]), cvRoundF((float)hgt * h->isf[l]));
    *cap = sum;
    return ORBX_OK;
}

int orbx_debug_check_guards(orbx_handle h) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    if (!guardOn()) return fail(ORBX_ERR_STATE, "guard bands are off (set ORBX_GUARD=1 before the first call)");
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaDeviceSynchronize());
    struct { const char* name; const DevMem* m; } all[] = {
        {"img", &h->mImg}, {"pyr", &h->mPyr}, {"blur", &h->mBlur}, {"cand", &h->mCand}, {"cellCnt", &h->mCellCnt}, {"gath", &h->mGath},
        {"nid", &h->mNid}, {"lkp", &h->mLkp}, {"lkpCnt", &h->mLkpCnt}, {"slot", &h->mSlot}, {"outK", &h->mOutK}, {"outD", &h->mOutD},
        {"outN", &h->mOutN}, {"outM", &h->mOutM}, {"tab", &h->mTab}, {"err", &h->mErr}, {"keys1", &h->mKeys1}, {"keys2", &h->mKeys2},
        {"match", &h->mMatch}, {"fastTab", &h->mFastTab}, {"blurTab", &h->mBlurTab}, {"stripTab", &h->mStripTab}, {"fb", &h->mFb},
        {"groupTab", &h->mGroupTab}, {"kpsUn", &h->mKpsUn}, {"raw", &h->mRaw}, {"stageTab", &h->mStageTab}};
    for (const auto& e : all) {
        const int r = e.m->guardIntact();
        if (r < 0) return fail(ORBX_ERR_CUDA, std::string("reading the guard band of ") + e.name);
        if (r > 0) return fail(ORBX_ERR_CAPACITY, std::string("guard band behind device buffer '") + e.name + "' was overwritten");
    }
    return ORBX_OK;
}

void orbx_destroy(orbx_handle h) {
    if (!h) return;
    DeviceGuard guard(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    DevMem* all[] = {&h->mImg, &h->mPyr, &h->mBlur, &h->mCand, &h->mCellCnt, &h->mGath, &h->mNid, &h->mLkp, &h->mLkpCnt,
                     &h->mSlot, &h->mOutK, &h->mOutD, &h->mOutN, &h->mOutM, &h->mTab, &h->mErr, &h->mKeys1, &h->mKeys2,
                     &h->mMatch, &h->mRaw, &h->mStageTab, &h->mFastTab, &h->mBlurTab, &h->mKpsUn, &h->mStripTab, &h->mFb, &h->mGroupTab};
    for (DevMem* m : all) m->release();
    if (h->hostStage) cudaFreeHost(h->hostStage);
    for (int i = 0; i < 12; ++i)
        if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    for (int p = 0; p < 8; ++p) {
        if (h->subS[p]) cudaStreamDestroy(h->subS[p]);
        if (h->subX[p]) cudaStreamDestroy(h->subX[p]);
        if (h->subBlur0[p]) cudaEventDestroy(h->subBlur0[p]);
        if (h->subBlur1[p]) cudaEventDestroy(h->subBlur1[p]);
        if (h->subDone[p]) cudaEventDestroy(h->subDone[p]);
        if (h->subPyr[p]) cudaEventDestroy(h->subPyr[p]);
        if (h->subWide[p]) cudaEventDestroy(h->subWide[p]);
    }
    if (h->subFork) cudaEventDestroy(h->subFork);
    if (h->aux) cudaStreamDestroy(h->aux);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int orbx_get_tables(orbx_handle h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* nfeat) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    for (int i = 0; i < h->prm.nlevels; ++i) {
        if (scale) scale[i] = h->sf[i];
        if (inv_scale) inv_scale[i] = h->isf[i];
        if (sigma2) sigma2[i] = h->sig2[i];
        if (inv_sigma2) inv_sigma2[i] = h->isig2[i];
        if (nfeat) nfeat[i] = h->nPerLevel[i];
    }
    return ORBX_OK;
}

int orbx_extract_batch_device(orbx_handle h, const uint8_t* d_imgs, int nframes, int w, int hgt, int stride,
                              size_t frame_stride, int lap0, int lap1) {
    if (!h || !d_imgs) return fail(ORBX_ERR_ARG, "null argument");
    if (stride < w) return fail(ORBX_ERR_ARG, "stride < width");
    DeviceGuard guard(h->device);
    int rc = ensureCapacity(h, w, hgt, nframes);
    if (rc != ORBX_OK) return rc;
    return runPipeline(h, d_imgs, stride, frame_stride, nframes, lap0, lap1);
}

// H2D of the frames + the pipeline, enqueued on the handle's stream (no host synchronisation)
static int enqueueHostBatch(orbx_extractor* h, const uint8_t* imgs, int nframes, int w, int hgt, int stride,
                            size_t frame_stride, int lap0, int lap1) {
    int rc = ensureCapacity(h, w, hgt, nframes);
    if (rc != ORBX_OK) return rc;
    const size_t frameBytes = (size_t)w * hgt;
    ORBX_CUDA(h->mImg.reserve((size_t)nframes * frameBytes + 64));
    h->buf.img = (uint8_t*)h->mImg.p;
    if (stride == w && frame_stride == frameBytes) {
        ORBX_CUDA(cudaMemcpyAsync(h->mImg.p, imgs, (size_t)nframes * frameBytes, cudaMemcpyHostToDevice, h->stream));
    } else {
        for (int f = 0; f < nframes; ++f)
            ORBX_CUDA(cudaMemcpy2DAsync((uint8_t*)h->mImg.p + (size_t)f * frameBytes, w, imgs + (size_t)f * frame_stride,
                                        stride, w, hgt, cudaMemcpyHostToDevice, h->stream));
    }
    return runPipeline(h, (const uint8_t*)h->mImg.p, w, frameBytes, nframes, lap0, lap1);
}

// D2H of full per-frame slots (cap_per_frame >= the handle's capacity), enqueued
static int enqueueResultCopies(orbx_extractor* h, int nframes, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* n_out,
                               int* mono_index) {
    const FramePlan& P = h->plan;
    if (kps)
        ORBX_CUDA(cudaMemcpy2DAsync(kps, (size_t)cap_per_frame * 28, h->buf.outKps, (size_t)P.outCap * 28,
                                    (size_t)P.outCap * 28, nframes, cudaMemcpyDeviceToHost, h->stream));
    if (desc)
        ORBX_CUDA(cudaMemcpy2DAsync(desc, (size_t)cap_per_frame * 32, h->buf.outDesc, (size_t)P.outCap * 32,
                                    (size_t)P.outCap * 32, nframes, cudaMemcpyDeviceToHost, h->stream));
    if (n_out) ORBX_CUDA(cudaMemcpyAsync(n_out, h->buf.outN, (size_t)nframes * 4, cudaMemcpyDeviceToHost, h->stream));
    if (mono_index)
        ORBX_CUDA(cudaMemcpyAsync(mono_index, h->buf.outMono, (size_t)nframes * 4, cudaMemcpyDeviceToHost, h->stream));
    return ORBX_OK;
}

int orbx_extract_batch_async(orbx_handle h, const uint8_t* imgs, int nframes, int w, int hgt, int stride, size_t frame_stride,
                             int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* n_out,
                             int* mono_index) {
    if (!h || !imgs) return fail(ORBX_ERR_ARG, "null argument");
    if (stride < w) return fail(ORBX_ERR_ARG, "stride < width");
    DeviceGuard guard(h->device);
    int rc = enqueueHostBatch(h, imgs, nframes, w, hgt, stride, frame_stride, lap0, lap1);
    if (rc != ORBX_OK) return rc;
    if ((kps || desc) && cap_per_frame < h->plan.outCap)
        return fail(ORBX_ERR_CAPACITY, "the asynchronous form needs cap_per_frame >= the extractor's capacity (sum of N_l + 4)");
    return enqueueResultCopies(h, nframes, kps, desc, cap_per_frame, n_out, mono_index);
}

int orbx_check(orbx_handle h) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    DeviceGuard guard(h->device);
    return checkDeviceFlags(h);
}

int orbx_extract_batch(orbx_handle h, const uint8_t* imgs, int nframes, int w, int hgt, int stride, size_t frame_stride,
                       int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* n_out,
                       int* mono_index) {
    if (!h || !imgs) return fail(ORBX_ERR_ARG, "null argument");
    if (stride < w) return fail(ORBX_ERR_ARG, "stride < width");
    DeviceGuard guard(h->device);
    int rc = enqueueHostBatch(h, imgs, nframes, w, hgt, stride, frame_stride, lap0, lap1);
    if (rc != ORBX_OK) return rc;
    const FramePlan& P = h->plan;
    if ((kps || desc) && cap_per_frame < P.outCap) {
        // the compact copy below needs the per-frame counts first
        std::vector<int> n(nframes);
        ORBX_CUDA(cudaMemcpyAsync(n.data(), h->buf.outN, (size_t)nframes * 4, cudaMemcpyDeviceToHost, h->stream));
        ORBX_CUDA(cudaStreamSynchronize(h->stream));
        for (int f = 0; f < nframes; ++f)
            if (n[f] > cap_per_frame) {
                if (n_out) std::memcpy(n_out, n.data(), (size_t)nframes * 4);
                return fail(ORBX_ERR_CAPACITY, "cap_per_frame smaller than a frame's keypoint count");
            }
        if (kps)
            ORBX_CUDA(cudaMemcpy2DAsync(kps, (size_t)cap_per_frame * 28, h->buf.outKps, (size_t)P.outCap * 28,
                                        (size_t)cap_per_frame * 28, nframes, cudaMemcpyDeviceToHost, h->stream));
        if (desc)
            ORBX_CUDA(cudaMemcpy2DAsync(desc, (size_t)cap_per_frame * 32, h->buf.outDesc, (size_t)P.outCap * 32,
                                        (size_t)cap_per_frame * 32, nframes, cudaMemcpyDeviceToHost, h->stream));
        rc = enqueueResultCopies(h, nframes, nullptr, nullptr, cap_per_frame, n_out, mono_index);
    } else {
        rc = enqueueResultCopies(h, nframes, kps, desc, cap_per_frame, n_out, mono_index);
    }
    if (rc != ORBX_OK) return rc;
    return checkDeviceFlags(h);
}

int orbx_extract(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, orbx_keypoint* kps,
                 uint8_t* desc, int cap, int* n_out, int* mono_index) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    if (n_out) *n_out = 0;
    if (!img || w <= 0 || hgt <= 0) {  // ORBextractor.cc:1090-1091
        if (mono_index) *mono_index = -1;
        return ORBX_OK;
    }
    DeviceGuard guard(h->device);
    int n = 0, mono = 0;
    int rc = orbx_extract_batch(h, img, 1, w, hgt, stride, (size_t)stride * hgt, lap0, lap1, nullptr, nullptr, 0, &n, &mono);
    if (rc != ORBX_OK) return rc;
    if (n_out) *n_out = n;
    if (mono_index) *mono_index = mono;
    if (n > cap) return fail(ORBX_ERR_CAPACITY, "keypoint buffer too small");
    if (kps && n) ORBX_CUDA(cudaMemcpyAsync(kps, h->buf.outKps, (size_t)n * 28, cudaMemcpyDeviceToHost, h->stream));
    if (desc && n) ORBX_CUDA(cudaMemcpyAsync(desc, h->buf.outDesc, (size_t)n * 32, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_extract_frame(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, int want_pyramid,
                       orbx_frame_result* out) {
    if (!h || !out) return fail(ORBX_ERR_ARG, "null argument");
    std::memset(out, 0, sizeof(*out));
    if (!img || w <= 0 || hgt <= 0) {  // ORBextractor.cc:1090-1091
        out->mono_index = -1;
        return ORBX_OK;
    }
    if (stride < w) return fail(ORBX_ERR_ARG, "stride < width");
    DeviceGuard guard(h->device);
    int rc = ensureCapacity(h, w, hgt, 1);
    if (rc != ORBX_OK) return rc;
    const FramePlan& P = h->plan;
    const size_t imgBytes = (size_t)w * hgt;
    const size_t oImg = 0, oCnt = alignUp((unsigned)imgBytes, 256), oKps = oCnt + 256, oDesc = oKps + alignUp((unsigned)P.outCap * 28, 256);
    const size_t oPyr = oDesc + alignUp((unsigned)P.outCap * 32, 256), need = oPyr + P.frameBytes;
    if (need > h->hostStageBytes) {
        if (h->hostStage) cudaFreeHost(h->hostStage);
        h->hostStage = nullptr;
        h->hostStageBytes = 0;
        ORBX_CUDA(cudaHostAlloc(&h->hostStage, need, cudaHostAllocDefault));
        h->hostStageBytes = need;
    }
    uint8_t* hs = static_cast<uint8_t*>(h->hostStage);
    if (stride == w) std::memcpy(hs + oImg, img, imgBytes);
    else
        for (int y = 0; y < hgt; ++y) std::memcpy(hs + oImg + (size_t)y * w, img + (size_t)y * stride, (size_t)w);
    ORBX_CUDA(h->mImg.reserve(imgBytes + 64));
    h->buf.img = (uint8_t*)h->mImg.p;
    ORBX_CUDA(cudaMemcpyAsync(h->mImg.p, hs + oImg, imgBytes, cudaMemcpyHostToDevice, h->stream));
    rc = runPipeline(h, (const uint8_t*)h->mImg.p, w, imgBytes, 1, lap0, lap1);
    if (rc != ORBX_OK) return rc;
    int* cnt = reinterpret_cast<int*>(hs + oCnt);  // [0] n, [1] monoIndex, [2] device-side flags
    ORBX_CUDA(cudaMemcpyAsync(cnt, h->buf.outN, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(cnt + 1, h->buf.outMono, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(cnt + 2, h->mErr.p, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(hs + oKps, h->buf.outKps, (size_t)P.outCap * 28, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(hs + oDesc, h->buf.outDesc, (size_t)P.outCap * 32, cudaMemcpyDeviceToHost, h->stream));
    if (want_pyramid) ORBX_CUDA(cudaMemcpyAsync(hs + oPyr, h->buf.pyr, P.frameBytes, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (cnt[2]) {
        cudaMemsetAsync(h->mErr.p, 0, 16, h->stream);
        char msg[96];
        snprintf(msg, sizeof msg, "device-side capacity invariant violated (flags 0x%x)", cnt[2]);
        return fail(ORBX_ERR_CAPACITY, msg);
    }
    out->kps = reinterpret_cast<const orbx_keypoint*>(hs + oKps);
    out->desc = hs + oDesc;
    out->n = cnt[0];
    out->mono_index = cnt[1];
    out->pyramid = want_pyramid ? hs + oPyr : nullptr;
    out->nlevels = P.nlevels;
    for (int l = 0; l < P.nlevels; ++l) {
        out->level[l].w = P.lv[l].w; out->level[l].h = P.lv[l].h;
        out->level[l].pitch = P.lv[l].pitch; out->level[l].offset = P.lv[l].planeOff;
    }
    return ORBX_OK;
}

int orbx_extract_color(orbx_handle h, const uint8_t* img, int w, int hgt, int stride, int channels, int rgb, int new_w, int new_h,
                       int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    if (n_out) *n_out = 0;
    if (!img || w <= 0 || hgt <= 0) {  // ORBextractor.cc:1090-1091
        if (mono_index) *mono_index = -1;
        return ORBX_OK;
    }
    if (channels != 1 && channels != 3 && channels != 4) return fail(ORBX_ERR_ARG, "channels must be 1, 3 or 4");
    if (stride < w * channels) return fail(ORBX_ERR_ARG, "stride < width * channels");
    if ((new_w > 0) != (new_h > 0)) return fail(ORBX_ERR_ARG, "new_w / new_h must both be set or both be 0");
    const bool resize = new_w > 0 && (new_w != w || new_h != hgt);
    const int dw = resize ? new_w : w, dh = resize ? new_h : hgt;
    if (channels == 1 && !resize) return orbx_extract(h, img, w, hgt, stride, lap0, lap1, kps, desc, cap, n_out, mono_index);
    DeviceGuard guard(h->device);
    int rc = ensureCapacity(h, dw, dh, 1);
    if (rc != ORBX_OK) return rc;
    const size_t rowBytes = (size_t)w * channels;
    ORBX_CUDA(h->mRaw.reserve(rowBytes * hgt + 64));
    ORBX_CUDA(h->mImg.reserve((size_t)dw * dh + 64));
    h->buf.img = (uint8_t*)h->mImg.p;
    ORBX_CUDA(cudaMemcpy2DAsync(h->mRaw.p, rowBytes, img, stride, rowBytes, hgt, cudaMemcpyHostToDevice, h->stream));
    const int16_t *tx = nullptr, *ty = nullptr;
    if (resize) {
        std::vector<int16_t> tab;
        resizeTable(w, dw, true, tab);
        resizeTable(hgt, dh, false, tab);
        ORBX_CUDA(h->mStageTab.reserve(tab.size() * sizeof(int16_t)));
        ORBX_CUDA(cudaMemcpyAsync(h->mStageTab.p, tab.data(), tab.size() * sizeof(int16_t), cudaMemcpyHostToDevice, h->stream));
        ORBX_CUDA(cudaStreamSynchronize(h->stream));  // `tab` is pageable host memory about to go out of scope
        tx = (const int16_t*)h->mStageTab.p;
        ty = tx + (size_t)dw * 4;
    }
    launchStage((const uint8_t*)h->mRaw.p, w, hgt, (int)rowBytes, channels, rgb, (uint8_t*)h->mImg.p, dw, dh, tx, ty, h->stream);
    h->stagedW = dw; h->stagedH = dh;
    rc = runPipeline(h, (const uint8_t*)h->mImg.p, dw, (size_t)dw * dh, 1, lap0, lap1);
    if (rc != ORBX_OK) return rc;
    int nm[2] = {0, 0};
    ORBX_CUDA(cudaMemcpyAsync(&nm[0], h->buf.outN, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(&nm[1], h->buf.outMono, 4, cudaMemcpyDeviceToHost, h->stream));
    rc = checkDeviceFlags(h);
    if (rc != ORBX_OK) return rc;
    if (n_out) *n_out = nm[0];
    if (mono_index) *mono_index = nm[1];
    if (nm[0] > cap) return fail(ORBX_ERR_CAPACITY, "keypoint buffer too small");
    if (kps && nm[0]) ORBX_CUDA(cudaMemcpyAsync(kps, h->buf.outKps, (size_t)nm[0] * 28, cudaMemcpyDeviceToHost, h->stream));
    if (desc && nm[0]) ORBX_CUDA(cudaMemcpyAsync(desc, h->buf.outDesc, (size_t)nm[0] * 32, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_get_staged_gray(orbx_handle h, uint8_t* out, int* w, int* hgt) {
    if (!h || h->stagedW == 0) return fail(ORBX_ERR_STATE, "no orbx_extract_color yet");
    if (w) *w = h->stagedW;
    if (hgt) *hgt = h->stagedH;
    if (!out) return ORBX_OK;
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaMemcpyAsync(out, h->mImg.p, (size_t)h->stagedW * h->stagedH, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_set_subbatches(orbx_handle h, int n) {
    if (!h || n < 1 || n > 8) return fail(ORBX_ERR_ARG, "sub-batch count must be in [1, 8]");
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (!h->subFork) ORBX_CUDA(cudaEventCreateWithFlags(&h->subFork, cudaEventDisableTiming));
    for (int p = 0; p < n; ++p) {
        if (h->subS[p]) continue;
        ORBX_CUDA(cudaStreamCreateWithFlags(&h->subS[p], cudaStreamNonBlocking));
        ORBX_CUDA(cudaStreamCreateWithFlags(&h->subX[p], cudaStreamNonBlocking));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->subBlur0[p], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->subBlur1[p], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->subDone[p], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->subPyr[p], cudaEventDisableTiming));
        ORBX_CUDA(cudaEventCreateWithFlags(&h->subWide[p], cudaEventDisableTiming));
    }
    h->nSub = n;
    return ORBX_OK;
}

int orbx_sync(orbx_handle h) {
    if (!h) return fail(ORBX_ERR_ARG, "null handle");
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

void* orbx_get_stream(orbx_handle h) { return h ? (void*)h->stream : nullptr; }

int orbx_last_timing(orbx_handle h, float* total_ms, float stage_ms[5]) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaEventSynchronize(h->ev[5]));
    if (total_ms) ORBX_CUDA(cudaEventElapsedTime(total_ms, h->ev[0], h->ev[5]));
    if (stage_ms && h->lastSplit) {  // the stages of different sub-batches overlap: there is no per-stage time
        for (int i = 0; i < 5; ++i) stage_ms[i] = 0.f;
    } else if (stage_ms) {
        // pyramid, fast, quadtree on the main stream; blur from its own events on the aux stream (it overlaps fast and
        // quadtree, so the stage times no longer add up to the total); describe after the join
        ORBX_CUDA(cudaEventElapsedTime(&stage_ms[0], h->ev[0], h->ev[1]));
        ORBX_CUDA(cudaEventElapsedTime(&stage_ms[1], h->ev[1], h->ev[2]));
        ORBX_CUDA(cudaEventElapsedTime(&stage_ms[2], h->ev[2], h->ev[3]));
        ORBX_CUDA(cudaEventElapsedTime(&stage_ms[3], h->ev[7], h->ev[8]));
        ORBX_CUDA(cudaEventElapsedTime(&stage_ms[4], h->ev[4], h->ev[5]));
    }
    return ORBX_OK;
}

int orbx_batch_results(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc, const int** d_n,
                       const int** d_mono, int* cap_per_frame) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (d_kps) *d_kps = (const orbx_keypoint*)h->buf.outKps;
    if (d_desc) *d_desc = h->buf.outDesc;
    if (d_n) *d_n = h->buf.outN;
    if (d_mono) *d_mono = h->buf.outMono;
    if (cap_per_frame) *cap_per_frame = h->plan.outCap;
    return ORBX_OK;
}

int orbx_batch_counts(orbx_handle h, int32_t* n_out, int32_t* mono_out) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    DeviceGuard guard(h->device);
    const size_t bytes = (size_t)h->lastFrames * 4;
    if (n_out) ORBX_CUDA(cudaMemcpyAsync(n_out, h->buf.outN, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (mono_out) ORBX_CUDA(cudaMemcpyAsync(mono_out, h->buf.outMono, bytes, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return checkDeviceFlags(h);
}

int orbx_batch_download(orbx_handle h, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (frame < 0 || frame >= h->lastFrames) return fail(ORBX_ERR_ARG, "frame out of range");
    DeviceGuard guard(h->device);
    int nm[2] = {0, 0};
    ORBX_CUDA(cudaMemcpyAsync(&nm[0], h->buf.outN + frame, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaMemcpyAsync(&nm[1], h->buf.outMono + frame, 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    if (n_out) *n_out = nm[0];
    if (mono_index) *mono_index = nm[1];
    if (nm[0] > cap) return fail(ORBX_ERR_CAPACITY, "keypoint buffer too small");
    const size_t o = (size_t)frame * h->plan.outCap;
    if (kps && nm[0]) ORBX_CUDA(cudaMemcpyAsync(kps, (uint8_t*)h->buf.outKps + o * 28, (size_t)nm[0] * 28, cudaMemcpyDeviceToHost, h->stream));
    if (desc && nm[0]) ORBX_CUDA(cudaMemcpyAsync(desc, h->buf.outDesc + o * 32, (size_t)nm[0] * 32, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return checkDeviceFlags(h);
}

int orbx_get_pyramid(orbx_handle h, int frame, int level, uint8_t* out, int* w, int* hgt) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (frame < 0 || frame >= h->lastFrames || level < 0 || level >= h->plan.nlevels) return fail(ORBX_ERR_ARG, "frame/level out of range");
    const LevelPlan& L = h->plan.lv[level];
    if (w) *w = L.w;
    if (hgt) *hgt = L.h;
    if (!out) return ORBX_OK;
    DeviceGuard guard(h->device);
    const int pw = L.w + 2 * kEdge, ph = L.h + 2 * kEdge;
    ORBX_CUDA(cudaMemcpy2DAsync(out, pw, h->buf.pyr + (size_t)frame * h->plan.frameBytes + L.planeOff, L.pitch, pw, ph,
                                cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_get_blurred(orbx_handle h, int frame, int level, uint8_t* out, int* w, int* hgt) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (frame < 0 || frame >= h->lastFrames || level < 0 || level >= h->plan.nlevels) return fail(ORBX_ERR_ARG, "frame/level out of range");
    const LevelPlan& L = h->plan.lv[level];
    if (w) *w = L.w;
    if (hgt) *hgt = L.h;
    if (!out) return ORBX_OK;
    DeviceGuard guard(h->device);
    ORBX_CUDA(cudaMemcpy2DAsync(out, L.w, h->buf.blur + (size_t)frame * h->plan.blurBytes + L.blurOff, L.bpitch, L.w, L.h,
                                cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_get_candidates(orbx_handle h, int frame, int level, int32_t* out, int cap, int* n_out) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (frame < 0 || frame >= h->lastFrames || level < 0 || level >= h->plan.nlevels) return fail(ORBX_ERR_ARG, "frame/level out of range");
    DeviceGuard guard(h->device);
    const FramePlan& P = h->plan;
    const LevelPlan& L = P.lv[level];
    const int nCells = L.nCols * L.nRows;
    std::vector<int> cnt(std::max(nCells, 1));
    if (nCells) ORBX_CUDA(cudaMemcpyAsync(cnt.data(), h->buf.cellCnt + (size_t)frame * P.nCellsTotal + L.cellBase, (size_t)nCells * 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    int n = 0;
    for (int c = 0; c < nCells; ++c) n += cnt[c];
    if (n_out) *n_out = n;
    if (!out) return ORBX_OK;
    if (n > cap) return fail(ORBX_ERR_CAPACITY, "candidate buffer too small");
    std::vector<uint32_t> slots((size_t)std::max(L.maxCand, 1));
    if (L.maxCand) ORBX_CUDA(cudaMemcpyAsync(slots.data(), h->buf.cand + (size_t)frame * P.candWords + L.candOff, (size_t)L.maxCand * 4, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA(cudaStreamSynchronize(h->stream));
    int o = 0;
    for (int c = 0; c < nCells; ++c)
        for (int k = 0; k < cnt[c]; ++k) {
            const uint32_t p = slots[(size_t)c * L.cellCap + k];
            out[3 * o] = p & 0xfff; out[3 * o + 1] = (p >> 12) & 0xfff; out[3 * o + 2] = p >> 24;
            ++o;
        }
    return ORBX_OK;
}

int orbx_synth_frames_device(uint8_t* d_out, int k0, int nframes, int w, int hgt, uint64_t seed, int right, int device) {
    if (!d_out || nframes < 1 || w < 1 || hgt < 1) return fail(ORBX_ERR_ARG, "bad argument");
    DeviceGuard guard(device);
    launchSynth(d_out, k0, nframes, w, hgt, seed, right, nullptr);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaDeviceSynchronize());
    return ORBX_OK;
}

}  // extern "C"
// ======================================================================================================
// Matching
// ======================================================================================================
#include <dlfcn.h>
// NCCL is loaded with dlopen (the extractor half must work on hosts without it), so the header is optional too: without it the few
// types and enumerators the sharded entry points use are declared here as NCCL 2.x defines them (nccl.h: ncclUniqueId is 128 opaque
// bytes, ncclComm_t an opaque pointer, ncclSuccess = 0, ncclUint64 = 5).
#if defined(__has_include) && !defined(ORBX_NO_NCCL_H)
#if __has_include(<nccl.h>)
#include <nccl.h>
#define ORBX_HAVE_NCCL_H 1
#endif
#endif
#ifndef ORBX_HAVE_NCCL_H
extern "C" {
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0 } ncclResult_t;
typedef enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5 } ncclDataType_t;
}
#endif

namespace orbx {

// scratch for the non-handle matching entry points (per thread, per device)
struct MatchScratch {
    int device = -1;
    DevMem q, db, k1, k2, i1, i2, d1, d2, misc;
    cudaStream_t stream = nullptr;
};
static thread_local MatchScratch tlsMatch;

static int matchScratch(int device, MatchScratch** out) {
    MatchScratch& S = tlsMatch;
    if (S.device != device) {
        if (S.device >= 0) {
            DeviceGuard g(S.device);
            DevMem* all[] = {&S.q, &S.db, &S.k1, &S.k2, &S.i1, &S.i2, &S.d1, &S.d2, &S.misc};
            for (DevMem* m : all) m->release();
            if (S.stream) cudaStreamDestroy(S.stream);
            S.stream = nullptr;
        }
        S.device = device;
    }
    if (!S.stream) ORBX_CUDA(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking));
    *out = &S;
    return ORBX_OK;
}

}  // namespace orbx
extern "C" void orbx_thread_cleanup(void) {
    MatchScratch& S = tlsMatch;
    if (S.device < 0) return;
    DeviceGuard g(S.device);
    if (S.stream) cudaStreamSynchronize(S.stream);
    DevMem* all[] = {&S.q, &S.db, &S.k1, &S.k2, &S.i1, &S.i2, &S.d1, &S.d2, &S.misc};
    for (DevMem* m : all) m->release();
    if (S.stream) cudaStreamDestroy(S.stream);
    S.stream = nullptr;
    S.device = -1;
}
namespace orbx {

static int smCount() {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms;
}

// local top-2 keys (one candidate set: all chunks already merged) of q x db on the current device
static int localTop2Keys(const uint8_t* dq, int nq, const uint8_t* ddb, int64_t nd, int64_t base, DevMem& k1, DevMem& k2,
                         int* nsets, cudaStream_t s) {
    const int chunks = nd > 0 ? matchNumChunks(nq, nd, smCount()) : 0;
    ORBX_CUDA(k1.reserve(std::max<size_t>((size_t)std::max(chunks, 1) * nq * 8, 16)));
    ORBX_CUDA(k2.reserve(std::max<size_t>((size_t)std::max(chunks, 1) * nq * 8, 16)));
    if (chunks == 0) {
        ORBX_CUDA(cudaMemsetAsync(k1.p, 0xff, (size_t)nq * 8, s));
        ORBX_CUDA(cudaMemsetAsync(k2.p, 0xff, (size_t)nq * 8, s));
        *nsets = 1;
        return ORBX_OK;
    }
    launchMatchTop2(dq, nq, ddb, nd, base, (uint64_t*)k1.p, (uint64_t*)k2.p, s);
    *nsets = chunks;
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

// ---- NCCL through dlopen: the extractor half of the library must load on hosts without NCCL -------------------
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi gNccl;
static std::mutex gNcclMutex;

static int loadNccl() {
    std::lock_guard<std::mutex> lock(gNcclMutex);
    if (gNccl.lib) return ORBX_OK;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return fail(ORBX_ERR_NCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
    gNccl.GetUniqueId = (decltype(gNccl.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    gNccl.CommInitRank = (decltype(gNccl.CommInitRank))dlsym(lib, "ncclCommInitRank");
    gNccl.CommDestroy = (decltype(gNccl.CommDestroy))dlsym(lib, "ncclCommDestroy");
    gNccl.AllGather = (decltype(gNccl.AllGather))dlsym(lib, "ncclAllGather");
    gNccl.GetErrorString = (decltype(gNccl.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!gNccl.GetUniqueId || !gNccl.CommInitRank || !gNccl.CommDestroy || !gNccl.AllGather || !gNccl.GetErrorString)
        return fail(ORBX_ERR_NCCL, "libnccl lacks a required symbol");
    gNccl.lib = lib;
    return ORBX_OK;
}
#define ORBX_NCCL(expr)                                                                               \
    do {                                                                                              \
        ncclResult_t r_ = (expr);                                                                     \
        if (r_ != ncclSuccess) return fail(ORBX_ERR_NCCL, std::string(#expr) + ": " + gNccl.GetErrorString(r_)); \
    } while (0)

}  // namespace orbx

struct orbx_comm_s {
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1, device = 0;
    cudaStream_t stream = nullptr;
    DevMem k1, k2, loc, gathered;
};

extern "C" {

// ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074): sum of popcounts of eight 32-bit XOR words
int orbx_hamming(const uint8_t* a, const uint8_t* b) {
    uint32_t pa[8], pb[8];
    std::memcpy(pa, a, 32);
    std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; ++i) dist += __builtin_popcount(pa[i] ^ pb[i]);
    return dist;
}

int orbx_match_top2_device(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t nd, int64_t db_index_base,
                           int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream) {
    if (nq < 0 || nd < 0 || (nq > 0 && !d_q) || (nd > 0 && !d_db)) return fail(ORBX_ERR_ARG, "bad argument");
    if (nq == 0) return ORBX_OK;
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    int rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    int nsets = 1;
    rc = localTop2Keys(d_q, nq, d_db, nd, db_index_base, S->k1, S->k2, &nsets, s);
    if (rc != ORBX_OK) return rc;
    launchUnpackTop2((const uint64_t*)S->k1.p, (const uint64_t*)S->k2.p, nsets, nq, d_idx1, d_d1, d_idx2, d_d2, s);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int orbx_match_top2(const uint8_t* q, int nq, const uint8_t* db, int64_t nd, int64_t* idx1, int32_t* d1, int64_t* idx2,
                    int32_t* d2, int device) {
    if (nq < 0 || nd < 0 || (nq > 0 && !q) || (nd > 0 && !db) || !idx1 || !d1 || !idx2 || !d2) return fail(ORBX_ERR_ARG, "bad argument");
    if (nq == 0) return ORBX_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    int rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA(S->q.reserve((size_t)nq * 32));
    ORBX_CUDA(S->db.reserve(std::max<size_t>((size_t)nd * 32, 16)));
    ORBX_CUDA(S->i1.reserve((size_t)nq * 8));
    ORBX_CUDA(S->i2.reserve((size_t)nq * 8));
    ORBX_CUDA(S->d1.reserve((size_t)nq * 4));
    ORBX_CUDA(S->d2.reserve((size_t)nq * 4));
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(S->q.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
    if (nd) ORBX_CUDA(cudaMemcpyAsync(S->db.p, db, (size_t)nd * 32, cudaMemcpyHostToDevice, s));
    rc = orbx_match_top2_device((const uint8_t*)S->q.p, nq, (const uint8_t*)S->db.p, nd, 0, (int64_t*)S->i1.p,
                                (int32_t*)S->d1.p, (int64_t*)S->i2.p, (int32_t*)S->d2.p, s);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA(cudaMemcpyAsync(idx1, S->i1.p, (size_t)nq * 8, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(idx2, S->i2.p, (size_t)nq * 8, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(d1, S->d1.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(d2, S->d2.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_merge_top2_device(const uint64_t* d_keys1, const uint64_t* d_keys2, int nranks, int nq, int64_t* d_idx1,
                           int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream) {
    if (!d_keys1 || !d_keys2 || nranks < 1 || nq < 0) return fail(ORBX_ERR_ARG, "bad argument");
    launchUnpackTop2(d_keys1, d_keys2, nranks, nq, d_idx1, d_d1, d_idx2, d_d2, (cudaStream_t)stream);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int orbx_match_batch_consecutive(orbx_handle h, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2, int to_host) {
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (!idx1 || !d1 || !idx2 || !d2) return fail(ORBX_ERR_ARG, "null output");
    const int F = h->lastFrames, cap = h->plan.outCap;
    if (F < 2) return ORBX_OK;
    DeviceGuard guard(h->device);
    const size_t n = (size_t)(F - 1) * cap;
    ORBX_CUDA(h->mKeys1.reserve(n * 8));
    ORBX_CUDA(h->mKeys2.reserve(n * 8));
    int32_t *o1 = idx1, *o2 = d1, *o3 = idx2, *o4 = d2;
    if (to_host) {
        ORBX_CUDA(h->mMatch.reserve(n * 16));
        o1 = (int32_t*)h->mMatch.p; o2 = o1 + n; o3 = o2 + n; o4 = o3 + n;
    }
    launchMatchConsecutive(h->buf.outDesc, h->buf.outN, F, cap, (uint64_t*)h->mKeys1.p, (uint64_t*)h->mKeys2.p, o1, o2, o3,
                           o4, h->stream);
    ORBX_CUDA(cudaGetLastError());
    if (to_host) {
        ORBX_CUDA(cudaMemcpyAsync(idx1, o1, n * 4, cudaMemcpyDeviceToHost, h->stream));
        ORBX_CUDA(cudaMemcpyAsync(d1, o2, n * 4, cudaMemcpyDeviceToHost, h->stream));
        ORBX_CUDA(cudaMemcpyAsync(idx2, o3, n * 4, cudaMemcpyDeviceToHost, h->stream));
        ORBX_CUDA(cudaMemcpyAsync(d2, o4, n * 4, cudaMemcpyDeviceToHost, h->stream));
        if (to_host != 2) ORBX_CUDA(cudaStreamSynchronize(h->stream));  // 2 = enqueue only (orbx_sync / orbx_check later)
    }
    return ORBX_OK;
}

int orbx_match_stereo_banded(const orbx_keypoint* kps_l, const uint8_t* desc_l, int nl, const orbx_keypoint* kps_r,
                             const uint8_t* desc_r, int nr, const float* scale_factors, int nlevels, int n_rows,
                             float max_d, int32_t* best_idx, int32_t* best_dist, int device) {
    if (nl < 0 || nr < 0 || !scale_factors || nlevels < 1 || !best_idx || !best_dist) return fail(ORBX_ERR_ARG, "bad argument");
    if (nl == 0) return ORBX_OK;
    // the kernel indexes scale_factors with the right keypoints' octaves (Frame.cc:831): reject what would read out of bounds
    for (int i = 0; i < nr; ++i)
        if (kps_r[i].octave < 0 || kps_r[i].octave >= nlevels) return fail(ORBX_ERR_ARG, "right keypoint octave outside [0, nlevels)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    int rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    const size_t bl = (size_t)nl * 60, br = (size_t)std::max(nr, 1) * 60;
    ORBX_CUDA(S->q.reserve(bl + 64));
    ORBX_CUDA(S->db.reserve(br + 64));
    ORBX_CUDA(S->misc.reserve((size_t)nlevels * 4 + (size_t)nl * 8 + 64));
    cudaStream_t s = S->stream;
    // layout: descriptors first (16-B aligned), keypoints after
    uint8_t* dl = (uint8_t*)S->q.p; uint8_t* kl = dl + (size_t)nl * 32;
    uint8_t* dr = (uint8_t*)S->db.p; uint8_t* kr = dr + (size_t)std::max(nr, 1) * 32;
    float* sf = (float*)S->misc.p;
    int32_t* bi = (int32_t*)((uint8_t*)S->misc.p + (((size_t)nlevels * 4 + 15) & ~(size_t)15));
    int32_t* bd = bi + nl;
    ORBX_CUDA(cudaMemcpyAsync(dl, desc_l, (size_t)nl * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(kl, kps_l, (size_t)nl * 28, cudaMemcpyHostToDevice, s));
    if (nr) {
        ORBX_CUDA(cudaMemcpyAsync(dr, desc_r, (size_t)nr * 32, cudaMemcpyHostToDevice, s));
        ORBX_CUDA(cudaMemcpyAsync(kr, kps_r, (size_t)nr * 28, cudaMemcpyHostToDevice, s));
    }
    ORBX_CUDA(cudaMemcpyAsync(sf, scale_factors, (size_t)nlevels * 4, cudaMemcpyHostToDevice, s));
    float sfMax = 1.f;
    for (int l = 0; l < nlevels; ++l) sfMax = std::max(sfMax, scale_factors[l]);
    const int bandMax = stereoBandMax(sfMax);
    ORBX_CUDA(S->k1.reserve(stereoRowScratchBytes(1, std::max(nr, 1), n_rows, bandMax)));
    launchStereoBanded(kl, dl, nl, kr, dr, nr, sf, n_rows, max_d, bi, bd, s, nullptr, S->k1.p, bandMax);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(best_idx, bi, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(best_dist, bd, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

static int needDevice() {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    return ORBX_OK;
}

int orbx_accept_top2_device(int rule, float nnratio, const int32_t* d_d1, const int32_t* d_d2, const int64_t* d_idx2,
                            const int32_t* d_level1, const int32_t* d_level2, int n, uint8_t* d_accept, void* stream) {
    if (rule < 0 || rule > ORBX_ACCEPT_LOW_RATIO || n < 0 || (n > 0 && (!d_d1 || !d_d2 || !d_accept))) return fail(ORBX_ERR_ARG, "bad argument");
    launchAccept(rule, nnratio, d_d1, d_d2, d_idx2, d_level1, d_level2, n, d_accept, (cudaStream_t)stream);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

int orbx_accept_top2(int rule, float nnratio, const int32_t* d1, const int32_t* d2, const int64_t* idx2, const int32_t* level1,
                     const int32_t* level2, int n, uint8_t* accept, int device) {
    if (rule < 0 || rule > ORBX_ACCEPT_LOW_RATIO || n < 0 || (n > 0 && (!d1 || !d2 || !accept))) return fail(ORBX_ERR_ARG, "bad argument");
    if (n == 0) return ORBX_OK;
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    const size_t n4 = ((size_t)n * 4 + 15) & ~(size_t)15, n8 = (size_t)n * 8;
    ORBX_CUDA(S->misc.reserve(4 * n4 + n8 + n + 64));
    uint8_t* base = (uint8_t*)S->misc.p;
    int32_t *a = (int32_t*)base, *b = (int32_t*)(base + n4), *l1 = (int32_t*)(base + 2 * n4), *l2 = (int32_t*)(base + 3 * n4);
    int64_t* i2 = (int64_t*)(base + 4 * n4);
    uint8_t* acc = base + 4 * n4 + n8;
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(a, d1, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(b, d2, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    if (level1) ORBX_CUDA(cudaMemcpyAsync(l1, level1, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    if (level2) ORBX_CUDA(cudaMemcpyAsync(l2, level2, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    if (idx2) ORBX_CUDA(cudaMemcpyAsync(i2, idx2, n8, cudaMemcpyHostToDevice, s));
    rc = orbx_accept_top2_device(rule, nnratio, a, b, idx2 ? i2 : nullptr, level1 ? l1 : nullptr, level2 ? l2 : nullptr, n, acc, s);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA(cudaMemcpyAsync(accept, acc, (size_t)n, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_grid_top2(const orbx_keypoint* kps_t, const uint8_t* desc_t, int nt, float min_x, float max_x, float min_y,
                          float max_y, const uint8_t* skip_t, const float* u_right_t, const orbx_grid_query* queries,
                          const uint8_t* desc_q, int nq, int32_t* idx1, int32_t* d1, int32_t* idx2, int32_t* d2, int32_t* lvl1,
                          int32_t* lvl2, int device) {
    if (nt < 0 || nq < 0 || (nt > 0 && (!kps_t || !desc_t)) || (nq > 0 && (!queries || !desc_q || !idx1 || !d1 || !idx2 || !d2)))
        return fail(ORBX_ERR_ARG, "bad argument");
    if (!(max_x > min_x) || !(max_y > min_y)) return fail(ORBX_ERR_ARG, "empty image bounds");
    if (nt > gridSortCapacity()) return fail(ORBX_ERR_ARG, "more than 32768 target keypoints");
    if (nq == 0) return ORBX_OK;
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    static_assert(sizeof(orbx_grid_query) == 24, "orbx_grid_query layout");
    const size_t T = (size_t)std::max(nt, 1), Q = (size_t)nq;
    // db: [desc_t | kps_t | skip | uRight | sorted | cellStart]   q: [desc_q | queries | 6 outputs]
    const size_t oKt = T * 32, oSk = oKt + ((T * 28 + 15) & ~(size_t)15), oUr = oSk + ((T + 15) & ~(size_t)15),
                 oSo = oUr + T * 4 + 16, oCs = oSo + T * 4 + 16, dbBytes = oCs + (size_t)(64 * 48 + 1) * 4 + 64;
    const size_t oQs = Q * 32, oOut = oQs + ((Q * 24 + 15) & ~(size_t)15), qBytes = oOut + 6 * Q * 4 + 64;
    ORBX_CUDA(S->db.reserve(dbBytes));
    ORBX_CUDA(S->q.reserve(qBytes));
    uint8_t* D = (uint8_t*)S->db.p;
    uint8_t* Qb = (uint8_t*)S->q.p;
    cudaStream_t s = S->stream;
    if (nt) {
        ORBX_CUDA(cudaMemcpyAsync(D, desc_t, (size_t)nt * 32, cudaMemcpyHostToDevice, s));
        ORBX_CUDA(cudaMemcpyAsync(D + oKt, kps_t, (size_t)nt * 28, cudaMemcpyHostToDevice, s));
        if (skip_t) ORBX_CUDA(cudaMemcpyAsync(D + oSk, skip_t, (size_t)nt, cudaMemcpyHostToDevice, s));
        if (u_right_t) ORBX_CUDA(cudaMemcpyAsync(D + oUr, u_right_t, (size_t)nt * 4, cudaMemcpyHostToDevice, s));
    }
    ORBX_CUDA(cudaMemcpyAsync(Qb, desc_q, Q * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Qb + oQs, queries, Q * 24, cudaMemcpyHostToDevice, s));
    int32_t* out = (int32_t*)(Qb + oOut);
    launchGridSearch(D + oKt, D, nt, min_x, max_x, min_y, max_y, skip_t ? D + oSk : nullptr, u_right_t ? (const float*)(D + oUr) : nullptr,
                     Qb + oQs, Qb, nq, (uint32_t*)(D + oSo), (int*)(D + oCs), out, out + Q, out + 2 * Q, out + 3 * Q, out + 4 * Q,
                     out + 5 * Q, s);
    ORBX_CUDA(cudaGetLastError());
    int32_t* host[6] = {idx1, d1, idx2, d2, lvl1, lvl2};
    for (int k = 0; k < 6; ++k)
        if (host[k]) ORBX_CUDA(cudaMemcpyAsync(host[k], out + k * Q, Q * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_compute_stereo_matches(orbx_handle left, orbx_handle right, int frame, const orbx_keypoint* kps_l, const uint8_t* desc_l,
                                int nl, const orbx_keypoint* kps_r, const uint8_t* desc_r, int nr, float bf, float b,
                                float* u_right, float* depth) {
    if (!left || !right || nl < 0 || nr < 0 || !u_right || !depth) return fail(ORBX_ERR_ARG, "bad argument");
    if (left->lastFrames == 0 || right->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (frame < 0 || frame >= left->lastFrames || frame >= right->lastFrames) return fail(ORBX_ERR_ARG, "frame out of range");
    if (left->device != right->device) return fail(ORBX_ERR_ARG, "the two extractors live on different devices");
    const FramePlan& P = left->plan;
    if (P.W != right->plan.W || P.H != right->plan.H || P.nlevels != right->plan.nlevels || left->sf != right->sf)
        return fail(ORBX_ERR_ARG, "left and right extractors differ in image size or pyramid");
    if (nl == 0) return ORBX_OK;
    for (int i = 0; i < nl; ++i)
        if (kps_l[i].octave < 0 || kps_l[i].octave >= P.nlevels) return fail(ORBX_ERR_ARG, "left keypoint octave out of range");
    for (int i = 0; i < nr; ++i)
        if (kps_r[i].octave < 0 || kps_r[i].octave >= P.nlevels) return fail(ORBX_ERR_ARG, "right keypoint octave out of range");
    DeviceGuard guard(left->device);
    ORBX_CUDA(cudaStreamSynchronize(left->stream));
    ORBX_CUDA(cudaStreamSynchronize(right->stream));
    MatchScratch* S = nullptr;
    int rc = matchScratch(left->device, &S);
    if (rc != ORBX_OK) return rc;
    const size_t L = (size_t)nl, R = (size_t)std::max(nr, 1);
    ORBX_CUDA(S->q.reserve(L * 60 + 64));
    ORBX_CUDA(S->db.reserve(R * 60 + 64));
    ORBX_CUDA(S->misc.reserve((size_t)kMaxLevels * 4 + L * 20 + 128));
    cudaStream_t s = S->stream;
    uint8_t* dl = (uint8_t*)S->q.p; uint8_t* kl = dl + L * 32;
    uint8_t* dr = (uint8_t*)S->db.p; uint8_t* kr = dr + R * 32;
    float* sf = (float*)S->misc.p;
    int32_t* bi = (int32_t*)((uint8_t*)S->misc.p + kMaxLevels * 4);
    int32_t* bd = bi + L;
    int32_t* sad = bd + L;
    float* ur = (float*)(sad + L);
    float* dp = ur + L;
    ORBX_CUDA(cudaMemcpyAsync(dl, desc_l, L * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(kl, kps_l, L * 28, cudaMemcpyHostToDevice, s));
    if (nr) {
        ORBX_CUDA(cudaMemcpyAsync(dr, desc_r, (size_t)nr * 32, cudaMemcpyHostToDevice, s));
        ORBX_CUDA(cudaMemcpyAsync(kr, kps_r, (size_t)nr * 28, cudaMemcpyHostToDevice, s));
    }
    ORBX_CUDA(cudaMemcpyAsync(sf, left->sf.data(), (size_t)P.nlevels * 4, cudaMemcpyHostToDevice, s));
    const float maxD = bf / b;  // Frame.cc:841
    const int bandMax = stereoBandMax(*std::max_element(left->sf.begin(), left->sf.end()));
    ORBX_CUDA(S->k1.reserve(stereoRowScratchBytes(1, std::max(nr, 1), P.H, bandMax)));
    launchStereoBanded(kl, dl, nl, kr, dr, nr, sf, P.H, maxD, bi, bd, s, nullptr, S->k1.p, bandMax);  // nRows = mvImagePyramid[0].rows (:818)
    launchStereoRefine(P, left->sf.data(), left->isf.data(), left->buf.pyr + (size_t)frame * P.frameBytes,
                       right->buf.pyr + (size_t)frame * right->plan.frameBytes, kl, kr, nl, bi, bd, bf, maxD, ur, dp, sad, s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(u_right, ur, L * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(depth, dp, L * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

// All pairs of the two handles' last batches at once, everything read in place on the device (config 3 of BASELINE.json).
int orbx_compute_stereo_matches_batch(orbx_handle left, orbx_handle right, float bf, float b, float* u_right, float* depth) {
    if (!left || !right || !u_right || !depth) return fail(ORBX_ERR_ARG, "bad argument");
    if (left->lastFrames == 0 || right->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (left->lastFrames != right->lastFrames) return fail(ORBX_ERR_ARG, "left and right batches differ in length");
    if (left->device != right->device) return fail(ORBX_ERR_ARG, "the two extractors live on different devices");
    const FramePlan& P = left->plan;
    if (P.W != right->plan.W || P.H != right->plan.H || P.nlevels != right->plan.nlevels || left->sf != right->sf ||
        P.outCap != right->plan.outCap)
        return fail(ORBX_ERR_ARG, "left and right extractors differ in image size, pyramid or capacity");
    DeviceGuard guard(left->device);
    MatchScratch* S = nullptr;
    int rc = matchScratch(left->device, &S);
    if (rc != ORBX_OK) return rc;
    const int F = left->lastFrames, cap = P.outCap;
    const size_t T = (size_t)F * cap;
    ORBX_CUDA(S->misc.reserve((size_t)kMaxLevels * 4 + T * 20 + 128));
    cudaStream_t s = S->stream;
    // both extractions may still be running: order the scratch stream behind the two handles
    ORBX_CUDA(cudaEventRecord(left->ev[6], left->stream));
    ORBX_CUDA(cudaStreamWaitEvent(s, left->ev[6], 0));
    ORBX_CUDA(cudaEventRecord(right->ev[6], right->stream));
    ORBX_CUDA(cudaStreamWaitEvent(s, right->ev[6], 0));
    float* sf = (float*)S->misc.p;
    int32_t* bi = (int32_t*)((uint8_t*)S->misc.p + kMaxLevels * 4);
    int32_t* bd = bi + T;
    int32_t* sad = bd + T;
    float* ur = (float*)(sad + T);
    float* dp = ur + T;
    ORBX_CUDA(cudaMemcpyAsync(sf, left->sf.data(), (size_t)P.nlevels * 4, cudaMemcpyHostToDevice, s));
    const StereoBatch SB{left->buf.outN, right->buf.outN, cap, F, (size_t)P.frameBytes, (size_t)right->plan.frameBytes};
    const float maxD = bf / b;  // Frame.cc:841
    const int bandMax = stereoBandMax(*std::max_element(left->sf.begin(), left->sf.end()));
    ORBX_CUDA(S->k1.reserve(stereoRowScratchBytes(F, cap, P.H, bandMax)));
    launchStereoBanded(left->buf.outKps, left->buf.outDesc, 0, right->buf.outKps, right->buf.outDesc, 0, sf, P.H, maxD, bi, bd, s, &SB,
                       S->k1.p, bandMax);
    launchStereoRefine(P, left->sf.data(), left->isf.data(), left->buf.pyr, right->buf.pyr, left->buf.outKps, right->buf.outKps, 0, bi,
                       bd, bf, maxD, ur, dp, sad, s, &SB);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(u_right, ur, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(depth, dp, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

// ---- Frame::UndistortKeyPoints / ComputeImageBounds (row N1, Frame.cc:747-809) ----------------------------------------
static int undistortCam(const float* K, const float* dist, int ndist, const float* P, UndistortCam* C) {
    if (!K || !P || ndist < 0 || ndist > 12 || (ndist > 0 && !dist)) return fail(ORBX_ERR_ARG, "bad camera");
    if (ndist != 0 && ndist != 4 && ndist != 5 && ndist != 8 && ndist != 12)
        return fail(ORBX_ERR_ARG, "distortion vector must have 4, 5, 8 or 12 coefficients");  // cv::undistortPoints' own check
    if (K[0] == 0.0f || K[4] == 0.0f) return fail(ORBX_ERR_ARG, "camera matrix with a zero focal length");
    C->ifx = 1. / (double)K[0];
    C->ify = 1. / (double)K[4];
    C->cx = (double)K[2];
    C->cy = (double)K[5];
    for (int i = 0; i < 12; ++i) C->k[i] = i < ndist ? (double)dist[i] : 0.0;
    for (int i = 0; i < 9; ++i) C->rr[i] = (double)P[i];
    return ORBX_OK;
}

int orbx_undistort_keypoints(const orbx_keypoint* kps, int n, const float* K, const float* dist, int ndist, const float* P,
                             orbx_keypoint* kps_un, int device) {
    if (n < 0 || (n > 0 && (!kps || !kps_un))) return fail(ORBX_ERR_ARG, "bad argument");
    UndistortCam C;
    int rc = undistortCam(K, dist, ndist, P, &C);
    if (rc != ORBX_OK) return rc;
    rc = needDevice();
    if (rc != ORBX_OK) return rc;
    if (n == 0) return ORBX_OK;
    if (ndist == 0 || dist[0] == 0.0) {  // Frame.cc:749-753 mvKeysUn = mvKeys
        if (kps_un != kps) std::memcpy(kps_un, kps, (size_t)n * sizeof(orbx_keypoint));
        return ORBX_OK;
    }
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA(S->misc.reserve((size_t)n * sizeof(orbx_keypoint)));
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(S->misc.p, kps, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, s));
    launchUndistort(S->misc.p, S->misc.p, nullptr, n, 1, n, C, s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(kps_un, S->misc.p, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_image_bounds(int cols, int rows, const float* K, const float* dist, int ndist, const float* P, float bounds[4], int device) {
    if (!bounds || cols <= 0 || rows <= 0) return fail(ORBX_ERR_ARG, "bad argument");
    UndistortCam C;
    int rc = undistortCam(K, dist, ndist, P, &C);
    if (rc != ORBX_OK) return rc;
    if (ndist == 0 || dist[0] == 0.0) {  // Frame.cc:804-809
        bounds[0] = 0.0f; bounds[1] = (float)cols; bounds[2] = 0.0f; bounds[3] = (float)rows;
        return ORBX_OK;
    }
    orbx_keypoint c[4] = {};
    c[1].x = (float)cols; c[2].y = (float)rows; c[3].x = (float)cols; c[3].y = (float)rows;  // :786-790
    rc = orbx_undistort_keypoints(c, 4, K, dist, ndist, P, c, device);
    if (rc != ORBX_OK) return rc;
    bounds[0] = std::min(c[0].x, c[2].x);  // :799-802
    bounds[1] = std::max(c[1].x, c[3].x);
    bounds[2] = std::min(c[0].y, c[1].y);
    bounds[3] = std::max(c[2].y, c[3].y);
    return ORBX_OK;
}

int orbx_undistort_batch(orbx_handle h, const float* K, const float* dist, int ndist, const float* P, orbx_keypoint* kps_un,
                         const orbx_keypoint** d_kps_un) {
    if (!h) return fail(ORBX_ERR_ARG, "bad argument");
    if (h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    UndistortCam C;
    int rc = undistortCam(K, dist, ndist, P, &C);
    if (rc != ORBX_OK) return rc;
    DeviceGuard guard(h->device);
    const int F = h->lastFrames, cap = h->plan.outCap;
    const size_t bytes = (size_t)F * cap * sizeof(orbx_keypoint);
    ORBX_CUDA(h->mKpsUn.reserve(bytes));
    cudaStream_t s = h->stream;  // behind the extraction that produced the keypoints
    if (ndist == 0 || dist[0] == 0.0)
        ORBX_CUDA(cudaMemcpyAsync(h->mKpsUn.p, h->buf.outKps, bytes, cudaMemcpyDeviceToDevice, s));
    else
        launchUndistort(h->buf.outKps, h->mKpsUn.p, h->buf.outN, 0, F, cap, C, s);
    ORBX_CUDA(cudaGetLastError());
    h->kpsUnValid = true;
    if (d_kps_un) *d_kps_un = (const orbx_keypoint*)h->mKpsUn.p;
    if (kps_un) {
        ORBX_CUDA(cudaMemcpyAsync(kps_un, h->mKpsUn.p, bytes, cudaMemcpyDeviceToHost, s));
        ORBX_CUDA(cudaStreamSynchronize(s));
        return checkDeviceFlags(h);
    }
    return ORBX_OK;
}

// ---- ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763) ---------------------------------------------------------
int orbx_search_for_initialization(const orbx_keypoint* kps1, const uint8_t* desc1, int n1, const orbx_keypoint* kps2,
                                   const uint8_t* desc2, int n2, float min_x, float max_x, float min_y, float max_y, float* prev_matched,
                                   int window_size, float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                   int device) {
    if (n1 < 0 || n2 < 0 || !nmatches || (n1 > 0 && (!kps1 || !desc1 || !prev_matched || !matches12)) || (n2 > 0 && (!kps2 || !desc2)))
        return fail(ORBX_ERR_ARG, "bad argument");
    if (!(max_x > min_x) || !(max_y > min_y)) return fail(ORBX_ERR_ARG, "empty image bounds");
    if (n2 > gridSortCapacity()) return fail(ORBX_ERR_ARG, "more than 32768 keypoints in the second frame");
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    *nmatches = 0;
    if (n1 == 0) return ORBX_OK;
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    const size_t A = (size_t)n1, B = (size_t)std::max(n2, 1), cap = std::max(A, B);
    auto al = [](size_t v) { return (v + 15) & ~(size_t)15; };
    // q: [desc1 | kps1 | prev | m12 | n]   db: [desc2 | kps2 | sorted | cellStart | 3 scratch arrays of cap ints]
    const size_t oK1 = A * 32, oPrev = oK1 + al(A * 28), oM12 = oPrev + A * 8, oN = oM12 + al(A * 4), qBytes = oN + 64;
    const size_t oK2 = B * 32, oSo = oK2 + al(B * 28), oCs = oSo + al(B * 4), oScr = oCs + al((size_t)(64 * 48 + 1) * 4),
                 dbBytes = oScr + 3 * cap * 4 + 64;
    ORBX_CUDA(S->q.reserve(qBytes));
    ORBX_CUDA(S->db.reserve(dbBytes));
    uint8_t* Q = (uint8_t*)S->q.p;
    uint8_t* D = (uint8_t*)S->db.p;
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(Q, desc1, A * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oK1, kps1, A * 28, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oPrev, prev_matched, A * 8, cudaMemcpyHostToDevice, s));
    if (n2) {
        ORBX_CUDA(cudaMemcpyAsync(D, desc2, (size_t)n2 * 32, cudaMemcpyHostToDevice, s));
        ORBX_CUDA(cudaMemcpyAsync(D + oK2, kps2, (size_t)n2 * 28, cudaMemcpyHostToDevice, s));
    }
    // candidate lists of phase A (ORBX_INIT_LISTS=0: phase B walks every window itself, A/B measurements)
    static const bool useLists = [] { const char* e = std::getenv("ORBX_INIT_LISTS"); return !e || std::atoi(e) != 0; }();
    uint32_t* candList = nullptr;
    int* candCount = nullptr;
    if (useLists) {
        ORBX_CUDA(S->k1.reserve(cap * (size_t)(kSearchInitListCap + 1) * 4 + 64));
        candList = (uint32_t*)S->k1.p;
        candCount = (int*)(candList + cap * (size_t)kSearchInitListCap);
    }
    // stride = cap so that the three scratch arrays (pairs x stride each, pairs = 1) are laid out back to back
    launchSearchInit(Q + oK1, Q, n1, D + oK2, D, n2, nullptr, 1, (int)cap, min_x, max_x, min_y, max_y, window_size, nnratio,
                     check_orientation != 0, 0, (uint32_t*)(D + oSo), (int*)(D + oCs), (float*)(Q + oPrev), (int*)(Q + oM12),
                     (int*)(D + oScr), (int*)(Q + oN), s, candList, candCount);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(matches12, Q + oM12, A * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(prev_matched, Q + oPrev, A * 8, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(nmatches, Q + oN, 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_by_projection(const orbx_keypoint* kps_last, const uint8_t* last_flags, const float* world, const uint8_t* mp_desc,
                              const int32_t* n_last, const orbx_keypoint* kps_cur, const uint8_t* desc_cur, const float* cur_u_right,
                              const uint8_t* cur_taken, const int32_t* n_cur, const float* poses, int pairs, int cap_last, int cap_cur,
                              const orbx_proj_params* params, int32_t* match_cur, int32_t* nmatches, int device) {
    if (pairs < 0 || cap_last < 0 || cap_cur < 0 || !params || !n_last || !n_cur || !poses || !nmatches ||
        (cap_last > 0 && (!kps_last || !last_flags || !world || !mp_desc)) || (cap_cur > 0 && (!kps_cur || !desc_cur || !match_cur)))
        return fail(ORBX_ERR_ARG, "bad argument");
    if (!(params->max_x > params->min_x) || !(params->max_y > params->min_y)) return fail(ORBX_ERR_ARG, "empty image bounds");
    if (params->nlevels < 1 || params->nlevels > kMaxLevels) return fail(ORBX_ERR_ARG, "nlevels must be in [1,16]");
    if (cap_cur > gridSortCapacity()) return fail(ORBX_ERR_ARG, "more than 32768 keypoints in the current frame");
    for (int p = 0; p < pairs; ++p) {
        if (n_last[p] < 0 || n_last[p] > cap_last || n_cur[p] < 0 || n_cur[p] > cap_cur) return fail(ORBX_ERR_ARG, "count exceeds capacity");
        nmatches[p] = 0;
        // octaves index mvScaleFactors (:1723): reject what the reference would read out of bounds
        for (int i = 0; i < n_last[p]; ++i) {
            const int oct = kps_last[(size_t)p * cap_last + i].octave;
            if (oct < 0 || oct >= params->nlevels) return fail(ORBX_ERR_ARG, "last-frame keypoint octave outside [0, nlevels)");
        }
    }
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    if (pairs == 0 || cap_last == 0 || cap_cur == 0) {
        for (size_t i = 0; i < (size_t)pairs * cap_cur; ++i) match_cur[i] = -1;
        return ORBX_OK;
    }
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t P = (size_t)pairs, L = P * cap_last, Cn = P * cap_cur;
    // q (last side): [mpDesc | kps | world | flags | nLast | nCur | poses | nMatches]
    const size_t oKl = al(L * 32), oW = oKl + al(L * 28), oF = oW + al(L * 12), oNl = oF + al(L), oNc = oNl + al(P * 4), oPo = oNc + al(P * 4),
                 oNm = oPo + al(P * 56), qBytes = oNm + al(P * 4);
    // db (current side): [desc | kps | uRight | taken | sorted | cellStart | match]
    const size_t oKc = al(Cn * 32), oUr = oKc + al(Cn * 28), oTk = oUr + al(Cn * 4), oSo = oTk + al(Cn), oCs = oSo + al(Cn * 4),
                 oMa = oCs + al(P * (size_t)(64 * 48 + 1) * 4), dbBytes = oMa + al(Cn * 4);
    ORBX_CUDA(S->q.reserve(qBytes));
    ORBX_CUDA(S->db.reserve(dbBytes));
    ORBX_CUDA(S->k1.reserve(L * projScratchWordsPerLast() * 4 + 256));
    uint8_t* Q = (uint8_t*)S->q.p;
    uint8_t* D = (uint8_t*)S->db.p;
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(Q, mp_desc, L * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oKl, kps_last, L * 28, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oW, world, L * 12, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oF, last_flags, L, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oNl, n_last, P * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oNc, n_cur, P * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oPo, poses, P * 56, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(D, desc_cur, Cn * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(D + oKc, kps_cur, Cn * 28, cudaMemcpyHostToDevice, s));
    if (cur_u_right) ORBX_CUDA(cudaMemcpyAsync(D + oUr, cur_u_right, Cn * 4, cudaMemcpyHostToDevice, s));
    if (cur_taken) ORBX_CUDA(cudaMemcpyAsync(D + oTk, cur_taken, Cn, cudaMemcpyHostToDevice, s));
    launchGridSortBatch(D + oKc, (const int*)(Q + oNc), pairs, cap_cur, params->min_x, params->max_x, params->min_y, params->max_y,
                        (uint32_t*)(D + oSo), (int*)(D + oCs), s);
    ProjCam C{};
    C.fx = params->fx; C.fy = params->fy; C.cx = params->cx; C.cy = params->cy;
    C.minX = params->min_x; C.maxX = params->max_x; C.minY = params->min_y; C.maxY = params->max_y;
    C.mbf = params->mbf; C.mb = params->mb; C.th = params->th; C.mono = params->mono != 0; C.checkOri = params->check_orientation != 0;
    C.nlevels = params->nlevels;
    for (int l = 0; l < params->nlevels; ++l) C.sf[l] = params->scale_factors[l];
    launchSearchByProjection(Q + oKl, Q + oF, (const float*)(Q + oW), Q, (const int*)(Q + oNl), D + oKc, D,
                             cur_u_right ? (const float*)(D + oUr) : nullptr, cur_taken ? D + oTk : nullptr, (const int*)(Q + oNc),
                             (const float*)(Q + oPo), pairs, cap_last, cap_cur, C, (uint32_t*)(D + oSo), (int*)(D + oCs), (uint32_t*)S->k1.p,
                             (int*)(D + oMa), (int*)(Q + oNm), s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(match_cur, D + oMa, Cn * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(nmatches, Q + oNm, P * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_by_projection_map(const float* track, const int32_t* level, const uint8_t* flags, const uint8_t* mp_desc, const int32_t* n_mp,
                                  const orbx_keypoint* kps_cur, const uint8_t* desc_cur, const float* cur_u_right, const uint8_t* cur_taken,
                                  const int32_t* n_cur, int pairs, int cap_mp, int cap_cur, const orbx_proj_params* params, float nnratio,
                                  int far_points, float th_far_points, int32_t* match_cur, int32_t* nmatches, int device) {
    if (pairs < 0 || cap_mp < 0 || cap_cur < 0 || !params || !n_mp || !n_cur || !nmatches ||
        (cap_mp > 0 && (!track || !level || !flags || !mp_desc)) || (cap_cur > 0 && (!kps_cur || !desc_cur || !match_cur)))
        return fail(ORBX_ERR_ARG, "bad argument");
    if (!(params->max_x > params->min_x) || !(params->max_y > params->min_y)) return fail(ORBX_ERR_ARG, "empty image bounds");
    if (params->nlevels < 1 || params->nlevels > kMaxLevels) return fail(ORBX_ERR_ARG, "nlevels must be in [1,16]");
    if (cap_cur > gridSortCapacity()) return fail(ORBX_ERR_ARG, "more than 32768 keypoints in the frame");
    for (int p = 0; p < pairs; ++p) {
        if (n_mp[p] < 0 || n_mp[p] > cap_mp || n_cur[p] < 0 || n_cur[p] > cap_cur) return fail(ORBX_ERR_ARG, "count exceeds capacity");
        nmatches[p] = 0;
        for (int i = 0; i < n_mp[p]; ++i) {  // mnTrackScaleLevel indexes mvScaleFactors (:71)
            const int l = level[(size_t)p * cap_mp + i];
            if ((flags[(size_t)p * cap_mp + i] & 1) && (l < 0 || l >= params->nlevels))
                return fail(ORBX_ERR_ARG, "predicted level of a map point in view outside [0, nlevels)");
        }
    }
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    if (pairs == 0 || cap_mp == 0 || cap_cur == 0) {
        for (size_t i = 0; i < (size_t)pairs * cap_cur; ++i) match_cur[i] = -1;
        return ORBX_OK;
    }
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    auto al = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t P = (size_t)pairs, L = P * cap_mp, Cn = P * cap_cur;
    // q (map point side): [mpDesc | track | level | flags | nMp | nCur | nMatches]
    const size_t oTr = al(L * 32), oLv = oTr + al(L * 20), oF = oLv + al(L * 4), oNm = oF + al(L), oNc = oNm + al(P * 4), oNr = oNc + al(P * 4),
                 qBytes = oNr + al(P * 4);
    const size_t oKc = al(Cn * 32), oUr = oKc + al(Cn * 28), oTk = oUr + al(Cn * 4), oSo = oTk + al(Cn), oCs = oSo + al(Cn * 4),
                 oMa = oCs + al(P * (size_t)(64 * 48 + 1) * 4), dbBytes = oMa + al(Cn * 4);
    ORBX_CUDA(S->q.reserve(qBytes));
    ORBX_CUDA(S->db.reserve(dbBytes));
    ORBX_CUDA(S->k1.reserve(L * mapScratchWordsPerPoint() * 4 + 256));
    uint8_t* Q = (uint8_t*)S->q.p;
    uint8_t* D = (uint8_t*)S->db.p;
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaMemcpyAsync(Q, mp_desc, L * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oTr, track, L * 20, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oLv, level, L * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oF, flags, L, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oNm, n_mp, P * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(Q + oNc, n_cur, P * 4, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(D, desc_cur, Cn * 32, cudaMemcpyHostToDevice, s));
    ORBX_CUDA(cudaMemcpyAsync(D + oKc, kps_cur, Cn * 28, cudaMemcpyHostToDevice, s));
    if (cur_u_right) ORBX_CUDA(cudaMemcpyAsync(D + oUr, cur_u_right, Cn * 4, cudaMemcpyHostToDevice, s));
    if (cur_taken) ORBX_CUDA(cudaMemcpyAsync(D + oTk, cur_taken, Cn, cudaMemcpyHostToDevice, s));
    launchGridSortBatch(D + oKc, (const int*)(Q + oNc), pairs, cap_cur, params->min_x, params->max_x, params->min_y, params->max_y,
                        (uint32_t*)(D + oSo), (int*)(D + oCs), s);
    ProjCam C{};
    C.minX = params->min_x; C.maxX = params->max_x; C.minY = params->min_y; C.maxY = params->max_y;
    C.th = params->th; C.nlevels = params->nlevels;
    for (int l = 0; l < params->nlevels; ++l) C.sf[l] = params->scale_factors[l];
    launchSearchByProjectionMap((const float*)(Q + oTr), (const int*)(Q + oLv), Q + oF, Q, (const int*)(Q + oNm), D + oKc, D,
                                cur_u_right ? (const float*)(D + oUr) : nullptr, cur_taken ? D + oTk : nullptr, (const int*)(Q + oNc), pairs,
                                cap_mp, cap_cur, C, nnratio, far_points != 0, th_far_points, (uint32_t*)(D + oSo), (int*)(D + oCs),
                                (uint32_t*)S->k1.p, (int*)(D + oMa), (int*)(Q + oNr), s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(match_cur, D + oMa, Cn * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(nmatches, Q + oNr, P * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_for_initialization_batch(orbx_handle h, int use_undistorted, float min_x, float max_x, float min_y, float max_y,
                                         int window_size, float nnratio, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                         float* prev_matched) {
    if (!h || !matches12 || !nmatches) return fail(ORBX_ERR_ARG, "bad argument");
    if (h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (!(max_x > min_x) || !(max_y > min_y)) return fail(ORBX_ERR_ARG, "empty image bounds");
    const int F = h->lastFrames, cap = h->plan.outCap;
    if (cap > gridSortCapacity()) return fail(ORBX_ERR_ARG, "more than 32768 keypoints per frame");
    if (use_undistorted && (!h->kpsUnValid || h->mKpsUn.bytes < (size_t)F * cap * sizeof(orbx_keypoint)))
        return fail(ORBX_ERR_STATE, "orbx_undistort_batch has not run on this batch");
    if (F < 2) return ORBX_OK;
    DeviceGuard guard(h->device);
    MatchScratch* S = nullptr;
    int rc = matchScratch(h->device, &S);
    if (rc != ORBX_OK) return rc;
    const size_t P = (size_t)(F - 1), T = P * cap;
    // misc: [sorted F x cap | cellStart F x 3073 | prev P x cap x 2 | m12 P x cap | 3 scratch arrays P x cap | n P]
    const size_t oCs = (size_t)F * cap * 4, oPrev = oCs + (((size_t)F * (64 * 48 + 1) * 4 + 15) & ~(size_t)15), oM12 = oPrev + T * 8,
                 oScr = oM12 + T * 4, oN = oScr + 3 * T * 4, bytes = oN + P * 4 + 64;
    ORBX_CUDA(S->misc.reserve(bytes));
    uint8_t* M = (uint8_t*)S->misc.p;
    cudaStream_t s = S->stream;
    ORBX_CUDA(cudaEventRecord(h->ev[6], h->stream));  // behind the extraction (and the undistortion) on the handle's stream
    ORBX_CUDA(cudaStreamWaitEvent(s, h->ev[6], 0));
    const void* kps = use_undistorted ? h->mKpsUn.p : h->buf.outKps;
    static const bool useLists = [] { const char* e = std::getenv("ORBX_INIT_LISTS"); return !e || std::atoi(e) != 0; }();
    uint32_t* candList = nullptr;
    int* candCount = nullptr;
    if (useLists) {
        ORBX_CUDA(S->k1.reserve(T * (size_t)(kSearchInitListCap + 1) * 4 + 64));
        candList = (uint32_t*)S->k1.p;
        candCount = (int*)(candList + T * (size_t)kSearchInitListCap);
    }
    launchSearchInit(kps, h->buf.outDesc, 0, nullptr, nullptr, 0, h->buf.outN, F, cap, min_x, max_x, min_y, max_y, window_size, nnratio,
                     check_orientation != 0, 1, (uint32_t*)M, (int*)(M + oCs), (float*)(M + oPrev), (int*)(M + oM12), (int*)(M + oScr),
                     (int*)(M + oN), s, candList, candCount);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(matches12, M + oM12, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(nmatches, M + oN, P * 4, cudaMemcpyDeviceToHost, s));
    if (prev_matched) ORBX_CUDA(cudaMemcpyAsync(prev_matched, M + oPrev, T * 8, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

// ---- Frame::ComputeBoW (row N4) ------------------------------------------------------------------------------------
struct orbx_vocabulary {
    int device = 0;
    int nNodes = 0, nWords = 0, maxChildren = 0, L = 0;
    DevMem mBegin, mId, mDesc, mWeight, mWord;   // the flat tree
    DevMem mLeaf, mNid, mIn, mN, mOut;           // per-call scratch (grown on demand)
    cudaStream_t stream = nullptr;
    BowVocDev dev{};
};

int orbx_voc_create(int n_nodes, const int32_t* parent, const uint8_t* node_desc, const double* weight, int L, int device,
                    orbx_voc* out) {
    if (!out || n_nodes < 2 || !parent || !node_desc || !weight || L < 1) return fail(ORBX_ERR_ARG, "bad argument");
    // children in increasing node id = the order loadFromTextFile appends them (TemplatedVocabulary.h:1373-1388)
    std::vector<int> begin((size_t)n_nodes + 1, 0), ids((size_t)n_nodes - 1), wordOf((size_t)n_nodes, -1);
    for (int i = 1; i < n_nodes; ++i) {
        if (parent[i] < 0 || parent[i] >= i) return fail(ORBX_ERR_ARG, "vocabulary: parent[i] must be in [0, i)");
        ++begin[(size_t)parent[i] + 1];
    }
    int maxC = 0;
    for (int i = 0; i < n_nodes; ++i) {
        maxC = std::max(maxC, begin[(size_t)i + 1]);
        begin[(size_t)i + 1] += begin[i];
    }
    if (maxC >= (1 << 20)) return fail(ORBX_ERR_ARG, "vocabulary: more than 2^20 children under one node");
    std::vector<int> fill(begin.begin(), begin.end() - 1);
    std::vector<uint8_t> cdesc((size_t)(n_nodes - 1) * 32);
    for (int i = 1; i < n_nodes; ++i) {
        const int pos = fill[parent[i]]++;
        ids[pos] = i;
        std::memcpy(&cdesc[(size_t)pos * 32], node_desc + (size_t)i * 32, 32);
    }
    int nWords = 0;
    for (int i = 1; i < n_nodes; ++i)
        if (begin[(size_t)i + 1] == begin[i]) wordOf[i] = nWords++;  // words numbered in node order (:1405-1411)
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    int dev = device;
    if (dev < 0) ORBX_CUDA(cudaGetDevice(&dev));
    DeviceGuard guard(dev);
    orbx_vocabulary* v = new orbx_vocabulary();
    v->device = dev; v->nNodes = n_nodes; v->nWords = nWords; v->maxChildren = maxC; v->L = L;
    cudaError_t e = cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = v->mBegin.reserve(begin.size() * 4);
    if (e == cudaSuccess) e = v->mId.reserve(ids.size() * 4);
    if (e == cudaSuccess) e = v->mDesc.reserve(cdesc.size());
    if (e == cudaSuccess) e = v->mWeight.reserve((size_t)n_nodes * 8);
    if (e == cudaSuccess) e = v->mWord.reserve((size_t)n_nodes * 4);
    if (e == cudaSuccess) e = cudaMemcpy(v->mBegin.p, begin.data(), begin.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->mId.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->mDesc.p, cdesc.data(), cdesc.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->mWeight.p, weight, (size_t)n_nodes * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(v->mWord.p, wordOf.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        orbx_voc_destroy(v);
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, std::string("vocabulary upload: ") + cudaGetErrorString(e));
    }
    v->dev = BowVocDev{n_nodes, L, (const int*)v->mBegin.p, (const int*)v->mId.p, (const uint8_t*)v->mDesc.p,
                       (const double*)v->mWeight.p, (const int*)v->mWord.p};
    *out = v;
    return ORBX_OK;
}

void orbx_voc_destroy(orbx_voc v) {
    if (!v) return;
    DeviceGuard guard(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    DevMem* all[] = {&v->mBegin, &v->mId, &v->mDesc, &v->mWeight, &v->mWord, &v->mLeaf, &v->mNid, &v->mIn, &v->mN, &v->mOut};
    for (DevMem* m : all) m->release();
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orbx_voc_info(orbx_voc v, int* n_nodes, int* n_words, int* max_children, int* L) {
    if (!v) return fail(ORBX_ERR_ARG, "null vocabulary");
    if (n_nodes) *n_nodes = v->nNodes;
    if (n_words) *n_words = v->nWords;
    if (max_children) *max_children = v->maxChildren;
    if (L) *L = v->L;
    return ORBX_OK;
}

// descriptors [F][cap][32] and counts already on the device (on stream `after`'s work, joined through an event)
static int bowRun(orbx_vocabulary* v, const uint8_t* dDesc, const int* dN, int F, int cap, int levelsup, int32_t* word_ids,
                  double* word_vals, int32_t* n_words, int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx,
                  int32_t* n_nodes) {
    const size_t T = (size_t)F * cap;
    ORBX_CUDA(v->mLeaf.reserve(T * 4 + 16));
    ORBX_CUDA(v->mNid.reserve(T * 4 + 16));
    // out: [wordVals f64 T | wordIds T | nodeIds T | featIdx T | nodeBegin F*(cap+1) | nWords F | nNodes F]
    const size_t oIds = T * 8, oNode = oIds + T * 4, oFeat = oNode + T * 4, oBeg = oFeat + T * 4,
                 oNw = oBeg + (size_t)F * (cap + 1) * 4, oNn = oNw + (size_t)F * 4, bytes = oNn + (size_t)F * 4;
    ORBX_CUDA(v->mOut.reserve(bytes + 16));
    uint8_t* O = (uint8_t*)v->mOut.p;
    cudaStream_t s = v->stream;
    launchBowTransform(v->dev, dDesc, dN, cap, F, levelsup, (int*)v->mLeaf.p, (int*)v->mNid.p, (int32_t*)(O + oIds), (double*)O,
                       (int32_t*)(O + oNw), (int32_t*)(O + oNode), (int32_t*)(O + oBeg), (int32_t*)(O + oFeat), (int32_t*)(O + oNn), s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_CUDA(cudaMemcpyAsync(word_vals, O, T * 8, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(word_ids, O + oIds, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(node_ids, O + oNode, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(feat_idx, O + oFeat, T * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(node_begin, O + oBeg, (size_t)F * (cap + 1) * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(n_words, O + oNw, (size_t)F * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(n_nodes, O + oNn, (size_t)F * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

static int bowArgs(orbx_voc v, int levelsup, const void* a, const void* b, const void* c, const void* d, const void* e,
                   const void* f, const void* g) {
    if (!v) return fail(ORBX_ERR_ARG, "null vocabulary");
    if (levelsup < 0) return fail(ORBX_ERR_ARG, "levelsup must be >= 0");
    if (!a || !b || !c || !d || !e || !f || !g) return fail(ORBX_ERR_ARG, "null output");
    return ORBX_OK;
}

int orbx_bow_transform(orbx_voc v, const uint8_t* desc, const int32_t* n_per_frame, int nframes, int cap, int levelsup,
                       int32_t* word_ids, double* word_vals, int32_t* n_words, int32_t* node_ids, int32_t* node_begin,
                       int32_t* feat_idx, int32_t* n_nodes) {
    int rc = bowArgs(v, levelsup, word_ids, word_vals, n_words, node_ids, node_begin, feat_idx, n_nodes);
    if (rc != ORBX_OK) return rc;
    if (!desc || !n_per_frame || nframes < 1 || cap < 1) return fail(ORBX_ERR_ARG, "bad argument");
    if (cap > bowSortCapacity()) return fail(ORBX_ERR_ARG, "more than 16384 descriptors per frame");
    for (int f = 0; f < nframes; ++f)
        if (n_per_frame[f] < 0 || n_per_frame[f] > cap) return fail(ORBX_ERR_ARG, "n_per_frame out of range");
    DeviceGuard guard(v->device);
    const size_t T = (size_t)nframes * cap;
    ORBX_CUDA(v->mIn.reserve(T * 32));
    ORBX_CUDA(v->mN.reserve((size_t)nframes * 4));
    ORBX_CUDA(cudaMemcpyAsync(v->mIn.p, desc, T * 32, cudaMemcpyHostToDevice, v->stream));
    ORBX_CUDA(cudaMemcpyAsync(v->mN.p, n_per_frame, (size_t)nframes * 4, cudaMemcpyHostToDevice, v->stream));
    return bowRun(v, (const uint8_t*)v->mIn.p, (const int*)v->mN.p, nframes, cap, levelsup, word_ids, word_vals, n_words, node_ids,
                  node_begin, feat_idx, n_nodes);
}

int orbx_bow_transform_extracted(orbx_voc v, orbx_handle h, int levelsup, int32_t* word_ids, double* word_vals, int32_t* n_words,
                                 int32_t* node_ids, int32_t* node_begin, int32_t* feat_idx, int32_t* n_nodes) {
    int rc = bowArgs(v, levelsup, word_ids, word_vals, n_words, node_ids, node_begin, feat_idx, n_nodes);
    if (rc != ORBX_OK) return rc;
    if (!h || h->lastFrames == 0) return fail(ORBX_ERR_STATE, "no extraction yet");
    if (h->device != v->device) return fail(ORBX_ERR_ARG, "vocabulary and extractor live on different devices");
    if (h->plan.outCap > bowSortCapacity()) return fail(ORBX_ERR_ARG, "more than 16384 keypoint slots per frame");
    DeviceGuard guard(v->device);
    // the extraction may still be running on the handle's stream: order the vocabulary's stream behind it
    ORBX_CUDA(cudaEventRecord(h->ev[6], h->stream));
    ORBX_CUDA(cudaStreamWaitEvent(v->stream, h->ev[6], 0));
    return bowRun(v, h->buf.outDesc, h->buf.outN, h->lastFrames, h->plan.outCap, levelsup, word_ids, word_vals, n_words, node_ids,
                  node_begin, feat_idx, n_nodes);
}

// shared by the keyframe -> frame form (valid_f == NULL) and the keyframe <-> keyframe form (valid_f given, strict TH_LOW)
static int searchByBowImpl(const uint8_t* desc_kf, int n_kf, const uint8_t* valid_kf, const float* angle_kf, const int32_t* kf_node_ids,
                           const int32_t* kf_node_begin, const int32_t* kf_feat_idx, int kf_nodes, const uint8_t* desc_f, int n_f,
                           const uint8_t* valid_f, const float* angle_f, const int32_t* f_node_ids, const int32_t* f_node_begin,
                           const int32_t* f_feat_idx, int f_nodes, float nnratio, int check_orientation, int strict_low, int32_t* match_f,
                           int32_t* n_matches, int device) {
    if (n_kf < 0 || n_f < 0 || kf_nodes < 0 || f_nodes < 0 || !match_f || !n_matches) return fail(ORBX_ERR_ARG, "bad argument");
    if (n_f >= 65536) return fail(ORBX_ERR_ARG, "more than 65535 frame features");
    if ((n_kf > 0 && (!desc_kf || !valid_kf || !angle_kf)) || (n_f > 0 && (!desc_f || !angle_f)) ||
        (kf_nodes > 0 && (!kf_node_ids || !kf_node_begin || !kf_feat_idx)) || (f_nodes > 0 && (!f_node_ids || !f_node_begin || !f_feat_idx)))
        return fail(ORBX_ERR_ARG, "null input");
    const int kfEntries = kf_nodes ? kf_node_begin[kf_nodes] : 0, fEntries = f_nodes ? f_node_begin[f_nodes] : 0;
    if (kfEntries < 0 || kfEntries > n_kf || fEntries < 0 || fEntries > n_f) return fail(ORBX_ERR_ARG, "FeatureVector does not fit the descriptor sets");
    for (int p = 0; p < kfEntries; ++p)
        if (kf_feat_idx[p] < 0 || kf_feat_idx[p] >= n_kf) return fail(ORBX_ERR_ARG, "keyframe feature index out of range");
    for (int p = 0; p < fEntries; ++p)
        if (f_feat_idx[p] < 0 || f_feat_idx[p] >= n_f) return fail(ORBX_ERR_ARG, "frame feature index out of range");
    int rc = needDevice();
    if (rc != ORBX_OK) return rc;
    DeviceGuard guard(device);
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    MatchScratch* S = nullptr;
    rc = matchScratch(dev, &S);
    if (rc != ORBX_OK) return rc;
    // one upload buffer: [descKF | descF | angleKF | angleF | kf ids, begin, idx | f ids, begin, idx | validKF], then [matchF | binF | n]
    const size_t K = (size_t)std::max(n_kf, 1), Fn = (size_t)std::max(n_f, 1), KN = (size_t)kf_nodes, FN = (size_t)f_nodes;
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o += (bytes + 15) & ~(size_t)15; return at; };
    const size_t oDK = take(K * 32), oDF = take(Fn * 32), oAK = take(K * 4), oAF = take(Fn * 4), oKI = take(KN * 4 + 4), oKB = take(KN * 4 + 4),
                 oKX = take(K * 4), oFI = take(FN * 4 + 4), oFB = take(FN * 4 + 4), oFX = take(Fn * 4), oV = take(K), oVF = take(Fn);
    ORBX_CUDA(S->db.reserve(o + 64));
    ORBX_CUDA(S->misc.reserve(Fn * 8 + 64));
    uint8_t* D = (uint8_t*)S->db.p;
    int* matchF = (int*)S->misc.p;
    int* binF = matchF + Fn;
    int* nM = binF + Fn;
    cudaStream_t s = S->stream;
    auto up = [&](size_t at, const void* src, size_t bytes) {
        return bytes ? cudaMemcpyAsync(D + at, src, bytes, cudaMemcpyHostToDevice, s) : cudaSuccess;
    };
    ORBX_CUDA(up(oDK, desc_kf, (size_t)n_kf * 32)); ORBX_CUDA(up(oDF, desc_f, (size_t)n_f * 32));
    ORBX_CUDA(up(oAK, angle_kf, (size_t)n_kf * 4)); ORBX_CUDA(up(oAF, angle_f, (size_t)n_f * 4));
    ORBX_CUDA(up(oKI, kf_node_ids, KN * 4)); ORBX_CUDA(up(oKB, kf_node_begin, kf_nodes ? (KN + 1) * 4 : 0)); ORBX_CUDA(up(oKX, kf_feat_idx, (size_t)kfEntries * 4));
    ORBX_CUDA(up(oFI, f_node_ids, FN * 4)); ORBX_CUDA(up(oFB, f_node_begin, f_nodes ? (FN + 1) * 4 : 0)); ORBX_CUDA(up(oFX, f_feat_idx, (size_t)fEntries * 4));
    ORBX_CUDA(up(oV, valid_kf, (size_t)n_kf));
    if (valid_f) ORBX_CUDA(up(oVF, valid_f, (size_t)n_f));
    launchBowSearch(D + oDK, D + oV, (const float*)(D + oAK), (const int*)(D + oKI), (const int*)(D + oKB), (const int*)(D + oKX), kf_nodes,
                    D + oDF, n_f, (const float*)(D + oAF), (const int*)(D + oFI), (const int*)(D + oFB), (const int*)(D + oFX), f_nodes, nnratio,
                    check_orientation != 0, matchF, binF, nM, s, valid_f ? D + oVF : nullptr, strict_low);
    ORBX_CUDA(cudaGetLastError());
    if (n_f) ORBX_CUDA(cudaMemcpyAsync(match_f, matchF, (size_t)n_f * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaMemcpyAsync(n_matches, nM, 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA(cudaStreamSynchronize(s));
    return ORBX_OK;
}

int orbx_search_by_bow(const uint8_t* desc_kf, int n_kf, const uint8_t* valid_kf, const float* angle_kf, const int32_t* kf_node_ids,
                       const int32_t* kf_node_begin, const int32_t* kf_feat_idx, int kf_nodes, const uint8_t* desc_f, int n_f,
                       const float* angle_f, const int32_t* f_node_ids, const int32_t* f_node_begin, const int32_t* f_feat_idx,
                       int f_nodes, float nnratio, int check_orientation, int32_t* match_f, int32_t* n_matches, int device) {
    return searchByBowImpl(desc_kf, n_kf, valid_kf, angle_kf, kf_node_ids, kf_node_begin, kf_feat_idx, kf_nodes, desc_f, n_f, nullptr,
                           angle_f, f_node_ids, f_node_begin, f_feat_idx, f_nodes, nnratio, check_orientation, 0, match_f, n_matches, device);
}

int orbx_search_by_bow_kf(const uint8_t* desc1, int n1, const uint8_t* valid1, const float* angle1, const int32_t* node_ids1,
                          const int32_t* node_begin1, const int32_t* feat_idx1, int nodes1, const uint8_t* desc2, int n2,
                          const uint8_t* valid2, const float* angle2, const int32_t* node_ids2, const int32_t* node_begin2,
                          const int32_t* feat_idx2, int nodes2, float nnratio, int check_orientation, int32_t* match12,
                          int32_t* n_matches, int device) {
    if (n1 < 0 || n2 < 0 || !match12 || !n_matches || (n2 > 0 && !valid2)) return fail(ORBX_ERR_ARG, "bad argument");
    // the device keeps the match per feature of keyframe 2 (= vbMatched2 and the partner in one array); a point of keyframe 2 is
    // given away at most once (:826, :853), so the map inverts into vpMatches12
    std::vector<int32_t> m21((size_t)std::max(n2, 1), -1);
    const int rc = searchByBowImpl(desc1, n1, valid1, angle1, node_ids1, node_begin1, feat_idx1, nodes1, desc2, n2, valid2, angle2,
                                   node_ids2, node_begin2, feat_idx2, nodes2, nnratio, check_orientation, 1, m21.data(), n_matches, device);
    if (rc != ORBX_OK) return rc;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    for (int j = 0; j < n2; ++j)
        if (m21[j] >= 0 && m21[j] < n1) match12[m21[j]] = j;
    return ORBX_OK;
}

int orbx_bench_popc(int device, double* gpopc_per_s) {
    if (!gpopc_per_s) return fail(ORBX_ERR_ARG, "null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(ORBX_ERR_CUDA, "no CUDA device: this library has no CPU path");
    }
    DeviceGuard guard(device);
    const double r = benchPopcRate(nullptr);
    if (r < 0) return fail(ORBX_ERR_CUDA, "POPC microbenchmark failed");
    *gpopc_per_s = r;
    return ORBX_OK;
}

// ---- sharded database ------------------------------------------------------------------------------------
int orbx_comm_unique_id(uint8_t id[128]) {
    int rc = loadNccl();
    if (rc != ORBX_OK) return rc;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId u;
    ORBX_NCCL(gNccl.GetUniqueId(&u));
    std::memcpy(id, &u, 128);
    return ORBX_OK;
}

int orbx_comm_init(int rank, int nranks, const uint8_t id[128], int device, orbx_comm* out) {
    if (!out || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(ORBX_ERR_ARG, "bad argument");
    int rc = loadNccl();
    if (rc != ORBX_OK) return rc;
    if (device >= 0) ORBX_CUDA(cudaSetDevice(device));
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    orbx_comm_s* c = new orbx_comm_s();
    c->rank = rank; c->nranks = nranks; c->device = dev;
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    ncclResult_t r = gNccl.CommInitRank(&c->comm, nranks, u, rank);
    if (r != ncclSuccess) {
        delete c;
        return fail(ORBX_ERR_NCCL, std::string("ncclCommInitRank: ") + gNccl.GetErrorString(r));
    }
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        gNccl.CommDestroy(c->comm);
        delete c;
        return fail(ORBX_ERR_CUDA, cudaGetErrorString(e));
    }
    *out = c;
    return ORBX_OK;
}

void orbx_comm_destroy(orbx_comm c) {
    if (!c) return;
    DeviceGuard guard(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) gNccl.CommDestroy(c->comm);
    c->k1.release(); c->k2.release(); c->loc.release(); c->gathered.release();
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

// local top-2 on this rank's shard -> pack (k1[nq] | k2[nq]) -> ncclAllGather over NVLink -> merge on every rank, all enqueued on
// the communicator's stream: no host synchronisation (orbx_comm_sync waits).  Calls on one communicator are ordered by that stream,
// so the scratch buffers can be reused; nq must not grow between un-synchronised calls (a reallocation would free memory in use).
static int shardedEnqueue(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local, int64_t db_index_base,
                          int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2);

int orbx_match_top2_sharded_async(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local,
                                  int64_t db_index_base, int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2) {
    if (!c || nq < 1 || !d_q || nd_local < 0) return fail(ORBX_ERR_ARG, "bad argument");
    DeviceGuard guard(c->device);
    return shardedEnqueue(c, d_q, nq, d_db_local, nd_local, db_index_base, d_idx1, d_d1, d_idx2, d_d2);
}

int orbx_comm_sync(orbx_comm c) {
    if (!c) return fail(ORBX_ERR_ARG, "null communicator");
    DeviceGuard guard(c->device);
    ORBX_CUDA(cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

// Synchronous with respect to the host on return.
int orbx_match_top2_sharded(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local,
                            int64_t db_index_base, int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2) {
    if (!c || nq < 1 || !d_q || nd_local < 0) return fail(ORBX_ERR_ARG, "bad argument");
    DeviceGuard guard(c->device);
    const int rc = shardedEnqueue(c, d_q, nq, d_db_local, nd_local, db_index_base, d_idx1, d_d1, d_idx2, d_d2);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA(cudaStreamSynchronize(c->stream));
    return ORBX_OK;
}

static int shardedEnqueue(orbx_comm c, const uint8_t* d_q, int nq, const uint8_t* d_db_local, int64_t nd_local, int64_t db_index_base,
                          int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2) {
    cudaStream_t s = c->stream;
    int nsets = 1;
    int rc = localTop2Keys(d_q, nq, d_db_local, nd_local, db_index_base, c->k1, c->k2, &nsets, s);
    if (rc != ORBX_OK) return rc;
    // reduce this rank's chunks to one candidate set, written as keys into loc = [k1 | k2]
    ORBX_CUDA(c->loc.reserve((size_t)nq * 16));
    ORBX_CUDA(c->gathered.reserve((size_t)c->nranks * nq * 16));
    orbx::launchReduceKeys((const uint64_t*)c->k1.p, (const uint64_t*)c->k2.p, nsets, nq, (uint64_t*)c->loc.p,
                           (uint64_t*)c->loc.p + nq, s);
    ORBX_CUDA(cudaGetLastError());
    ORBX_NCCL(gNccl.AllGather(c->loc.p, c->gathered.p, (size_t)nq * 2, ncclUint64, c->comm, s));
    // gathered layout: [rank][k1[nq] | k2[nq]]  -> strided candidate sets
    orbx::launchUnpackTop2Strided((const uint64_t*)c->gathered.p, (const uint64_t*)c->gathered.p + nq, c->nranks, nq,
                                  (size_t)nq * 2, d_idx1, d_d1, d_idx2, d_d2, s);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}

}  // extern "C"

// ---- orbx_match_top2_device_noalloc: the top-2 matcher for callers that may not allocate ------------------------------------
// No counterpart in the reference; added (CUDA graph capture, tight serving loops and
// harnesses that forbid cudaMalloc).  Same kernels as orbx_match_top2_device (matchKernel<4> + mergeKernel); the per-chunk
// partial keys live in a file-scope device array instead of the per-thread DevMem scratch.  2 x 2^20 keys x 8 B = 16 MiB per
// GPU that loads the library, held until exit.  Every call shares the array: calls that overlap on the GPU race.
// matchKernel writes every [chunk][query < nq] slot mergeKernel reads (k_match.cu:106-115), so nothing has to start cleared;
// the nd = 0 case fills the first nq slots with the "no key" value itself.
namespace orbx {
constexpr size_t kNoAllocKeys = ORBX_NOALLOC_MAX_KEYS;
static __device__ __align__(256) unsigned long long gNoAllocKeys[2 * kNoAllocKeys];

constexpr int kNoAllocMaxDevices = 64;
static uint64_t* gNoAllocPtr[kNoAllocMaxDevices];  // device address of gNoAllocKeys per device (the symbol itself is a host-side stand-in)
static std::mutex gNoAllocMutex;

// once per device, with that device current
static int noAllocScratch(int dev, uint64_t** out) {
    if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(ORBX_ERR_ARG, "device ordinal beyond the no-alloc table");
    std::lock_guard<std::mutex> lock(gNoAllocMutex);
    if (!gNoAllocPtr[dev]) {
        void* p = nullptr;
        ORBX_CUDA(cudaGetSymbolAddress(&p, gNoAllocKeys));
        if (!p) return fail(ORBX_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        gNoAllocPtr[dev] = (uint64_t*)p;
    }
    *out = gNoAllocPtr[dev];
    return ORBX_OK;
}
}  // namespace orbx

extern "C" int orbx_match_top2_device_noalloc(const uint8_t* d_q, int nq, const uint8_t* d_db, int64_t nd, int64_t db_index_base,
                                              int64_t* d_idx1, int32_t* d_d1, int64_t* d_idx2, int32_t* d_d2, void* stream) {
    if (nq < 0 || nd < 0 || (nq > 0 && !d_q) || (nd > 0 && !d_db)) return fail(ORBX_ERR_ARG, "bad argument");
    if (nq == 0) return ORBX_OK;
    if (!d_idx1 || !d_d1 || !d_idx2 || !d_d2) return fail(ORBX_ERR_ARG, "bad argument");
    int dev = 0;
    ORBX_CUDA(cudaGetDevice(&dev));
    const int chunks = nd > 0 ? matchNumChunks(nq, nd, smCount()) : 0;
    const size_t need = (size_t)std::max(chunks, 1) * (size_t)nq;
    if (need > kNoAllocKeys) {
        char msg[192];
        std::snprintf(msg, sizeof msg, "orbx_match_top2_device_noalloc: nq (%d) x database chunks (%d) = %zu partial keys exceeds "
                      "ORBX_NOALLOC_MAX_KEYS (%zu); use orbx_match_top2_device", nq, chunks, need, kNoAllocKeys);
        return fail(ORBX_ERR_CAPACITY, msg);
    }
    uint64_t* keys = nullptr;
    const int rc = noAllocScratch(dev, &keys);
    if (rc != ORBX_OK) return rc;
    uint64_t *k1 = keys, *k2 = keys + kNoAllocKeys;
    cudaStream_t s = (cudaStream_t)stream;
    if (chunks == 0) {
        ORBX_CUDA(cudaMemsetAsync(k1, 0xff, (size_t)nq * 8, s));
        ORBX_CUDA(cudaMemsetAsync(k2, 0xff, (size_t)nq * 8, s));
    } else {
        launchMatchTop2(d_q, nq, d_db, nd, db_index_base, k1, k2, s);
        ORBX_CUDA(cudaGetLastError());
    }
    launchUnpackTop2(k1, k2, std::max(chunks, 1), nq, d_idx1, d_d1, d_idx2, d_d2, s);
    ORBX_CUDA(cudaGetLastError());
    return ORBX_OK;
}
