// k_fast_strip.cu — K-b, second generation: FAST-9/16 over STRIPS of the reference's cell grid.
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (ORB_SLAM3/src/ORBextractor.cc:787-872) and the
// cv::FAST(cell, kps, th, true) calls inside it (:826, :845), including the minThFAST redo of a cell that stayed empty at
// iniThFAST (:843-846).
//
// Why strips: cv::FAST's score is a function of the LEVEL image; the cell grid only decides (1) which
// threshold a cell ends up with, (2) that the 3x3 non-max suppression sees nothing across a cell seam and (3) the output
// order.  The tested rectangles of the cells tile the level exactly: with x' = x - 19, y' = y - 19 the pixel (x', y') in
// [0, width-6) x [0, height-6) (width = w - 32, the reference's maxBorderX - minBorderX) belongs to cell
// (y' / hCell, x' / wCell).  The per-cell kernel (k_fast.cu, one warp per 35-px cell) spends most of its instructions on
// per-cell fixed costs and on 16-byte chunks that straddle its 42-px tile (measured 1.49 warp-instructions per level
// pixel, the byte-SIMD quick test itself being 0.25).  Here one CTA owns a strip = several neighbouring cells of one cell
// row (up to 150 tested columns = 4 cells at the usual 36-px cells, one TMA box; up to 235 columns fit a box), so
//   * the quick test walks dense rows of 16-byte chunks (94 % of the lanes' pixels are tested pixels, was 56-75 %);
//   * the flagged pixels of the whole strip go to ONE list (order-free, shared-memory atomics) and the exact arc strength is
//     computed with every lane busy (was: per-cell lists of ~130 pixels in 64-pixel trips);
//   * corners (arc > t) go to a second list; the suppression compares each with its 8 neighbours in the strip's arc map,
//     ignoring the columns beyond the corner's own cell (rows beyond the cell row are never computed and stay 0);
//   * survivors get their place in the reference's output order (cell by cell, row-major inside the cell, :863-868) by
//     RANK: with at most 128 survivors per strip each one counts the smaller (cell, row, x) keys of the list; busier strips
//     let one warp per cell read its rectangle of a survivor bitmap row by row;
//   * a cell that stayed empty is redone by the same CTA at minThFAST on the tile it already holds (the arc map is
//     threshold-free and stays valid).
// A strip whose lists overflow (more than kListCap flagged pixels / kCornerCap corners: white noise) hands its cells to the
// per-cell kernel through a device-side list (k_fast.cu, list mode).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "orbx_internal.cuh"
#include "tma.cuh"
#include "fast_common.cuh"

namespace orbx {
namespace {

// Strip geometry and occupancy (measured, FAST stage ms / frames per s of the whole step; extra shared memory per CTA shows the
// sensitivity: 6 resident CTAs per SM 0.459 ms, 5 0.516, 4 0.518, 3 0.623, 2 0.846):
//   235 columns, 4096-entry list, 64 registers (6 CTAs / SM)   0.459-0.465 / 158.4 k
//   150 columns, 2560-entry list, 48 registers (9 CTAs / SM)   0.439 / 160.1 k        <- default
//   150 columns, 2048-entry list, 48 registers                 0.445 / 161.4 k
//   115 columns, 2048-entry list, 48 / 40 registers            0.473 / 0.498 (more strips: more fixed cost, small-level strips overflow)
// flagged pixels per strip (u16 entries: tile row << 8 | tile x) and corners per strip (3/8 of that): ORBX_STRIP_LIST
static int stripListCap() {
    static const int v = [] {
        const char* e = std::getenv("ORBX_STRIP_LIST");
        const int n = e ? std::atoi(e) : 0;
        return n >= 256 && n <= 16384 ? (n & ~63) : 2560;
    }();
    return v;
}
// widest strip in tested columns (one TMA box: 15 + width + 6 <= 256): ORBX_STRIP_MAXW
static int stripMaxWidth() {
    static const int v = [] {
        const char* e = std::getenv("ORBX_STRIP_MAXW");
        const int n = e ? std::atoi(e) : 0;
        return n >= 16 && n <= 235 ? n : 150;
    }();
    return v;
}
constexpr int kKeptCap = 128;     // survivors per strip ranked by counting; more = bitmap walk
constexpr int kBmWords = 8;       // bitmap words per row (tile pitch <= 256)
constexpr int kMaxStripCells = 32;

// one strip, everything derived on the host (16 ints = four 128-bit loads per thread instead of ~80 integer instructions)
struct StripDesc {
    int level;      // level | class << 8 (class 1 = the per-cell kernel's wide class) | (hCell + 6) << 16
    int pitch;      // tile pitch TP (bytes, multiple of 16, <= 256)
    int wCell;
    unsigned divW;  // 65536 / wCell + 1: (x * divW) >> 16 == x / wCell for x < 256
    int nCells;     // cells of the strip
    int testedW;    // tested columns of the strip (<= 0: nothing to test)
    int nRy;        // tested rows (<= 0: nothing to test)
    int shift;      // tile x of sub-image x = 0 (plane column of the strip's iniX, & 15)
    int gx, gy;     // plane column (16-byte aligned) / row of tile (0, 0)
    int cell0;      // index of the strip's first cell in the frame's cell arrays
    int cand0;      // word offset of that cell's slot in the frame's candidate block
    int cellCap;
    int yOut, xOut; // tile (row - 3, x) -> cell-grid coordinates of ORBextractor.cc:865-866
    int entry0;     // level << 24 | cell row << 12 | first cell column (fallback list entry)
};
static_assert(sizeof(StripDesc) == 64, "four 128-bit loads");

struct StripLayout {
    int tileBytes, arcBytes, bmBytes, total;
};
__host__ __device__ inline StripLayout stripLayout(int rowsMax, int pitchMax, int kListCap) {
    const int kCornerCap = kListCap * 3 / 8;
    StripLayout s;
    s.tileBytes = (rowsMax * pitchMax + 127) & ~127;
    s.arcBytes = ((rowsMax - 4) * pitchMax + 15) & ~15;
    s.bmBytes = (rowsMax - 6) * kBmWords * 4;
    s.total = s.tileBytes + s.arcBytes + s.bmBytes + 2 * kListCap + 2 * kCornerCap + 4 * kKeptCap + 4 * kMaxStripCells + 32 + 8;
    return s;
}

// FMAQ: the quick test's per-byte threshold adds and its mask packing as integer multiply-adds (IMAD, FMA pipe) instead of IADD / LEA /
// SHF (ALU pipe, the busier one: 56 % against 14 %); `one` is a kernel argument (= 1) so that the compiler keeps the multiply.
template <int NT, bool FMAQ, int MINB>
__global__ void __launch_bounds__(NT, MINB)
    fastStripKernel(const __grid_constant__ LevelMaps M, const StripDesc* __restrict__ stripTab, uint32_t* __restrict__ cand,
                    int* __restrict__ cellCnt, unsigned long long* __restrict__ fbList, int* __restrict__ fbCount, int fbCap,
                    int candWords, int nCellsTotal, int tIni, int tMin, int rowsMax, int pitchMax, int frame0, int kListCap, uint32_t one) {
    const int kCornerCap = kListCap * 3 / 8;
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int f = blockIdx.y;
    StripDesc D;
    {
        const int4* dp = reinterpret_cast<const int4*>(stripTab + blockIdx.x);
        int4* dd = reinterpret_cast<int4*>(&D);
#pragma unroll
        for (int k = 0; k < 4; ++k) dd[k] = __ldg(dp + k);
    }
    const int TP = D.pitch, P16 = TP >> 4, wCell = D.wCell;
    const int nCells = D.nCells, testedW = D.testedW, nRy = D.nRy, shift = D.shift;
    int* cntBase = cellCnt + (size_t)f * nCellsTotal + D.cell0;
    uint32_t* candBase = cand + (size_t)f * candWords + D.cand0;

    if (testedW <= 0 || nRy <= 0) {  // nothing to test: cv::FAST gets a sub-image below 7 x 7 or the cell is skipped (:813-823)
        if (tid < nCells) cntBase[tid] = 0;
        return;
    }

    const StripLayout lay = stripLayout(rowsMax, pitchMax, kListCap);
    uint8_t* tile = smem;                                   // [hCell + 6][TP]; tile x = sub-image x + shift
    uint8_t* arcStore = tile + lay.tileBytes;               // [nRy + 2][TP]: arc strength of tile rows 2 .. nRy + 3 (first / last stay 0)
    uint32_t* bitmap = reinterpret_cast<uint32_t*>(arcStore + lay.arcBytes);  // [nRy][kBmWords]: survivors, bit = tile x
    uint16_t* list = reinterpret_cast<uint16_t*>(bitmap + (rowsMax - 6) * kBmWords);
    uint16_t* corners = list + kListCap;
    uint32_t* keptList = reinterpret_cast<uint32_t*>(corners + kCornerCap);  // cell << 24 | row << 16 | x << 8 | arc
    int* cellKept = reinterpret_cast<int*>(keptList + kKeptCap);             // survivors per cell of the strip
    int* ctr = cellKept + kMaxStripCells;  // [0] flagged pixels, [1] corners, [2] overflow, [3] survivors
    uint64_t* bar = reinterpret_cast<uint64_t*>(ctr + 8);

    // ---- stage the strip: one TMA box, first byte on a 16-byte boundary of the plane row (the tile keeps the shift) ----
    // thread 0 alone owns the mbarrier (init, expect, wait); the block barrier behind its wait publishes the tile
    if (tid == 0) {
        mbarInit(bar, 1);
        mbarInitFence();
        mbarExpectTx(bar, (uint32_t)(TP * (D.level >> 16)));
        tmaLoadTile(tile, &M.m[D.level & 0xff], D.gx, D.gy, frame0 + f, bar);
    }
    {
        uint4* a4 = reinterpret_cast<uint4*>(arcStore);
        const int nA = (nRy + 2) * P16;
        for (int i = tid; i < nA; i += NT) a4[i] = make_uint4(0, 0, 0, 0);
        uint4* b4 = reinterpret_cast<uint4*>(bitmap);
        const int nB = nRy * (kBmWords / 4);
        for (int i = tid; i < nB; i += NT) b4[i] = make_uint4(0, 0, 0, 0);
        if (tid < kMaxStripCells + 8) cellKept[tid] = 0;  // + the counters behind it
    }
    if (tid == 0) mbarWait(bar, 0);
    __syncthreads();

    const int xs = shift + 3, xe = xs + testedW;  // tested tile columns; tested tile rows are [3, 3 + nRy)
    const uint32_t* tile32 = reinterpret_cast<const uint32_t*>(tile);
    const uint4* tile128 = reinterpret_cast<const uint4*>(tile);
    const unsigned divW = D.divW;

    // ---- quick reject of the tile columns [rxa, rxb) at threshold t, 16 pixels per lane (any 9-arc holds
    // ring[0] or ring[8] and ring[4] or ring[12]); the flagged pixels are appended to `list` ----------------------------------
    auto quickPhase = [&](int rxa, int rxb, int t) {
        const int c0 = rxa >> 4, nC = ((rxb - 1) >> 4) - c0 + 1;
        const int nItems = nRy * nC;
        const unsigned divC = 65536u / (unsigned)nC + 1u;  // (i * divC) >> 16 == i / nC for i < 2048, nC <= 32
        // per byte: ((d & 0x7f) + K) | d has bit 7 set iff d > min(t, 126) (a NECESSARY condition for t > 126)
        const uint32_t K = (uint32_t)(0x7f - min(t, 126)) * 0x01010101u;
        for (int i = tid; i < nItems; i += NT) {
            const int r = (int)(((unsigned)i * divC) >> 16);
            const int cx = c0 + (i - r * nC);
            const int idx = (r + 3) * P16 + cx;
            const uint4 cc = tile128[idx], up = tile128[idx - 3 * P16], dn = tile128[idx + 3 * P16];
            const uint32_t lw = tile32[4 * idx - 1], rw = tile32[4 * idx + 4];
            const uint32_t cw4[6] = {lw, cc.x, cc.y, cc.z, cc.w, rw};
            const uint32_t uw[4] = {up.x, up.y, up.z, up.w}, dw4[4] = {dn.x, dn.y, dn.z, dn.w};
            uint32_t m = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t ce = cw4[k + 1];
                const uint32_t lf = __byte_perm(cw4[k], ce, 0x4321);      // pixels x-3 .. x
                const uint32_t rt = __byte_perm(ce, cw4[k + 2], 0x6543);  // pixels x+3 .. x+6
                const uint32_t d0 = __vabsdiffu4(uw[k], ce), d8 = __vabsdiffu4(dw4[k], ce);
                const uint32_t d4 = __vabsdiffu4(rt, ce), d12 = __vabsdiffu4(lf, ce);
                if (FMAQ) {
                    const uint32_t v = (((d0 & 0x7f7f7f7fu) * one + K) | d0) | (((d8 & 0x7f7f7f7fu) * one + K) | d8);
                    const uint32_t h = (((d4 & 0x7f7f7f7fu) * one + K) | d4) | (((d12 & 0x7f7f7f7fu) * one + K) | d12);
                    const uint32_t F = v & h & 0x80808080u;
                    // flags at bits 7, 15, 23, 31 -> bits 28 .. 31: F * (2^21 + 2^14 + 2^7 + 1); the cross terms stay below bit 24
                    m = ((F * 0x00204081u) >> 28) * (one << (4 * k)) + m;
                } else {
                const uint32_t v = (((d0 & 0x7f7f7f7fu) + K) | d0) | (((d8 & 0x7f7f7f7fu) + K) | d8);
                const uint32_t h = (((d4 & 0x7f7f7f7fu) + K) | d4) | (((d12 & 0x7f7f7f7fu) + K) | d12);
                const uint32_t F = v & h & 0x80808080u;
                m |= (((F >> 7) * 0x01020408u) >> 24) << (4 * k);  // bit 4k+j = pixel 4k+j of the chunk
                }
            }
            const int x0 = cx << 4;
            const int lo = min(max(rxa - x0, 0), 16), hi = min(max(rxb - x0, 0), 16);
            m &= ((1u << hi) - 1u) & ~((1u << lo) - 1u);
            if (m) {
                const int n = __popc(m);
                const int base = atomicAdd(&ctr[0], n);
                if (base + n <= kListCap) {
                    // two pixels per trip (lowest / highest set bit); the order inside the list does not matter
                    int pos = base, end = base + n - 1;
                    const uint32_t yb = (uint32_t)((r + 3) << 8) | (uint32_t)x0;
                    while (m) {
                        const int jl = __ffs(m) - 1, jh = 31 - __clz(m);
                        list[pos++] = (uint16_t)(yb + jl);
                        list[end--] = (uint16_t)(yb + jh);
                        m &= m - 1;
                        m &= ~(1u << jh);
                    }
                } else {
                    ctr[2] = 1;
                }
            }
        }
    };

    // ---- exact arc strength of the listed pixels, two per lane, into the arc map; corners (arc > t) go to `corners` -------
    auto arcPhase = [&](int t) {
        const int nList = ctr[0];
        uint8_t* arc = arcStore - 2 * TP;  // indexed like the tile
        int rowOff[7];
#pragma unroll
        for (int d = 0; d < 7; ++d) rowOff[d] = (d - 3) * TP;
        for (int b = 0; b < nList; b += 2 * NT) {
            const int rem = min(nList - b, 2 * NT), half = (rem + 1) >> 1;
            if (tid >= half) break;  // the remaining pixels are split evenly: whole warps drop out of a short last trip
            const int j0 = b + tid, j1 = j0 + half;
            const bool v1 = tid + half < rem;
            const int p0 = list[j0], p1 = list[v1 ? j1 : j0];
            const int o0 = (p0 >> 8) * TP + (p0 & 255), o1 = (p1 >> 8) * TP + (p1 & 255);
            constexpr int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
            constexpr int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
            uint32_t q[16];
#pragma unroll
            for (int k = 0; k < 16; ++k)
                q[k] = (uint32_t)tile[o0 + rowOff[dy[k] + 3] + dx[k]] | ((uint32_t)tile[o1 + rowOff[dy[k] + 3] + dx[k]] << 16);
            const uint32_t cen = (uint32_t)tile[o0] | ((uint32_t)tile[o1] << 16);
            const uint32_t a = arcStrength2(q, cen);
            const int a0 = (int)(a & 0xffffu), a1 = (int)(a >> 16);
            arc[o0] = (uint8_t)a0;
            if (v1) arc[o1] = (uint8_t)a1;
            const bool c0k = a0 > t, c1k = v1 && a1 > t;
            if (c0k || c1k) {
                const int n = (int)c0k + (int)c1k;
                int k = atomicAdd(&ctr[1], n);
                if (k + n <= kCornerCap) {
                    if (c0k) corners[k++] = (uint16_t)p0;
                    if (c1k) corners[k] = (uint16_t)p1;
                } else {
                    ctr[2] = 1;
                }
            }
        }
    };

    // ---- strict 3x3 non-max suppression among the corners; nothing is seen across a cell seam (FAST ran on the cell) ----
    auto nmsPhase = [&]() {
        const int nCor = ctr[1];
        for (int j = tid; j < nCor; j += NT) {
            const int p = corners[j];
            const int row = p >> 8, x = p & 255;
            const uint8_t* ap = arcStore + (row - 2) * TP + x;
            const int a = ap[0];
            const int xrel = x - xs;
            const int cl = (int)(((unsigned)xrel * divW) >> 16);
            const int xin = xrel - cl * wCell;
            int m = max((int)ap[-TP], (int)ap[TP]);
            if (xin > 0) m = __vimax3_s32(m, __vimax3_s32(ap[-TP - 1], ap[-1], ap[TP - 1]), 0);
            if (xin < wCell - 1) m = __vimax3_s32(m, __vimax3_s32(ap[-TP + 1], ap[1], ap[TP + 1]), 0);
            if (a > m) {
                atomicOr(&bitmap[(row - 3) * kBmWords + (x >> 5)], 1u << (x & 31));
                atomicAdd(&cellKept[cl], 1);
                const int k = atomicAdd(&ctr[3], 1);
                if (k < kKeptCap) keptList[k] = ((uint32_t)cl << 24) | ((uint32_t)row << 16) | ((uint32_t)x << 8) | (uint32_t)a;
            }
        }
    };

    // ---- ordered output of the cells [clA, clB) (ORBextractor.cc:863-868: row-major inside the cell) -----------------------
    auto emitPhase = [&](int clA, int clB) {
        const int nK = ctr[3];
        if (nK <= kKeptCap) {
            // few survivors: the place of each = the number of smaller (cell, row, x) keys, minus those of the cells before its own
            for (int j = tid; j < nK; j += NT) {
                const uint32_t key = keptList[j];
                int rank = 0;
                for (int i = 0; i < nK; ++i) rank += keptList[i] < key ? 1 : 0;
                const int cl = (int)(key >> 24);
                for (int c = clA; c < cl; ++c) rank -= cellKept[c];
                const int row = (int)((key >> 16) & 0xffu), x = (int)((key >> 8) & 0xffu), a = (int)(key & 0xffu);
                if (rank < D.cellCap)
                    candBase[(size_t)cl * D.cellCap + rank] =
                        ((uint32_t)(a - 1) << 24) | ((uint32_t)(row - 3 + D.yOut) << 12) | (uint32_t)(x + D.xOut);
            }
            return;
        }
        const uint8_t* arcRows = arcStore + TP;  // arc strength of tested row r (tile row r + 3) at arcRows[r * TP + x]
        for (int cl = clA + warp; cl < clB; cl += NT / 32) {
            const int xa = xs + cl * wCell, xb = min(xa + wCell, xe);
            if (xb <= xa) continue;
            uint32_t* slot = candBase + (size_t)cl * D.cellCap;
            int kept = 0;
            const int wa = xa >> 5, wb = (xb - 1) >> 5;
            for (int r0 = 0; r0 < nRy; r0 += 32) {
                const int r = r0 + lane;
                uint32_t bits[4] = {0, 0, 0, 0};
                int n = 0;
                if (r < nRy) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int wi = wa + k;
                        if (wi <= wb) {
                            const int lo = max(xa - wi * 32, 0), hi = min(xb - wi * 32, 32);
                            const uint32_t mask = (hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u);
                            bits[k] = bitmap[r * kBmWords + wi] & mask;
                            n += __popc(bits[k]);
                        }
                    }
                }
                const unsigned any = __ballot_sync(0xffffffffu, n != 0);
                if (any == 0) continue;
                int incl = n;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += v;
                }
                int pos = kept + incl - n;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    uint32_t bk = bits[k];
                    while (bk) {
                        const int x = (wa + k) * 32 + __ffs(bk) - 1;
                        bk &= bk - 1;
                        const int a = arcRows[r * TP + x];
                        if (pos < D.cellCap)
                            slot[pos] = ((uint32_t)(a - 1) << 24) | ((uint32_t)(r + D.yOut) << 12) | (uint32_t)(x + D.xOut);
                        ++pos;
                    }
                }
                kept += __shfl_sync(0xffffffffu, incl, 31);
            }
        }
    };

    const int cls = (D.level >> 8) & 1;  // bits 16.. hold the tile rows
    auto handOver = [&](int clA, int clB, uint32_t flag) {  // cells [clA, clB) go to the per-cell kernel (it writes their counts)
        const int cl = clA + tid;
        if (cl < clB) {
            const int k = atomicAdd(&fbCount[cls], 1);
            if (k < fbCap) fbList[(size_t)cls * fbCap + k] = ((unsigned long long)f << 32) | flag | (uint32_t)(D.entry0 + cl);
        }
    };

    // ---- pass 1: the whole strip at iniThFAST -------------------------------------------------------------------------------
    quickPhase(xs, xe, tIni);
    __syncthreads();
    if (ctr[2]) {  // more flagged pixels than the list holds: the per-cell kernel takes the strip (both passes)
        handOver(0, nCells, 0u);
        return;
    }
    arcPhase(tIni);
    __syncthreads();
    if (ctr[2]) {
        handOver(0, nCells, 0u);
        return;
    }
    nmsPhase();
    __syncthreads();
    emitPhase(0, nCells);
    // counts; a cell with tested pixels but no keypoint is redone at minThFAST (:843-846)
    bool redo = false;
    if (tid < nCells) {
        const int xa = xs + tid * wCell;
        redo = cellKept[tid] == 0 && xa < xe && tMin < tIni;
        if (!redo) cntBase[tid] = cellKept[tid];
    }
    if (!__syncthreads_or(redo)) return;

    // ---- pass 2: the empty cells, one after the other, at minThFAST on the same tile ---------------------------------------
    for (int cl = 0; cl < nCells; ++cl) {
        const int xa = xs + cl * wCell, xb = min(xa + wCell, xe);
        if (cellKept[cl] != 0 || xa >= xe) continue;  // block-uniform (cellKept only changes between barriers)
        __syncthreads();  // the previous round's readers are done
        if (tid < 4) ctr[tid] = 0;
        __syncthreads();
        quickPhase(xa, xb, tMin);
        __syncthreads();
        bool over = ctr[2] != 0;
        if (!over) {
            arcPhase(tMin);
            __syncthreads();
            over = ctr[2] != 0;
        }
        if (over) {  // a single cell overflowing the lists (wide cells of tiny images): the per-cell kernel redoes it at minThFAST
            handOver(cl, cl + 1, 0x80000000u);
            continue;
        }
        nmsPhase();
        __syncthreads();
        emitPhase(cl, cl + 1);
        if (tid == 0) cntBase[cl] = cellKept[cl];
    }
}

int stripThreads() {
    static const int nt = [] {
        const char* e = std::getenv("ORBX_STRIP_THREADS");
        const int v = e ? std::atoi(e) : 0;
        return v == 128 || v == 256 ? v : 128;
    }();
    return nt;
}

}  // namespace

// cells per strip: as many as one TMA box (<= 256 bytes per row, first byte 16-byte aligned, 3-px apron on both sides) holds,
// then evened out over the strips of a cell row
void planFastStrips(FramePlan& P) {
    int tiles = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        LevelPlan& L = P.lv[l];
        L.stripCells = L.stripSegs = L.stripPitch = 0;
        if (L.nCols <= 0 || L.nRows <= 0) continue;
        const int maxCells = std::min(kMaxStripCells, std::max(1, stripMaxWidth() / L.wCell));
        L.stripSegs = (L.nCols + maxCells - 1) / maxCells;
        L.stripCells = (L.nCols + L.stripSegs - 1) / L.stripSegs;
        L.stripSegs = (L.nCols + L.stripCells - 1) / L.stripCells;
        L.stripPitch = (15 + L.stripCells * L.wCell + 6 + 15) & ~15;
        tiles += L.stripSegs * L.nRows;
    }
    P.stripTiles = tiles;
}

size_t fastStripDescBytes() { return sizeof(StripDesc); }

// the strips of a frame, the small levels first (their strips carry the most flagged pixels per tested pixel); 16 ints each
void buildFastStripTable(const FramePlan& P, std::vector<uint32_t>& tab) {
    tab.clear();
    for (int l = P.nlevels - 1; l >= 0; --l) {
        const LevelPlan& L = P.lv[l];
        for (int ci = 0; ci < L.nRows; ++ci)
            for (int sg = 0; sg < L.stripSegs; ++sg) {
                StripDesc D;
                const int cj0 = sg * L.stripCells;
                D.level = l | (fastWideLevel(L) ? 0x100 : 0) | ((L.hCell + 6) << 16);
                D.pitch = L.stripPitch;
                D.wCell = L.wCell;
                D.divW = 65536u / (unsigned)L.wCell + 1u;
                D.nCells = std::min(L.stripCells, L.nCols - cj0);
                // tested pixels relative to the level's tested origin (19, 19): columns [cj0 * wCell, .. + testedW), rows
                // [ci * hCell, .. + nRy); the last cell column / row is clipped by the reference's maxBorder (:809-812, :817-819)
                D.testedW = std::min(D.nCells * L.wCell, (L.w - 2 * kMinBorder - 6) - cj0 * L.wCell);
                D.nRy = std::min(L.hCell, (L.h - 2 * kMinBorder - 6) - ci * L.hCell);
                const int gx0 = kEdge + kMinBorder + cj0 * L.wCell;  // plane column of sub-image x = 0 (iniX of the first cell, :809)
                D.shift = gx0 & 15;
                D.gx = gx0 - D.shift;
                D.gy = kEdge + kMinBorder + ci * L.hCell;
                D.cell0 = L.cellBase + ci * L.nCols + cj0;
                D.cand0 = (int)(L.candOff + (unsigned)(ci * L.nCols + cj0) * (unsigned)L.cellCap);
                D.cellCap = L.cellCap;
                D.yOut = ci * L.hCell + 3;
                D.xOut = cj0 * L.wCell - D.shift;
                D.entry0 = (int)(((uint32_t)l << 24) | ((uint32_t)ci << 12) | (uint32_t)cj0);
                const uint32_t* w = reinterpret_cast<const uint32_t*>(&D);
                tab.insert(tab.end(), w, w + sizeof(StripDesc) / 4);
            }
    }
}

int buildStripMaps(const FramePlan& P, const uint8_t* pyrBase, int capFrames, void* out) {
    LevelMaps* M = reinterpret_cast<LevelMaps*>(out);
    std::memset(M, 0, sizeof(LevelMaps));
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelPlan& L = P.lv[l];
        if (L.stripSegs <= 0) continue;
        if (L.stripPitch > 256 || L.hCell + 6 > 256) return -2;
        const int rc = encodeLevelMap(&M->m[l], P, l, pyrBase, capFrames, L.stripPitch, L.hCell + 6);
        if (rc != 0) return rc;
    }
    return 0;
}

template <int NT, bool FMAQ, int MINB>
static void launchStripT(const FramePlan& P, const Buffers& B, int F, int rowsMax, int pitchMax, cudaStream_t s, int tile0, int nTiles) {
    StripLayout lay = stripLayout(rowsMax, pitchMax, stripListCap());
    // ORBX_STRIP_PAD_KB (measurements only): extra dynamic shared memory per CTA = fewer resident CTAs per SM (occupancy sensitivity)
    static const int padKb = [] { const char* e = std::getenv("ORBX_STRIP_PAD_KB"); return e ? std::atoi(e) : 0; }();
    lay.total += padKb * 1024;
    auto kern = fastStripKernel<NT, FMAQ, MINB>;
    if (lay.total > 40 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, lay.total);
    kern<<<dim3(nTiles, F), NT, lay.total, s>>>(*reinterpret_cast<const LevelMaps*>(B.stripMaps),
                                                     reinterpret_cast<const StripDesc*>(B.stripTab) + tile0, B.cand, B.cellCnt, B.fbList, B.fbCount,
                                                     B.fbCap, (int)P.candWords, P.nCellsTotal, std::max(P.iniTh, 1), std::max(P.minTh, 1),
                                                     rowsMax, pitchMax, B.frame0, stripListCap(), 1u);
    countLaunch();
}

void launchFastStrips(const FramePlan& P, const Buffers& B, int F, cudaStream_t s, int part, int splitLevel) {
    int rowsMax = 7, pitchMax = 16;
    // the strip table lists the levels from the smallest down to level 0: the strips of the levels > splitLevel come first
    int upper = 0;
    for (int l = P.nlevels - 1; l > splitLevel; --l) upper += P.lv[l].stripSegs * P.lv[l].nRows;
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelPlan& L = P.lv[l];
        if (L.stripSegs <= 0) continue;
        rowsMax = std::max(rowsMax, L.hCell + 6);
        pitchMax = std::max(pitchMax, L.stripPitch);
    }
    if (part == 0 || part == 4) {  // the list counts [2] and the per-cell kernel's two work counters
        cudaMemsetAsync(B.fbCount, 0, 2 * sizeof(int), s);
        cudaMemsetAsync(B.workCtr, 0, 2 * sizeof(int), s);
    }
    const int tile0 = part == 1 ? upper : 0;
    const int nTiles = part == 0 ? P.stripTiles : (part == 1 ? P.stripTiles - upper : (part == 2 ? upper : 0));  // parts 3, 4: none
    if (nTiles > 0) {
        static const bool fmaq = [] { const char* e = std::getenv("ORBX_STRIP_FMAQ"); return e && e[0] == '1'; }();
        // ORBX_STRIP_MINB: resident CTAs per SM the register allocation aims at (8 = 64 registers, 10 = 48, 12 = 40 with 44 bytes of spills)
        static const int minb = [] { const char* e = std::getenv("ORBX_STRIP_MINB"); return e ? std::atoi(e) : 10; }();
        if (stripThreads() == 256) launchStripT<256, false, 4>(P, B, F, rowsMax, pitchMax, s, tile0, nTiles);
        else if (fmaq) launchStripT<128, true, 8>(P, B, F, rowsMax, pitchMax, s, tile0, nTiles);
        else if (minb >= 12) launchStripT<128, false, 12>(P, B, F, rowsMax, pitchMax, s, tile0, nTiles);
        else if (minb >= 10) launchStripT<128, false, 10>(P, B, F, rowsMax, pitchMax, s, tile0, nTiles);
        else launchStripT<128, false, 8>(P, B, F, rowsMax, pitchMax, s, tile0, nTiles);
    }
    if (part == 0 || part == 3) launchFastFallback(P, B, F, B.fbList, B.fbCount, B.fbCap, s);
}

}  // namespace orbx
