// C ABI of the B200-native ORB extractor: handle management, geometry tables, launch sequence.
// Host-side logic mirrors ORBextractor's constructor and operator() (src/ORBextractor.cc:409-469, 1086-1168).
#include "orbx_kernels_desc.cuh"
#include "orbx_kernels_fast.cuh"
#include "orbx_kernels_stereo.cuh"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <algorithm>

using namespace orbx;

namespace {

const int kPatternTable[1024] = {
#include "../../include/orb_pattern_31.inc"
};

struct Extractor {
    orbx_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr, stream2 = nullptr;
    cudaStream_t aux[2] = {nullptr, nullptr};          // side streams: the blur runs beside FAST + quadtree
    cudaEvent_t evFork[2] = {nullptr, nullptr}, evJoin[2] = {nullptr, nullptr};
    // completion of the last pass of each slot and the stream it ran on: a pass on ANOTHER stream (a caller's stream handed to
    // orbx_extract_device) first waits for it - the handle's buffers and side streams are shared - and every entry point that
    // reads results waits for it, whatever stream the pass ran on (the handle's streams are non-blocking: the legacy-stream
    // copies would not order against them)
    cudaEvent_t evDone[2] = {nullptr, nullptr};
    cudaStream_t lastSt[2] = {nullptr, nullptr};
    bool overlap = true;
    // CUDA graphs of the launch sequence of one pass, keyed by everything the kernels' arguments depend on
    struct PassGraph { const uint8_t* base; size_t frameBytes; int pitch, n, lap0, lap1, fb, slot; cudaGraphExec_t exec; int kernels; };
    std::vector<PassGraph> graphs;
    bool useGraphs = true;
    bool blurEarly = true;           // blur forked right after the pyramid (beside FAST + quadtree); ORBX_BLUR_EARLY=0: beside the quadtree only
    int prio = 0;                    // launch priority of the latency-bound kernels (0 = none)
    int chunk = 0;                   // frames per pipelined sub-batch of the host path (0 = auto)
    std::string err;
    long long launches = 0;

    // ctor tables (host)
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> perLevel;

    // shape-dependent state
    int W = 0, H = 0;
    DevGeom geom;
    DevGeom* dGeom = nullptr;
    std::vector<FastCell> cells;
    FastCell* dCells = nullptr;
    ResizeTap* dTaps = nullptr;
    BlurTiles blurTiles;
    int blurTilesTotal = 0;
    int fastTilePitch = 0, fastTileRows = 0, fastThreads = kFastThreads;
    bool fastPitchConst = false;
    bool pyrFast[kMaxLevels] = {};                  // level may use k_pyr_level_fast (4 pixels read <= 8 source bytes)
    size_t fastSmem = 0, treeSmem = 0;
    int pyrRoots = 0;                // quadtree count pyramid: roots it is sized for (0 = off)
    // FAST cells in two launch groups: [0, fastNA) of dCellOrder = cells of at most fastRowsA tile rows (shared memory fastSmemA)
    int* dCellOrder = nullptr;
    int fastNA = 0, fastRowsA = 0, fastNA0 = 0;     // [0, fastNA0): the level-0 cells of group A
    bool fastEarly = false;          // ORBX_FAST_EARLY=1: FAST of level 0 launched beside the pyramid (e2e +1.7 %, device-resident -2 %)
    cudaEvent_t evStart[2] = {nullptr, nullptr}, evF0[2] = {nullptr, nullptr};
    size_t fastSmemA = 0;
    // second-generation FAST (orbx_kernels_fast.cuh): image-wide strength pass + per-cell select pass
    int fastGen = 2;                 // ORBX_FAST_GEN=1: the first-generation one-CTA-per-cell kernel (A/B switch)
    bool fastCompass = false;        // ORBX_FAST_COMPASS=1: warp-uniform compass early-out in the strength pass
    FsPlan fsPlan{};
    FsMaps fsMaps{}, scoreMaps{};    // tensor maps of the pyramid levels (level 0 per pass) and of the score planes
    FsMaps blurMaps{};               // the pyramid levels again, with the blur's 144 x 36 box
    FsMaps descMaps{};               // the BLURRED planes with the descriptor's 64 x 37 box
    FsMaps oriMaps{};                // the pyramid levels with the orientation disc's 48 x 31 box (level 0 per pass)
    FsMaps pyrMaps{};                // m[l] = SOURCE level l-1 with the box of the resize to level l (m[1] per pass)
    PyrBox pyrBox[kMaxLevels] = {};  // boxW = 0: level l keeps k_pyr_level_fast / k_pyr_level
    bool pyrTma = true;              // ORBX_PYR_TMA=0 (A/B switch)
    int smCount = 148;
    int treeThreadsSmall = 512, treeThreadsLarge = 128;   // ORBX_TREE_THREADS_SMALL / _LARGE (A/B switches; 128, 256 or 512)
    bool pyrChainOk = false;         // all levels of a few frames in ONE launch (k_pyr_chain_tma); ORBX_PYR_CHAIN=0 switches it off
    PyrChain pyrChain{};
    size_t pyrChainSmem = 0;
    unsigned int* dPyrArrive = nullptr;   // arrival counters [max_batch][kMaxLevels], zero at rest
    bool descTma = true;             // ORBX_DESC_TMA=0: k_orient_brief with global gathers (A/B switch)
    bool blurTma = true;             // ORBX_BLUR_TMA=0: the register-only k_blur7 (A/B switch; also the fallback when level 0 is not TMA-addressable)
    NmsTabs nmsTabs{};               // same-cell column / row masks of the non-max suppression
    uint2* dCbits = nullptr;         // survivor bits (k_fast_nms -> k_fast_select)
    int poolCap = 0, scanCap = 0;
    int maxLevelW4 = 0;

    std::vector<void*> allocs;       // everything cudaMalloc'ed for the current shape
    uint8_t* dLvl0 = nullptr;        // B x plane(0): staging for host images
    uint32_t* dCand = nullptr;
    int* dCellCount = nullptr;
    uint64_t *dQtA = nullptr, *dQtB = nullptr, *dLvlKp = nullptr;
    int* dLvlCnt = nullptr;
    orbx_keypoint* dOutKp = nullptr;
    uint8_t* dOutDesc = nullptr;
    int* dOutCnt = nullptr;
    WorkItem* dWork = nullptr;
    // pinned host staging
    uint8_t* hImg = nullptr;
    orbx_keypoint* hKp = nullptr;
    uint8_t* hDesc = nullptr;
    int* hCnt = nullptr;
    size_t outBlockBytes = 0;          // dOutKp | dOutDesc | dOutCnt (and hKp | hDesc | hCnt) are one allocation of this size
    int lastFrames = 0;
    Level0 lastL0{};

    bool timing = false;
    cudaEvent_t ev[8] = {};
    float stageMs[7] = {};
};

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            E->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return ORBX_ERR_CUDA;                                                                  \
        }                                                                                          \
    } while (0)

int cvRoundF(float v) { return (int)lrintf(v); }

// cuTensorMapEncodeTiled through the runtime (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn tensorMapEncoder() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// 3-D u8 tensor {x, y, frame} of one pyramid level for k_fast_score's tile loads.  false: TMA cannot address it (base, row
// pitch or frame stride not a multiple of 16 bytes) and the kernel stages that level with plain loads.
bool encodeLevelMap(CUtensorMap* m, const uint8_t* base, int w, int h, int frames, size_t pitch, size_t frameBytes, int boxH = kFsBoxH,
                    int boxW = kFsBoxW) {
    EncodeTiledFn enc = tensorMapEncoder();
    static const bool off = getenv("ORBX_FAST_TMA") && atoi(getenv("ORBX_FAST_TMA")) == 0;      // A/B switch: plain-load staging
    if (off || !enc || (((uintptr_t)base | pitch | frameBytes) & 15) != 0 || frames < 1) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)frames};
    const cuuint64_t strides[2] = {(cuuint64_t)pitch, (cuuint64_t)frameBytes};
    const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

void freeShape(Extractor* E) {
    for (auto& G : E->graphs) cudaGraphExecDestroy(G.exec);      // they hold pointers into the buffers freed below
    E->graphs.clear();
    for (void* p : E->allocs) cudaFree(p);
    E->allocs.clear();
    if (E->hImg) cudaFreeHost(E->hImg);
    if (E->hKp) cudaFreeHost(E->hKp);           // one block: hDesc and hCnt point into it
    E->hImg = nullptr; E->hKp = nullptr; E->hDesc = nullptr; E->hCnt = nullptr;
    E->W = E->H = 0;
}

template <class T>
int devAlloc(Extractor* E, T** p, size_t count) {
    void* q = nullptr;
    CK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    E->allocs.push_back(q);
    *p = (T*)q;
    return ORBX_OK;
}

// 11-bit bilinear taps of cv::resize; `rows` keeps the fraction at the ends (rows are clamped instead)
void makeTaps(int ssize, int dsize, bool rows, std::vector<ResizeTap>& out) {
    const double invScale = (double)dsize / ssize;
    const double sc = 1.0 / invScale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * sc - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (!rows) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        }
        ResizeTap t;
        t.ofs = s;
        t.c0 = (short)cvRoundF((1.f - f) * 2048.f);
        t.c1 = (short)cvRoundF(f * 2048.f);
        out.push_back(t);
    }
}

// (Re)build every shape-dependent table and buffer.
int configureShape(Extractor* E, int W, int H) {
    if (E->W == W && E->H == H) return ORBX_OK;
    freeShape(E);
    const int L = E->cfg.nlevels, B = E->cfg.max_batch;
    DevGeom& g = E->geom;
    memset(&g, 0, sizeof(g));
    g.nlevels = L; g.iniTh = E->cfg.ini_th_fast; g.minTh = E->cfg.min_th_fast;
    E->cells.clear();
    std::vector<ResizeTap> taps;
    size_t qtOff = 0; int kpOff = 0, candOff = 0, maxCells = 0, maxCellW = 0, maxCellH = 0, maxCellPx = 0;
    E->poolCap = 0;
    int tileStart = 0;
    for (int l = 0; l < L; ++l) {
        LevelGeom& G = g.lv[l];
        G.w = cvRoundF((float)W * E->invScale[l]);           // :1174-1175
        G.h = cvRoundF((float)H * E->invScale[l]);
        G.pitch = (G.w + 127) & ~127;
        G.planeBytes = ((size_t)G.pitch * G.h + 255) & ~(size_t)255;
        G.scale = E->scale[l];
        G.sizeF = (float)(int)(31 * E->scale[l]);            // scaledPatchSize (:876)
        G.quota = E->perLevel[l];
        const int maxBX = G.w - kEdge + 3, maxBY = G.h - kEdge + 3;
        const float width = (float)(maxBX - kMinBorder), height = (float)(maxBY - kMinBorder);
        if (width < 35.f || height < 35.f) { E->err = "pyramid level smaller than one 35-px FAST cell"; return ORBX_ERR_UNSUPPORTED; }
        G.nCols = (int)(width / 35.f); G.nRows = (int)(height / 35.f);
        G.wCell = (int)ceilf(width / G.nCols); G.hCell = (int)ceilf(height / G.nRows);
        G.treeW = maxBX - kMinBorder; G.treeH = maxBY - kMinBorder;
        G.nIni = (int)roundf((float)G.treeW / (float)G.treeH);   // :559
        if (G.nIni < 1) { E->err = "image too tall for the quadtree root layout (nIni = 0)"; return ORBX_ERR_UNSUPPORTED; }
        G.hX = (float)G.treeW / G.nIni;
        if (G.w > 65535 || G.h > 65535) { E->err = "level larger than 65535"; return ORBX_ERR_UNSUPPORTED; }
        G.cellBegin = (int)E->cells.size();
        G.candCap = 0;
        for (int i = 0; i < G.nRows; ++i) {
            const float iniY = (float)(kMinBorder + i * G.hCell);
            float maxY = iniY + G.hCell + 6;
            if (iniY >= maxBY - 3) continue;
            if (maxY > maxBY) maxY = (float)maxBY;
            for (int j = 0; j < G.nCols; ++j) {
                const float iniX = (float)(kMinBorder + j * G.wCell);
                float maxX = iniX + G.wCell + 6;
                if (iniX >= maxBX - 6) continue;
                if (maxX > maxBX) maxX = (float)maxBX;
                FastCell c;
                c.level = (short)l;
                c.x0 = (short)((int)iniX + 3); c.y0 = (short)((int)iniY + 3);
                c.w = (short)((int)maxX - (int)iniX - 6); c.h = (short)((int)maxY - (int)iniY - 6);
                if (c.w <= 0 || c.h <= 0) continue;           // cv::FAST tests nothing in a < 7 px wide window
                c.relX = (short)(j * G.wCell + 3); c.relY = (short)(i * G.hCell + 3);
                c.pad = 0;
                c.cap = ((c.w + 1) / 2) * ((c.h + 1) / 2);    // 3x3 strict maxima cannot be denser
                c.outOffset = candOff;
                candOff += c.cap; G.candCap += c.cap;
                if (c.w > 255 || c.h > 255 || ((c.w + 1) & ~1) * c.h > 32 * kFastThreads) { E->err = "FAST cell too large"; return ORBX_ERR_UNSUPPORTED; }
                maxCellPx = std::max<int>(maxCellPx, ((c.w + 1) & ~1) * c.h);      // pixel slots: rows padded to whole pairs
                maxCellW = std::max<int>(maxCellW, c.w); maxCellH = std::max<int>(maxCellH, c.h);
                E->cells.push_back(c);
            }
        }
        G.cellEnd = (int)E->cells.size();
        maxCells = std::max(maxCells, G.cellEnd - G.cellBegin);
        G.poolCap = std::max(G.quota + 3, 4 * G.nIni) + 5;
        E->poolCap = std::max(E->poolCap, G.poolCap);
        G.kpCap = G.poolCap; G.kpOffset = kpOff; kpOff += G.kpCap;
        G.qtOffset = qtOff; qtOff += (size_t)G.candCap + 32;
        if (l > 0) {
            const LevelGeom& P = g.lv[l - 1];
            G.areaFast = (P.w == 2 * G.w && P.h == 2 * G.h) ? 1 : 0;
            G.tabX = (int)taps.size(); makeTaps(P.w, G.w, false, taps);
            G.tabY = (int)taps.size(); makeTaps(P.h, G.h, true, taps);
            bool ok = !G.areaFast && (P.pitch & 3) == 0;
            for (int x = 0; x < G.w && ok; x += 4) {
                const int last = std::min(x + 3, G.w - 1);
                if (std::min(taps[G.tabX + last].ofs + 1, P.w - 1) - taps[G.tabX + x].ofs > 7) ok = false;
            }
            E->pyrFast[l] = ok;
            // box of the TMA-staged resize: the exact source extent of a 128 x 64 destination tile, from the tap tables
            E->pyrBox[l] = PyrBox{0, 0};
            if (ok) {
                int ex = 0, ey = 0;
                for (int X0 = 0; X0 < G.w; X0 += 128) {
                    const int lastGroup = std::min(X0 + 124, (G.w - 1) & ~3);
                    const int start = taps[G.tabX + X0].ofs & ~15;
                    const int end = (taps[G.tabX + lastGroup].ofs & ~3) + 11;       // the third word a thread reads
                    ex = std::max(ex, end - start + 1);
                }
                for (int Y0 = 0; Y0 < G.h; Y0 += kPrTileH) {
                    const int Yl = std::min(Y0 + kPrTileH - 1, G.h - 1);
                    const int gy = std::min(std::max(taps[G.tabY + Y0].ofs, 0), P.h - 1);
                    const int last = std::min(std::max(taps[G.tabY + Yl].ofs + 1, 0), P.h - 1);
                    ey = std::max(ey, last - gy + 1);
                }
                bool rowsPlain = true;      // k_pyr_level_tma walks the source rows s0, s0 + 1: no clamped row taps (never for a scale > 1)
                for (int y = 0; y < G.h; ++y)
                    if (taps[G.tabY + y].ofs < 0 || taps[G.tabY + y].ofs + 1 > P.h - 1 || (y && taps[G.tabY + y].ofs < taps[G.tabY + y - 1].ofs)) rowsPlain = false;
                const int bw = (ex + 15) & ~15;
                if (rowsPlain && bw <= 256 && ey <= 256 && (size_t)bw * ey <= 40 * 1024) E->pyrBox[l] = PyrBox{bw, ey};
            }
        }
        E->blurTiles.start[l] = tileStart;
        E->blurTiles.tilesX[l] = (G.w + kBlurStripW - 1) / kBlurStripW;
        tileStart += E->blurTiles.tilesX[l] * ((G.h + kBlurStripH - 1) / kBlurStripH);
    }
    E->blurTiles.start[L] = tileStart;
    E->blurTilesTotal = tileStart;
    g.nCells = (int)E->cells.size();
    g.candCapFrame = candOff; g.qtCapFrame = qtOff; g.kpCapFrame = kpOff; g.outCap = kpOff;
    if (E->poolCap > 60000) { E->err = "nfeatures too large"; return ORBX_ERR_UNSUPPORTED; }
    E->scanCap = std::max(std::max(E->poolCap, maxCells), 64) + 8;
    E->treeSmem = tree_smem_layout(E->poolCap, E->scanCap, nullptr, nullptr);
    E->fastTilePitch = (((maxCellW + 6 + 3 + 3) >> 2) + 1) * 2 + 1;      // words per row of each u16x2 pair copy (odd: spreads banks)
    E->fastPitchConst = E->fastTilePitch <= 33;      // the usual 35-px cells: compile-time pitch (k_fast_cells<., 33>)
    if (E->fastPitchConst) E->fastTilePitch = 33;
    E->fastTileRows = maxCellH + 6;
    E->fastSmem = (size_t)2 * E->fastTileRows * E->fastTilePitch * 4 + (((size_t)(maxCellH + 2) * (maxCellW + 4) + 3) & ~(size_t)3) + 16;
    {
        // The tallest cell (a coarse level with two or three cell rows) sets fastSmem; at EuRoC that is 18.7 KB per CTA = 12
        // CTAs per SM, although 86 % of the cells would fit 16 (the register limit of the 64-thread kernel).  Pick the largest
        // CTA count per SM k for which at least 60 % of the cells fit the per-CTA budget and launch those cells separately.
        auto need = [&](int h) {
            return (size_t)2 * (h + 6) * E->fastTilePitch * 4 + ((((size_t)(h + 2) * (maxCellW + 4)) + 3) & ~(size_t)3) + 16;
        };
        int hA = 0;
        for (int k = 16; k >= 13 && !hA; --k) {
            const size_t budget = (size_t)(233472 / k) / 128 * 128 - 1024 - 512;       // driver reserve + static, 128-B granules
            int h = maxCellH;
            while (h > 0 && need(h) > budget) --h;
            size_t fit = 0;
            for (const FastCell& c : E->cells) fit += c.h <= h;
            if (h > 0 && h < maxCellH && fit * 10 >= E->cells.size() * 6) hA = h;
            if (h >= maxCellH) break;                                                  // everything fits already
        }
        std::vector<int> order;
        // group A starts with its level-0 cells: they need no pyramid and are launched beside it (launchPass)
        for (size_t i = 0; i < E->cells.size(); ++i) if ((!hA || E->cells[i].h <= hA) && E->cells[i].level == 0) order.push_back((int)i);
        E->fastNA0 = (int)order.size();
        for (size_t i = 0; i < E->cells.size(); ++i) if ((!hA || E->cells[i].h <= hA) && E->cells[i].level != 0) order.push_back((int)i);
        E->fastNA = (int)order.size();
        for (size_t i = 0; i < E->cells.size(); ++i) if (hA && E->cells[i].h > hA) order.push_back((int)i);
        E->fastRowsA = hA ? hA + 6 : E->fastTileRows;
        E->fastSmemA = hA ? need(hA) : E->fastSmem;
        if (getenv("ORBX_FAST_GROUPS") && atoi(getenv("ORBX_FAST_GROUPS")) == 0) {     // A/B switch: one launch sized for the tallest cell
            E->fastNA = (int)order.size(); E->fastRowsA = E->fastTileRows; E->fastSmemA = E->fastSmem; E->fastNA0 = 0;
            for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
        }
        int rc2;
        if ((rc2 = devAlloc(E, &E->dCellOrder, order.size()))) return rc2;
        CK(cudaMemcpy(E->dCellOrder, order.data(), order.size() * sizeof(int), cudaMemcpyHostToDevice));
    }
    if (E->treeSmem > 220 * 1024) { E->err = "nfeatures too large for on-chip quadtree state"; return ORBX_ERR_UNSUPPORTED; }
    {   // count-pyramid form of the quadtree (orbx_kernels_tree.cuh): 8 bytes x 1365 cells per root on top of the list state
        int roots = 0;
        for (int l = 0; l < L; ++l) roots = std::max(roots, g.lv[l].nIni);
        const char* pe = getenv("ORBX_TREE_PYRAMID");                     // A/B switch
        const size_t withPyr = align16(E->treeSmem) + (size_t)roots * kPyrCellsPerRoot * 8;
        E->pyrRoots = (roots <= kPyrMaxRoots && withPyr <= 100 * 1024 && !(pe && atoi(pe) == 0)) ? roots : 0;
        if (E->pyrRoots) E->treeSmem = withPyr;
    }
    // A quadtree CTA is 512 threads x 64 registers: two of them hold an SM's whole register file for ~100 us while issuing at
    // ~40 %, and no CTA of another handle's kernels can co-reside.  ORBX_TREE_CTAS=k caps the resident quadtree CTAs per SM at k
    // by padding the dynamic shared memory request (the occupancy limiter that needs no recompilation).
    if (const char* tc = getenv("ORBX_TREE_CTAS")) {
        const int k = atoi(tc);
        if (k >= 1 && k <= 4) E->treeSmem = std::max<size_t>(E->treeSmem, (size_t)(227 * 1024 / (k + 1)) + 1024);
        if (E->treeSmem > 220 * 1024) E->treeSmem = 220 * 1024;
    }
    CK(raise_dyn_smem((const void*)k_octree, E->device, E->treeSmem));
    CK(raise_dyn_smem((const void*)k_debug_distribute, E->device, 200 * 1024));
    CK(raise_dyn_smem((const void*)k_orient_brief_tma, E->device, kDtSmemBytes));
    // fewer threads per cell when the cells are small: the per-thread fixed cost (prologue, scans, barriers) is spread over
    // more pixels; every thread owns a raster chunk of at most 32 pixels
    const char* ft = getenv("ORBX_FAST_THREADS");
    E->fastThreads = maxCellPx <= 32 * 64 ? 64 : maxCellPx <= 32 * 128 ? 128 : 256;
    if (ft && (atoi(ft) == 64 || atoi(ft) == 128 || atoi(ft) == 256) && maxCellPx <= 32 * atoi(ft)) E->fastThreads = atoi(ft);
    CK(raise_dyn_smem((const void*)k_fast_cells<64, 33>, E->device, E->fastSmem));
    CK(raise_dyn_smem((const void*)k_fast_cells<64, 0>, E->device, E->fastSmem));
    CK(raise_dyn_smem((const void*)k_fast_cells<128, 33>, E->device, E->fastSmem));
    CK(raise_dyn_smem((const void*)k_fast_cells<128, 0>, E->device, E->fastSmem));
    CK(raise_dyn_smem((const void*)k_fast_cells<256, 33>, E->device, E->fastSmem));
    CK(raise_dyn_smem((const void*)k_fast_cells<256, 0>, E->device, E->fastSmem));

    // strength-pass plan: per level ceil(testedW / 64) x ceil(testedH / 28) warp jobs, fine levels first
    {
        FsPlan& P = E->fsPlan;
        memset(&P, 0, sizeof(P));
        int job = 0, bitsOff = 0;
        for (int l = 0; l < L; ++l) {
            const LevelGeom& G = g.lv[l];
            P.lv[l].jobBegin = job;
            P.lv[l].nCx = (G.w - kEdge - kFsX0 + kFsCols - 1) / kFsCols;
            P.lv[l].nSy = (G.h - 2 * kEdge + 2 * kFsRows - 1) / (2 * kFsRows);
            P.lv[l].bitsOffset = bitsOff;
            job += P.lv[l].nCx * P.lv[l].nSy;
            bitsOff += 2 * P.lv[l].nSy * 16 * P.lv[l].nCx;
        }
        for (int l = L; l < kMaxLevels; ++l) P.lv[l].jobBegin = job;
        P.nJobs = job; P.bitsPerFrame = bitsOff;
        for (const FastCell& c : E->cells)
            if (c.w * c.h > 32 * kSelMaxWords) { E->err = "FAST cell too large"; return ORBX_ERR_UNSUPPORTED; }
    }

    // device buffers
    for (int l = 0; l < L; ++l) {
        int rc;
        if (l > 0) { if ((rc = devAlloc(E, &g.lvl[l], g.lv[l].planeBytes * B + 256))) return rc; }
        if ((rc = devAlloc(E, &g.blur[l], g.lv[l].planeBytes * B + 256))) return rc;
        if ((rc = devAlloc(E, &g.score[l], g.lv[l].planeBytes * B + 256))) return rc;
        CK(cudaMemset(g.score[l], 0, g.lv[l].planeBytes * B + 256));
    }
    int rc;
    if ((rc = devAlloc(E, &E->dLvl0, g.lv[0].planeBytes * B + 256))) return rc;
    g.lvl[0] = E->dLvl0;
    if ((rc = devAlloc(E, &E->dCand, (size_t)g.candCapFrame * B))) return rc;
    if ((rc = devAlloc(E, &E->dCellCount, (size_t)g.nCells * B))) return rc;
    if ((rc = devAlloc(E, &E->dCbits, (size_t)E->fsPlan.bitsPerFrame * B))) return rc;
    if ((rc = devAlloc(E, &E->dPyrArrive, (size_t)B * kMaxLevels))) return rc;
    CK(cudaMemset(E->dPyrArrive, 0, (size_t)B * kMaxLevels * sizeof(unsigned int)));
    memset(&E->fsMaps, 0, sizeof(E->fsMaps));
    memset(&E->scoreMaps, 0, sizeof(E->scoreMaps));
    memset(&E->blurMaps, 0, sizeof(E->blurMaps));
    memset(&E->descMaps, 0, sizeof(E->descMaps));
    memset(&E->oriMaps, 0, sizeof(E->oriMaps));
    memset(&E->pyrMaps, 0, sizeof(E->pyrMaps));
    for (int l = 0; l < L; ++l) {
        if (l > 0) E->fsMaps.tmaOk[l] = encodeLevelMap(&E->fsMaps.m[l], g.lvl[l], g.lv[l].w, g.lv[l].h, B, (size_t)g.lv[l].pitch, g.lv[l].planeBytes) ? 1 : 0;
        E->scoreMaps.tmaOk[l] = encodeLevelMap(&E->scoreMaps.m[l], g.score[l], g.lv[l].w, g.lv[l].h, B, (size_t)g.lv[l].pitch, g.lv[l].planeBytes, kNmsBoxH) ? 1 : 0;
        E->descMaps.tmaOk[l] = encodeLevelMap(&E->descMaps.m[l], g.blur[l], g.lv[l].w, g.lv[l].h, B, (size_t)g.lv[l].pitch, g.lv[l].planeBytes, kDtBoxH, kDtBoxW) ? 1 : 0;
        if (l > 1 && E->pyrBox[l].boxW)
            E->pyrMaps.tmaOk[l] = encodeLevelMap(&E->pyrMaps.m[l], g.lvl[l - 1], g.lv[l - 1].w, g.lv[l - 1].h, B, (size_t)g.lv[l - 1].pitch,
                                                 g.lv[l - 1].planeBytes, E->pyrBox[l].boxH, E->pyrBox[l].boxW) ? 1 : 0;
        if (l > 0) E->oriMaps.tmaOk[l] = encodeLevelMap(&E->oriMaps.m[l], g.lvl[l], g.lv[l].w, g.lv[l].h, B, (size_t)g.lv[l].pitch, g.lv[l].planeBytes, kOtBoxH, kOtBoxW) ? 1 : 0;
        if (l > 0) E->blurMaps.tmaOk[l] = encodeLevelMap(&E->blurMaps.m[l], g.lvl[l], g.lv[l].w, g.lv[l].h, B, (size_t)g.lv[l].pitch, g.lv[l].planeBytes, kBtBoxH, kBtBoxW) ? 1 : 0;
    }
    cudaDeviceGetAttribute(&E->smCount, cudaDevAttrMultiProcessorCount, E->device);
    for (int which = 0; which < 2; ++which)
        if (const char* tt = getenv(which ? "ORBX_TREE_THREADS_LARGE" : "ORBX_TREE_THREADS_SMALL")) {
            const int v = atoi(tt);
            if (v == 128 || v == 256 || v == 512) (which ? E->treeThreadsLarge : E->treeThreadsSmall) = v;
        }
    {   // chained pyramid: usable when every level takes the TMA kernel
        const char* ce = getenv("ORBX_PYR_CHAIN");
        bool ok = E->pyrTma && L > 1 && !(ce && atoi(ce) == 0);
        int tiles = 0; size_t smem = 0;
        for (int l = 1; l < L && ok; ++l) {
            if (!E->pyrFast[l] || E->pyrBox[l].boxW == 0 || (l > 1 && !E->pyrMaps.tmaOk[l])) { ok = false; break; }
            E->pyrChain.tileBegin[l] = tiles;
            E->pyrChain.tilesX[l] = (g.lv[l].w + 127) / 128;
            E->pyrChain.box[l] = E->pyrBox[l];
            tiles += E->pyrChain.tilesX[l] * ((g.lv[l].h + kPrTileH - 1) / kPrTileH);
            smem = std::max(smem, (size_t)E->pyrBox[l].boxW * E->pyrBox[l].boxH);
        }
        E->pyrChain.tileBegin[L] = tiles;
        E->pyrChainOk = ok && tiles > 0;
        E->pyrChainSmem = smem;
    }
    {
        // same-cell masks of the 3x3 suppression.  The cells of a level form a regular grid (src/ORBextractor.cc:805-821: the
        // column range depends only on j, the row range only on i), so "neighbour in the same cell" = column test AND row test.
        // col[x]: bit 0 = x-1 is in x's cell column, bit 1 = x+1 is, bit 2 = x is in some cell column.
        // row[y]: {0 / ~0: the row above counts, the row below counts, 0 / 0x80008000: y is in some cell row, 0}
        std::vector<uint8_t> col; std::vector<uint4> row;
        NmsTabs& T = E->nmsTabs;
        for (int l = 0; l < L; ++l) {
            const int nc = kFsX0 + kFsCols * E->fsPlan.lv[l].nCx + 8, nr = kFsY0 + 2 * kFsRows * E->fsPlan.lv[l].nSy + 8;
            T.colOff[l] = (int)col.size(); T.rowOff[l] = (int)row.size();
            col.resize(col.size() + ((nc + 3) & ~3), 0);
            row.resize(row.size() + nr, make_uint4(0, 0, 0, 0));
            uint8_t* C = col.data() + T.colOff[l]; uint4* R = row.data() + T.rowOff[l];
            for (int ci = g.lv[l].cellBegin; ci < g.lv[l].cellEnd; ++ci) {
                const FastCell& c = E->cells[ci];
                for (int x = c.x0; x < c.x0 + c.w; ++x) C[x] = (uint8_t)(4 | (x > c.x0 ? 1 : 0) | (x < c.x0 + c.w - 1 ? 2 : 0));
                for (int y = c.y0; y < c.y0 + c.h; ++y) R[y] = make_uint4(y > c.y0 ? ~0u : 0u, y < c.y0 + c.h - 1 ? ~0u : 0u, 0x80008000u, 0u);
            }
        }
        uint8_t* dCol = nullptr; uint4* dRow = nullptr;
        if ((rc = devAlloc(E, &dCol, col.size()))) return rc;
        if ((rc = devAlloc(E, &dRow, row.size()))) return rc;
        CK(cudaMemcpy(dCol, col.data(), col.size(), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dRow, row.data(), row.size() * sizeof(uint4), cudaMemcpyHostToDevice));
        T.col = dCol; T.row = dRow;
    }
    if ((rc = devAlloc(E, &E->dQtA, g.qtCapFrame * B))) return rc;
    if ((rc = devAlloc(E, &E->dQtB, g.qtCapFrame * B))) return rc;
    if ((rc = devAlloc(E, &E->dLvlKp, (size_t)g.kpCapFrame * B))) return rc;
    if ((rc = devAlloc(E, &E->dLvlCnt, (size_t)L * B))) return rc;
    // keypoints | descriptors | {n, monoIndex} of all frames in ONE block (device and page-locked host): a single frame's
    // results come back in one copy instead of three
    const size_t outKpBytes = ((size_t)g.outCap * B * sizeof(orbx_keypoint) + 255) & ~(size_t)255;
    const size_t outDsBytes = ((size_t)g.outCap * B * 32 + 255) & ~(size_t)255;
    E->outBlockBytes = outKpBytes + outDsBytes + (size_t)2 * B * sizeof(int);
    uint8_t* outBlock = nullptr;
    if ((rc = devAlloc(E, &outBlock, E->outBlockBytes))) return rc;
    E->dOutKp = reinterpret_cast<orbx_keypoint*>(outBlock);
    E->dOutDesc = outBlock + outKpBytes;
    E->dOutCnt = reinterpret_cast<int*>(outBlock + outKpBytes + outDsBytes);
    if ((rc = devAlloc(E, &E->dWork, (size_t)g.outCap * B))) return rc;
    if ((rc = devAlloc(E, &E->dGeom, 1))) return rc;
    if ((rc = devAlloc(E, &E->dCells, E->cells.size()))) return rc;
    if ((rc = devAlloc(E, &E->dTaps, taps.size()))) return rc;
    CK(cudaMemcpy(E->dGeom, &g, sizeof(g), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(E->dCells, E->cells.data(), E->cells.size() * sizeof(FastCell), cudaMemcpyHostToDevice));
    if (!taps.empty()) CK(cudaMemcpy(E->dTaps, taps.data(), taps.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
    CK(cudaMallocHost((void**)&E->hImg, g.lv[0].planeBytes * B));
    uint8_t* hostBlock = nullptr;
    CK(cudaMallocHost((void**)&hostBlock, E->outBlockBytes));
    E->hKp = reinterpret_cast<orbx_keypoint*>(hostBlock);
    E->hDesc = hostBlock + outKpBytes;
    E->hCnt = reinterpret_cast<int*>(hostBlock + outKpBytes + outDsBytes);
    E->W = W; E->H = H;
    E->lastFrames = 0;
    return ORBX_OK;
}

// The launch sequence of one pass over frames [fb, fb+n) of the handle's buffers; level 0 of frame fb+i lives at
// l0.base + i*l0.frameBytes.
int launchPass(Extractor* E, Level0 l0, int n, int lap0, int lap1, cudaStream_t st, int fb) {
    const DevGeom& g = E->geom;
    const int L = g.nlevels;
    const bool tm = E->timing && fb == 0;
    int evi = 0;
    auto mark = [&]() { if (tm) cudaEventRecord(E->ev[evi++], st); };
    l0.base -= (size_t)fb * l0.frameBytes;          // kernels index level 0 by the global frame slot fb+i
    mark();
    const int slot = st == E->stream2 ? 1 : 0;
    const bool side = E->overlap && !E->timing && E->aux[slot];
    // the blur of all levels: TMA-staged tiles when every level (incl. the caller's level 0) is TMA-addressable
    const size_t l0stride = (fb + n == 1) ? (((size_t)l0.pitch * g.lv[0].h + 15) & ~(size_t)15) : l0.frameBytes;
    auto launchBlur = [&](cudaStream_t bs) {
        const dim3 bgrid((E->blurTilesTotal + 7) / 8, n);
        bool tma = E->blurTma;
        FsMaps bm = E->blurMaps;
        if (tma) bm.tmaOk[0] = encodeLevelMap(&bm.m[0], l0.base, g.lv[0].w, g.lv[0].h, fb + n, (size_t)l0.pitch, l0stride, kBtBoxH, kBtBoxW) ? 1 : 0;
        for (int l = 0; l < L && tma; ++l) tma = bm.tmaOk[l] != 0;
        if (tma) k_blur7_tma<<<bgrid, 256, 0, bs>>>(bm, E->dGeom, E->blurTiles, fb);
        else k_blur7<<<bgrid, 256, 0, bs>>>(E->dGeom, l0, E->blurTiles, fb);
    };
    // FAST on level 0 reads only the input image: its cells go first, on the side stream, beside the (latency-bound) chain
    // of pyramid launches
    const int nA0 = side && E->fastEarly && E->fastGen < 2 ? E->fastNA0 : 0;
#define ORBX_FAST_GROUP(T, P, STREAM, BASE, COUNT, ROWS, SMEM)                                                                      \
    k_fast_cells<T, P><<<dim3(COUNT, n), T, SMEM, STREAM>>>(E->dGeom, l0, E->dCells, E->dCand, E->dCellCount, E->fastTilePitch,   \
                                                            ROWS, fb, E->dCellOrder, BASE)
#define ORBX_FAST_DISPATCH(STREAM, BASE, COUNT, ROWS, SMEM)                                                                        \
    do {                                                                                                                           \
        if (E->fastPitchConst) {                                                                                                   \
            if (E->fastThreads == 64) ORBX_FAST_GROUP(64, 33, STREAM, BASE, COUNT, ROWS, SMEM);                                    \
            else if (E->fastThreads == 128) ORBX_FAST_GROUP(128, 33, STREAM, BASE, COUNT, ROWS, SMEM);                             \
            else ORBX_FAST_GROUP(256, 33, STREAM, BASE, COUNT, ROWS, SMEM);                                                        \
        } else {                                                                                                                   \
            if (E->fastThreads == 64) ORBX_FAST_GROUP(64, 0, STREAM, BASE, COUNT, ROWS, SMEM);                                     \
            else if (E->fastThreads == 128) ORBX_FAST_GROUP(128, 0, STREAM, BASE, COUNT, ROWS, SMEM);                              \
            else ORBX_FAST_GROUP(256, 0, STREAM, BASE, COUNT, ROWS, SMEM);                                                         \
        }                                                                                                                          \
        ++E->launches;                                                                                                             \
    } while (0)
    if (nA0 > 0) {
        CK(cudaEventRecord(E->evStart[slot], st));
        CK(cudaStreamWaitEvent(E->aux[slot], E->evStart[slot], 0));
        ORBX_FAST_DISPATCH(E->aux[slot], 0, nA0, E->fastRowsA, E->fastSmemA);
        CK(cudaEventRecord(E->evF0[slot], E->aux[slot]));
    }
    // few frames: the whole pyramid in one launch (tiles of a level wait on the arrival counter of the level below); all tiles must
    // be resident at once - 4 CTAs per SM is a conservative bound for 256 threads x 40 registers x <= 16 KB
    bool chained = false;
    if (E->pyrChainOk && (long long)n * E->pyrChain.tileBegin[L] <= (long long)E->smCount * 4 &&
        ((((uintptr_t)l0.base | (uintptr_t)l0.pitch | (uintptr_t)l0.frameBytes) & 3) == 0)) {
        FsMaps cm = E->pyrMaps;
        if (encodeLevelMap(&cm.m[1], l0.base, g.lv[0].w, g.lv[0].h, fb + n, (size_t)l0.pitch, l0stride, E->pyrBox[1].boxH, E->pyrBox[1].boxW)) {
            k_pyr_chain_tma<<<dim3(E->pyrChain.tileBegin[L], n), 256, E->pyrChainSmem, st>>>(cm, E->dGeom, E->dTaps, E->pyrChain, fb,
                                                                                             E->dPyrArrive + (size_t)fb * kMaxLevels);
            ++E->launches;
            chained = true;
        }
    }
    for (int l = 1; l < L && !chained; ++l) {
        dim3 grid((g.lv[l].w + 127) / 128, (g.lv[l].h + 7) / 8, n), block(32, 8);
        const bool srcAligned = l > 1 || ((((uintptr_t)l0.base | (uintptr_t)l0.pitch | (uintptr_t)l0.frameBytes) & 3) == 0);
        bool ptma = E->pyrTma && E->pyrFast[l] && srcAligned && E->pyrBox[l].boxW > 0;
        CUtensorMap pm = E->pyrMaps.m[l];
        if (ptma && l == 1) ptma = encodeLevelMap(&pm, l0.base, g.lv[0].w, g.lv[0].h, fb + n, (size_t)l0.pitch, l0stride, E->pyrBox[1].boxH, E->pyrBox[1].boxW);
        else if (ptma) ptma = E->pyrMaps.tmaOk[l] != 0;
        if (ptma) {
            dim3 fgrid((g.lv[l].w + 127) / 128, (g.lv[l].h + kPrTileH - 1) / kPrTileH, n);
            k_pyr_level_tma<<<fgrid, 256, (size_t)E->pyrBox[l].boxW * E->pyrBox[l].boxH, st>>>(pm, E->dGeom, l, E->dTaps, E->pyrBox[l], fb);
        } else if (E->pyrFast[l] && srcAligned) {
            dim3 fgrid((g.lv[l].w + 127) / 128, (g.lv[l].h + 8 * kPyrRows - 1) / (8 * kPyrRows), n);
            k_pyr_level_fast<<<fgrid, block, 0, st>>>(E->dGeom, l0, l, E->dTaps, fb);
        } else {
            k_pyr_level<<<grid, block, 0, st>>>(E->dGeom, l0, l, E->dTaps, fb);
        }
        ++E->launches;
    }
    mark();
    // The blur needs only the pyramid, and the quadtree (latency-bound, ~1/3 of the issue slots) leaves most of the SMs'
    // issue bandwidth unused: run the blur on a side stream beside the quadtree and join before the descriptors.
    // With stage timing on, everything stays on one stream so that the per-stage events mean what they say.
    const bool early = side && E->blurEarly;       // fork right after the pyramid: blur || (FAST -> quadtree)
    if (early) {
        CK(cudaEventRecord(E->evFork[slot], st));
        CK(cudaStreamWaitEvent(E->aux[slot], E->evFork[slot], 0));
        launchBlur(E->aux[slot]);
        CK(cudaEventRecord(E->evJoin[slot], E->aux[slot]));
    }
    if (E->fastGen >= 2) {
        // level 0 is the caller's image: its tensor map depends on this pass's pointer / pitch / frame stride
        FsMaps maps = E->fsMaps;
        maps.tmaOk[0] = encodeLevelMap(&maps.m[0], l0.base, g.lv[0].w, g.lv[0].h, fb + n, (size_t)l0.pitch, l0stride) ? 1 : 0;
        const dim3 sgrid((E->fsPlan.nJobs + kFsWarps - 1) / kFsWarps, n);
        if (E->fastCompass) k_fast_score<true><<<sgrid, kFsWarps * 32, 0, st>>>(maps, E->dGeom, l0, E->fsPlan, fb);
        else k_fast_score<false><<<sgrid, kFsWarps * 32, 0, st>>>(maps, E->dGeom, l0, E->fsPlan, fb);
        k_fast_nms<<<sgrid, kFsWarps * 32, 0, st>>>(E->scoreMaps, E->dGeom, E->fsPlan, E->nmsTabs, E->dCbits, fb);
        k_fast_select<<<dim3((g.nCells + kSelWarps - 1) / kSelWarps, n), kSelWarps * 32, 0, st>>>(E->dGeom, E->dCells, E->fsPlan, E->dCbits,
                                                                                               E->dCand, E->dCellCount, fb);
        E->launches += 3;
    } else {
        if (E->fastNA > nA0) ORBX_FAST_DISPATCH(st, nA0, E->fastNA - nA0, E->fastRowsA, E->fastSmemA);
        if (g.nCells > E->fastNA) ORBX_FAST_DISPATCH(st, E->fastNA, g.nCells - E->fastNA, E->fastTileRows, E->fastSmem);
    }
    if (nA0 > 0) CK(cudaStreamWaitEvent(st, E->evF0[slot], 0));
#undef ORBX_FAST_DISPATCH
#undef ORBX_FAST_GROUP
    mark();
    if (side && !early) {
        CK(cudaEventRecord(E->evFork[slot], st));
        CK(cudaStreamWaitEvent(E->aux[slot], E->evFork[slot], 0));
        launchBlur(E->aux[slot]);
        CK(cudaEventRecord(E->evJoin[slot], E->aux[slot]));
    }
    // few CTAs in flight: wide CTAs (latency of one frame); a full batch: narrow ones (see orbx_kernels_tree.cuh).  "Few" = all wide
    // CTAs resident at once (two 512 x 64-register CTAs per SM): one frame, or the 32 KITTI frames of the bench (51 k frames/s with 256
    // threads, 54 k with 512, 48 k with 128)
    const int treeThreads = n * L <= 2 * E->smCount ? E->treeThreadsSmall : E->treeThreadsLarge;
    CK(launch_prio(k_octree, dim3(n, L), dim3(treeThreads), E->treeSmem, st, E->prio, E->dGeom, E->dCells, E->dCand, E->dCellCount,
                   E->dQtA, E->dQtB, E->dLvlKp, E->dLvlCnt, E->poolCap, E->scanCap, fb, E->pyrRoots));
    ++E->launches;
    mark();
    if (!side) launchBlur(st);
    ++E->launches;
    mark();
    CK(launch_prio(k_finalize, dim3(n), dim3(kFinalThreads), 0, st, E->prio, E->dGeom, E->dLvlKp, E->dLvlCnt, lap0, lap1, E->dOutKp,
                   E->dOutCnt, E->dWork, fb));
    ++E->launches;
    mark();
    if (side) CK(cudaStreamWaitEvent(st, E->evJoin[slot], 0));
    bool dtma = E->descTma;
    FsMaps om = E->oriMaps;
    if (dtma) om.tmaOk[0] = encodeLevelMap(&om.m[0], l0.base, g.lv[0].w, g.lv[0].h, fb + n, (size_t)l0.pitch, l0stride, kOtBoxH, kOtBoxW) ? 1 : 0;
    for (int l = 0; l < L && dtma; ++l) dtma = E->descMaps.tmaOk[l] != 0 && om.tmaOk[l] != 0;
    if (dtma)
        k_orient_brief_tma<<<dim3((g.outCap + kDtWarps * kDescGroup - 1) / (kDtWarps * kDescGroup), n), kDtWarps * 32, kDtSmemBytes, st>>>(
            E->descMaps, om, E->dGeom, E->dWork, E->dOutCnt, E->dOutKp, E->dOutDesc, fb);
    else
        k_orient_brief<<<dim3((g.outCap + kDescWarps * kDescGroup - 1) / (kDescWarps * kDescGroup), n), kDescWarps * 32, 0, st>>>(
            E->dGeom, l0, E->dWork, E->dOutCnt, E->dOutKp, E->dOutDesc, fb);
    ++E->launches;
    mark();
    CK(cudaGetLastError());
    return ORBX_OK;
}

// One pass = 13 kernel launches, 14 with two FAST groups (+ the side-stream fork / join).  At 16 frames per sub-batch that is ~0.3 ms of GPU work
// behind ~20 driver calls, so the host path was launch-bound; the sequence is captured once per argument set into a CUDA
// graph and replayed with a single call.  Stage timing (events between the kernels) uses the plain launches.
int runPass(Extractor* E, Level0 l0, int n, int lap0, int lap1, cudaStream_t st, int fb = 0) {
    Level0 l0abs = l0;
    int rc = ORBX_OK;
    const int slot = st == E->stream2 ? 1 : 0;
    for (int s2 = 0; s2 < 2; ++s2)          // order this pass after earlier passes of the handle that ran on other streams
        if (E->lastSt[s2] && E->lastSt[s2] != st && (s2 == slot || fb == 0)) CK(cudaStreamWaitEvent(st, E->evDone[s2], 0));
    if (!E->useGraphs || E->timing) {
        rc = launchPass(E, l0, n, lap0, lap1, st, fb);
    } else {
        Extractor::PassGraph* hit = nullptr;
        for (auto& G : E->graphs)
            if (G.base == l0.base && G.frameBytes == l0.frameBytes && G.pitch == l0.pitch && G.n == n && G.lap0 == lap0 && G.lap1 == lap1 &&
                G.fb == fb && G.slot == slot) { hit = &G; break; }
        if (!hit) {
            if (E->graphs.size() >= 64) { for (auto& G : E->graphs) cudaGraphExecDestroy(G.exec); E->graphs.clear(); }
            const long long before = E->launches;
            cudaGraph_t graph = nullptr;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            rc = launchPass(E, l0, n, lap0, lap1, st, fb);
            cudaError_t ce = cudaStreamEndCapture(st, &graph);
            if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
            CK(ce);
            Extractor::PassGraph G{l0.base, l0.frameBytes, l0.pitch, n, lap0, lap1, fb, slot, nullptr, (int)(E->launches - before)};
            E->launches = before;
            ce = cudaGraphInstantiate(&G.exec, graph, 0);
            cudaGraphDestroy(graph);
            CK(ce);
            E->graphs.push_back(G);
            hit = &E->graphs.back();
        }
        CK(cudaGraphLaunch(hit->exec, st));
        E->launches += hit->kernels;
    }
    if (rc) return rc;
    CK(cudaEventRecord(E->evDone[slot], st));
    E->lastSt[slot] = st;
    if (fb == 0) { E->lastFrames = n; E->lastL0 = l0abs; }
    else E->lastFrames = std::max(E->lastFrames, fb + n);
    return ORBX_OK;
}

// wait for every pass launched on this handle, whatever stream it ran on
int syncPasses(Extractor* E) {
    for (int s = 0; s < 2; ++s)
        if (E->lastSt[s]) CK(cudaEventSynchronize(E->evDone[s]));
    return ORBX_OK;
}

}  // namespace

struct orbx_extractor : Extractor {};
#define E_FROM(h) Extractor* E = static_cast<Extractor*>(h)

extern "C" {

int orbx_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int orbx_device_pci_bus_id(int device, char* out, int len) {
    if (!out || len < 16) return ORBX_ERR_INVALID;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return ORBX_ERR_NO_DEVICE;
    if (device < 0 || device >= n) return ORBX_ERR_INVALID;
    return cudaDeviceGetPCIBusId(out, len, device) == cudaSuccess ? ORBX_OK : ORBX_ERR_CUDA;
}

// Page-lock a host range the caller keeps its frames (or its output arrays) in, e.g. the camera ring buffer: pinned pointers
// are used in place by orbx_extract / orbx_extract_batch, pageable ones go through the handle's staging buffer (one host copy
// of the frame per call).  Thin wrappers over cudaHostRegister / cudaHostUnregister so that a C or C++ host needs no CUDA headers.
int orbx_pin_host(void* ptr, size_t bytes) {
    if (!ptr || !bytes) return ORBX_ERR_INVALID;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return ORBX_ERR_NO_DEVICE;
    const cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) { cudaGetLastError(); return ORBX_ERR_CUDA; }
    return ORBX_OK;
}
int orbx_unpin_host(void* ptr) {
    if (!ptr) return ORBX_ERR_INVALID;
    const cudaError_t e = cudaHostUnregister(ptr);
    if (e != cudaSuccess) { cudaGetLastError(); return ORBX_ERR_CUDA; }
    return ORBX_OK;
}

const char* orbx_strerror(int s) {
    switch (s) {
        case ORBX_OK: return "ok";
        case ORBX_ERR_INVALID: return "invalid argument";
        case ORBX_ERR_CUDA: return "CUDA error";
        case ORBX_ERR_CAPACITY: return "output buffer too small";
        case ORBX_ERR_UNSUPPORTED: return "unsupported configuration";
        case ORBX_ERR_NO_DEVICE: return "no CUDA device (no CPU fallback exists)";
    }
    return "unknown";
}

const char* orbx_last_error(orbx_handle h) { return h ? h->err.c_str() : ""; }

int orbx_create(const orbx_config* cfg, orbx_handle* out) {
    if (!cfg || !out || cfg->nfeatures <= 0 || cfg->nlevels <= 0 || cfg->nlevels > kMaxLevels || cfg->scale_factor < 1.f ||
        cfg->max_batch < 1)
        return ORBX_ERR_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return ORBX_ERR_NO_DEVICE;
    if (cfg->device < 0 || cfg->device >= ndev) return ORBX_ERR_INVALID;
    orbx_extractor* E = new orbx_extractor();
    E->cfg = *cfg;
    E->device = cfg->device;
    const int L = cfg->nlevels;
    // scale tables and per-level quotas with the constructor's float arithmetic (:414-445)
    const double sf = (double)cfg->scale_factor;
    E->scale.assign(L, 1.f); E->sigma2.assign(L, 1.f); E->invScale.resize(L); E->invSigma2.resize(L);
    for (int i = 1; i < L; ++i) {
        E->scale[i] = (float)(E->scale[i - 1] * sf);
        E->sigma2[i] = E->scale[i] * E->scale[i];
    }
    for (int i = 0; i < L; ++i) { E->invScale[i] = 1.0f / E->scale[i]; E->invSigma2[i] = 1.0f / E->sigma2[i]; }
    E->perLevel.resize(L);
    const float factor = (float)(1.0f / sf);
    float want = cfg->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; ++l) { E->perLevel[l] = cvRoundF(want); sum += E->perLevel[l]; want *= factor; }
    E->perLevel[L - 1] = std::max(cfg->nfeatures - sum, 0);

    auto fail = [&](cudaError_t e) { fprintf(stderr, "orbx_create: %s\n", cudaGetErrorString(e)); delete E; return ORBX_ERR_CUDA; };
    cudaError_t e;
    if ((e = cudaSetDevice(E->device)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamCreateWithFlags(&E->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail(e);
    if ((e = cudaStreamCreateWithFlags(&E->stream2, cudaStreamNonBlocking)) != cudaSuccess) return fail(e);
    for (int i = 0; i < 2; ++i) {
        if ((e = cudaStreamCreateWithFlags(&E->aux[i], cudaStreamNonBlocking)) != cudaSuccess) return fail(e);
        if ((e = cudaEventCreateWithFlags(&E->evFork[i], cudaEventDisableTiming)) != cudaSuccess) return fail(e);
        if ((e = cudaEventCreateWithFlags(&E->evJoin[i], cudaEventDisableTiming)) != cudaSuccess) return fail(e);
        if ((e = cudaEventCreateWithFlags(&E->evStart[i], cudaEventDisableTiming)) != cudaSuccess) return fail(e);
        if ((e = cudaEventCreateWithFlags(&E->evDone[i], cudaEventDisableTiming)) != cudaSuccess) return fail(e);
        if ((e = cudaEventCreateWithFlags(&E->evF0[i], cudaEventDisableTiming)) != cudaSuccess) return fail(e);
    }
    if (const char* ov = getenv("ORBX_OVERLAP")) E->overlap = atoi(ov) != 0;
    if (const char* gr = getenv("ORBX_GRAPHS")) E->useGraphs = atoi(gr) != 0;
    if (const char* be = getenv("ORBX_BLUR_EARLY")) E->blurEarly = atoi(be) != 0;
    if (const char* fe = getenv("ORBX_FAST_EARLY")) E->fastEarly = atoi(fe) != 0;
    if (const char* fg = getenv("ORBX_FAST_GEN")) E->fastGen = atoi(fg) == 1 ? 1 : 2;
    if (const char* fc = getenv("ORBX_FAST_COMPASS")) E->fastCompass = atoi(fc) != 0;
    if (const char* bt = getenv("ORBX_BLUR_TMA")) E->blurTma = atoi(bt) != 0;
    if (const char* dt = getenv("ORBX_DESC_TMA")) E->descTma = atoi(dt) != 0;
    if (const char* pt = getenv("ORBX_PYR_TMA")) E->pyrTma = atoi(pt) != 0;
    {
        int least = 0, greatest = 0;
        cudaDeviceGetStreamPriorityRange(&least, &greatest);
        E->prio = greatest;
        if (const char* pr = getenv("ORBX_PRIO")) E->prio = atoi(pr) ? greatest : 0;
    }
    if (const char* c = getenv("ORBX_CHUNK")) E->chunk = atoi(c);
    for (auto& ev : E->ev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail(e);
    // algorithm constants: circular-patch half-widths (:453-468) and the BRIEF pattern
    int umax[16];
    const int vmax = (int)floorf(kHalfPatch * sqrtf(2.f) / 2 + 1), vmin = (int)ceilf(kHalfPatch * sqrtf(2.f) / 2);
    for (int v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt((double)kHalfPatch * kHalfPatch - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0; ++v0;
    }
    char4 pat[256];
    for (int i = 0; i < 256; ++i)
        pat[i] = make_char4((char)kPatternTable[4 * i], (char)kPatternTable[4 * i + 1], (char)kPatternTable[4 * i + 2], (char)kPatternTable[4 * i + 3]);
    if ((e = cudaMemcpyToSymbol(c_umax, umax, sizeof(umax))) != cudaSuccess) return fail(e);
    if ((e = cudaMemcpyToSymbol(c_pattern, pat, sizeof(pat))) != cudaSuccess) return fail(e);
    *out = E;
    return ORBX_OK;
}

void orbx_destroy(orbx_handle h) {
    if (!h) return;
    E_FROM(h);
    cudaSetDevice(E->device);
    cudaStreamSynchronize(E->stream);
    cudaStreamSynchronize(E->stream2);
    freeShape(E);
    for (auto& ev : E->ev) if (ev) cudaEventDestroy(ev);
    if (E->stream) cudaStreamDestroy(E->stream);
    if (E->stream2) cudaStreamDestroy(E->stream2);
    for (int i = 0; i < 2; ++i) {
        if (E->aux[i]) { cudaStreamSynchronize(E->aux[i]); cudaStreamDestroy(E->aux[i]); }
        if (E->evFork[i]) cudaEventDestroy(E->evFork[i]);
        if (E->evJoin[i]) cudaEventDestroy(E->evJoin[i]);
        if (E->evStart[i]) cudaEventDestroy(E->evStart[i]);
        if (E->evDone[i]) { cudaEventSynchronize(E->evDone[i]); cudaEventDestroy(E->evDone[i]); }
        if (E->evF0[i]) cudaEventDestroy(E->evF0[i]);
    }
    delete h;
}

int orbx_get_tables(orbx_handle h, float* scale, float* inv, float* sigma2, float* invSigma2, int* perLevel) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    for (int i = 0; i < E->cfg.nlevels; ++i) {
        if (scale) scale[i] = E->scale[i];
        if (inv) inv[i] = E->invScale[i];
        if (sigma2) sigma2[i] = E->sigma2[i];
        if (invSigma2) invSigma2[i] = E->invSigma2[i];
        if (perLevel) perLevel[i] = E->perLevel[i];
    }
    return ORBX_OK;
}

int orbx_max_keypoints(orbx_handle h) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (E->W) return E->geom.outCap;
    int s = 0;
    for (int q : E->perLevel) s += q + 24;     // shape-independent bound; the exact bound needs nIni per level
    return s;
}

}  // extern "C"

// ---- extraction entry points -------------------------------------------------------------------------------------
namespace {

int collectStageTimes(Extractor* E) {
    if (!E->timing) return ORBX_OK;
    float tot = 0;
    for (int i = 0; i < 6; ++i) {
        CK(cudaEventElapsedTime(&E->stageMs[i], E->ev[i], E->ev[i + 1]));
        tot += E->stageMs[i];
    }
    E->stageMs[6] = tot;
    return ORBX_OK;
}

// copy the results of frames [0,n) of the last pass to the pinned staging area and wait
int stageResults(Extractor* E, int n, cudaStream_t st) {
    const DevGeom& g = E->geom;
    CK(cudaMemcpyAsync(E->hCnt, E->dOutCnt, (size_t)2 * n * sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(E->hKp, E->dOutKp, (size_t)n * g.outCap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(E->hDesc, E->dOutDesc, (size_t)n * g.outCap * 32, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return collectStageTimes(E);
}

}  // namespace

extern "C" {

// true iff p points into page-locked host memory (cudaMallocHost / cudaHostRegister): copies from/to it are truly async
static bool isPinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// Host path.  A batch is cut into sub-batches that alternate between the handle's two streams, so that the H2D copy of
// one sub-batch overlaps the kernels of the previous one and its D2H the kernels of the next.  Page-locked caller buffers
// are used in place (no staging copy); pageable ones go through the pinned staging area.
int orbx_extract_batch(orbx_handle h, const uint8_t* const* imgs, int nimg, int width, int height, int stride,
                       int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_out) {
    if (!h || !imgs || nimg <= 0 || !n_out || !mono_out) return ORBX_ERR_INVALID;
    if (width <= 0 || height <= 0) return ORBX_ERR_INVALID;    // the reference returns -1 for an empty image (:1090)
    if (stride < width) return ORBX_ERR_INVALID;
    E_FROM(h);
    CK(cudaSetDevice(E->device));
    int rc = configureShape(E, width, height);
    if (rc) return rc;
    const DevGeom& g = E->geom;
    const int B = E->cfg.max_batch;
    const int p0 = g.lv[0].pitch;
    const size_t plane = g.lv[0].planeBytes;
    int C = E->chunk > 0 ? E->chunk : (B >= 8 ? B / 2 : B);      // two sub-batches: smaller ones lose more to kernel tails than they hide
    C = std::max(1, std::min(C, B));
    const bool inPinned = isPinned(imgs[0]) && isPinned(imgs[nimg - 1]);
    const bool kpPinned = kps && isPinned(kps) && isPinned(kps + (size_t)nimg * cap - 1);
    const bool dsPinned = desc && isPinned(desc) && isPinned(desc + (size_t)nimg * cap * 32 - 1);
    const int rowsOut = std::min(cap, g.outCap);
    cudaStream_t streams[2] = {E->stream, E->stream2};
    int status = ORBX_OK;
    for (int base = 0; base < nimg; base += B) {
        const int n = std::min(B, nimg - base);
        for (int c0 = 0, ci = 0; c0 < n; c0 += C, ++ci) {
            const int nc = std::min(C, n - c0);
            cudaStream_t st = streams[ci & 1];
            // Rows are copied with the caller's stride (one linear DMA per frame, or per sub-batch when the frames are
            // contiguous) and the kernels read level 0 with that pitch: a pitched 2-D copy of 752-byte rows runs far below
            // PCIe speed.  Strides that do not fit the slot fall back to re-pitching.
            const size_t frameLin = (size_t)stride * (height - 1) + width;
            const bool linear = (size_t)stride * height <= plane;
            bool contiguous = linear;
            for (int f = 0; f < nc; ++f) {
                if (!imgs[base + c0 + f]) return ORBX_ERR_INVALID;
                if (f && imgs[base + c0 + f] != imgs[base + c0 + f - 1] + (size_t)stride * height) contiguous = false;
            }
            const size_t slot = linear ? (size_t)stride * height : plane;         // distance between frames on the device
            const int pitch0 = linear ? stride : p0;
            uint8_t* dchunk = E->dLvl0 + (size_t)c0 * slot;
            if (inPinned) {
                if (contiguous) {
                    CK(cudaMemcpyAsync(dchunk, imgs[base + c0], slot * (nc - 1) + frameLin, cudaMemcpyHostToDevice, st));
                } else {
                    for (int f = 0; f < nc; ++f) {
                        if (linear) CK(cudaMemcpyAsync(dchunk + (size_t)f * slot, imgs[base + c0 + f], frameLin, cudaMemcpyHostToDevice, st));
                        else CK(cudaMemcpy2DAsync(dchunk + (size_t)f * slot, p0, imgs[base + c0 + f], stride, width, height, cudaMemcpyHostToDevice, st));
                    }
                }
            } else {
                uint8_t* hchunk = E->hImg + (size_t)c0 * slot;
                for (int f = 0; f < nc; ++f) {
                    const uint8_t* src = imgs[base + c0 + f];
                    uint8_t* hdst = hchunk + (size_t)f * slot;
                    if (linear) memcpy(hdst, src, frameLin);
                    else for (int y = 0; y < height; ++y) memcpy(hdst + (size_t)y * p0, src + (size_t)y * stride, width);
                }
                CK(cudaMemcpyAsync(dchunk, hchunk, slot * (nc - 1) + (linear ? frameLin : plane), cudaMemcpyHostToDevice, st));
            }
            Level0 l0{dchunk, slot, pitch0};
            if ((rc = runPass(E, l0, nc, lap0, lap1, st, c0))) return rc;
            if (B == 1 && kps && desc && !kpPinned && !dsPinned) {
                // one frame per pass (an agent's Tracking thread): the whole result block in one copy
                CK(cudaMemcpyAsync(E->hKp, E->dOutKp, E->outBlockBytes, cudaMemcpyDeviceToHost, st));
                continue;
            }
            CK(cudaMemcpyAsync(E->hCnt + 2 * c0, E->dOutCnt + 2 * c0, (size_t)2 * nc * sizeof(int), cudaMemcpyDeviceToHost, st));
            if (kps) {
                if (kpPinned)
                    CK(cudaMemcpy2DAsync(kps + (size_t)(base + c0) * cap, (size_t)cap * sizeof(orbx_keypoint), E->dOutKp + (size_t)c0 * g.outCap,
                                         (size_t)g.outCap * sizeof(orbx_keypoint), (size_t)rowsOut * sizeof(orbx_keypoint), nc, cudaMemcpyDeviceToHost, st));
                else
                    CK(cudaMemcpyAsync(E->hKp + (size_t)c0 * g.outCap, E->dOutKp + (size_t)c0 * g.outCap, (size_t)nc * g.outCap * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, st));
            }
            if (desc) {
                if (dsPinned)
                    CK(cudaMemcpy2DAsync(desc + (size_t)(base + c0) * cap * 32, (size_t)cap * 32, E->dOutDesc + (size_t)c0 * g.outCap * 32,
                                         (size_t)g.outCap * 32, (size_t)rowsOut * 32, nc, cudaMemcpyDeviceToHost, st));
                else
                    CK(cudaMemcpyAsync(E->hDesc + (size_t)c0 * g.outCap * 32, E->dOutDesc + (size_t)c0 * g.outCap * 32, (size_t)nc * g.outCap * 32, cudaMemcpyDeviceToHost, st));
            }
        }
        CK(cudaStreamSynchronize(E->stream));
        if (n > C) CK(cudaStreamSynchronize(E->stream2));        // a second sub-batch ran there
        E->lastFrames = n;
        if ((rc = collectStageTimes(E))) return rc;
        for (int f = 0; f < n; ++f) {
            const int cnt = E->hCnt[2 * f];
            n_out[base + f] = cnt;
            mono_out[base + f] = E->hCnt[2 * f + 1];
            if (cnt > cap) { status = ORBX_ERR_CAPACITY; continue; }
            if (kps && !kpPinned) memcpy(kps + (size_t)(base + f) * cap, E->hKp + (size_t)f * g.outCap, (size_t)cnt * sizeof(orbx_keypoint));
            if (desc && !dsPinned) memcpy(desc + (size_t)(base + f) * cap * 32, E->hDesc + (size_t)f * g.outCap * 32, (size_t)cnt * 32);
        }
    }
    return status;
}

int orbx_extract(orbx_handle h, const uint8_t* img, int width, int height, int stride, int lap0, int lap1,
                 orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_out) {
    if (!img) return ORBX_ERR_INVALID;
    const uint8_t* one[1] = {img};
    return orbx_extract_batch(h, one, 1, width, height, stride, lap0, lap1, kps, desc, cap, n_out, mono_out);
}

int orbx_extract_device(orbx_handle h, const uint8_t* d_imgs, size_t image_bytes, int nimg, int width, int height,
                        int stride, int lap0, int lap1, void* stream) {
    if (!h || !d_imgs || nimg <= 0 || width <= 0 || height <= 0 || stride < width) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (nimg > E->cfg.max_batch) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    int rc = configureShape(E, width, height);
    if (rc) return rc;
    Level0 l0{d_imgs, image_bytes, stride};
    return runPass(E, l0, nimg, lap0, lap1, stream ? (cudaStream_t)stream : E->stream);
}

int orbx_synchronize(orbx_handle h) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    CK(cudaSetDevice(E->device));
    int rc = syncPasses(E);
    if (rc) return rc;
    CK(cudaStreamSynchronize(E->stream));
    return collectStageTimes(E);
}

int orbx_results_device(orbx_handle h, int frame, const orbx_keypoint** d_kps, const uint32_t** d_desc, int* n_out, int* mono_out) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (frame < 0 || frame >= E->lastFrames) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    int cnt[2];
    { int rc = syncPasses(E); if (rc) return rc; }
    CK(cudaMemcpyAsync(cnt, E->dOutCnt + 2 * frame, sizeof(cnt), cudaMemcpyDeviceToHost, E->stream));
    CK(cudaStreamSynchronize(E->stream));
    if (d_kps) *d_kps = E->dOutKp + (size_t)frame * E->geom.outCap;
    if (d_desc) *d_desc = reinterpret_cast<const uint32_t*>(E->dOutDesc + (size_t)frame * E->geom.outCap * 32);
    if (n_out) *n_out = cnt[0];
    if (mono_out) *mono_out = cnt[1];
    return ORBX_OK;
}

int orbx_results_batch_device(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc, const int** d_counts, int* frame_stride) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (!E->W) return ORBX_ERR_INVALID;
    if (d_kps) *d_kps = E->dOutKp;
    if (d_desc) *d_desc = E->dOutDesc;
    if (d_counts) *d_counts = E->dOutCnt;
    if (frame_stride) *frame_stride = E->geom.outCap;
    return ORBX_OK;
}

int orbx_fetch_results(orbx_handle h, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_out) {
    const orbx_keypoint* dk; const uint32_t* dd; int n, mono;
    int rc = orbx_results_device(h, frame, &dk, &dd, &n, &mono);
    if (rc) return rc;
    E_FROM(h);
    if (n_out) *n_out = n;
    if (mono_out) *mono_out = mono;
    if (n > cap) return ORBX_ERR_CAPACITY;
    if (kps && n) CK(cudaMemcpy(kps, dk, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost));
    if (desc && n) CK(cudaMemcpy(desc, dd, (size_t)n * 32, cudaMemcpyDeviceToHost));
    return ORBX_OK;
}

int orbx_level_size(orbx_handle h, int level, int* width, int* height) {
    if (!h) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (!E->W || level < 0 || level >= E->cfg.nlevels) return ORBX_ERR_INVALID;
    if (width) *width = E->geom.lv[level].w;
    if (height) *height = E->geom.lv[level].h;
    return ORBX_OK;
}

static int downloadPlane(Extractor* E, const uint8_t* dsrc, int pitch, int w, int hgt, int padded, uint8_t* dst, int dst_stride) {
    std::vector<uint8_t> tmp((size_t)w * hgt);
    { int rc = syncPasses(E); if (rc) return rc; }
    CK(cudaStreamSynchronize(E->stream));
    CK(cudaMemcpy2D(tmp.data(), w, dsrc, pitch, w, hgt, cudaMemcpyDeviceToHost));
    if (!padded) {
        for (int y = 0; y < hgt; ++y) memcpy(dst + (size_t)y * dst_stride, &tmp[(size_t)y * w], w);
        return ORBX_OK;
    }
    // mvImagePyramid's backing buffer: 19-px BORDER_REFLECT_101 ring synthesised from the level itself (:1185,1190).
    // Nothing on the mono hot path reads the ring (FAST, IC_Angle and BRIEF stay >= 1 px inside the level), so the
    // device never stores it; it is produced here for callers that read the padded pyramid (stereo matching).
    auto refl = [](int p, int n) { if (n == 1) return 0; while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p; return p; };
    for (int y = -kEdge; y < hgt + kEdge; ++y) {
        const uint8_t* S = &tmp[(size_t)refl(y, hgt) * w];
        uint8_t* D = dst + (size_t)(y + kEdge) * dst_stride;
        for (int x = -kEdge; x < w + kEdge; ++x) D[x + kEdge] = S[refl(x, w)];
    }
    return ORBX_OK;
}

int orbx_download_level(orbx_handle h, int frame, int level, int padded, uint8_t* dst, int dst_stride) {
    if (!h || !dst) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (frame < 0 || frame >= E->lastFrames || level < 0 || level >= E->cfg.nlevels) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    const LevelGeom& G = E->geom.lv[level];
    const uint8_t* src; int pitch;
    if (level == 0) { src = E->lastL0.base + (size_t)frame * E->lastL0.frameBytes; pitch = E->lastL0.pitch; }
    else { src = E->geom.lvl[level] + (size_t)frame * G.planeBytes; pitch = G.pitch; }
    return downloadPlane(E, src, pitch, G.w, G.h, padded, dst, dst_stride);
}

int orbx_download_blurred(orbx_handle h, int frame, int level, uint8_t* dst, int dst_stride) {
    if (!h || !dst) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (frame < 0 || frame >= E->lastFrames || level < 0 || level >= E->cfg.nlevels) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    const LevelGeom& G = E->geom.lv[level];
    return downloadPlane(E, E->geom.blur[level] + (size_t)frame * G.planeBytes, G.pitch, G.w, G.h, 0, dst, dst_stride);
}

int orbx_download_candidates(orbx_handle h, int frame, int level, int* out, int cap, int* n_out) {
    if (!h || !n_out) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (frame < 0 || frame >= E->lastFrames || level < 0 || level >= E->cfg.nlevels) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    { int rc = syncPasses(E); if (rc) return rc; }
    CK(cudaStreamSynchronize(E->stream));
    const DevGeom& g = E->geom;
    const LevelGeom& G = g.lv[level];
    std::vector<int> counts(g.nCells);
    std::vector<uint32_t> cand(g.candCapFrame);
    CK(cudaMemcpy(counts.data(), E->dCellCount + (size_t)frame * g.nCells, counts.size() * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cand.data(), E->dCand + (size_t)frame * g.candCapFrame, cand.size() * 4, cudaMemcpyDeviceToHost));
    int n = 0;
    for (int c = G.cellBegin; c < G.cellEnd; ++c) {
        const FastCell& fc = E->cells[c];
        for (int k = 0; k < counts[c]; ++k, ++n) {
            if (n >= cap || !out) continue;
            const uint32_t e = cand[fc.outOffset + k];
            out[3 * n] = (int)(e & 255u) + fc.relX; out[3 * n + 1] = (int)((e >> 8) & 255u) + fc.relY; out[3 * n + 2] = (int)(e >> 16);
        }
    }
    *n_out = n;
    return n > cap ? ORBX_ERR_CAPACITY : ORBX_OK;
}

int orbx_download_level_keypoints(orbx_handle h, int frame, int level, int* out, int cap, int* n_out) {
    if (!h || !n_out) return ORBX_ERR_INVALID;
    E_FROM(h);
    if (frame < 0 || frame >= E->lastFrames || level < 0 || level >= E->cfg.nlevels) return ORBX_ERR_INVALID;
    CK(cudaSetDevice(E->device));
    { int rc = syncPasses(E); if (rc) return rc; }
    CK(cudaStreamSynchronize(E->stream));
    const DevGeom& g = E->geom;
    const LevelGeom& G = g.lv[level];
    int n = 0;
    CK(cudaMemcpy(&n, E->dLvlCnt + frame * g.nlevels + level, 4, cudaMemcpyDeviceToHost));
    *n_out = n;
    if (n > cap) return ORBX_ERR_CAPACITY;
    std::vector<uint64_t> keys(n);
    if (n) CK(cudaMemcpy(keys.data(), E->dLvlKp + (size_t)frame * g.kpCapFrame + G.kpOffset, (size_t)n * 8, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && out; ++i) {
        out[3 * i] = (int)(keys[i] & 0xffff); out[3 * i + 1] = (int)((keys[i] >> 16) & 0xffff); out[3 * i + 2] = (int)((keys[i] >> 32) & 0xff);
    }
    return ORBX_OK;
}

int orbx_debug_sort_nodes(orbx_handle h, const int* size, const int* ul_x, int n, int* perm, int mode) {
    if (!h || n < 0 || (mode == 2 && n > 65535)) return ORBX_ERR_INVALID;
    E_FROM(h);
    CK(cudaSetDevice(E->device));
    std::vector<ise::Elem> e(std::max(n, 1));
    for (int i = 0; i < n; ++i) { e[i].key = ((uint64_t)(uint32_t)size[i] << 32) | (uint32_t)ul_x[i]; e[i].id = i; e[i].pad = 0; }
    ise::Elem* d = nullptr;
    CK(cudaMalloc((void**)&d, e.size() * sizeof(ise::Elem)));
    cudaError_t er = cudaMemcpy(d, e.data(), e.size() * sizeof(ise::Elem), cudaMemcpyHostToDevice);
    uint32_t* scratch = nullptr;
    if (er == cudaSuccess) er = cudaMalloc((void**)&scratch, ((size_t)3 * n + 16) * sizeof(uint32_t));
    if (er == cudaSuccess) { k_debug_sort<<<1, 32, 0, E->stream>>>(d, n, mode, scratch); ++E->launches; er = cudaStreamSynchronize(E->stream); }
    cudaFree(scratch);
    if (er == cudaSuccess) er = cudaMemcpy(e.data(), d, e.size() * sizeof(ise::Elem), cudaMemcpyDeviceToHost);
    cudaFree(d);
    CK(er);
    for (int i = 0; i < n; ++i) perm[i] = e[i].id;
    return ORBX_OK;
}

int orbx_debug_distribute(orbx_handle h, const int* keys, int n, int min_x, int max_x, int min_y, int max_y,
                          int n_features, int* out, int cap, int* n_out) {
    return orbx_debug_distribute2(h, keys, n, min_x, max_x, min_y, max_y, n_features, 0, out, cap, n_out, nullptr);
}

int orbx_debug_distribute2(orbx_handle h, const int* keys, int n, int min_x, int max_x, int min_y, int max_y,
                           int n_features, int use_pyramid, int* out, int cap, int* n_out, int* used_pyramid) {
    if (!h || n < 0 || !n_out || max_x <= min_x || max_y <= min_y || n_features < 1) return ORBX_ERR_INVALID;
    E_FROM(h);
    CK(cudaSetDevice(E->device));
    const int treeW = max_x - min_x, treeH = max_y - min_y;
    const int nIni = (int)roundf((float)treeW / (float)treeH);
    if (nIni < 1) return ORBX_ERR_UNSUPPORTED;
    const float hX = (float)treeW / nIni;
    const int P = std::max(n_features + 3, 4 * nIni) + 5;
    const int scanCap = std::max(P, 64) + 8;
    const size_t smem = align16(tree_smem_layout(P, scanCap, nullptr, nullptr)) + (size_t)kPyrMaxRoots * kPyrCellsPerRoot * 8;
    if (smem > 200 * 1024) return ORBX_ERR_UNSUPPORTED;
    CK(raise_dyn_smem((const void*)k_debug_distribute, E->device, 200 * 1024));
    std::vector<uint64_t> k(std::max(n, 1));
    for (int i = 0; i < n; ++i)
        k[i] = (uint64_t)(uint32_t)keys[3 * i] | ((uint64_t)(uint32_t)keys[3 * i + 1] << 16) | ((uint64_t)(uint32_t)keys[3 * i + 2] << 32);
    uint64_t *dA = nullptr, *dB = nullptr, *dOut = nullptr; int* dCnt = nullptr;
    cudaError_t er = cudaMalloc((void**)&dA, k.size() * 8);
    if (er == cudaSuccess) er = cudaMalloc((void**)&dB, k.size() * 8);
    if (er == cudaSuccess) er = cudaMalloc((void**)&dOut, (size_t)P * 8);
    if (er == cudaSuccess) er = cudaMalloc((void**)&dCnt, 8);
    if (er == cudaSuccess) er = cudaMemcpy(dA, k.data(), k.size() * 8, cudaMemcpyHostToDevice);
    int cnt = 0;
    std::vector<uint64_t> res(P);
    if (er == cudaSuccess) {
        k_debug_distribute<<<1, E->treeThreadsSmall, smem, E->stream>>>(dA, dB, n, n_features, nIni, hX, treeH, dOut, dCnt, P, scanCap, use_pyramid);
        ++E->launches;
        er = cudaStreamSynchronize(E->stream);
    }
    int cnt2[2] = {0, 0};
    if (er == cudaSuccess) er = cudaMemcpy(cnt2, dCnt, 8, cudaMemcpyDeviceToHost);
    cnt = cnt2[0];
    if (used_pyramid) *used_pyramid = cnt2[1];
    if (er == cudaSuccess) er = cudaMemcpy(res.data(), dOut, (size_t)P * 8, cudaMemcpyDeviceToHost);
    cudaFree(dA); cudaFree(dB); cudaFree(dOut); cudaFree(dCnt);
    CK(er);
    *n_out = cnt;
    if (cnt > cap) return ORBX_ERR_CAPACITY;
    for (int i = 0; i < cnt && out; ++i) {
        out[3 * i] = (int)(res[i] & 0xffff); out[3 * i + 1] = (int)((res[i] >> 16) & 0xffff); out[3 * i + 2] = (int)((res[i] >> 32) & 0xff);
    }
    return ORBX_OK;
}

// Frame::ComputeStereoMatches (src/Frame.cc:811-981)
int orbx_stereo_matches(orbx_handle left, orbx_handle right, float mb, float mbf, float* u_right, float* depth, int cap,
                        int* n_left, int* n_kept) {
    if (!left || !right || !u_right || !depth || !n_left || !n_kept || !(mb > 0.f)) return ORBX_ERR_INVALID;
    Extractor* EL = static_cast<Extractor*>(left);
    Extractor* ER = static_cast<Extractor*>(right);
    Extractor* E = EL;
    if (EL->device != ER->device || EL->lastFrames < 1 || ER->lastFrames < 1 || EL->cfg.nlevels != ER->cfg.nlevels ||
        EL->cfg.scale_factor != ER->cfg.scale_factor) {
        EL->err = "orbx_stereo_matches: both extractors must hold one extracted frame, on the same device, with the same pyramid";
        return ORBX_ERR_INVALID;
    }
    CK(cudaSetDevice(EL->device));
    { Extractor* E = ER; int rc = syncPasses(E); if (rc) return rc; }
    { int rc = syncPasses(EL); if (rc) return rc; }
    CK(cudaStreamSynchronize(ER->stream));
    const int L = EL->cfg.nlevels, outCap = EL->geom.outCap;
    StereoParams P;
    for (int l = 0; l < L; ++l) { P.scale[l] = EL->scale[l]; P.invScale[l] = EL->invScale[l]; P.colsR[l] = ER->geom.lv[l].w; }
    P.mb = mb; P.mbf = mbf;
    float* dU = nullptr;                       // uRight | depth | sad | kept
    CK(cudaMalloc((void**)&dU, (size_t)outCap * 12 + 16));
    float* dD = dU + outCap;
    int* dSad = (int*)(dD + outCap);
    int* dKept = dSad + outCap;
    k_stereo_search<<<(outCap + kStereoWarps - 1) / kStereoWarps, kStereoWarps * 32, 0, EL->stream>>>(
        EL->dGeom, EL->lastL0, ER->dGeom, ER->lastL0, EL->dOutKp, (const uint32_t*)EL->dOutDesc, EL->dOutCnt, ER->dOutKp,
        (const uint32_t*)ER->dOutDesc, ER->dOutCnt, P, dU, dD, dSad);
    k_stereo_filter<<<1, 1024, 0, EL->stream>>>(EL->dOutCnt, dSad, dU, dD, dKept);
    EL->launches += 2;
    int hn[2] = {0, 0};
    cudaError_t e = cudaMemcpyAsync(&hn[0], EL->dOutCnt, 4, cudaMemcpyDeviceToHost, EL->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&hn[1], dKept, 4, cudaMemcpyDeviceToHost, EL->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(EL->stream);
    if (e == cudaSuccess && hn[0] <= cap && hn[0] > 0) {
        e = cudaMemcpy(u_right, dU, (size_t)hn[0] * 4, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(depth, dD, (size_t)hn[0] * 4, cudaMemcpyDeviceToHost);
    }
    cudaFree(dU);
    CK(e);
    *n_left = hn[0]; *n_kept = hn[1];
    return hn[0] <= cap ? ORBX_OK : ORBX_ERR_CAPACITY;
}

#ifdef ORBX_TREE_PROFILE
extern "C" int orbx_debug_tree_profile(long long* out16, int reset) {
    if (reset) { long long z[16] = {}; cudaMemcpyToSymbol(g_treeProf, z, sizeof(z)); return 0; }
    cudaDeviceSynchronize();
    return cudaMemcpyFromSymbol(out16, g_treeProf, 16 * sizeof(long long)) == cudaSuccess ? 0 : -1;
}
#endif

long long orbx_launch_count(orbx_handle h) { return h ? static_cast<Extractor*>(h)->launches : 0; }
void* orbx_stream(orbx_handle h) { return h ? (void*)static_cast<Extractor*>(h)->stream : nullptr; }

int orbx_set_stage_timing(orbx_handle h, int enable) {
    if (!h) return ORBX_ERR_INVALID;
    static_cast<Extractor*>(h)->timing = enable != 0;
    return ORBX_OK;
}
int orbx_stage_times(orbx_handle h, float* ms7) {
    if (!h || !ms7) return ORBX_ERR_INVALID;
    E_FROM(h);
    for (int i = 0; i < 7; ++i) ms7[i] = E->stageMs[i];
    return ORBX_OK;
}

}  // extern "C"
