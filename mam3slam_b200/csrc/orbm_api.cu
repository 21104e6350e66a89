// C ABI of the B200-native ORB matcher (include/orbm.h): handle / frame management, query marshalling, launches.
// Host-side logic mirrors what each ORBmatcher method does before and after its Hamming loop (src/ORBmatcher.cc).
#include "orbm_kernels.cuh"
#include "orbm_kernels_tc.cuh"
#include "orbm_kernels_rounds.cuh"
#include "orbx_internal.h"
#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

// every allocating CUDA call of the matcher goes through these, so that tests can assert that a steady-state call sequence (the
// adapter's pooled objects, the allocation-free entry point) performs none: orbm_debug_alloc_count()
static std::atomic<long long> g_allocCalls{0};
template <class T> static cudaError_t cMalloc(T** p, size_t n) { ++g_allocCalls; return cudaMalloc((void**)p, n); }
template <class T> static cudaError_t cMallocHost(T** p, size_t n) { ++g_allocCalls; return cudaMallocHost((void**)p, n); }
static cudaError_t cStreamCreate(cudaStream_t* s, unsigned flags) { ++g_allocCalls; return cudaStreamCreateWithFlags(s, flags); }

using namespace orbm;

struct orbm_frame_s {
    int n = 0, cap = 0;
    bool ownsInput = false;            // kpsDev / descDev were allocated here (host upload) rather than borrowed
    orbx_keypoint* kpsDev = nullptr;
    uint32_t* descDev = nullptr;
    float4* kp4 = nullptr;
    int* cellOf = nullptr;
    int* cellOff = nullptr;
    int* cellIdx = nullptr;
    float minX = 0, minY = 0, maxX = 0, maxY = 0, invW = 0, invH = 0;
    std::vector<orbx_keypoint> kpsHost;   // kept for the host-side marshalling of BoW / initialization searches
    bool haveHost = false;
    // host uploads: keypoints and descriptors share one device block and go up in ONE copy from a page-locked staging buffer
    // of the slot (two copies from the caller's pageable arrays cost ~8 us of host time each)
    uint8_t* stage = nullptr; size_t descOff = 0;
    const void* uploadedBy = nullptr;     // handle whose stream carried the last upload from `stage`, and its syncEpoch then
    unsigned long long uploadEpoch = 0;
};

struct orbm_db_s {
    uint32_t* rows = nullptr;
    bool owns = false;
    long long n = 0, rowOffset = 0;
};

struct orbm_matcher {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    long long launches = 0;
    long long lastEvals = 0;
    int smCount = 148;
    int prio = 0;                    // launch priority of the tracking step (latency-bound; 0 = none)
    // growable device scratch for queries / results
    void* scratch = nullptr; size_t scratchBytes = 0;
    void* pinned = nullptr; size_t pinnedBytes = 0;
    unsigned long long* dEvals = nullptr;
    unsigned long long* knnPart = nullptr; size_t knnPartBytes = 0;
    unsigned int* knnBound = nullptr; int knnBoundCap = 0;         // shared pruning bound per query (all ones at rest)
    unsigned int* knnCounter = nullptr; int knnCounterCap = 0;     // chunks-done counter per query block (fused merge), zero at rest
    void* trackBuf = nullptr; size_t trackBytes = 0;
    unsigned long long syncEpoch = 1;          // completed waits on `stream` (syncStream)
    int lastRounds = 0, lastSerial = 0;          // claim rounds of the last search (0 = in-order kernel), 1 = finished in order
};

namespace {

#define CKM(call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            M->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return ORBM_ERR_CUDA;                                                                  \
        }                                                                                          \
    } while (0)
#define CKM_(MM, call)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            (MM)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                           \
            return ORBM_ERR_CUDA;                                                                  \
        }                                                                                          \
    } while (0)

int ensureScratch(orbm_matcher* M, size_t bytes) {
    if (bytes <= M->scratchBytes) return ORBM_OK;
    if (M->scratch) cudaFree(M->scratch);
    if (M->pinned) cudaFreeHost(M->pinned);
    M->scratch = nullptr; M->pinned = nullptr; M->scratchBytes = M->pinnedBytes = 0;
    const size_t want = std::max<size_t>(bytes * 2, 1 << 20);
    CKM(cMalloc(&M->scratch, want));
    CKM(cMallocHost(&M->pinned, want));
    M->scratchBytes = M->pinnedBytes = want;
    return ORBM_OK;
}

// every wait on the handle's stream goes through here: frame slots use the count to know that their last upload has left the
// page-locked staging buffer (a search always waits, so in steady state the next upload needs no CUDA call to be safe)
static inline cudaError_t syncStream(orbm_matcher* M) {
    const cudaError_t e = cudaStreamSynchronize(M->stream);
    if (e == cudaSuccess) ++M->syncEpoch;
    return e;
}

// bump allocator over the pinned staging buffer and its device twin (same offsets)
struct Stage {
    orbm_matcher* M; size_t off = 0;
    explicit Stage(orbm_matcher* m) : M(m) {}
    template <class T> size_t take(size_t count) { off = (off + 255) & ~(size_t)255; size_t at = off; off += count * sizeof(T); return at; }
    template <class T> T* host(size_t at) { return reinterpret_cast<T*>((uint8_t*)M->pinned + at); }
    template <class T> T* dev(size_t at) { return reinterpret_cast<T*>((uint8_t*)M->scratch + at); }
};

void freeFrameBuffers(orbm_frame_s* f) {
    if (f->ownsInput) cudaFree(f->kpsDev);         // one block: descDev points into it
    if (f->stage) { cudaFreeHost(f->stage); f->stage = nullptr; }
    cudaFree(f->kp4); cudaFree(f->cellOf); cudaFree(f->cellOff); cudaFree(f->cellIdx);
    f->kpsDev = nullptr; f->descDev = nullptr; f->kp4 = nullptr; f->cellOf = f->cellOff = f->cellIdx = nullptr;
    f->cap = 0; f->ownsInput = false;
}

int setFrame(orbm_matcher* M, orbm_frame_s* f, const orbm_frame_view* v) {
    if (!v || v->n < 0 || (v->n > 0 && (!v->kps || !v->desc)) || !(v->max_x > v->min_x) || !(v->max_y > v->min_y)) return ORBM_ERR_INVALID;
    if (v->n >= (1 << 19)) { M->err = "more than 2^19 keypoints per frame"; return ORBM_ERR_UNSUPPORTED; }
    CKM(cudaSetDevice(M->device));
    const int n = v->n;
    const bool needOwn = !v->on_device;
    if (n > f->cap || needOwn != f->ownsInput || !f->cellOff) {
        freeFrameBuffers(f);
        const int cap = std::max(n, 1024);
        if (needOwn) {
            f->descOff = ((size_t)cap * sizeof(orbx_keypoint) + 255) & ~(size_t)255;
            CKM(cMalloc((void**)&f->kpsDev, f->descOff + (size_t)cap * 32));
            f->descDev = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(f->kpsDev) + f->descOff);
            CKM(cudaHostAlloc((void**)&f->stage, f->descOff + (size_t)cap * 32, cudaHostAllocDefault));
        }
        CKM(cMalloc((void**)&f->kp4, (size_t)cap * sizeof(float4)));
        CKM(cMalloc((void**)&f->cellOf, (size_t)cap * sizeof(int)));
        CKM(cMalloc((void**)&f->cellOff, (size_t)(kGridCells + 1) * sizeof(int)));
        CKM(cMalloc((void**)&f->cellIdx, (size_t)cap * sizeof(int)));
        f->cap = cap; f->ownsInput = needOwn;
    }
    f->n = n;
    f->minX = v->min_x; f->minY = v->min_y; f->maxX = v->max_x; f->maxY = v->max_y;
    // mfGridElementWidthInv / HeightInv (src/Frame.cc:251-252)
    f->invW = static_cast<float>(kGridCols) / static_cast<float>(f->maxX - f->minX);
    f->invH = static_cast<float>(kGridRows) / static_cast<float>(f->maxY - f->minY);
    if (needOwn) {
        if (n) {
            // a previous upload of this slot may still be reading the staging buffer: not if its stream has been waited on since
            // (every search does); otherwise wait now - on the same handle its stream, on another one the device (rare)
            if (f->uploadedBy == M) { if (M->syncEpoch == f->uploadEpoch) CKM(syncStream(M)); }
            else if (f->uploadedBy) CKM(cudaDeviceSynchronize());
            memcpy(f->stage, v->kps, (size_t)n * sizeof(orbx_keypoint));
            memcpy(f->stage + f->descOff, v->desc, (size_t)n * 32);
            if ((size_t)n * 60 * 2 >= f->descOff + (size_t)n * 32) {        // the gap between the two arrays is small: one copy
                CKM(cudaMemcpyAsync(f->kpsDev, f->stage, f->descOff + (size_t)n * 32, cudaMemcpyHostToDevice, M->stream));
            } else {
                CKM(cudaMemcpyAsync(f->kpsDev, f->stage, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, M->stream));
                CKM(cudaMemcpyAsync(f->descDev, f->stage + f->descOff, (size_t)n * 32, cudaMemcpyHostToDevice, M->stream));
            }
            f->uploadedBy = M; f->uploadEpoch = M->syncEpoch;
        }
        f->kpsHost.assign(v->kps, v->kps + n);
        f->haveHost = true;
    } else {
        f->kpsDev = const_cast<orbx_keypoint*>(v->kps);
        f->descDev = reinterpret_cast<uint32_t*>(const_cast<uint8_t*>(v->desc));
        f->haveHost = false;
    }
    if (n <= 8 * kGridThreads) {
        // cell of every keypoint + the CSR grid in one launch
        k_frame_grid<<<1, kGridThreads, 0, M->stream>>>(n, f->kpsDev, f->kp4, f->cellOf, f->cellOff, f->cellIdx, f->minX, f->minY, f->invW, f->invH);
        ++M->launches;
    } else {
        k_grid_prepare<<<(n + 255) / 256, 256, 0, M->stream>>>(n, f->kpsDev, f->kp4, f->cellOf, f->minX, f->minY, f->invW, f->invH);
        k_grid_build<<<1, kGridThreads, 0, M->stream>>>(n, f->cellOf, f->cellOff, f->cellIdx);
        M->launches += 2;
    }
    CKM(cudaGetLastError());
    return ORBM_OK;
}

int ensureHostKps(orbm_matcher* M, orbm_frame_s* f) {
    if (f->haveHost) return ORBM_OK;
    f->kpsHost.resize(f->n);
    if (f->n) {
        CKM(cudaMemcpyAsync(f->kpsHost.data(), f->kpsDev, (size_t)f->n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, M->stream));
        CKM(syncStream(M));
    }
    f->haveHost = true;
    return ORBM_OK;
}

FrameDev devView(const orbm_frame_s* f) {
    FrameDev d;
    d.n = f->n; d.kp4 = f->kp4; d.desc = f->descDev; d.cellOff = f->cellOff; d.cellIdx = f->cellIdx;
    d.minX = f->minX; d.minY = f->minY; d.invW = f->invW; d.invH = f->invH;
    return d;
}

}  // namespace

extern "C" {

const char* orbm_strerror(int s) {
    switch (s) {
        case ORBM_OK: return "ok";
        case ORBM_ERR_INVALID: return "invalid argument";
        case ORBM_ERR_CUDA: return "CUDA error";
        case ORBM_ERR_CAPACITY: return "output buffer too small";
        case ORBM_ERR_UNSUPPORTED: return "unsupported configuration";
        case ORBM_ERR_NO_DEVICE: return "no CUDA device (no CPU fallback exists)";
    }
    return "unknown";
}
const char* orbm_last_error(orbm_handle h) { return h ? h->err.c_str() : ""; }

int orbm_create(int device, orbm_handle* out) {
    if (!out) return ORBM_ERR_INVALID;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return ORBM_ERR_NO_DEVICE;
    if (device < 0 || device >= ndev) return ORBM_ERR_INVALID;
    orbm_matcher* M = new orbm_matcher();
    M->device = device;
    if (cudaSetDevice(device) != cudaSuccess || cStreamCreate(&M->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cMalloc((void**)&M->dEvals, 128) != cudaSuccess) {
        fprintf(stderr, "orbm_create: %s\n", cudaGetErrorString(cudaGetLastError()));
        delete M;
        return ORBM_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&M->smCount, cudaDevAttrMultiProcessorCount, device);
    {
        int least = 0, greatest = 0;
        cudaDeviceGetStreamPriorityRange(&least, &greatest);
        M->prio = greatest;
        if (const char* pr = getenv("ORBX_PRIO")) M->prio = atoi(pr) ? greatest : 0;
    }
    // per (function, device): set for THIS handle's device (orbx_internal.h raise_dyn_smem)
    orbx::raise_dyn_smem((const void*)k_search<true>, device, 200 * 1024);
    orbx::raise_dyn_smem((const void*)k_search<false>, device, 200 * 1024);
    orbx::raise_dyn_smem((const void*)k_track_search, device, 200 * 1024);
    orbx::raise_dyn_smem((const void*)k_search_rounds, device, kRoundsSmemMax);
    orbx::raise_dyn_smem((const void*)k_track_resolve, device, kRoundsSmemMax);
    *out = M;
    return ORBM_OK;
}

void orbm_destroy(orbm_handle M) {
    if (!M) return;
    cudaSetDevice(M->device);
    syncStream(M);
    if (M->scratch) cudaFree(M->scratch);
    if (M->pinned) cudaFreeHost(M->pinned);
    if (M->dEvals) cudaFree(M->dEvals);
    if (M->knnPart) cudaFree(M->knnPart);
    if (M->knnCounter) cudaFree(M->knnCounter);
    if (M->knnBound) cudaFree(M->knnBound);
    if (M->trackBuf) cudaFree(M->trackBuf);
    cudaStreamDestroy(M->stream);
    delete M;
}

long long orbm_launch_count(orbm_handle h) { return h ? h->launches : 0; }
void* orbm_stream(orbm_handle h) { return h ? (void*)h->stream : nullptr; }
long long orbm_last_evaluations(orbm_handle h) { return h ? h->lastEvals : 0; }

// DescriptorDistance (src/ORBmatcher.cc:2058-2074), host side
int orbm_distance(const uint8_t* a, const uint8_t* b) {
    uint32_t pa[8], pb[8];
    memcpy(pa, a, 32); memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
    }
    return dist;
}

int orbm_frame_create(orbm_handle M, const orbm_frame_view* view, orbm_frame* out) {
    if (!M || !out) return ORBM_ERR_INVALID;
    orbm_frame_s* f = new orbm_frame_s();
    int rc = setFrame(M, f, view);
    if (rc) { freeFrameBuffers(f); delete f; return rc; }
    *out = f;
    return ORBM_OK;
}
int orbm_frame_update(orbm_handle M, orbm_frame f, const orbm_frame_view* view) {
    if (!M || !f) return ORBM_ERR_INVALID;
    return setFrame(M, f, view);
}
void orbm_frame_destroy(orbm_handle M, orbm_frame f) {
    if (!M || !f) return;
    cudaSetDevice(M->device);
    syncStream(M);
    freeFrameBuffers(f);
    delete f;
}

int orbm_frame_grid(orbm_handle M, orbm_frame f, int* off, int* idx) {
    if (!M || !f || !off || !idx) return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    CKM(cudaMemcpyAsync(off, f->cellOff, (size_t)(kGridCells + 1) * sizeof(int), cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    if (off[kGridCells]) CKM(cudaMemcpy(idx, f->cellIdx, (size_t)off[kGridCells] * sizeof(int), cudaMemcpyDeviceToHost));
    return ORBM_OK;
}

}  // extern "C"

// ---- generic search driver -------------------------------------------------------------------------------------------
namespace {

// parity tap kernel: the candidate list of one window query, in the reference's order
__global__ void k_area_list(SearchParams P, int* out, int cap, int* nOut) {
    const int lane = threadIdx.x & 31;
    uint32_t qd[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    unsigned long long ne = 0;
    int count = 0;
    walk_candidates(P, P.tgt, 0, qd, ne, [&](bool ok, uint32_t entry) {
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (ok) { const int pos = count + __popc(m & ((1u << lane) - 1)); if (pos < cap) out[pos] = (int)(entry & 0x7ffff); }
        count += __popc(m);
    });
    if (lane == 0) *nOut = count;
}

struct SearchIO {
    int nq = 0;
    std::vector<float4> qwin;
    std::vector<int2> qlev, qrun;
    std::vector<float> qangle;
    const uint8_t* qdescSrc = nullptr;      // rows gathered through qsrc
    std::vector<int> qsrc;                  // original query index of each compacted query
    const int* bowIdx = nullptr; int nBow = 0;
    const uint8_t* elig = nullptr;
    uint8_t* occ = nullptr;                 // in/out, n bytes
    const uint8_t* qocc = nullptr;          // per ORIGINAL query (indexed through qsrc): its match occupies the keypoint; NULL = always
    std::vector<int2> nodeQ;                // BoW: query range of every shared node (k_search_nodes), empty = single-CTA path
    const uint8_t* qdescDev = nullptr;      // BoW, node-parallel: the source frame's descriptors in HBM (rows gathered on the device)
    std::vector<int> result, dist;
    void reserve(int nq) { qwin.reserve(nq); qlev.reserve(nq); qangle.reserve(nq); qsrc.reserve(nq); }
};

// uploads the compacted queries, runs k_search (+ k_rot_filter), downloads per-query results
int runSearch(orbm_matcher* M, const orbm_frame_s* T, SearchParams P, SearchIO& io) {
    CKM(cudaSetDevice(M->device));
    const int nq = io.nq, n = T->n;
    io.result.assign(nq, -1); io.dist.assign(nq, 256);
    M->lastEvals = 0;
    if (nq == 0) return ORBM_OK;
    Stage S(M);
    // layout: [uploaded: queries ... occupancy, counters] [downloaded with the last two: results] [device only: lists]
    // so that one copy goes up and one contiguous copy (occupancy, counters, results) comes back
    const size_t oWin = S.take<float4>(nq), oLev = S.take<int2>(nq), oRun = S.take<int2>(nq), oDesc = S.take<uint32_t>((size_t)nq * 8);
    const size_t oAng = S.take<float>(nq), oBow = S.take<int>(std::max(io.nBow, 1)), oElig = S.take<uint8_t>(std::max(n, 1));
    const size_t oQocc = S.take<uint8_t>(nq);
    const bool byNode = !io.nodeQ.empty() && io.qdescDev && P.source == 1 && P.state == 1;
    const int nNodes = (int)io.nodeQ.size();
    const size_t oNode = S.take<int2>(std::max(nNodes, 1)), oSrc = S.take<int>(byNode ? nq : 1), oTicket = S.take<unsigned int>(4);
    const size_t oOcc = S.take<uint8_t>(std::max(n, 1)), oEv = S.take<unsigned long long>(16);
    const size_t inEnd = S.off;
    const size_t oRes = S.take<int>(nq), oDist = S.take<int>(nq), oBin = S.take<int>(nq);
    const size_t outEnd = S.off;
    // window searches under "best unblocked candidate <= th" against an occupancy state: claim rounds over many CTAs
    // (orbm_kernels_rounds.cuh); ORBM_ROUNDS=0 keeps them on the single-CTA in-order resolver (A/B)
    static const bool roundsOk = [] { const char* e = getenv("ORBM_ROUNDS"); return !(e && e[0] == '0'); }();
    const bool byRounds = roundsOk && P.source == 0 && P.state == 1 && rounds_rule(P.rule) && io.occ && nq <= kRoundsMaxQueries &&
                          rounds_smem_bytes(n, nq) <= kRoundsSmemMax;
    const size_t oLists = S.take<uint4>(byRounds ? nq : 1), oCnt = S.take<int>(byRounds ? nq : 1);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    if (!io.qwin.empty()) memcpy(S.host<float4>(oWin), io.qwin.data(), (size_t)nq * sizeof(float4));
    if (!io.qlev.empty()) memcpy(S.host<int2>(oLev), io.qlev.data(), (size_t)nq * sizeof(int2));
    if (!io.qrun.empty()) memcpy(S.host<int2>(oRun), io.qrun.data(), (size_t)nq * sizeof(int2));
    if (!io.qangle.empty()) memcpy(S.host<float>(oAng), io.qangle.data(), (size_t)nq * sizeof(float));
    if (byNode) {
        memcpy(S.host<int2>(oNode), io.nodeQ.data(), (size_t)nNodes * sizeof(int2));
        memcpy(S.host<int>(oSrc), io.qsrc.data(), (size_t)nq * sizeof(int));
        *S.host<unsigned int>(oTicket) = 0;
    } else {
        *S.host<unsigned int>(oTicket) = 0;
        uint8_t* hd = S.host<uint8_t>(oDesc);
        for (int q = 0; q < nq; ++q) memcpy(hd + 32 * (size_t)q, io.qdescSrc + 32 * (size_t)io.qsrc[q], 32);
    }
    if (io.nBow) memcpy(S.host<int>(oBow), io.bowIdx, (size_t)io.nBow * sizeof(int));
    if (io.elig && n) memcpy(S.host<uint8_t>(oElig), io.elig, n);
    if (io.occ && n) memcpy(S.host<uint8_t>(oOcc), io.occ, n);
    if (io.qocc) for (int q = 0; q < nq; ++q) S.host<uint8_t>(oQocc)[q] = io.qocc[io.qsrc[q]];
    memset(S.host<unsigned long long>(oEv), 0, 128);
    CKM(cudaMemcpyAsync(M->scratch, M->pinned, inEnd, cudaMemcpyHostToDevice, M->stream));
    P.tgt = devView(T);
    P.nq = nq;
    P.qwin = S.dev<float4>(oWin); P.qlev = S.dev<int2>(oLev); P.qrun = S.dev<int2>(oRun);
    P.qdesc = S.dev<uint32_t>(oDesc); P.qangle = S.dev<float>(oAng); P.bowIdx = S.dev<int>(oBow);
    P.elig = io.elig ? S.dev<uint8_t>(oElig) : nullptr;
    P.occ = io.occ ? S.dev<uint8_t>(oOcc) : nullptr;
    P.qocc = io.qocc ? S.dev<uint8_t>(oQocc) : nullptr;
    P.qResult = S.dev<int>(oRes); P.qDist = S.dev<int>(oDist); P.qBin = S.dev<int>(oBin);
    P.evals = S.dev<unsigned long long>(oEv);
    // stateful searches run on one CTA: stage the target frame on chip; stateless ones spread over many CTAs instead
    P.stageTgt = (P.state != 0 && search_smem_bytes(n, P.state, true, P.source == 0) <= kSearchSmemMax && nq >= 64) ? 1 : 0;
    const size_t smem = search_smem_bytes(n, P.state, P.stageTgt != 0, P.source == 0);
    if (!byRounds && smem > kSearchSmemMax) { M->err = "target frame too large for the on-chip match state"; return ORBM_ERR_UNSUPPORTED; }
    // window searches walk one query per lane (32 per warp), BoW / triangulation runs one query per warp
    const int units = P.source == 0 ? (nq + 31) / 32 : nq;
    const int grid = P.state != 0 ? 1 : std::max(1, std::min(units, M->smCount * 2));
    if (byRounds) {
        P.ticket = S.dev<unsigned int>(oTicket);
        // the resolving CTA writes [occupancy, counters, results] straight into the page-locked mirror (UVA: the pinned buffer is
        // addressable from the device), so the call is one copy up, one launch, one wait
        const size_t outBytes = (outEnd - oOcc + 15) & ~(size_t)15;
        k_search_rounds<<<(nq + kRoundQueries - 1) / kRoundQueries, kRoundThreads, rounds_smem_bytes(n, nq), M->stream>>>(
            P, S.dev<uint4>(oLists), S.dev<int>(oCnt), S.host<uint4>(oOcc), S.dev<uint4>(oOcc), (int)(outBytes / 16));
        ++M->launches;
    } else if (byNode) {
        // BoW: a warp per shared node, rotation filter fused into the last CTA (orbm_kernels.cuh, k_search_nodes)
        P.nodeQ = S.dev<int2>(oNode); P.nNodes = nNodes; P.qsrc = S.dev<int>(oSrc);
        P.qdescBase = reinterpret_cast<const uint32_t*>(io.qdescDev); P.ticket = S.dev<unsigned int>(oTicket);
        k_search_nodes<<<(nNodes + kNodeWarps - 1) / kNodeWarps, kNodeWarps * 32, 0, M->stream>>>(P);
        ++M->launches;
    } else {
        if (P.source == 0) k_search<true><<<grid, kSearchThreads, smem, M->stream>>>(P);
        else k_search<false><<<grid, kSearchThreads, smem, M->stream>>>(P);
        ++M->launches;
        if (P.checkOri) { k_rot_filter<<<1, 256, 0, M->stream>>>(nq, P.qBin, P.qResult); ++M->launches; }
    }
    CKM(cudaGetLastError());
    if (!byRounds) CKM(cudaMemcpyAsync((uint8_t*)M->pinned + oOcc, (uint8_t*)M->scratch + oOcc, outEnd - oOcc, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    memcpy(io.result.data(), S.host<int>(oRes), (size_t)nq * sizeof(int));
    memcpy(io.dist.data(), S.host<int>(oDist), (size_t)nq * sizeof(int));
    if (io.occ && n && P.state == 1) memcpy(io.occ, S.host<uint8_t>(oOcc), n);
    M->lastEvals = (long long)*S.host<unsigned long long>(oEv);
    M->lastRounds = byRounds ? (int)S.host<unsigned long long>(oEv)[1] : 0;
    M->lastSerial = byRounds ? (int)S.host<unsigned long long>(oEv)[2] : 0;
#ifdef ORBM_STATS
    if (byRounds) fprintf(stderr, "rounds: nq %d evals %llu rounds %llu incomplete %llu resolve cycles %llu walk cycles (max CTA) %llu; re-walks %llu in %llu rounds, %llu cycles\n", nq,
                          S.host<unsigned long long>(oEv)[0], S.host<unsigned long long>(oEv)[1], S.host<unsigned long long>(oEv)[12],
                          S.host<unsigned long long>(oEv)[3], S.host<unsigned long long>(oEv)[4], S.host<unsigned long long>(oEv)[5],
                          S.host<unsigned long long>(oEv)[6], S.host<unsigned long long>(oEv)[7]);
    if (byRounds) { fprintf(stderr, "  cycles in step 1 / 1r / 1b / 2:"); for (int i = 8; i < 12; ++i) fprintf(stderr, " %llu", S.host<unsigned long long>(oEv)[i]); fprintf(stderr, "\n"); }
#endif
    return ORBM_OK;
}

SearchParams baseParams() {
    SearchParams P;
    memset(&P, 0, sizeof(P));
    P.none2 = 256;
    return P;
}

// level bounds of Frame::GetFeaturesInArea (src/Frame.cc:689-707): bCheckLevels = minLevel > 0 || maxLevel >= 0
int2 levelBounds(int minLevel, int maxLevel) {
    const bool check = (minLevel > 0) || (maxLevel >= 0);
    if (!check) return make_int2(INT_MIN, INT_MAX);
    return make_int2(minLevel, maxLevel >= 0 ? maxLevel : INT_MAX);
}

}  // namespace

extern "C" {

int orbm_features_in_area(orbm_handle M, orbm_frame f, float x, float y, float r, int min_level, int max_level, int* out,
                          int cap, int* n_out) {
    if (!M || !f || !n_out || cap < 0) return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    Stage S(M);
    const size_t oWin = S.take<float4>(1), oLev = S.take<int2>(1), oOut = S.take<int>(std::max(cap, 1)), oN = S.take<int>(1);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    *S.host<float4>(oWin) = make_float4(x, y, r, 0);
    *S.host<int2>(oLev) = levelBounds(min_level, max_level);
    CKM(cudaMemcpyAsync(M->scratch, M->pinned, oOut, cudaMemcpyHostToDevice, M->stream));
    SearchParams P = baseParams();
    P.tgt = devView(f); P.nq = 1; P.source = 0;
    P.qwin = S.dev<float4>(oWin); P.qlev = S.dev<int2>(oLev);
    // the walk also evaluates a distance per candidate: give it a valid descriptor row to read (the frame's first)
    if (f->n == 0) { *n_out = 0; return ORBM_OK; }
    k_area_list<<<1, 32, 0, M->stream>>>(P, S.dev<int>(oOut), cap, S.dev<int>(oN));
    ++M->launches;
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync((uint8_t*)M->pinned + oOut, (uint8_t*)M->scratch + oOut, S.off - oOut, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    const int n = *S.host<int>(oN);
    *n_out = n;
    if (out) memcpy(out, S.host<int>(oOut), (size_t)std::min(n, cap) * sizeof(int));
    return n > cap ? ORBM_ERR_CAPACITY : ORBM_OK;
}

// M1: src/ORBmatcher.cc:43-213 (mono branch)
int orbm_search_local_map(orbm_handle M, orbm_frame f, uint8_t* occupied, const float* scale_factors, int nlevels, int nq,
                          const uint8_t* valid, const float* proj_x, const float* proj_y, const int* level,
                          const float* view_cos, const uint8_t* qdesc, float th, float nnratio, int* kp_match, int* nmatches,
                          const uint8_t* q_occupies) {
    if (!M || !f || !occupied || !scale_factors || nq < 0 || !kp_match || !nmatches) return ORBM_ERR_INVALID;
    const bool bFactor = th != 1.0;
    SearchIO io;
    io.reserve(nq);
    for (int q = 0; q < nq; ++q) {
        if (!valid[q]) continue;
        const int l = level[q];
        if (l < 0 || l >= nlevels) return ORBM_ERR_INVALID;
        float r = (view_cos[q] > 0.998) ? 2.5f : 4.0f;           // RadiusByViewingCos (:215-221)
        if (bFactor) r *= th;
        io.qwin.push_back(make_float4(proj_x[q], proj_y[q], r * scale_factors[l], 0));
        io.qlev.push_back(levelBounds(l - 1, l));
        io.qsrc.push_back(q);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = qdesc; io.occ = occupied; io.qocc = q_occupies;
    SearchParams P = baseParams();
    P.source = 0; P.state = 1; P.rule = RULE_LOCAL_MAP; P.nnratio = nnratio;
    int rc = runSearch(M, f, P, io);
    if (rc) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) if (io.result[q] >= 0) { kp_match[io.result[q]] = io.qsrc[q]; ++nm; }
    *nmatches = nm;
    return ORBM_OK;
}

// M2 / M3: src/ORBmatcher.cc:1676-1887 (mono), :1889-2010
int orbm_search_frame(orbm_handle M, orbm_frame f, uint8_t* occupied, const float* scale_factors, int nlevels, int nq,
                      const uint8_t* valid, const float* u, const float* v, const int* octave, const float* qangle,
                      const uint8_t* qdesc, float th, int th_dist, int check_orientation, int* kp_match, int* nmatches,
                      const uint8_t* q_occupies) {
    if (!M || !f || !occupied || !scale_factors || nq < 0 || !kp_match || !nmatches) return ORBM_ERR_INVALID;
    SearchIO io;
    io.reserve(nq);
    for (int q = 0; q < nq; ++q) {
        if (!valid[q]) continue;
        const int o = octave[q];
        if (o < 0 || o >= nlevels) return ORBM_ERR_INVALID;
        const float radius = th * scale_factors[o];
        io.qwin.push_back(make_float4(u[q], v[q], radius, 0));
        io.qlev.push_back(levelBounds(o - 1, o + 1));
        io.qangle.push_back(qangle ? qangle[q] : 0.f);
        io.qsrc.push_back(q);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = qdesc; io.occ = occupied; io.qocc = q_occupies;
    SearchParams P = baseParams();
    P.source = 0; P.state = 1; P.rule = RULE_BEST_LE; P.thF = (float)th_dist; P.checkOri = check_orientation && qangle;
    int rc = runSearch(M, f, P, io);
    if (rc) return rc;
    // assignments in query order (a later one overwrites: only possible after a non-occupying match), then the keypoints of
    // the matches the rotation histogram dropped are cleared (:1862-1884); nmatches = assignments - dropped
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) {
        const int r = io.result[q];
        if (r >= 0) { kp_match[r] = io.qsrc[q]; ++nm; }
        else if (r <= -2) kp_match[-2 - r] = io.qsrc[q];
    }
    for (int q = 0; q < io.nq; ++q) if (io.result[q] <= -2) kp_match[-2 - io.result[q]] = -1;
    *nmatches = nm;
    return ORBM_OK;
}

// M7: src/ORBmatcher.cc:427-532, :534-646
int orbm_search_sim3(orbm_handle M, orbm_frame kf, uint8_t* matched, const float* scale_factors, int nlevels, int nq,
                     const uint8_t* valid, const float* u, const float* v, const int* level, const uint8_t* qdesc, int th,
                     float ratio_hamming, int* kp_match, int* nmatches) {
    if (!M || !kf || !matched || !scale_factors || nq < 0 || !kp_match || !nmatches) return ORBM_ERR_INVALID;
    SearchIO io;
    io.reserve(nq);
    for (int q = 0; q < nq; ++q) {
        if (!valid[q]) continue;
        const int l = level[q];
        if (l < 0 || l >= nlevels) return ORBM_ERR_INVALID;
        const float radius = th * scale_factors[l];
        io.qwin.push_back(make_float4(u[q], v[q], radius, 0));
        io.qlev.push_back(make_int2(l - 1, l));                   // the loop's own level test (:505)
        io.qsrc.push_back(q);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = qdesc; io.occ = matched;
    SearchParams P = baseParams();
    P.source = 0; P.state = 1; P.rule = RULE_BEST_LE; P.thF = ORBM_TH_LOW * ratio_hamming;
    int rc = runSearch(M, kf, P, io);
    if (rc) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) if (io.result[q] >= 0) { kp_match[io.result[q]] = io.qsrc[q]; ++nm; }
    *nmatches = nm;
    return ORBM_OK;
}

// M9 (candidate part): src/ORBmatcher.cc:1233-1291, :1400-1437
int orbm_fuse_candidates(orbm_handle M, orbm_frame kf, const float* scale_factors, const float* inv_sigma2, int nlevels, int nq,
                         const uint8_t* valid, const float* u, const float* v, const int* level, const uint8_t* qdesc, float th,
                         int chi2_gate, int* best_idx, int* best_dist) {
    if (!M || !kf || !scale_factors || nq < 0 || !best_idx || !best_dist || (chi2_gate && !inv_sigma2)) return ORBM_ERR_INVALID;
    SearchIO io;
    for (int q = 0; q < nq; ++q) {
        best_idx[q] = -1; best_dist[q] = 256;
        if (!valid[q]) continue;
        const int l = level[q];
        if (l < 0 || l >= nlevels) return ORBM_ERR_INVALID;
        io.qwin.push_back(make_float4(u[q], v[q], th * scale_factors[l], 0));
        io.qlev.push_back(make_int2(l - 1, l));
        io.qsrc.push_back(q);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = qdesc;
    SearchParams P = baseParams();
    P.source = 0; P.state = 0; P.rule = RULE_NONE; P.chi2 = chi2_gate;
    // the chi2 table rides in the (otherwise unused) elig slot of the staging area: upload it separately
    float* dSig = nullptr;
    if (chi2_gate) {
        CKM(cudaSetDevice(M->device));
        CKM(cMalloc((void**)&dSig, (size_t)nlevels * sizeof(float)));
        cudaError_t e = cudaMemcpyAsync(dSig, inv_sigma2, (size_t)nlevels * sizeof(float), cudaMemcpyHostToDevice, M->stream);
        if (e != cudaSuccess) { cudaFree(dSig); CKM(e); }
        P.invSigma2 = dSig;
    }
    int rc = runSearch(M, kf, P, io);
    if (dSig) cudaFree(dSig);
    if (rc) return rc;
    for (int q = 0; q < io.nq; ++q) { best_idx[io.qsrc[q]] = io.result[q]; best_dist[io.qsrc[q]] = io.dist[q]; }
    return ORBM_OK;
}

// M8: src/ORBmatcher.cc:1457-1674 — one stateless window search per direction, then the agreement check
static int sim3OneWay(orbm_matcher* M, const orbm_frame_s* T, const float* sf, int nlevels, int nq, const uint8_t* valid,
                      const float* u, const float* v, const int* level, const uint8_t* qdesc, float th, std::vector<int>& match) {
    SearchIO io;
    match.assign(nq, -1);
    for (int q = 0; q < nq; ++q) {
        if (!valid[q]) continue;
        const int l = level[q];
        if (l < 0 || l >= nlevels) return ORBM_ERR_INVALID;
        io.qwin.push_back(make_float4(u[q], v[q], th * sf[l], 0));
        io.qlev.push_back(make_int2(l - 1, l));
        io.qsrc.push_back(q);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = qdesc;
    SearchParams P = baseParams();
    P.source = 0; P.state = 0; P.rule = RULE_NONE;
    const int rc = runSearch(M, T, P, io);
    if (rc) return rc;
    for (int q = 0; q < io.nq; ++q)
        if (io.result[q] >= 0 && io.dist[q] <= ORBM_TH_HIGH) match[io.qsrc[q]] = io.result[q];   // bestDist<=TH_HIGH (:1569,1649)
    return ORBM_OK;
}
int orbm_search_by_sim3(orbm_handle M, orbm_frame kf1, orbm_frame kf2, const float* scale_factors1, const float* scale_factors2,
                        int nlevels, const uint8_t* valid1, const float* u12, const float* v12, const int* level12,
                        const uint8_t* qdesc1, const uint8_t* valid2, const float* u21, const float* v21, const int* level21,
                        const uint8_t* qdesc2, float th, int* matches12, int* nfound) {
    if (!M || !kf1 || !kf2 || !scale_factors1 || !scale_factors2 || !matches12 || !nfound) return ORBM_ERR_INVALID;
    if ((kf1->n && (!valid1 || !u12 || !v12 || !level12 || !qdesc1)) || (kf2->n && (!valid2 || !u21 || !v21 || !level21 || !qdesc2)))
        return ORBM_ERR_INVALID;
    std::vector<int> m1, m2;
    int rc = sim3OneWay(M, kf2, scale_factors2, nlevels, kf1->n, valid1, u12, v12, level12, qdesc1, th, m1);
    long long ev = M->lastEvals;
    if (rc) return rc;
    if ((rc = sim3OneWay(M, kf1, scale_factors1, nlevels, kf2->n, valid2, u21, v21, level21, qdesc2, th, m2))) return rc;
    M->lastEvals += ev;
    int n = 0;
    for (int i1 = 0; i1 < kf1->n; ++i1) {                      // "check agreement" (:1655-1671)
        const int idx2 = m1[i1];
        matches12[i1] = (idx2 >= 0 && m2[idx2] == i1) ? idx2 : -1;
        n += matches12[i1] >= 0;
    }
    *nfound = n;
    return ORBM_OK;
}

// M5: src/ORBmatcher.cc:648-763
int orbm_search_initialization(orbm_handle M, orbm_frame f1, orbm_frame f2, float* prev_matched, int window_size, float nnratio,
                               int check_orientation, int* matches12, int* nmatches) {
    if (!M || !f1 || !f2 || !prev_matched || !matches12 || !nmatches) return ORBM_ERR_INVALID;
    int rc = ensureHostKps(M, f1);
    if (rc) return rc;
    if ((rc = ensureHostKps(M, f2))) return rc;
    const int n1 = f1->n;
    // query descriptors = F1's rows: fetch them to the host once (F1 may be device-resident)
    std::vector<uint8_t> d1((size_t)std::max(n1, 1) * 32);
    if (n1) {
        CKM(cudaSetDevice(M->device));
        CKM(cudaMemcpyAsync(d1.data(), f1->descDev, (size_t)n1 * 32, cudaMemcpyDeviceToHost, M->stream));
        CKM(syncStream(M));
    }
    SearchIO io;
    for (int i1 = 0; i1 < n1; ++i1) {
        matches12[i1] = -1;
        const int level1 = f1->kpsHost[i1].octave;
        if (level1 > 0) continue;
        io.qwin.push_back(make_float4(prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window_size, 0));
        io.qlev.push_back(levelBounds(level1, level1));
        io.qangle.push_back(f1->kpsHost[i1].angle);
        io.qsrc.push_back(i1);
    }
    io.nq = (int)io.qsrc.size(); io.qdescSrc = d1.data();
    SearchParams P = baseParams();
    P.source = 0; P.state = 2; P.rule = RULE_INIT; P.thI = ORBM_TH_LOW; P.nnratio = nnratio; P.none2 = INT_MAX;
    P.checkOri = check_orientation;
    if ((rc = runSearch(M, f2, P, io))) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q)
        if (io.result[q] >= 0) {
            const int i1 = io.qsrc[q];
            matches12[i1] = io.result[q];
            prev_matched[2 * i1] = f2->kpsHost[io.result[q]].x;       // "update prev matched" (:758-760)
            prev_matched[2 * i1 + 1] = f2->kpsHost[io.result[q]].y;
            ++nm;
        }
    *nmatches = nm;
    return ORBM_OK;
}

}  // extern "C"

// ---- BoW-guided searches ---------------------------------------------------------------------------------------------
namespace {

bool validFeatVec(const orbm_featvec* fv, int n) {
    if (!fv || fv->n_nodes < 0) return false;
    if (fv->n_nodes == 0) return true;
    if (!fv->node || !fv->off || !fv->idx) return false;
    for (int i = 1; i < fv->n_nodes; ++i) if (fv->node[i] <= fv->node[i - 1]) return false;
    for (int i = 0; i < fv->n_nodes; ++i) if (fv->off[i + 1] < fv->off[i]) return false;
    for (int i = fv->off[0]; i < fv->off[fv->n_nodes]; ++i) if (fv->idx[i] < 0 || fv->idx[i] >= n) return false;
    return true;
}

// A DBoW2 FeatureVector lists every feature under exactly one node; a hand-built orbm_featvec need not.  The node-parallel
// search (k_search_nodes) relies on it for side 2.  ORBM_BOW_NODES=0 forces the single-CTA path (A/B switch).
bool listsEachFeatureOnce(const orbm_featvec* fv, int n) {
    static const bool enabled = [] { const char* e = getenv("ORBM_BOW_NODES"); return !(e && e[0] == '0'); }();
    if (!enabled) return false;
    if (fv->n_nodes == 0) return true;
    std::vector<uint8_t> seen((size_t)std::max(n, 1), 0);
    for (int i = fv->off[0]; i < fv->off[fv->n_nodes]; ++i) { if (seen[fv->idx[i]]) return false; seen[fv->idx[i]] = 1; }
    return true;
}

// The merge-join of two DBoW2::FeatureVector maps (e.g. src/ORBmatcher.cc:241-392): for every shared node, every
// eligible feature of side 1 becomes a query (in node order, then list order) whose candidates are the node's run of
// side 2.
void joinQueries(const orbm_featvec* a, const orbm_featvec* b, const uint8_t* elig1, SearchIO& io, bool nodeRanges = false) {
    int ia = 0, ib = 0;
    while (ia < a->n_nodes && ib < b->n_nodes) {
        if (a->node[ia] == b->node[ib]) {
            const int q0 = (int)io.qsrc.size();
            for (int p = a->off[ia]; p < a->off[ia + 1]; ++p) {
                const int i1 = a->idx[p];
                if (elig1 && !elig1[i1]) continue;
                io.qrun.push_back(make_int2(b->off[ib] - b->off[0], b->off[ib + 1] - b->off[0]));
                io.qsrc.push_back(i1);
            }
            if (nodeRanges && (int)io.qsrc.size() > q0) io.nodeQ.push_back(make_int2(q0, (int)io.qsrc.size()));
            ++ia; ++ib;
        } else if (a->node[ia] < b->node[ib]) {
            ia = (int)(std::lower_bound(a->node, a->node + a->n_nodes, b->node[ib]) - a->node);
        } else {
            ib = (int)(std::lower_bound(b->node, b->node + b->n_nodes, a->node[ia]) - b->node);
        }
    }
    io.nq = (int)io.qsrc.size();
    io.bowIdx = b->n_nodes ? b->idx + b->off[0] : nullptr;
    io.nBow = b->n_nodes ? b->off[b->n_nodes] - b->off[0] : 0;
}

int fetchDescHost(orbm_matcher* M, orbm_frame_s* f, std::vector<uint8_t>& out) {
    out.resize((size_t)std::max(f->n, 1) * 32);
    if (f->n) {
        CKM(cudaSetDevice(M->device));
        CKM(cudaMemcpyAsync(out.data(), f->descDev, (size_t)f->n * 32, cudaMemcpyDeviceToHost, M->stream));
        CKM(syncStream(M));
    }
    return ORBM_OK;
}

}  // namespace

extern "C" {

// M4a: src/ORBmatcher.cc:223-425
int orbm_search_bow_kf_frame(orbm_handle M, orbm_frame kf, const orbm_featvec* fv_kf, const uint8_t* elig1, orbm_frame f,
                             const orbm_featvec* fv_f, float nnratio, int check_orientation, int* match_f, int* nmatches) {
    if (!M || !kf || !f || !elig1 || !match_f || !nmatches || !validFeatVec(fv_kf, kf->n) || !validFeatVec(fv_f, f->n)) return ORBM_ERR_INVALID;
    int rc = ensureHostKps(M, kf);
    if (rc) return rc;
    std::vector<uint8_t> d1;
    const bool byNode = listsEachFeatureOnce(fv_f, f->n);
    if (!byNode && (rc = fetchDescHost(M, kf, d1))) return rc;
    for (int j = 0; j < f->n; ++j) match_f[j] = -1;
    SearchIO io;
    joinQueries(fv_kf, fv_f, elig1, io, byNode);
    for (int q = 0; q < io.nq; ++q) io.qangle.push_back(kf->kpsHost[io.qsrc[q]].angle);
    io.qdescSrc = d1.data();
    if (byNode) io.qdescDev = reinterpret_cast<const uint8_t*>(kf->descDev);
    std::vector<uint8_t> occ((size_t)std::max(f->n, 1), 0);        // vpMapPointMatches starts all NULL (:227)
    io.occ = occ.data();
    SearchParams P = baseParams();
    P.source = 1; P.state = 1; P.rule = RULE_RATIO_LE; P.thI = ORBM_TH_LOW; P.nnratio = nnratio; P.checkOri = check_orientation;
    if ((rc = runSearch(M, f, P, io))) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) if (io.result[q] >= 0) { match_f[io.result[q]] = io.qsrc[q]; ++nm; }
    *nmatches = nm;
    return ORBM_OK;
}

// M4b: src/ORBmatcher.cc:765-905
int orbm_search_bow_kf_kf(orbm_handle M, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* elig1, orbm_frame kf2,
                          const orbm_featvec* fv2, const uint8_t* elig2, float nnratio, int check_orientation, int* matches12,
                          int* nmatches) {
    if (!M || !kf1 || !kf2 || !elig1 || !elig2 || !matches12 || !nmatches || !validFeatVec(fv1, kf1->n) || !validFeatVec(fv2, kf2->n))
        return ORBM_ERR_INVALID;
    int rc = ensureHostKps(M, kf1);
    if (rc) return rc;
    std::vector<uint8_t> d1;
    const bool byNode = listsEachFeatureOnce(fv2, kf2->n);
    if (!byNode && (rc = fetchDescHost(M, kf1, d1))) return rc;
    for (int i = 0; i < kf1->n; ++i) matches12[i] = -1;
    SearchIO io;
    joinQueries(fv1, fv2, elig1, io, byNode);
    for (int q = 0; q < io.nq; ++q) io.qangle.push_back(kf1->kpsHost[io.qsrc[q]].angle);
    io.qdescSrc = d1.data();
    if (byNode) io.qdescDev = reinterpret_cast<const uint8_t*>(kf1->descDev);
    io.elig = elig2;
    std::vector<uint8_t> occ((size_t)std::max(kf2->n, 1), 0);      // vbMatched2 (:778)
    io.occ = occ.data();
    SearchParams P = baseParams();
    P.source = 1; P.state = 1; P.rule = RULE_RATIO_LT; P.thI = ORBM_TH_LOW; P.nnratio = nnratio; P.checkOri = check_orientation;
    if ((rc = runSearch(M, kf2, P, io))) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) if (io.result[q] >= 0) { matches12[io.qsrc[q]] = io.result[q]; ++nm; }
    *nmatches = nm;
    return ORBM_OK;
}

// M6: src/ORBmatcher.cc:907-1146 (mono, Pinhole)
int orbm_search_triangulation(orbm_handle M, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* free1, orbm_frame kf2,
                              const orbm_featvec* fv2, const uint8_t* free2, const float* ep, const float* scale_factors2,
                              const float* F12, const float* sigma2, int nlevels, int coarse, int check_orientation,
                              int* matches12, int* nmatches) {
    if (!M || !kf1 || !kf2 || !free1 || !free2 || !ep || !scale_factors2 || !matches12 || !nmatches || nlevels < 1 ||
        (!coarse && (!F12 || !sigma2)) || !validFeatVec(fv1, kf1->n) || !validFeatVec(fv2, kf2->n))
        return ORBM_ERR_INVALID;
    int rc = ensureHostKps(M, kf1);
    if (rc) return rc;
    std::vector<uint8_t> d1;
    if ((rc = fetchDescHost(M, kf1, d1))) return rc;
    for (int i = 0; i < kf1->n; ++i) matches12[i] = -1;
    SearchIO io;
    joinQueries(fv1, fv2, free1, io);
    for (int q = 0; q < io.nq; ++q) {
        const orbx_keypoint& k = kf1->kpsHost[io.qsrc[q]];
        io.qwin.push_back(make_float4(k.x, k.y, 0, 0));
        io.qangle.push_back(k.angle);
    }
    io.qdescSrc = d1.data();
    io.elig = free2;
    SearchParams P = baseParams();
    P.source = 1; P.state = 0; P.rule = RULE_LAST_MIN; P.tri = 1; P.coarse = coarse; P.checkOri = check_orientation;
    P.ep0 = ep[0]; P.ep1 = ep[1];
    if (F12) memcpy(P.F, F12, sizeof(P.F));
    CKM(cudaSetDevice(M->device));
    float* dTab = nullptr;
    CKM(cMalloc((void**)&dTab, (size_t)2 * nlevels * sizeof(float)));
    std::vector<float> tab(2 * (size_t)nlevels, 0.f);
    for (int l = 0; l < nlevels; ++l) { tab[l] = scale_factors2[l]; tab[nlevels + l] = sigma2 ? sigma2[l] : 0.f; }
    cudaError_t e = cudaMemcpyAsync(dTab, tab.data(), tab.size() * sizeof(float), cudaMemcpyHostToDevice, M->stream);
    if (e != cudaSuccess) { cudaFree(dTab); CKM(e); }
    P.scale2 = dTab; P.sigma2 = dTab + nlevels;
    rc = runSearch(M, kf2, P, io);
    cudaFree(dTab);
    if (rc) return rc;
    int nm = 0;
    for (int q = 0; q < io.nq; ++q) if (io.result[q] >= 0) { matches12[io.qsrc[q]] = io.result[q]; ++nm; }
    *nmatches = nm;
    return ORBM_OK;
}

// SearchForTriangulation for cameras whose epipolar predicate stays on the host (any GeometricCamera; the reference ships only
// KannalaBrandt8 settings, whose predicate triangulates through an Eigen SVD: src/CameraModels/KannalaBrandt8.cpp:216-220,
// 306-380).  Returns, per keypoint i1 of KF1, the candidates the reference's inner loop would hand to
// pCamera1->epipolarConstrain if its running `bestDist` were ignored: KF2 keypoints of the same vocabulary node, without a map
// point, DescriptorDistance <= TH_LOW, outside the epipole's 10-px-per-scale disc - in the reference's scan order.
int orbm_triangulation_candidates(orbm_handle M, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* free1, orbm_frame kf2,
                                  const orbm_featvec* fv2, const uint8_t* free2, const float* ep, const float* scale_factors2,
                                  int nlevels, int cap, int* cand_idx2, int* cand_dist, int* cand_count) {
    if (!M || !kf1 || !kf2 || !free1 || !free2 || !ep || !scale_factors2 || nlevels < 1 || cap < 1 || !cand_idx2 || !cand_dist ||
        !cand_count || !validFeatVec(fv1, kf1->n) || !validFeatVec(fv2, kf2->n))
        return ORBM_ERR_INVALID;
    for (int i = 0; i < kf1->n; ++i) cand_count[i] = 0;
    SearchIO io;
    joinQueries(fv1, fv2, free1, io);
    M->lastEvals = 0;
    if (io.nq == 0) return ORBM_OK;
    CKM(cudaSetDevice(M->device));
    const int nq = io.nq, n2 = kf2->n;
    Stage S(M);
    const size_t oSrc = S.take<int>(nq), oRun = S.take<int2>(nq), oBow = S.take<int>(std::max(io.nBow, 1)), oFree = S.take<uint8_t>(std::max(n2, 1));
    const size_t oTab = S.take<float>(nlevels), oEv = S.take<unsigned long long>(1);
    const size_t inEnd = S.off;
    const size_t oCnt = S.take<int>(nq), oIdx = S.take<int>((size_t)nq * cap), oDist = S.take<int>((size_t)nq * cap);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    memcpy(S.host<int>(oSrc), io.qsrc.data(), (size_t)nq * sizeof(int));
    memcpy(S.host<int2>(oRun), io.qrun.data(), (size_t)nq * sizeof(int2));
    if (io.nBow) memcpy(S.host<int>(oBow), io.bowIdx, (size_t)io.nBow * sizeof(int));
    if (n2) memcpy(S.host<uint8_t>(oFree), free2, (size_t)n2);
    memcpy(S.host<float>(oTab), scale_factors2, (size_t)nlevels * sizeof(float));
    *S.host<unsigned long long>(oEv) = 0;
    CKM(cudaMemcpyAsync(M->scratch, M->pinned, inEnd, cudaMemcpyHostToDevice, M->stream));
    k_tri_candidates<<<(nq + 7) / 8, 256, 0, M->stream>>>(nq, S.dev<int>(oSrc), S.dev<int2>(oRun), S.dev<int>(oBow), kf1->descDev,
                                                          kf2->descDev, kf2->kp4, S.dev<uint8_t>(oFree), S.dev<float>(oTab), ep[0], ep[1],
                                                          ORBM_TH_LOW, cap, S.dev<int>(oIdx), S.dev<int>(oDist), S.dev<int>(oCnt),
                                                          S.dev<unsigned long long>(oEv));
    ++M->launches;
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync((uint8_t*)M->pinned + oEv, (uint8_t*)M->scratch + oEv, S.off - oEv, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    M->lastEvals = (long long)*S.host<unsigned long long>(oEv);
    const int* hc = S.host<int>(oCnt); const int* hi = S.host<int>(oIdx); const int* hd = S.host<int>(oDist);
    int status = ORBM_OK;
    for (int q = 0; q < nq; ++q) {
        const int i1 = io.qsrc[q];
        cand_count[i1] = hc[q];                                 // the true number: > cap means the lists were cut
        if (hc[q] > cap) status = ORBM_ERR_CAPACITY;
        const int m = std::min(hc[q], cap);
        memcpy(cand_idx2 + (size_t)i1 * cap, hi + (size_t)q * cap, (size_t)m * sizeof(int));
        memcpy(cand_dist + (size_t)i1 * cap, hd + (size_t)q * cap, (size_t)m * sizeof(int));
    }
    if (status == ORBM_ERR_CAPACITY) M->err = "orbm_triangulation_candidates: a keypoint has more candidates than `cap`; call again with a larger cap";
    return status;
}

// ---- brute-force k-NN ----------------------------------------------------------------------------------------------------
int orbm_db_create(orbm_handle M, const uint8_t* rows, long long ndb, int rows_on_device, long long row_offset, orbm_db* out) {
    if (!M || !out || ndb < 0 || (ndb > 0 && !rows)) return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    orbm_db_s* db = new orbm_db_s();
    db->n = ndb; db->rowOffset = row_offset;
    if (rows_on_device) {
        db->rows = reinterpret_cast<uint32_t*>(const_cast<uint8_t*>(rows)); db->owns = false;
    } else {
        cudaError_t e = cMalloc((void**)&db->rows, (size_t)std::max<long long>(ndb, 1) * 32);
        if (e == cudaSuccess && ndb) e = cudaMemcpy(db->rows, rows, (size_t)ndb * 32, cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { if (db->rows) cudaFree(db->rows); delete db; CKM(e); }
        db->owns = true;
    }
    *out = db;
    return ORBM_OK;
}
void orbm_db_destroy(orbm_handle M, orbm_db db) {
    if (!M || !db) return;
    cudaSetDevice(M->device);
    syncStream(M);
    if (db->owns) cudaFree(db->rows);
    delete db;
}

namespace {

// chunking of a k-NN scan: about 4 resident CTAs per SM over all query blocks.  Chunks are NOT rounded to whole 128-row tiles
// (the last tile of a chunk may be partial; rows are 32 bytes, so any row offset keeps the 16-byte alignment of the bulk copies):
// rounding left e.g. 560 CTAs for 148 SMs on a 125 k-row shard - 3.8 per SM, the time of 4.
struct KnnPlan { int qblocks, nchunks; long long rowsPerChunk; };
KnnPlan knnPlan(const orbm_matcher* M, long long n, int nq, int qpt) {
    KnnPlan P;
    P.qblocks = (nq + kKnnThreads * qpt - 1) / (kKnnThreads * qpt);
    const long long target = std::max<long long>(1, ((long long)M->smCount * 4) / P.qblocks);
    long long rows = (n + target - 1) / target;
    rows = std::max<long long>(kKnnTileRows, (rows + 3) & ~3LL);
    rows = std::min<long long>(rows, 1LL << 22);
    P.rowsPerChunk = rows;
    P.nchunks = (int)std::max<long long>(1, (n + rows - 1) / rows);
    return P;
}

// The chunks of a query block share a pruning bound (k_knn, KnnFused::bound) when a chunk is short: the list warm-up it saves
// is a fixed ~20 us per chunk, 6 % of a 1700-row chunk (a 125 k-row shard: 354 -> 342 us) but nothing against the per-tile
// read it costs on a 13 k-row chunk (1 M rows on one GPU: 2.56 -> 2.59 ms).  ORBM_KNN_BOUND=0 / 1 forces it off / on.
bool knnShareBound(const KnnPlan& P) {
    static const int mode = [] { const char* e = getenv("ORBM_KNN_BOUND"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
    return mode < 0 ? P.rowsPerChunk <= 4096 : mode == 1;
}

int knnScratch(orbm_matcher* M, const KnnPlan& P, int nq, int k) {
    const size_t need = (size_t)P.nchunks * nq * k * sizeof(unsigned long long);
    if (need > M->knnPartBytes) {
        if (M->knnPart) cudaFree(M->knnPart);
        M->knnPart = nullptr; M->knnPartBytes = 0;
        CKM(cMalloc((void**)&M->knnPart, need));
        M->knnPartBytes = need;
    }
    if (nq > M->knnBoundCap) {
        if (M->knnBound) cudaFree(M->knnBound);
        M->knnBound = nullptr; M->knnBoundCap = 0;
        const int cap = std::max(nq, 4096);
        CKM(cMalloc((void**)&M->knnBound, (size_t)cap * sizeof(unsigned int)));
        CKM(cudaMemset(M->knnBound, 0xff, (size_t)cap * sizeof(unsigned int)));
        M->knnBoundCap = cap;
    }
    if (P.qblocks > M->knnCounterCap) {
        if (M->knnCounter) cudaFree(M->knnCounter);
        M->knnCounter = nullptr; M->knnCounterCap = 0;
        const int cap = std::max(P.qblocks, 256);
        CKM(cMalloc((void**)&M->knnCounter, (size_t)cap * sizeof(unsigned int)));
        CKM(cudaMemset(M->knnCounter, 0, (size_t)cap * sizeof(unsigned int)));
        M->knnCounterCap = cap;
    }
    return ORBM_OK;
}

// kK = compile-time bound of k: the sorted insertion is unrolled kK deep, so a k = 2 search pays for 2 steps, not 16
extern "C++" template <int kQ, int kK>
void launchKnnK(orbm_matcher* M, const KnnPlan& P, orbm_db db, const uint8_t* d_q, int nq, int k, const KnnFused& F, cudaStream_t st) {
    const dim3 grid(P.qblocks, P.nchunks);
    const uint32_t* q = reinterpret_cast<const uint32_t*>(d_q);
    if (F.counter) k_knn<kQ, kK, true><<<grid, kKnnThreads, 0, st>>>(q, nq, db->rows, db->n, db->rowOffset, (int)P.rowsPerChunk, k, M->knnPart, F);
    else k_knn<kQ, kK, false><<<grid, kKnnThreads, 0, st>>>(q, nq, db->rows, db->n, db->rowOffset, (int)P.rowsPerChunk, k, M->knnPart, F);
}

void launchKnn(orbm_matcher* M, const KnnPlan& P, int qpt, orbm_db db, const uint8_t* d_q, int nq, int k, const KnnFused& F, cudaStream_t st) {
    if (qpt == 2) launchKnnK<2, 4>(M, P, db, d_q, nq, k, F, st);
    else if (k <= 2) launchKnnK<1, 2>(M, P, db, d_q, nq, k, F, st);
    else if (k <= 4) launchKnnK<1, 4>(M, P, db, d_q, nq, k, F, st);
    else if (k <= 8) launchKnnK<1, 8>(M, P, db, d_q, nq, k, F, st);
    else launchKnnK<1, kKnnMaxK>(M, P, db, d_q, nq, k, F, st);
    ++M->launches;
}

}  // namespace

int orbm_knn_device(orbm_handle M, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx, void* stream) {
    if (!M || !db || !d_q || nq <= 0 || k < 1 || k > kKnnMaxK || !d_dist || !d_idx) return ORBM_ERR_INVALID;
    if (db->rowOffset + db->n > 0x7fffffffLL) { M->err = "global row ids beyond int32"; return ORBM_ERR_UNSUPPORTED; }
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    // two queries per thread when the list is short (registers) and there are enough queries to fill the blocks
    const char* qe = getenv("ORBM_KNN_QPT");
    const int qpt = (k <= 4 && nq >= 2 * kKnnThreads && qe && atoi(qe) == 2) ? 2 : 1;      // measured slower on B200; kept for experiments
    const KnnPlan P = knnPlan(M, db->n, nq, qpt);
    int rc = knnScratch(M, P, nq, k);
    if (rc) return rc;
    KnnFused F;
    memset(&F, 0, sizeof(F));
    // ORBM_KNN_FUSED=1: the CTA that finishes a query block's last chunk merges the block's lists itself (one launch instead of
    // two).  Off by default on one GPU: the epilogue costs the scan kernel 7 registers and measured ~2 % slower than scan +
    // k_knn_merge (7 us); the sharded search always uses it (it is what pushes the lists to the peers).
    static const bool fused = [] { const char* e = getenv("ORBM_KNN_FUSED"); return e && atoi(e) != 0; }();
    if (fused) { F.counter = M->knnCounter; F.dist = d_dist; F.idx = d_idx; }
    if (knnShareBound(P)) F.bound = M->knnBound;
    launchKnn(M, P, qpt, db, d_q, nq, k, F, st);
    if (!fused) { k_knn_merge<<<(nq + 7) / 8, 256, 0, st>>>(M->knnPart, P.nchunks, nq, k, d_dist, d_idx, F.bound); ++M->launches; }
    M->lastEvals = (long long)nq * db->n;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

// ---- k-NN on the tensor cores (A/B beside orbm_knn_device; orbm_kernels_tc.cuh) -----------------------------------------------
namespace {
extern "C++" template <int kK>
void launchKnnTc(orbm_matcher* M, int qtiles, int nchunks, long long rowsPerChunk, orbm_db db, const uint8_t* d_q, int nq, int k, cudaStream_t st) {
    k_knn_tc<kK><<<dim3(qtiles, nchunks), kTcThreads, kTcSmemBytes, st>>>(reinterpret_cast<const uint32_t*>(d_q), nq, db->rows, db->n, db->rowOffset,
                                                                          (int)rowsPerChunk, k, M->knnPart,
                                                                          getenv("ORBM_TC_STATS") ? reinterpret_cast<unsigned long long*>(strtoull(getenv("ORBM_TC_STATS"), nullptr, 10)) : nullptr);
}
}  // namespace

namespace {
// the scan half of the tensor-core search: per-(chunk, slice) lists in M->knnPart, *nLists of them per query
int knnTcScan(orbm_matcher* M, orbm_db db, const uint8_t* d_q, int nq, int k, cudaStream_t st, int* nLists) {
    // one CTA per SM (its two operand stages and the whole TMEM): query tiles of 128 x as many row chunks as fill the SMs
    const int qtiles = (nq + kTcM - 1) / kTcM;
    const long long want = std::max<long long>(1, M->smCount / qtiles);
    long long rows = (std::max<long long>(db->n, 1) + want - 1) / want;
    rows = std::max<long long>(kTcN, (rows + kTcN - 1) / kTcN * kTcN);
    rows = std::min<long long>(rows, (1LL << 22));
    const int nchunks = (int)std::max<long long>(1, (db->n + rows - 1) / rows);
    KnnPlan P; P.qblocks = 1; P.nchunks = nchunks * kTcSlices; P.rowsPerChunk = rows;      // kTcSlices lists per (chunk, query)
    int rc = knnScratch(M, P, nq, k);
    if (rc) return rc;
    CKM(orbx::raise_dyn_smem((const void*)k_knn_tc<2>, M->device, kTcSmemBytes));
    CKM(orbx::raise_dyn_smem((const void*)k_knn_tc<4>, M->device, kTcSmemBytes));
    CKM(orbx::raise_dyn_smem((const void*)k_knn_tc<8>, M->device, kTcSmemBytes));
    CKM(orbx::raise_dyn_smem((const void*)k_knn_tc<kKnnMaxK>, M->device, kTcSmemBytes));
    if (k <= 2) launchKnnTc<2>(M, qtiles, nchunks, rows, db, d_q, nq, k, st);
    else if (k <= 4) launchKnnTc<4>(M, qtiles, nchunks, rows, db, d_q, nq, k, st);
    else if (k <= 8) launchKnnTc<8>(M, qtiles, nchunks, rows, db, d_q, nq, k, st);
    else launchKnnTc<kKnnMaxK>(M, qtiles, nchunks, rows, db, d_q, nq, k, st);
    ++M->launches;
    *nLists = nchunks * kTcSlices;
    return ORBM_OK;
}
}  // namespace

int orbm_knn_device_tc(orbm_handle M, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx, void* stream) {
    if (!M || !db || !d_q || nq <= 0 || k < 1 || k > kKnnMaxK || !d_dist || !d_idx) return ORBM_ERR_INVALID;
    if (db->rowOffset + db->n > 0x7fffffffLL) { M->err = "global row ids beyond int32"; return ORBM_ERR_UNSUPPORTED; }
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    int nLists = 0;
    int rc = knnTcScan(M, db, d_q, nq, k, st, &nLists);
    if (rc) return rc;
    k_knn_merge<<<(nq + 7) / 8, 256, 0, st>>>(M->knnPart, nLists, nq, k, d_dist, d_idx, nullptr);
    ++M->launches;
    M->lastEvals = (long long)nq * db->n;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

// ---- sharded k-NN over peer memory (no counterpart in the reference, which has no GPU and one address space) ----
// Every rank scans its own shard of the database; the all-gather of the per-rank top-k lists and the final merge are part of
// the kernels (KnnFused / k_knn_xchg_merge in orbm_kernels.cuh): k_knn stores a finished query block straight into every rank's
// exchange buffer over NVLink and bumps that rank's arrival counter, k_knn_xchg_merge waits for the counters and merges.
// Buffer of a rank: keys[2 parities][world][slotKeys] | flags[2][kXchgMaxRanks] | status.  Two parities: a rank can be at most
// one call ahead of a peer (it needs the peer's arrival of call t to finish call t), so call t + 1 never overwrites what a
// peer still merges from call t... and call t + 2 reuses the slot of call t only after every peer has started call t + 1.
struct orbm_xchg_s {
    orbm_matcher* M = nullptr;
    int rank = 0, world = 1, maxNq = 0, maxK = 0;
    size_t slotKeys = 0, flagsOff = 0, bytes = 0;
    uint8_t* buf = nullptr;                       // this rank's buffer
    uint8_t* peer[kXchgMaxRanks] = {};            // every rank's buffer as mapped here (peer[rank] == buf)
    bool ipcOpened[kXchgMaxRanks] = {};
    unsigned int expect[2] = {0, 0};              // arrivals every counter of a parity has seen once the current call is complete
    unsigned long long calls = 0;
};

int orbm_xchg_create(orbm_handle M, int rank, int world, int max_nq, int max_k, orbm_xchg* out) {
    if (!M || !out || world < 1 || world > kXchgMaxRanks || rank < 0 || rank >= world || max_nq < 1 || max_k < 1 || max_k > kKnnMaxK)
        return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    orbm_xchg_s* X = new orbm_xchg_s();
    X->M = M; X->rank = rank; X->world = world; X->maxNq = max_nq; X->maxK = max_k;
    X->slotKeys = (size_t)max_nq * max_k;
    X->flagsOff = 2 * (size_t)world * X->slotKeys * sizeof(unsigned long long);
    X->bytes = X->flagsOff + (2 * kXchgMaxRanks + 4) * sizeof(unsigned int);
    if (cMalloc((void**)&X->buf, X->bytes) != cudaSuccess || cudaMemset(X->buf, 0, X->bytes) != cudaSuccess) {
        M->err = cudaGetErrorString(cudaGetLastError());
        delete X;
        return ORBM_ERR_CUDA;
    }
    CKM(cudaDeviceSynchronize());
    X->peer[rank] = X->buf;
    *out = X;
    return ORBM_OK;
}

int orbm_xchg_ipc_handle(orbm_xchg X, void* handle64) {
    if (!X || !handle64) return ORBM_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == ORBM_XCHG_HANDLE_BYTES, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    CKM_(X->M, cudaSetDevice(X->M->device));
    CKM_(X->M, cudaIpcGetMemHandle(&h, X->buf));
    memcpy(handle64, &h, sizeof(h));
    return ORBM_OK;
}

int orbm_xchg_open(orbm_xchg X, int peer_rank, const void* handle64) {
    if (!X || !handle64 || peer_rank < 0 || peer_rank >= X->world) return ORBM_ERR_INVALID;
    if (peer_rank == X->rank) return ORBM_OK;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    void* p = nullptr;
    CKM_(X->M, cudaSetDevice(X->M->device));
    CKM_(X->M, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    X->peer[peer_rank] = static_cast<uint8_t*>(p);
    X->ipcOpened[peer_rank] = true;
    return ORBM_OK;
}

int orbm_xchg_open_local(orbm_xchg X, int peer_rank, orbm_xchg peer) {
    if (!X || !peer || peer_rank < 0 || peer_rank >= X->world || peer->rank != peer_rank || peer->world != X->world ||
        peer->slotKeys != X->slotKeys)
        return ORBM_ERR_INVALID;
    if (peer->M->device != X->M->device) {
        CKM_(X->M, cudaSetDevice(X->M->device));
        const cudaError_t e = cudaDeviceEnablePeerAccess(peer->M->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { X->M->err = cudaGetErrorString(e); return ORBM_ERR_CUDA; }
        cudaGetLastError();
    }
    X->peer[peer_rank] = peer->buf;
    return ORBM_OK;
}

int orbm_xchg_status(orbm_xchg X) {
    if (!X) return ORBM_ERR_INVALID;
    unsigned int v = 0;
    CKM_(X->M, cudaSetDevice(X->M->device));
    CKM_(X->M, cudaMemcpy(&v, X->buf + X->flagsOff + 2 * kXchgMaxRanks * sizeof(unsigned int), sizeof(v), cudaMemcpyDeviceToHost));
    return (int)v;
}

void orbm_xchg_destroy(orbm_xchg X) {
    if (!X) return;
    cudaSetDevice(X->M->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < X->world; ++r) if (X->ipcOpened[r]) cudaIpcCloseMemHandle(X->peer[r]);
    cudaFree(X->buf);
    delete X;
}

int orbm_knn_sharded(orbm_handle M, orbm_xchg X, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx, void* stream) {
    if (!M || !X || X->M != M || !db || !d_q || nq <= 0 || k < 1 || k > kKnnMaxK || !d_dist || !d_idx) return ORBM_ERR_INVALID;
    if (nq > X->maxNq || k > X->maxK) { M->err = "orbm_knn_sharded: nq / k beyond the exchange buffer's capacity"; return ORBM_ERR_CAPACITY; }
    if (db->rowOffset + db->n > 0x7fffffffLL) { M->err = "global row ids beyond int32"; return ORBM_ERR_UNSUPPORTED; }
    for (int r = 0; r < X->world; ++r) if (!X->peer[r]) { M->err = "orbm_knn_sharded: a rank's exchange buffer has not been opened"; return ORBM_ERR_INVALID; }
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    // ORBM_KNN_TC=1 (A/B, every rank alike): the shard is scanned by the tensor-core kernel (orbm_kernels_tc.cuh); the exchange is the same
    static const bool tcScan = [] { const char* e = getenv("ORBM_KNN_TC"); return e && atoi(e) != 0; }();
    const KnnPlan P = knnPlan(M, std::max<long long>(db->n, 1), nq, 1);
    int rc = tcScan ? ORBM_OK : knnScratch(M, P, nq, k);
    if (rc) return rc;
    const int par = (int)(X->calls & 1);
    KnnFused F;
    memset(&F, 0, sizeof(F));
    F.world = X->world;
    if (!tcScan && knnShareBound(P)) F.bound = M->knnBound;
    for (int r = 0; r < X->world; ++r) {
        F.peerSlot[r] = reinterpret_cast<unsigned long long*>(X->peer[r]) + ((size_t)par * X->world + X->rank) * X->slotKeys;
        F.peerFlag[r] = reinterpret_cast<unsigned int*>(X->peer[r] + X->flagsOff) + par * kXchgMaxRanks + X->rank;
    }
    // ORBM_KNN_FUSED=1: the scan kernel's last CTA per query block merges and pushes (one launch less, but that CTA works alone
    // at the end of the scan); default: plain scan, then k_knn_merge_push with a warp per query.  Every rank must use the same
    // setting: the arrival counters count query blocks in one form and merge CTAs in the other.
    static const bool fused = [] { const char* e = getenv("ORBM_KNN_FUSED"); return e && atoi(e) != 0; }();
    if (tcScan) {
        int nLists = 0;
        if ((rc = knnTcScan(M, db, d_q, nq, k, st, &nLists))) return rc;
        k_knn_merge_push<<<(nq + 7) / 8, 256, 0, st>>>(M->knnPart, nLists, nq, k, F);
        ++M->launches;
        X->expect[par] += (unsigned int)((nq + 7) / 8);
    } else if (fused) {
        F.counter = M->knnCounter;
        launchKnn(M, P, 1, db, d_q, nq, k, F, st);
        X->expect[par] += (unsigned int)P.qblocks;     // nq is the same on every rank (the queries are replicated), so is qblocks
    } else {
        launchKnn(M, P, 1, db, d_q, nq, k, F, st);
        k_knn_merge_push<<<(nq + 7) / 8, 256, 0, st>>>(M->knnPart, P.nchunks, nq, k, F);
        ++M->launches;
        X->expect[par] += (unsigned int)((nq + 7) / 8);
    }
    static const unsigned long long timeoutNs = [] { const char* e = getenv("ORBM_XCHG_TIMEOUT_MS"); return (unsigned long long)(e ? atoll(e) : 4000) * 1000000ull; }();
    const unsigned int* flags = reinterpret_cast<const unsigned int*>(X->buf + X->flagsOff) + par * kXchgMaxRanks;
    unsigned int* status = reinterpret_cast<unsigned int*>(X->buf + X->flagsOff) + 2 * kXchgMaxRanks;
    k_knn_xchg_merge<<<(nq + 7) / 8, 256, 0, st>>>(reinterpret_cast<const unsigned long long*>(X->buf) + (size_t)par * X->world * X->slotKeys,
                                                   X->slotKeys, flags, X->expect[par], X->world, nq, k, d_dist, d_idx, status, timeoutNs);
    ++M->launches;
    ++X->calls;
    M->lastEvals = (long long)nq * db->n;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

// test aid: number of allocating CUDA calls (cudaMalloc / cudaMallocHost / cudaStreamCreate) the matcher library has made so far
long long orbm_debug_alloc_count(void) { return g_allocCalls.load(); }

// ---- allocation-free k-NN (no counterpart in the reference) ----------------
// Same search as orbm_knn_device, for callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The per-chunk partial top-k lists - the scratch orbm_knn_device keeps in a cudaMalloc'ed buffer - live in a
// file-scope device array of fixed size: ORBM_KNN_STATIC_MAX_ENTRIES 64-bit keys = 8 MiB, held by every GPU that loads the
// library until the process exits.  Host code must not hand the array's NAME to a kernel (on the host it is a stand-in, not a
// device address): the address is resolved once per device with cudaGetSymbolAddress, with that device current.
// Nothing in the array has to start at zero: k_knn writes every entry k_knn_merge reads (chunk < nchunks, query < nq, slot <
// k) in the same call, so the contents left by the previous call are never read.  All calls share the one array per device:
// calls that overlap on the GPU (different streams, or graphs replayed concurrently) race - serialise them.
static __device__ __align__(256) unsigned long long g_knnStaticPart[ORBM_KNN_STATIC_MAX_ENTRIES];

namespace {
constexpr int kMaxDevicesStatic = 64;
unsigned long long* g_knnStaticPtr[kMaxDevicesStatic] = {};
std::mutex g_knnStaticMu;

// device address of g_knnStaticPart on `device` (which must be current); resolved once per device
cudaError_t knnStaticAddress(int device, unsigned long long** out) {
    std::lock_guard<std::mutex> lk(g_knnStaticMu);
    if (device < 0 || device >= kMaxDevicesStatic) return cudaErrorInvalidDevice;
    if (!g_knnStaticPtr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_knnStaticPart);
        if (e != cudaSuccess) return e;
        g_knnStaticPtr[device] = static_cast<unsigned long long*>(p);
    }
    *out = g_knnStaticPtr[device];
    return cudaSuccess;
}
}  // namespace

int orbm_knn_device_static(orbm_handle M, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx, void* stream) {
    if (!M || !db || !d_q || nq <= 0 || k < 1 || k > kKnnMaxK || !d_dist || !d_idx) return ORBM_ERR_INVALID;
    if (db->rowOffset + db->n > 0x7fffffffLL) { M->err = "global row ids beyond int32"; return ORBM_ERR_UNSUPPORTED; }
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    // the chunking of orbm_knn_device (one query per thread): the result does not depend on it, the scratch size does
    const int qblocks = (nq + kKnnThreads - 1) / kKnnThreads;
    const long long target = std::max<long long>(1, ((long long)M->smCount * 4 + qblocks - 1) / qblocks);
    long long rowsPerChunk = (db->n + target - 1) / target;
    rowsPerChunk = std::max<long long>(kKnnTileRows, (rowsPerChunk + kKnnTileRows - 1) / kKnnTileRows * kKnnTileRows);
    rowsPerChunk = std::min<long long>(rowsPerChunk, 1LL << 22);
    const int nchunks = (int)std::max<long long>(1, (db->n + rowsPerChunk - 1) / rowsPerChunk);
    const unsigned long long need = (unsigned long long)nchunks * (unsigned long long)nq * (unsigned long long)k;
    if (need > (unsigned long long)ORBM_KNN_STATIC_MAX_ENTRIES) {
        char msg[256];
        snprintf(msg, sizeof(msg), "orbm_knn_device_static: %d chunks x %d queries x k=%d = %llu scratch entries exceed the static "
                 "maximum of %llu (ORBM_KNN_STATIC_MAX_ENTRIES); use fewer queries per call or orbm_knn_device",
                 nchunks, nq, k, need, (unsigned long long)ORBM_KNN_STATIC_MAX_ENTRIES);
        M->err = msg;
        return ORBM_ERR_CAPACITY;
    }
    unsigned long long* part = nullptr;
    CKM(knnStaticAddress(M->device, &part));
    KnnFused none;                                              // no fused merge here: its done-counters would be one more buffer
    memset(&none, 0, sizeof(none));
    k_knn<1, kKnnMaxK, false><<<dim3(qblocks, nchunks), kKnnThreads, 0, st>>>(reinterpret_cast<const uint32_t*>(d_q), nq, db->rows, db->n,
                                                                             db->rowOffset, (int)rowsPerChunk, k, part, none);
    k_knn_merge<<<(nq + 7) / 8, 256, 0, st>>>(part, nchunks, nq, k, d_dist, d_idx, nullptr);
    M->launches += 2;
    M->lastEvals = (long long)nq * db->n;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

int orbm_knn(orbm_handle M, orbm_db db, const uint8_t* q, int nq, int k, uint16_t* dist, int32_t* idx) {
    if (!M || !db || !q || nq <= 0 || k < 1 || k > kKnnMaxK || !dist || !idx) return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    Stage S(M);
    const size_t oQ = S.take<uint8_t>((size_t)nq * 32), oD = S.take<uint16_t>((size_t)nq * k), oI = S.take<int32_t>((size_t)nq * k);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    memcpy(S.host<uint8_t>(oQ), q, (size_t)nq * 32);
    CKM(cudaMemcpyAsync(S.dev<uint8_t>(oQ), S.host<uint8_t>(oQ), (size_t)nq * 32, cudaMemcpyHostToDevice, M->stream));
    if ((rc = orbm_knn_device(M, db, S.dev<uint8_t>(oQ), nq, k, S.dev<uint16_t>(oD), S.dev<int32_t>(oI), nullptr))) return rc;
    CKM(cudaMemcpyAsync(S.host<uint8_t>(oD), S.dev<uint8_t>(oD), S.off - oD, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    memcpy(dist, S.host<uint16_t>(oD), (size_t)nq * k * sizeof(uint16_t));
    memcpy(idx, S.host<int32_t>(oI), (size_t)nq * k * sizeof(int32_t));
    return ORBM_OK;
}

// k smallest (dist, global row) per query over nshards sorted lists (host side: 2000 x 8 x k entries at most)
int orbm_knn_merge(const uint16_t* dist, const int32_t* idx, int nshards, int nq, int k, uint16_t* out_dist, int32_t* out_idx) {
    if (!dist || !idx || nshards < 1 || nq < 0 || k < 1 || !out_dist || !out_idx) return ORBM_ERR_INVALID;
    std::vector<unsigned long long> keys((size_t)nshards * k);
    for (int q = 0; q < nq; ++q) {
        size_t m = 0;
        for (int s = 0; s < nshards; ++s)
            for (int t = 0; t < k; ++t) {
                const size_t at = ((size_t)s * nq + q) * k + t;
                if (idx[at] < 0) continue;
                keys[m++] = ((unsigned long long)dist[at] << 32) | (uint32_t)idx[at];
            }
        std::sort(keys.begin(), keys.begin() + m);
        for (int t = 0; t < k; ++t) {
            if ((size_t)t < m) { out_dist[(size_t)q * k + t] = (uint16_t)(keys[t] >> 32); out_idx[(size_t)q * k + t] = (int32_t)(keys[t] & 0xffffffffu); }
            else { out_dist[(size_t)q * k + t] = 65535; out_idx[(size_t)q * k + t] = -1; }
        }
    }
    return ORBM_OK;
}

int orbm_knn_merge_device(orbm_handle M, const void* d_dist, const void* d_idx, size_t shard_stride_bytes, int nshards, int nq,
                          int k, uint16_t* d_out_dist, int32_t* d_out_idx, void* stream) {
    if (!M || !d_dist || !d_idx || nshards < 1 || nq < 0 || k < 1 || !d_out_dist || !d_out_idx) return ORBM_ERR_INVALID;
    if (nq == 0) return ORBM_OK;
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    k_knn_merge_shards<<<(nq + 7) / 8, 256, 0, st>>>((const uint8_t*)d_dist, (const uint8_t*)d_idx, shard_stride_bytes, nshards, nq, k,
                                                     d_out_dist, d_out_idx);
    ++M->launches;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

// MapPoint::ComputeDistinctiveDescriptors selection (src/MapPoint.cc:366-396), batched over map points
int orbm_distinctive_descriptors(orbm_handle M, const uint8_t* desc, const int* off, int npoints, int* best) {
    if (!M || !off || npoints < 0 || !best || (npoints > 0 && off[npoints] > 0 && !desc)) return ORBM_ERR_INVALID;
    if (npoints == 0) return ORBM_OK;
    CKM(cudaSetDevice(M->device));
    const int rows = off[npoints];
    Stage S(M);
    const size_t oD = S.take<uint8_t>((size_t)std::max(rows, 1) * 32), oO = S.take<int>(npoints + 1), oB = S.take<int>(npoints);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    if (rows) memcpy(S.host<uint8_t>(oD), desc, (size_t)rows * 32);
    memcpy(S.host<int>(oO), off, (size_t)(npoints + 1) * sizeof(int));
    CKM(cudaMemcpyAsync(M->scratch, M->pinned, oB, cudaMemcpyHostToDevice, M->stream));
    k_distinctive<<<(npoints + 7) / 8, 256, 0, M->stream>>>(S.dev<uint32_t>(oD), S.dev<int>(oO), npoints, S.dev<int>(oB));
    ++M->launches;
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(S.host<int>(oB), S.dev<int>(oB), (size_t)npoints * sizeof(int), cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    memcpy(best, S.host<int>(oB), (size_t)npoints * sizeof(int));
    return ORBM_OK;
}

}  // extern "C"

// ---- device-resident tracking step -----------------------------------------------------------------------------------
extern "C" int orbm_track_batch_device(orbm_handle M, int nframes, const orbx_keypoint* d_kps, const uint8_t* d_desc,
                                       const int* d_counts, int frame_stride, int count_stride, const float* bounds4,
                                       const float* scale_factors, int nlevels, float th, int th_dist, int check_orientation,
                                       int* d_kp_match, int* d_nmatches, void* stream) {
    if (!M || nframes < 1 || !d_kps || !d_desc || !d_counts || frame_stride < 1 || count_stride < 1 || !bounds4 || !scale_factors ||
        nlevels < 1 || nlevels > 16 || !d_kp_match || !d_nmatches || !(bounds4[2] > bounds4[0]) || !(bounds4[3] > bounds4[1]))
        return ORBM_ERR_INVALID;
    if (frame_stride >= (1 << 19)) return ORBM_ERR_UNSUPPORTED;
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    const size_t NS = (size_t)nframes * frame_stride;
    size_t off = 0;
    auto take = [&](size_t bytes) { off = (off + 255) & ~(size_t)255; size_t at = off; off += bytes; return at; };
    const size_t oKp4 = take(NS * 16), oCellOf = take(NS * 4), oCellIdx = take(NS * 4), oCellOff = take((size_t)nframes * (kGridCells + 1) * 4);
    const size_t oWin = take(NS * 16), oLev = take(NS * 8), oAng = take(NS * 4), oOcc = take(NS);
    const size_t oRes = take(NS * 4), oDist = take(NS * 4), oBin = take(NS * 4);
    const size_t oLists = take(NS * 16), oListCnt = take(NS * 4);
    if (off > M->trackBytes) {
        if (M->trackBuf) { CKM(cudaStreamSynchronize(st)); cudaFree(M->trackBuf); }
        M->trackBuf = nullptr; M->trackBytes = 0;
        CKM(cMalloc(&M->trackBuf, off));
        M->trackBytes = off;
    }
    uint8_t* b = (uint8_t*)M->trackBuf;
    TrackBatch T;
    memset(&T, 0, sizeof(T));
    T.nframes = nframes; T.stride = frame_stride; T.countStride = count_stride;
    T.kps = d_kps; T.desc = reinterpret_cast<const uint32_t*>(d_desc); T.counts = d_counts;
    T.minX = bounds4[0]; T.minY = bounds4[1];
    T.invW = static_cast<float>(kGridCols) / static_cast<float>(bounds4[2] - bounds4[0]);
    T.invH = static_cast<float>(kGridRows) / static_cast<float>(bounds4[3] - bounds4[1]);
    T.th = th; T.thF = (float)th_dist; T.checkOri = check_orientation;
    for (int l = 0; l < 16; ++l) T.scale[l] = l < nlevels ? scale_factors[l] : scale_factors[nlevels - 1];
    T.kp4 = (float4*)(b + oKp4); T.cellOf = (int*)(b + oCellOf); T.cellIdx = (int*)(b + oCellIdx); T.cellOff = (int*)(b + oCellOff);
    T.qwin = (float4*)(b + oWin); T.qlev = (int2*)(b + oLev); T.qangle = (float*)(b + oAng); T.occ = b + oOcc;
    T.qResult = (int*)(b + oRes); T.qDist = (int*)(b + oDist); T.qBin = (int*)(b + oBin);
    T.kpMatch = d_kp_match; T.nMatches = d_nmatches; T.evals = M->dEvals;
    T.lists = (uint4*)(b + oLists); T.listCounts = (int*)(b + oListCnt);
    // Off by default for the batch: measured on 64-frame steps the stage is shorter by itself (0.124 against 0.135 ms) but the
    // overlapped step loses 8 % (113.8 k against 122.9 k frames/s): the 63 resolving CTAs of 1024 threads take a whole SM each
    // where the in-order kernel's 63 CTAs of 256 threads leave room for the other handles' kernels.  ORBM_TRACK_ROUNDS=1 turns
    // it on (A/B).
    static const bool roundsOk = [] { const char* e = getenv("ORBM_TRACK_ROUNDS"); return e && e[0] == '1'; }();
    const bool byRounds = roundsOk && frame_stride <= kRoundsMaxQueries && rounds_smem_bytes(frame_stride, frame_stride) <= kRoundsSmemMax;
    // ORBM_TRACK_STAGE=0 (A/B): the target frame stays in L2 instead of shared memory - a slower kernel with a third of the footprint
    static const bool stageOk = [] { const char* e = getenv("ORBM_TRACK_STAGE"); return !(e && e[0] == '0'); }();
    T.stageTgt = (stageOk && search_smem_bytes(frame_stride, 1, true, true) <= kSearchSmemMax) ? 1 : 0;
    const size_t smem = search_smem_bytes(frame_stride, 1, T.stageTgt != 0, true);
    if (!byRounds && smem > kSearchSmemMax) return ORBM_ERR_UNSUPPORTED;
    CKM(cudaMemsetAsync(M->dEvals, 0, 128, st));
    // the whole step is latency-bound (one CTA per frame pair): dispatched ahead of the queued CTAs of other handles' kernels
    CKM(orbx::launch_prio(k_track_prepare, dim3((frame_stride + 255) / 256, nframes), dim3(256), 0, st, M->prio, T));
    CKM(orbx::launch_prio(k_track_grid, dim3(nframes), dim3(kGridThreads), 0, st, M->prio, T));
    M->launches += 2;
    if (nframes > 1 && byRounds) {
        // claim rounds: a warp per query over the whole GPU (small CTAs), then one CTA per pair resolves, filters and scatters
        CKM(orbx::launch_prio(k_track_walk, dim3((frame_stride + 7) / 8, nframes - 1), dim3(256), 0, st, M->prio, T));
        CKM(orbx::launch_prio(k_track_resolve, dim3(nframes - 1), dim3(kRoundThreads), rounds_smem_bytes(frame_stride, frame_stride), st,
                              M->prio, T));
        M->launches += 2;
    } else if (nframes > 1) {
        CKM(orbx::launch_prio(k_track_search, dim3(nframes - 1), dim3(kSearchThreads), smem, st, M->prio, T));
        CKM(orbx::launch_prio(k_track_finish, dim3(nframes - 1), dim3(256), 0, st, M->prio, T));
        M->launches += 2;
    } else {
        CKM(cudaMemsetAsync(d_nmatches, 0, sizeof(int), st));
    }
    CKM(cudaGetLastError());
    return ORBM_OK;
}

// evaluations of the last asynchronous tracking step (synchronises the handle's device)
extern "C" long long orbm_track_evaluations(orbm_handle M) {
    if (!M) return 0;
    unsigned long long v = 0;
    cudaSetDevice(M->device);
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    if (cudaMemcpy(&v, M->dEvals, 8, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (long long)v;
}

// developer aid, not part of include/orbm.h: how the last single search was resolved - claim rounds run (0 = the in-order
// kernel) and whether it finished in query order on one warp (long chains of competing queries)
extern "C" int orbm_debug_last_rounds(orbm_handle M, int* rounds, int* serial) {
    if (!M) return ORBM_ERR_INVALID;
    if (rounds) *rounds = M->lastRounds;
    if (serial) *serial = M->lastSerial;
    return ORBM_OK;
}

// developer aid, not part of include/orbm.h: the 16 counters of the last tracking step ([0] = evaluations; the rest are
// cycle / replay counters that exist only in builds with -DORBM_STATS)
extern "C" int orbm_debug_track_counters(orbm_handle M, unsigned long long* out16) {
    if (!M || !out16) return ORBM_ERR_INVALID;
    cudaSetDevice(M->device);
    if (cudaDeviceSynchronize() != cudaSuccess) return ORBM_ERR_CUDA;
    return cudaMemcpy(out16, M->dEvals, 128, cudaMemcpyDeviceToHost) == cudaSuccess ? ORBM_OK : ORBM_ERR_CUDA;
}

// ---- Frame-side steps: undistortion and the DBoW2 transform ----------------------------------------------
#include "orbm_kernels_frame.cuh"

struct orbm_vocab_s {
    int* childOff = nullptr; int* children = nullptr; uint4* desc = nullptr; int* wordId = nullptr; double* weight = nullptr;
    int nnodes = 0, levels = 0, maxDepth = 0, weighting = 0, norm = 0, nwords = 0;
    std::vector<double> wordWeight;      // host copy, indexed by word id (BowVector assembly)
};

namespace {

int makeUndistortParams(orbm_matcher* M, const float* K4, const float* dist, int ndist, UndistortParams* P) {
    if (!K4 || !dist || !(ndist == 4 || ndist == 5 || ndist == 8 || ndist == 12)) return ORBM_ERR_INVALID;
    if (!(K4[0] != 0.f) || !(K4[1] != 0.f)) { M->err = "zero focal length"; return ORBM_ERR_INVALID; }
    memset(P, 0, sizeof(*P));
    for (int i = 0; i < ndist; ++i) P->k[i] = (double)dist[i];
    P->fx = K4[0]; P->fy = K4[1]; P->cx = K4[2]; P->cy = K4[3];
    P->ifx = 1. / P->fx; P->ify = 1. / P->fy;
    return ORBM_OK;
}

VocabDev vocabDev(const orbm_vocab_s* v) {
    VocabDev V;
    V.childOff = v->childOff; V.children = v->children; V.desc = v->desc; V.wordId = v->wordId; V.weight = v->weight;
    V.nnodes = v->nnodes; V.levels = v->levels; V.maxDepth = v->maxDepth;
    return V;
}

}  // namespace

extern "C" {

int orbm_undistort_points(orbm_handle M, const float* xy, int n, const float* K4, const float* dist, int ndist, float* out) {
    if (!M || n < 0 || (n > 0 && (!xy || !out))) return ORBM_ERR_INVALID;
    UndistortParams P;
    int rc = makeUndistortParams(M, K4, dist, ndist, &P);
    if (rc) return rc;
    if (n == 0) return ORBM_OK;
    CKM(cudaSetDevice(M->device));
    Stage S(M);
    const size_t oIn = S.take<float>((size_t)n * 2), oOut = S.take<float>((size_t)n * 2);
    rc = ensureScratch(M, S.off);
    if (rc) return rc;
    memcpy(S.host<float>(oIn), xy, (size_t)n * 8);
    CKM(cudaMemcpyAsync(S.dev<float>(oIn), S.host<float>(oIn), (size_t)n * 8, cudaMemcpyHostToDevice, M->stream));
    k_undistort<<<dim3((n + 255) / 256, 1), 256, 0, M->stream>>>(S.dev<uint8_t>(oIn), S.dev<uint8_t>(oOut), 8, nullptr, 0, n, n, 0, P);
    ++M->launches;
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(S.host<float>(oOut), S.dev<float>(oOut), (size_t)n * 8, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    memcpy(out, S.host<float>(oOut), (size_t)n * 8);
    return ORBM_OK;
}

int orbm_undistort_keypoints_device(orbm_handle M, int nframes, const orbx_keypoint* d_kps, const int* d_counts, int frame_stride,
                                    int count_stride, const float* K4, const float* dist, int ndist, orbx_keypoint* d_kps_un,
                                    void* stream) {
    if (!M || nframes < 1 || nframes > 65535 || !d_kps || !d_counts || frame_stride < 1 || count_stride < 1 || !d_kps_un)
        return ORBM_ERR_INVALID;
    UndistortParams P;
    int rc = makeUndistortParams(M, K4, dist, ndist, &P);
    if (rc) return rc;
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    static_assert(sizeof(orbx_keypoint) == 28, "cv::KeyPoint layout");
    const int bx = std::min((frame_stride + 255) / 256, 64);
    k_undistort<<<dim3(bx, nframes), 256, 0, st>>>(reinterpret_cast<const uint8_t*>(d_kps), reinterpret_cast<uint8_t*>(d_kps_un),
                                                   (int)sizeof(orbx_keypoint), d_counts, count_stride, frame_stride, 0, 1, P);
    ++M->launches;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

int orbm_vocab_create(orbm_handle M, const orbm_vocab_view* v, orbm_vocab* out) {
    if (!M || !v || !out || v->nnodes < 2 || v->levels < 1 || !v->child_off || !v->children || !v->node_desc || !v->word_id ||
        !v->weight || v->weighting < 0 || v->weighting > 3 || v->norm < 0 || v->norm > 2)
        return ORBM_ERR_INVALID;
    const int nn = v->nnodes;
    // the arrays must describe one tree rooted at node 0 (the device loop follows them without further checks)
    if (v->child_off[0] != 0) { M->err = "child_off[0] != 0"; return ORBM_ERR_INVALID; }
    for (int p = 0; p < nn; ++p)
        if (v->child_off[p + 1] < v->child_off[p]) { M->err = "child_off not monotone"; return ORBM_ERR_INVALID; }
    const int nch = v->child_off[nn];
    if (nch != nn - 1) { M->err = "children do not form a tree (expected nnodes - 1 edges)"; return ORBM_ERR_INVALID; }
    std::vector<int> depth(nn, -1);
    std::vector<int> order; order.reserve(nn);
    depth[0] = 0; order.push_back(0);
    int maxDepth = 0, nwords = 0;
    for (size_t h = 0; h < order.size(); ++h) {
        const int p = order[h];
        for (int c = v->child_off[p]; c < v->child_off[p + 1]; ++c) {
            const int id = v->children[c];
            if (id <= 0 || id >= nn || depth[id] >= 0) { M->err = "child id out of range or reached twice"; return ORBM_ERR_INVALID; }
            depth[id] = depth[p] + 1; maxDepth = std::max(maxDepth, depth[id]);
            order.push_back(id);
        }
        if (v->child_off[p] == v->child_off[p + 1]) {
            if (v->word_id[p] < 0) { M->err = "negative word id"; return ORBM_ERR_INVALID; }
            nwords = std::max(nwords, v->word_id[p] + 1);
        }
    }
    if ((int)order.size() != nn) { M->err = "nodes unreachable from the root"; return ORBM_ERR_INVALID; }
    CKM(cudaSetDevice(M->device));
    orbm_vocab_s* V = new orbm_vocab_s();
    V->nnodes = nn; V->levels = v->levels; V->maxDepth = maxDepth; V->weighting = v->weighting; V->norm = v->norm; V->nwords = nwords;
    V->wordWeight.assign(nwords, 0.0);
    for (int p = 0; p < nn; ++p)
        if (v->child_off[p] == v->child_off[p + 1]) V->wordWeight[v->word_id[p]] = v->weight[p];
    cudaError_t e = cudaSuccess;
    auto up = [&](void** dst, const void* src, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cMalloc(dst, std::max<size_t>(bytes, 16));
        if (e == cudaSuccess && bytes) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    up((void**)&V->childOff, v->child_off, (size_t)(nn + 1) * 4);
    up((void**)&V->children, v->children, (size_t)nch * 4);
    up((void**)&V->desc, v->node_desc, (size_t)nn * 32);
    up((void**)&V->wordId, v->word_id, (size_t)nn * 4);
    up((void**)&V->weight, v->weight, (size_t)nn * 8);
    if (e != cudaSuccess) {
        cudaFree(V->childOff); cudaFree(V->children); cudaFree(V->desc); cudaFree(V->wordId); cudaFree(V->weight);
        delete V;
        CKM(e);
    }
    *out = V;
    return ORBM_OK;
}

void orbm_vocab_destroy(orbm_handle M, orbm_vocab V) {
    if (!M || !V) return;
    cudaSetDevice(M->device);
    syncStream(M);
    cudaFree(V->childOff); cudaFree(V->children); cudaFree(V->desc); cudaFree(V->wordId); cudaFree(V->weight);
    delete V;
}

int orbm_bow_transform_device(orbm_handle M, orbm_vocab V, int nframes, const uint8_t* d_desc, const int* d_counts, int frame_stride,
                              int count_stride, int levelsup, int* d_feat_word, int* d_feat_node, void* stream) {
    if (!M || !V || nframes < 1 || nframes > 65535 || !d_desc || !d_counts || frame_stride < 1 || count_stride < 1 || levelsup < 0 ||
        !d_feat_word || !d_feat_node)
        return ORBM_ERR_INVALID;
    CKM(cudaSetDevice(M->device));
    cudaStream_t st = stream ? (cudaStream_t)stream : M->stream;
    const int bx = std::min((frame_stride + 127) / 128, 256);
    k_bow_descend<<<dim3(bx, nframes), 128, 0, st>>>(vocabDev(V), reinterpret_cast<const uint4*>(d_desc), d_counts, count_stride,
                                                     frame_stride, 0, levelsup, d_feat_word, d_feat_node);
    ++M->launches;
    CKM(cudaGetLastError());
    return ORBM_OK;
}

int orbm_bow_transform(orbm_handle M, orbm_vocab V, const uint8_t* desc, int n, int levelsup, int* feat_word, int* feat_node,
                       int* n_bow, int* bow_word, double* bow_value, int* n_fv, int* fv_node, int* fv_off, int* fv_idx) {
    if (!M || !V || n < 0 || (n > 0 && !desc) || levelsup < 0 || !n_bow || !n_fv || !fv_off ||
        (n > 0 && (!bow_word || !bow_value || !fv_node || !fv_idx)))
        return ORBM_ERR_INVALID;
    *n_bow = 0; *n_fv = 0; fv_off[0] = 0;
    if (n == 0) return ORBM_OK;
    CKM(cudaSetDevice(M->device));
    Stage S(M);
    const size_t oD = S.take<uint8_t>((size_t)n * 32), oW = S.take<int>(n), oN = S.take<int>(n);
    int rc = ensureScratch(M, S.off);
    if (rc) return rc;
    memcpy(S.host<uint8_t>(oD), desc, (size_t)n * 32);
    CKM(cudaMemcpyAsync(S.dev<uint8_t>(oD), S.host<uint8_t>(oD), (size_t)n * 32, cudaMemcpyHostToDevice, M->stream));
    k_bow_descend<<<dim3((n + 127) / 128, 1), 128, 0, M->stream>>>(vocabDev(V), S.dev<uint4>(oD), nullptr, 0, n, n, levelsup,
                                                                  S.dev<int>(oW), S.dev<int>(oN));
    ++M->launches;
    CKM(cudaGetLastError());
    CKM(cudaMemcpyAsync(S.host<int>(oW), S.dev<int>(oW), (size_t)(oN - oW) + (size_t)n * 4, cudaMemcpyDeviceToHost, M->stream));
    CKM(syncStream(M));
    const int* w = S.host<int>(oW);
    const int* nd = S.host<int>(oN);
    if (feat_word) memcpy(feat_word, w, (size_t)n * 4);
    if (feat_node) memcpy(feat_node, nd, (size_t)n * 4);
    // BowVector (BowVector.cpp:35-88): per word the weights are added in feature order, the norm runs over ascending words.
    // FeatureVector (FeatureVector.cpp:32-46): ascending node ids, feature indices in insertion (= ascending) order.
    std::vector<uint64_t> byWord, byNode;
    byWord.reserve(n); byNode.reserve(n);
    for (int i = 0; i < n; ++i) {
        if (w[i] < 0) continue;                                         // stopped word
        byWord.push_back(((uint64_t)(uint32_t)w[i] << 32) | (uint32_t)i);
        byNode.push_back(((uint64_t)(uint32_t)nd[i] << 32) | (uint32_t)i);
    }
    std::sort(byWord.begin(), byWord.end());
    std::sort(byNode.begin(), byNode.end());
    const bool tf = V->weighting == 0 || V->weighting == 1;
    int nb = 0;
    for (size_t a = 0; a < byWord.size();) {
        const int word = (int)(byWord[a] >> 32);
        const double wt = V->wordWeight[word];
        double val = wt;
        size_t b = a + 1;
        for (; b < byWord.size() && (int)(byWord[b] >> 32) == word; ++b)
            if (tf) val += wt;                                          // addWeight; addIfNotExist keeps the first
        bow_word[nb] = word; bow_value[nb] = val; ++nb;
        a = b;
    }
    if (tf && nb > 0 && V->norm == 0) {
        const double cnt = (double)nb;
        for (int a = 0; a < nb; ++a) bow_value[a] /= cnt;
    }
    if (V->norm != 0) {
        double norm = 0.0;
        if (V->norm == 1) { for (int a = 0; a < nb; ++a) norm += std::fabs(bow_value[a]); }
        else { for (int a = 0; a < nb; ++a) norm += bow_value[a] * bow_value[a]; norm = std::sqrt(norm); }
        if (norm > 0.0) for (int a = 0; a < nb; ++a) bow_value[a] /= norm;
    }
    *n_bow = nb;
    int nf = 0, o = 0;
    for (size_t a = 0; a < byNode.size(); ++a) {
        const int node = (int)(byNode[a] >> 32);
        if (nf == 0 || fv_node[nf - 1] != node) { fv_node[nf] = node; fv_off[nf] = o; ++nf; }
        fv_idx[o++] = (int)(uint32_t)byNode[a];
    }
    fv_off[nf] = o;
    *n_fv = nf;
    return ORBM_OK;
}

}  // extern "C"
