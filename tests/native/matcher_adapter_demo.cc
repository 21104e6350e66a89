// Runs every search / fuse member of the C++ drop-in ORBmatcher (mam3slam_b200/cpp/ORBmatcher.h) on mock Frame /
// KeyFrame / MapPoint objects (slam_shim.h) built from a scenario file written by tests/test_cpp_adapter_gpu.py, the way
// Tracking / LocalMapping / LoopClosing call the reference class, and prints the results ("name ret n v0 v1 ...") for the
// Python side to compare against the oracle.  argv[1] = scenario file.
#include "slam_shim.h"
#include "../../mam3slam_b200/cpp/ORBmatcher.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <string>

using namespace shim;

typedef std::map<std::string, std::vector<uint8_t> > Blob;
static Blob load(const char* path) {
    Blob b;
    FILE* f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(3); }
    int32_t n = 0;
    if (fread(&n, 4, 1, f) != 1) exit(3);
    for (int i = 0; i < n; i++) {
        char name[17] = {0};
        int64_t bytes = 0;
        if (fread(name, 1, 16, f) != 16 || fread(&bytes, 8, 1, f) != 1) exit(3);
        std::vector<uint8_t> v((size_t)bytes);
        if (bytes && fread(v.data(), 1, (size_t)bytes, f) != (size_t)bytes) exit(3);
        b[name] = v;
    }
    fclose(f);
    return b;
}
template <class T> static const T* arr(const Blob& b, const char* name, size_t* n = nullptr) {
    auto it = b.find(name);
    if (it == b.end()) { fprintf(stderr, "missing array %s\n", name); exit(4); }
    if (n) *n = it->second.size() / sizeof(T);
    return reinterpret_cast<const T*>(it->second.data());
}

struct World {
    Camera cam;
    std::vector<std::unique_ptr<MapPoint> > mp;
    KeyFrame KF1, KF2;
    Frame F1, F2;
    std::vector<MapPoint*> pts;
    std::set<MapPoint*> already;
    Blob blob;
    MapPoint* at(int id) { return id < 0 ? nullptr : mp[id].get(); }
};

static void fillBase(FrameBase& F, const Blob& b, const char* k, const char* d, const char* node, const float* t, Camera* cam) {
    size_t n = 0;
    const cv::KeyPoint* kp = arr<cv::KeyPoint>(b, k, &n);
    F.N = (int)n;
    F.mvKeys.assign(kp, kp + n);
    F.mvKeysUn = F.mvKeys;
    F.mDescriptors = cv::Mat((int)n, 32, CV_8UC1, (void*)arr<uint8_t>(b, d), 32);
    const float* bd = arr<float>(b, "bounds");
    F.mnMinX = bd[0]; F.mnMinY = bd[1]; F.mnMaxX = bd[2]; F.mnMaxY = bd[3];
    size_t nl = 0;
    const float* sf = arr<float>(b, "scale", &nl);
    F.mvScaleFactors.assign(sf, sf + nl);
    F.mvLevelSigma2.assign(arr<float>(b, "sigma2"), arr<float>(b, "sigma2") + nl);
    F.mvInvLevelSigma2.assign(arr<float>(b, "invsigma2"), arr<float>(b, "invsigma2") + nl);
    const int32_t* nd = arr<int32_t>(b, node);
    F.mFeatVec.clear();
    for (size_t i = 0; i < n; i++) F.mFeatVec[(unsigned)nd[i]].push_back((unsigned)i);      // FeatureVector::addFeature order
    F.mpCamera = cam;
    F.Tcw = SE3(Mat3(), Vec3(t[0], t[1], t[2]));
}

static std::unique_ptr<World> build(const char* path) {
    std::unique_ptr<World> W(new World);
    W->blob = load(path);
    const Blob& b = W->blob;
    size_t nmp = 0;
    const int32_t* lvl = arr<int32_t>(b, "mp_level", &nmp);
    const float* pos = arr<float>(b, "mp_pos");
    const float* nrm = arr<float>(b, "mp_normal");
    for (size_t i = 0; i < nmp; i++) {
        std::unique_ptr<MapPoint> p(new MapPoint);
        p->id = (int)i;
        p->pos = Vec3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        p->normal = Vec3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
        p->level = lvl[i];
        p->desc = cv::Mat(1, 32, CV_8UC1, (void*)(arr<uint8_t>(b, "mp_desc") + 32 * i), 32);
        p->bad = arr<uint8_t>(b, "mp_bad")[i] != 0;
        p->nObs = arr<int32_t>(b, "mp_nobs")[i];
        p->mbTrackInView = arr<uint8_t>(b, "mp_inview")[i] != 0;
        p->mTrackProjX = arr<float>(b, "mp_projx")[i];
        p->mTrackProjY = arr<float>(b, "mp_projy")[i];
        p->mTrackDepth = arr<float>(b, "mp_depth")[i];
        p->mTrackViewCos = arr<float>(b, "mp_viewcos")[i];
        p->mnTrackScaleLevel = arr<int32_t>(b, "mp_tracklevel")[i];
        W->mp.push_back(std::move(p));
    }
    const float* t1 = arr<float>(b, "t1");
    const float* t2 = arr<float>(b, "t2");
    fillBase(W->KF1, b, "k1", "d1", "node1", t1, &W->cam);
    fillBase(W->KF2, b, "k2", "d2", "node2", t2, &W->cam);
    fillBase(W->F1, b, "k1", "d1", "node1", t1, &W->cam);
    fillBase(W->F2, b, "k2", "d2", "node2", t2, &W->cam);
    const int32_t* m1 = arr<int32_t>(b, "kf1_mp");
    const int32_t* m2 = arr<int32_t>(b, "kf2_mp");
    for (int i = 0; i < W->KF1.N; i++) { W->KF1.mps.push_back(W->at(m1[i])); if (m1[i] >= 0) W->at(m1[i])->obs[&W->KF1] = i; }
    for (int i = 0; i < W->KF2.N; i++) { W->KF2.mps.push_back(W->at(m2[i])); if (m2[i] >= 0) W->at(m2[i])->obs[&W->KF2] = i; }
    W->F1.mvpMapPoints = W->KF1.mps;
    W->F2.mvpMapPoints = W->KF2.mps;
    const uint8_t* o1 = arr<uint8_t>(b, "outlier1");
    W->F1.mvbOutlier.assign(o1, o1 + W->F1.N);
    W->F2.mvbOutlier.assign(W->F2.N, false);
    size_t np = 0;
    const int32_t* pts = arr<int32_t>(b, "pts", &np);
    for (size_t i = 0; i < np; i++) W->pts.push_back(W->at(pts[i]));
    const int32_t* al = arr<int32_t>(b, "already", &np);
    for (size_t i = 0; i < np; i++) W->already.insert(W->at(al[i]));
    g_events.clear();
    return W;
}

static void emit(const char* name, int ret, const std::vector<int>& v) {
    printf("%s %d %zu", name, ret, v.size());
    for (int x : v) printf(" %d", x);
    printf("\n");
}
static std::vector<int> ids(const std::vector<MapPoint*>& v) {
    std::vector<int> o;
    for (MapPoint* p : v) o.push_back(p ? p->id : -1);
    return o;
}
static void emitEvents(const char* name) {
    std::vector<int> o;
    for (const Event& e : g_events) { o.push_back((int)e.kind); o.push_back(e.a); o.push_back(e.b); }
    emit(name, (int)g_events.size(), o);
}

// `pool` mode: the reference constructs an ORBmatcher per call (src/Tracking.cc:2789: `ORBmatcher matcher(0.9,true);` inside
// TrackWithMotionModel).  With the adapter's thread_local pool that must cost the same as reusing one object and must not
// allocate: N calls with a matcher constructed per call against N calls on one object, plus the library's allocation counter.
static int poolMode(const char* path) {
    const int N = 1000;
    auto W = build(path);
    const std::vector<MapPoint*> pristine = W->F2.mvpMapPoints;
    std::vector<int> first, last;
    {   // warm-up: the first object creates the pooled set and its device frame slot
        MAM3SLAM::ORBmatcher m(0.9f, true);
        m.SearchByProjection(W->F2, W->F1, 15.0f, true);
        first = ids(W->F2.mvpMapPoints);
    }
    const long long alloc0 = orbm_debug_alloc_count();
    auto t0 = std::chrono::steady_clock::now();
    int r = 0;
    for (int i = 0; i < N; i++) {
        W->F2.mvpMapPoints = pristine;
        MAM3SLAM::ORBmatcher matcher(0.9f, true);                            // per call, as the reference does
        r = matcher.SearchByProjection(W->F2, W->F1, 15.0f, true);
    }
    auto t1 = std::chrono::steady_clock::now();
    last = ids(W->F2.mvpMapPoints);
    const long long alloc1 = orbm_debug_alloc_count();
    double reusedUs = 0;
    {
        MAM3SLAM::ORBmatcher one(0.9f, true);
        auto t2 = std::chrono::steady_clock::now();
        for (int i = 0; i < N; i++) {
            W->F2.mvpMapPoints = pristine;
            r = one.SearchByProjection(W->F2, W->F1, 15.0f, true);
        }
        reusedUs = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t2).count() / N;
    }
    const double perCallUs = std::chrono::duration<double, std::micro>(t1 - t0).count() / N;
    std::vector<int> o = {N, (int)(perCallUs * 10), (int)(reusedUs * 10), (int)(alloc1 - alloc0), (int)MAM3SLAM::ORBmatcher::PooledSets(),
                          first == last ? 1 : 0};
    emit("pool", r, o);
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    const char* path = argv[1];
    if (argc >= 3 && std::string(argv[2]) == "pool") return poolMode(path);
    MAM3SLAM::ORBmatcher matcher(0.8f, true);            // Tracking::SearchLocalPoints (src/Tracking.cc:3146)
    {   // Tracking::SearchLocalPoints -> SearchByProjection(Frame, local map points, th)
        auto W = build(path);
        std::vector<MapPoint*> local;
        for (MapPoint* p : W->pts) if (p) local.push_back(p);
        const int r = matcher.SearchByProjection(W->F2, local, 3.0f, true, 40.0f);
        emit("local_map", r, ids(W->F2.mvpMapPoints));
    }
    MAM3SLAM::ORBmatcher m09(0.9f, true);                // Tracking::TrackWithMotionModel (src/Tracking.cc:2789)
    {
        auto W = build(path);
        const int r = m09.SearchByProjection(W->F2, W->F1, 15.0f, true);
        emit("frame", r, ids(W->F2.mvpMapPoints));
    }
    {   // Tracking::Relocalization (src/Tracking.cc:3457)
        auto W = build(path);
        const int r = m09.SearchByProjection(W->F2, &W->KF1, W->already, 10.0f, 100);
        emit("reloc", r, ids(W->F2.mvpMapPoints));
    }
    Sim3 Scw2;                                           // world -> KF2 with scale 2: Tcw = [R | t / s]
    {
        auto W = build(path);
        const float* t2 = arr<float>(W->blob, "t2");
        Scw2.s = 2.f; Scw2.t = Vec3(2 * t2[0], 2 * t2[1], 2 * t2[2]);
        std::vector<MapPoint*> pts;
        for (MapPoint* p : W->pts) if (p) pts.push_back(p);
        std::vector<MapPoint*> vpMatched = W->KF2.mps;
        int r = m09.SearchByProjection(&W->KF2, Scw2, pts, vpMatched, 4, 1.5f);     // src/LoopClosing.cc:1173
        emit("sim3proj", r, ids(vpMatched));
        std::vector<KeyFrame*> ptKFs, matchedKF(W->KF2.N, nullptr);
        for (size_t i = 0; i < pts.size(); i++) ptKFs.push_back(i % 2 ? &W->KF1 : &W->KF2);
        vpMatched = W->KF2.mps;
        r = m09.SearchByProjection(&W->KF2, Scw2, pts, ptKFs, vpMatched, matchedKF, 4, 1.0f);   // src/LoopClosing.cc:1310
        std::vector<int> o = ids(vpMatched);
        for (KeyFrame* k : matchedKF) o.push_back(k == &W->KF1 ? 1 : k == &W->KF2 ? 2 : 0);
        emit("sim3proj_kfs", r, o);
    }
    MAM3SLAM::ORBmatcher m07(0.7f, true);                // Tracking::TrackReferenceKeyFrame (src/Tracking.cc:2648)
    {
        auto W = build(path);
        std::vector<MapPoint*> out;
        int r = m07.SearchByBoW(&W->KF1, W->F2, out);
        emit("bow_kf_f", r, ids(out));
        r = m07.SearchByBoW(&W->KF1, &W->KF2, out);      // src/LoopClosing.cc:904
        emit("bow_kf_kf", r, ids(out));
    }
    {   // Tracking::MonocularInitialization (src/Tracking.cc:2393)
        auto W = build(path);
        std::vector<cv::Point2f> prev(W->F1.N);
        for (int i = 0; i < W->F1.N; i++) prev[i] = W->F1.mvKeysUn[i].pt;
        std::vector<int> m12;
        const int r = m09.SearchForInitialization(W->F1, W->F2, prev, m12, 100);
        for (int i = 0; i < W->F1.N; i++) {
            int32_t bx, by;
            std::memcpy(&bx, &prev[i].x, 4); std::memcpy(&by, &prev[i].y, 4);
            m12.push_back(bx); m12.push_back(by);
        }
        emit("init", r, m12);
    }
    MAM3SLAM::ORBmatcher m06(0.6f, false);               // LocalMapping::CreateNewMapPoints (src/LocalMapping.cc:528,582)
    {
        auto W = build(path);
        std::vector<std::pair<size_t, size_t> > pairs;
        int r = m06.SearchForTriangulation(&W->KF1, &W->KF2, pairs, false, false);
        std::vector<int> o;
        for (auto& p : pairs) { o.push_back((int)p.first); o.push_back((int)p.second); }
        emit("tri", r, o);
        r = m06.SearchForTriangulation(&W->KF1, &W->KF2, pairs, false, true);
        o.clear();
        for (auto& p : pairs) { o.push_back((int)p.first); o.push_back((int)p.second); }
        emit("tri_coarse", r, o);
        r = m06.SearchForTriangulation(&W->KF1, &W->KF2, pairs, true, false);
        emit("tri_stereo", r, std::vector<int>(1, (int)pairs.size()));
    }
    for (int ori = 0; ori < 2; ori++) {   // the same pair seen through a non-Pinhole camera: the predicate is the camera's own (:1072)
        auto W = build(path);
        W->KF1.mpCamera->mnType = 1;                     // CAM_FISHEYE (KF1 and KF2 share the camera object, like one agent's keyframes)
        MAM3SLAM::ORBmatcher mf(0.6f, ori != 0);
        std::vector<std::pair<size_t, size_t> > pairs;
        const int r = mf.SearchForTriangulation(&W->KF1, &W->KF2, pairs, false, false);
        std::vector<int> o;
        o.push_back((int)W->KF1.mpCamera->calls);
        o.push_back((int)(W->KF1.mpCamera->callHash & 0x7fffffffu));
        for (auto& p : pairs) { o.push_back((int)p.first); o.push_back((int)p.second); }
        emit(ori ? "tri_fisheye_ori" : "tri_fisheye", r, o);
    }
    {
        auto W = build(path);
        const float* t1 = arr<float>(W->blob, "t1");
        const float* t2 = arr<float>(W->blob, "t2");
        Sim3 S12;
        S12.t = Vec3(t1[0] - t2[0], t1[1] - t2[1], t1[2] - t2[2]);
        size_t n = 0;
        const int32_t* mi = arr<int32_t>(W->blob, "m12_init", &n);
        std::vector<MapPoint*> m12;
        for (size_t i = 0; i < n; i++) m12.push_back(W->at(mi[i]));
        const int r = m09.SearchBySim3(&W->KF1, &W->KF2, m12, S12, 7.5f);
        emit("sim3mutual", r, ids(m12));
    }
    {   // LocalMapping::SearchInNeighbors (src/LocalMapping.cc:888)
        auto W = build(path);
        const int r = m09.Fuse(&W->KF2, W->pts, 3.0f);
        emit("fuse", r, ids(W->KF2.mps));
        emitEvents("fuse_events");
    }
    {   // LoopClosing::SearchAndFuse (src/LoopClosing.cc:3201)
        auto W = build(path);
        std::vector<MapPoint*> pts;
        for (MapPoint* p : W->pts) if (p) pts.push_back(p);
        std::vector<MapPoint*> repl(pts.size(), nullptr);
        const int r = m09.Fuse(&W->KF2, Scw2, pts, 4.0f, repl);
        emit("fuse_sim3", r, ids(repl));
        emitEvents("fuse_sim3_events");
    }
    return 0;
}
