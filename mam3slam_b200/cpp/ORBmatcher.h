// Drop-in replacement of MAM3SLAM's ORBmatcher (reference include/ORBmatcher.h:36-103) on top of the C ABI
// (include/orbm.h).  Each method does on the host exactly what the reference does before and after its Hamming loop
// (projection, frustum / distance / viewing-angle tests, PredictScale — all MapPoint / pose dependent) and hands the
// loop itself (candidate gather through the frame grid, DescriptorDistance, best / second best, thresholds, ratio,
// occupancy, rotation histogram) to the CUDA library.  The methods are templates over the reference's own Frame /
// KeyFrame / MapPoint / Sophus types so that this header has no dependency of its own; in a MAM3SLAM tree include it
// instead of include/ORBmatcher.h after Frame.h / KeyFrame.h / MapPoint.h.  All twelve search / fuse members of the
// reference class are present (monocular branches; the stereo / fisheye `Nleft != -1` branches are out of scope).
// tests/native/slam_shim.h + matcher_adapter_demo.cc exercise every member against the oracle on a GPU box.
#ifndef MAM3SLAM_B200_ORBMATCHER_H
#define MAM3SLAM_B200_ORBMATCHER_H

#include <cmath>
#include <cstring>
#include <set>
#include <stdexcept>
#include <string>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>
#include "../../include/orbm.h"

namespace MAM3SLAM {

class ORBmatcher {
public:
    static const int TH_LOW = ORBM_TH_LOW;
    static const int TH_HIGH = ORBM_TH_HIGH;
    static const int HISTO_LENGTH = ORBM_HISTO_LENGTH;

    // The reference constructs an ORBmatcher as a stack local on every tracking / mapping / loop-closing call
    // (src/Tracking.cc:2393,2648,2789,3146,3359,3398, src/LocalMapping.cc:528,882, src/LoopClosing.cc:904-905,1541,3185,3225).
    // Construction must therefore be free: the CUDA-side state (orbm handle = stream + scratch, two device frame slots) comes
    // from a thread_local pool - the constructor pops a set for `device` (creating one only when the pool has none), the
    // destructor pushes it back.  In steady state neither makes a CUDA API call; the pool's sets are destroyed when the thread
    // exits.  A set is only ever used by the thread that owns it, like the reference's per-call objects.
    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0) : mfNNratio(nnratio), mbCheckOrientation(checkOri), device_(device) {
        std::vector<PoolSet>& free = pool().free;
        for (size_t i = free.size(); i-- > 0;)
            if (free[i].device == device) {
                h_ = free[i].h; a_ = free[i].a; b_ = free[i].b;
                free.erase(free.begin() + (long)i);
                return;
            }
        const int rc = orbm_create(device, &h_);
        if (rc != ORBM_OK) throw std::runtime_error(std::string("orbm_create: ") + orbm_strerror(rc));   // no CPU fallback
    }
    ~ORBmatcher() {
        PoolSet s; s.h = h_; s.a = a_; s.b = b_; s.device = device_;
        pool().free.push_back(s);
    }
    // sets waiting in the calling thread's pool (tests)
    static size_t PooledSets() { return pool().free.size(); }
    ORBmatcher(const ORBmatcher&) = delete;
    ORBmatcher& operator=(const ORBmatcher&) = delete;

    // src/ORBmatcher.cc:2058-2074
    template <class MatT> static int DescriptorDistance(const MatT& a, const MatT& b) {
        return orbm_distance(a.template ptr<uint8_t>(), b.template ptr<uint8_t>());
    }

    // SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono)
    // (src/ORBmatcher.cc:1676-1887), monocular.
    template <class FrameT> int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool /*bMono*/) {
        const auto Tcw = CurrentFrame.GetPose();
        const int N = LastFrame.N;
        std::vector<uint8_t> valid(N, 0), qdesc((size_t)N * 32), qocc(N, 0);
        std::vector<float> u(N, 0.f), v(N, 0.f), ang(N, 0.f);
        std::vector<int> oct(N, 0);
        for (int i = 0; i < N; i++) {                                         // :1696-1729
            auto* pMP = LastFrame.mvpMapPoints[i];
            if (!pMP || LastFrame.mvbOutlier[i]) continue;
            const auto x3Dc = Tcw * pMP->GetWorldPos();
            const float invzc = 1.0 / x3Dc(2);
            if (invzc < 0) continue;
            const auto uv = CurrentFrame.mpCamera->project(x3Dc);
            if (uv(0) < CurrentFrame.mnMinX || uv(0) > CurrentFrame.mnMaxX) continue;
            if (uv(1) < CurrentFrame.mnMinY || uv(1) > CurrentFrame.mnMaxY) continue;
            valid[i] = 1; u[i] = uv(0); v[i] = uv(1);
            qocc[i] = pMP->Observations() > 0;                                  // whether this point, once assigned, blocks its keypoint (:1745-1747)
            oct[i] = LastFrame.mvKeys[i].octave;
            ang[i] = LastFrame.mvKeysUn[i].angle;
            const auto d = pMP->GetDescriptor();
            std::memcpy(&qdesc[(size_t)i * 32], d.template ptr<uint8_t>(), 32);
        }
        bindFrame(CurrentFrame, a_);
        std::vector<uint8_t> occ(CurrentFrame.N, 0);
        for (int i = 0; i < CurrentFrame.N; i++)                               // :1745-1747
            occ[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
        std::vector<int> kpMatch(CurrentFrame.N, -1);
        int nmatches = 0;
        check(orbm_search_frame(h_, a_, occ.data(), CurrentFrame.mvScaleFactors.data(), (int)CurrentFrame.mvScaleFactors.size(), N,
                                valid.data(), u.data(), v.data(), oct.data(), ang.data(), qdesc.data(), th, TH_HIGH,
                                mbCheckOrientation, kpMatch.data(), &nmatches, qocc.data()), "orbm_search_frame");
        for (int i2 = 0; i2 < CurrentFrame.N; i2++)
            if (kpMatch[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = LastFrame.mvpMapPoints[kpMatch[i2]];
        return nmatches;
    }

    // SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, const float th, bool bFarPoints, float thFarPoints)
    // (src/ORBmatcher.cc:43-213), monocular.
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3, const bool bFarPoints = false,
                           const float thFarPoints = 50.0f) {
        const int nq = (int)vpMapPoints.size();
        std::vector<uint8_t> valid(nq, 0), qdesc((size_t)nq * 32), qocc(nq, 0);
        std::vector<float> px(nq, 0.f), py(nq, 0.f), vc(nq, 0.f);
        std::vector<int> lvl(nq, 0);
        for (int i = 0; i < nq; i++) {                                        // :51-63
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP->mbTrackInView) continue;
            if (bFarPoints && pMP->mTrackDepth > thFarPoints) continue;
            if (pMP->isBad()) continue;
            valid[i] = 1; px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; lvl[i] = pMP->mnTrackScaleLevel; vc[i] = pMP->mTrackViewCos;
            qocc[i] = pMP->Observations() > 0;                                 // once assigned, does it block its keypoint (:92-94)
            const auto d = pMP->GetDescriptor();
            std::memcpy(&qdesc[(size_t)i * 32], d.template ptr<uint8_t>(), 32);
        }
        bindFrame(F, a_);
        std::vector<uint8_t> occ(F.N, 0);
        for (int i = 0; i < F.N; i++) occ[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;   // :92-94
        std::vector<int> kpMatch(F.N, -1);
        int nmatches = 0;
        check(orbm_search_local_map(h_, a_, occ.data(), F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(), nq, valid.data(),
                                    px.data(), py.data(), lvl.data(), vc.data(), qdesc.data(), th, mfNNratio, kpMatch.data(),
                                    &nmatches, qocc.data()), "orbm_search_local_map");
        for (int i = 0; i < F.N; i++) if (kpMatch[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[kpMatch[i]];
        return nmatches;
    }

    // SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, const float th, const int ORBdist)
    // (src/ORBmatcher.cc:1889-2010) — relocalisation refinement.
    template <class FrameT, class KeyFrameT, class MapPointT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const std::set<MapPointT*>& sAlreadyFound, const float th, const int ORBdist) {
        const auto Tcw = CurrentFrame.GetPose();
        const auto Ow = Tcw.inverse().translation();
        const std::vector<MapPointT*> vpMPs = pKF->GetMapPointMatches();
        const int nq = (int)vpMPs.size();
        Queries Q(nq);
        std::vector<float> ang(nq, 0.f);
        for (int i = 0; i < nq; i++) {                                         // :1906-1934
            MapPointT* pMP = vpMPs[i];
            if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
            const auto x3Dw = pMP->GetWorldPos();
            const auto x3Dc = Tcw * x3Dw;
            const auto uv = CurrentFrame.mpCamera->project(x3Dc);
            if (uv(0) < CurrentFrame.mnMinX || uv(0) > CurrentFrame.mnMaxX) continue;
            if (uv(1) < CurrentFrame.mnMinY || uv(1) > CurrentFrame.mnMaxY) continue;
            const auto PO = x3Dw - Ow;
            float dist3D = PO.norm();
            const float maxDistance = pMP->GetMaxDistanceInvariance();
            const float minDistance = pMP->GetMinDistanceInvariance();
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            Q.set(i, uv(0), uv(1), pMP->PredictScale(dist3D, &CurrentFrame), pMP->GetDescriptor());
            ang[i] = pKF->mvKeysUn[i].angle;
        }
        bindFrame(CurrentFrame, a_);
        std::vector<uint8_t> occ(CurrentFrame.N, 0);
        for (int i = 0; i < CurrentFrame.N; i++) occ[i] = CurrentFrame.mvpMapPoints[i] != nullptr;   // :1955 (any map point blocks)
        std::vector<int> kpMatch(CurrentFrame.N, -1);
        int nmatches = 0;
        check(orbm_search_frame(h_, a_, occ.data(), CurrentFrame.mvScaleFactors.data(), (int)CurrentFrame.mvScaleFactors.size(), nq,
                                Q.valid.data(), Q.u.data(), Q.v.data(), Q.level.data(), ang.data(), Q.desc.data(), th, ORBdist,
                                mbCheckOrientation, kpMatch.data(), &nmatches, nullptr), "orbm_search_frame");   // :1955: any point blocks
        for (int i2 = 0; i2 < CurrentFrame.N; i2++)
            if (kpMatch[i2] >= 0) CurrentFrame.mvpMapPoints[i2] = vpMPs[kpMatch[i2]];
        return nmatches;
    }

    // SearchByProjection(KeyFrame* pKF, Sophus::Sim3f &Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*> &vpMatched,
    // int th, float ratioHamming) (src/ORBmatcher.cc:427-532) — loop detection.
    template <class KeyFrameT, class Sim3T, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched,
                           int th, float ratioHamming = 1.0) {
        std::vector<int> kpMatch;
        const int nmatches = sim3Projection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming, kpMatch);
        for (int i = 0; i < pKF->N; i++) if (kpMatch[i] >= 0) vpMatched[i] = vpPoints[kpMatch[i]];
        return nmatches;
    }
    // ... with vpPointsKFs / vpMatchedKF (src/ORBmatcher.cc:534-646) — place recognition and merging.
    template <class KeyFrameT, class Sim3T, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints, const std::vector<KeyFrameT*>& vpPointsKFs,
                           std::vector<MapPointT*>& vpMatched, std::vector<KeyFrameT*>& vpMatchedKF, int th, float ratioHamming = 1.0) {
        std::vector<int> kpMatch;
        const int nmatches = sim3Projection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming, kpMatch);
        for (int i = 0; i < pKF->N; i++)
            if (kpMatch[i] >= 0) { vpMatched[i] = vpPoints[kpMatch[i]]; vpMatchedKF[i] = vpPointsKFs[kpMatch[i]]; }
        return nmatches;
    }

    // SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (src/ORBmatcher.cc:223-425), monocular.
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
        const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(nullptr));
        std::vector<uint8_t> elig(vpMapPointsKF.size(), 0);
        for (size_t i = 0; i < elig.size(); i++) elig[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();   // :254-260
        FeatVecCSR fvKF(pKF->mFeatVec), fvF(F.mFeatVec);
        bindFrame(*pKF, a_);
        bindFrame(F, b_);
        std::vector<int> matchF(F.N, -1);
        int nmatches = 0;
        check(orbm_search_bow_kf_frame(h_, a_, &fvKF.view, elig.data(), b_, &fvF.view, mfNNratio, mbCheckOrientation, matchF.data(),
                                       &nmatches), "orbm_search_bow_kf_frame");
        for (int j = 0; j < F.N; j++) if (matchF[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[matchF[j]];
        return nmatches;
    }

    // SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (src/ORBmatcher.cc:765-905).
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches();
        const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(nullptr));
        std::vector<uint8_t> e1(vpMapPoints1.size(), 0), e2(vpMapPoints2.size(), 0);
        for (size_t i = 0; i < e1.size(); i++) e1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();         // :805-809
        for (size_t i = 0; i < e2.size(); i++) e2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();         // :825-829
        FeatVecCSR fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
        bindFrame(*pKF1, a_);
        bindFrame(*pKF2, b_);
        std::vector<int> m12(vpMapPoints1.size(), -1);
        int nmatches = 0;
        check(orbm_search_bow_kf_kf(h_, a_, &fv1.view, e1.data(), b_, &fv2.view, e2.data(), mfNNratio, mbCheckOrientation, m12.data(),
                                    &nmatches), "orbm_search_bow_kf_kf");
        for (size_t i = 0; i < m12.size(); i++) if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];
        return nmatches;
    }

    // SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f> &vbPrevMatched, vector<int> &vnMatches12, int windowSize)
    // (src/ORBmatcher.cc:648-763).  Point2fT = cv::Point2f (two packed floats).
    template <class FrameT, class Point2fT>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<Point2fT>& vbPrevMatched, std::vector<int>& vnMatches12,
                                int windowSize = 10) {
        static_assert(sizeof(Point2fT) == 2 * sizeof(float), "cv::Point2f layout");
        vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);
        bindFrame(F1, a_);
        bindFrame(F2, b_);
        int nmatches = 0;
        check(orbm_search_initialization(h_, a_, b_, reinterpret_cast<float*>(vbPrevMatched.data()), windowSize, mfNNratio,
                                         mbCheckOrientation, vnMatches12.data(), &nmatches), "orbm_search_initialization");
        return nmatches;
    }

    // SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, vector<pair<size_t,size_t>> &vMatchedPairs, bOnlyStereo, bCoarse)
    // (src/ORBmatcher.cc:907-1146), monocular.  The reference tests each candidate with the virtual
    // pCamera1->epipolarConstrain (:1072).  Pinhole pairs (GetType() == CAM_PINHOLE) take the device path, which evaluates
    // Pinhole's fundamental-matrix test inside the kernel; every other camera (KannalaBrandt8: the only settings the reference
    // ships) takes searchForTriangulationHostPredicate below, where the device produces the candidates and the caller's own
    // camera object judges them in the reference's order - the same matches as the reference for ANY GeometricCamera.
    template <class KeyFrameT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo, const bool bCoarse = false) {
        vMatchedPairs.clear();
        if (bOnlyStereo) return 0;                       // mono keypoints are never stereo (:977-981): every idx1 is skipped
        const auto T1w = pKF1->GetPose();
        const auto T2w = pKF2->GetPose();
        const auto Tw2 = pKF2->GetPoseInverse();
        const auto Cw = pKF1->GetCameraCenter();
        const auto C2 = T2w * Cw;
        const auto ep = pKF2->mpCamera->project(C2);                          // :915-920
        const auto T12 = T1w * Tw2;
        const auto R12 = T12.rotationMatrix();
        const auto t12 = T12.translation();
        // the fundamental matrix exactly as Pinhole::epipolarConstrain forms it (src/CameraModels/Pinhole.cpp:107-112)
        auto K1 = pKF1->mpCamera->toK_();
        auto K2 = pKF2->mpCamera->toK_();
        decltype(K1) t12x = K1;                                                // Sophus::SO3f::hat(t12)
        t12x(0, 0) = 0.f;      t12x(0, 1) = -t12(2);  t12x(0, 2) = t12(1);
        t12x(1, 0) = t12(2);   t12x(1, 1) = 0.f;      t12x(1, 2) = -t12(0);
        t12x(2, 0) = -t12(1);  t12x(2, 1) = t12(0);   t12x(2, 2) = 0.f;
        const decltype(K1) F12 = K1.transpose().inverse() * t12x * R12 * K2.inverse();
        float F[9], epf[2] = {ep(0), ep(1)};
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F[3 * r + c] = F12(r, c);
        std::vector<uint8_t> free1(pKF1->N, 0), free2(pKF2->N, 0);
        for (int i = 0; i < pKF1->N; i++) free1[i] = pKF1->GetMapPoint(i) == nullptr;   // :970-976
        for (int i = 0; i < pKF2->N; i++) free2[i] = pKF2->GetMapPoint(i) == nullptr;   // :1001-1005
        FeatVecCSR fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
        bindFrame(*pKF1, a_);
        bindFrame(*pKF2, b_);
        std::vector<int> vMatches12(pKF1->N, -1);
        int nmatches = 0;
        if (!bCoarse && (pKF1->mpCamera->GetType() != 0u || pKF2->mpCamera->GetType() != 0u)) {      // not CAM_PINHOLE (GeometricCamera.h:95)
            nmatches = searchForTriangulationHostPredicate(pKF1, pKF2, fv1, fv2, free1, free2, epf, R12, t12, vMatches12);
            vMatchedPairs.reserve(nmatches);
            for (size_t i = 0, iend = vMatches12.size(); i < iend; i++)
                if (vMatches12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)vMatches12[i]));
            return nmatches;
        }
        check(orbm_search_triangulation(h_, a_, &fv1.view, free1.data(), b_, &fv2.view, free2.data(), epf, pKF2->mvScaleFactors.data(),
                                        F, pKF2->mvLevelSigma2.data(), (int)pKF2->mvScaleFactors.size(), bCoarse, mbCheckOrientation,
                                        vMatches12.data(), &nmatches), "orbm_search_triangulation");
        vMatchedPairs.reserve(nmatches);
        for (size_t i = 0, iend = vMatches12.size(); i < iend; i++)           // :1135-1143
            if (vMatches12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)vMatches12[i]));
        return nmatches;
    }

    // The inner loop of SearchForTriangulation (:995-1078) over the device-made candidate lists, with the caller's cameras.
    // The lists hold, in scan order, exactly the candidates that reach `bCoarse || pCamera1->epipolarConstrain(...)` when the
    // running bestDist is ignored; applying `dist > bestDist -> continue` here restores the reference's sequence of predicate
    // calls and its "last minimum among those that pass" result.  Then the rotation histogram (:1092-1133).
    template <class KeyFrameT, class FeatVecT, class Mat3T, class Vec3T>
    int searchForTriangulationHostPredicate(KeyFrameT* pKF1, KeyFrameT* pKF2, FeatVecT& fv1, FeatVecT& fv2, const std::vector<uint8_t>& free1,
                                            const std::vector<uint8_t>& free2, const float* epf, const Mat3T& R12, const Vec3T& t12,
                                            std::vector<int>& vMatches12) {
        const int n1 = pKF1->N;
        int cap = 16;
        std::vector<int> ci, cd, cc(n1, 0);
        for (;;) {
            ci.assign((size_t)n1 * cap, -1); cd.assign((size_t)n1 * cap, 0);
            const int rc = orbm_triangulation_candidates(h_, a_, &fv1.view, free1.data(), b_, &fv2.view, free2.data(), epf,
                                                         pKF2->mvScaleFactors.data(), (int)pKF2->mvScaleFactors.size(), cap, ci.data(),
                                                         cd.data(), cc.data());
            if (rc == ORBM_ERR_CAPACITY) { for (int c : cc) if (c > cap) cap = c; continue; }
            check(rc, "orbm_triangulation_candidates");
            break;
        }
        auto* pCamera1 = pKF1->mpCamera;
        auto* pCamera2 = pKF2->mpCamera;
        const float factor = 1.0f / HISTO_LENGTH;                              // :943
        std::vector<int> binOf(n1, -1), binSize(HISTO_LENGTH, 0);
        int nmatches = 0;
        for (size_t p1 = 0; p1 < fv1.idx.size(); p1++) {                       // nodes in order, each node's list in order (:955-966)
            const int idx1 = fv1.idx[p1];                                       // (a keypoint belongs to exactly one node)
            if (!cc[idx1]) continue;
            const auto& kp1 = pKF1->mvKeysUn[idx1];
            int bestDist = TH_LOW, bestIdx2 = -1;                               // :995-996
            for (int c = 0; c < cc[idx1]; c++) {
                const int idx2 = ci[(size_t)idx1 * cap + c], dist = cd[(size_t)idx1 * cap + c];
                if (dist > bestDist) continue;                                  // :1017
                const auto& kp2 = pKF2->mvKeysUn[idx2];
                if (pCamera1->epipolarConstrain(pCamera2, kp1, kp2, R12, t12, pKF1->mvLevelSigma2[kp1.octave],
                                                pKF2->mvLevelSigma2[kp2.octave])) {                       // :1072
                    bestIdx2 = idx2; bestDist = dist;
                }
            }
            if (bestIdx2 < 0) continue;
            vMatches12[idx1] = bestIdx2;
            nmatches++;
            if (mbCheckOrientation) {                                           // :1092-1103
                float rot = kp1.angle - pKF2->mvKeysUn[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                binOf[idx1] = bin; binSize[bin]++;
            }
        }
        if (mbCheckOrientation) {                                               // :1112-1133, ComputeThreeMaxima :2012-2053
            int top[3] = {-1, -1, -1}, size[3] = {0, 0, 0};
            for (int i = 0; i < HISTO_LENGTH; i++) {
                int at = binSize[i] > size[0] ? 0 : binSize[i] > size[1] ? 1 : binSize[i] > size[2] ? 2 : 3;
                for (int k = 2; k > at; k--) { top[k] = top[k - 1]; size[k] = size[k - 1]; }
                if (at < 3) { top[at] = i; size[at] = binSize[i]; }
            }
            if (size[1] < 0.1f * (float)size[0]) top[1] = top[2] = -1;
            else if (size[2] < 0.1f * (float)size[0]) top[2] = -1;
            for (int idx1 = 0; idx1 < n1; idx1++) {
                const int b = binOf[idx1];
                if (b < 0 || b == top[0] || b == top[1] || b == top[2]) continue;
                vMatches12[idx1] = -1;
                nmatches--;
            }
        }
        return nmatches;
    }

    // SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*> &vpMatches12, const Sophus::Sim3f &S12, const float th)
    // (src/ORBmatcher.cc:1457-1674).
    template <class KeyFrameT, class MapPointT, class Sim3T>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const Sim3T& S12, const float th) {
        const float fx = pKF1->fx, fy = pKF1->fy, cx = pKF1->cx, cy = pKF1->cy;
        const auto T1w = pKF1->GetPose();
        const auto T2w = pKF2->GetPose();
        const auto S21 = S12.inverse();
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches();
        const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {                                         // :1480-1490
            MapPointT* pMP = vpMatches12[i];
            if (pMP) {
                vbAlreadyMatched1[i] = true;
                const int idx2 = std::get<0>(pMP->GetIndexInKeyFrame(pKF2));
                if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
            }
        }
        Queries Q1(N1), Q2(N2);
        for (int i1 = 0; i1 < N1; i1++) {                                      // :1496-1537
            MapPointT* pMP = vpMapPoints1[i1];
            if (!pMP || vbAlreadyMatched1[i1] || pMP->isBad()) continue;
            const auto p3Dc1 = T1w * pMP->GetWorldPos();
            const auto p3Dc2 = S21 * p3Dc1;
            if (p3Dc2(2) < 0.0) continue;
            const float invz = 1.0 / p3Dc2(2);
            const float x = p3Dc2(0) * invz, y = p3Dc2(1) * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF2->IsInImage(u, v)) continue;
            const float dist3D = p3Dc2.norm();
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            Q1.set(i1, u, v, pMP->PredictScale(dist3D, pKF2), pMP->GetDescriptor());
        }
        for (int i2 = 0; i2 < N2; i2++) {                                      // :1576-1617
            MapPointT* pMP = vpMapPoints2[i2];
            if (!pMP || vbAlreadyMatched2[i2] || pMP->isBad()) continue;
            const auto p3Dc2 = T2w * pMP->GetWorldPos();
            const auto p3Dc1 = S12 * p3Dc2;
            if (p3Dc1(2) < 0.0) continue;
            const float invz = 1.0 / p3Dc1(2);
            const float x = p3Dc1(0) * invz, y = p3Dc1(1) * invz;
            const float u = fx * x + cx, v = fy * y + cy;
            if (!pKF1->IsInImage(u, v)) continue;
            const float dist3D = p3Dc1.norm();
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            Q2.set(i2, u, v, pMP->PredictScale(dist3D, pKF1), pMP->GetDescriptor());
        }
        bindFrame(*pKF1, a_);
        bindFrame(*pKF2, b_);
        std::vector<int> m12(N1, -1);
        int nFound = 0;
        check(orbm_search_by_sim3(h_, a_, b_, pKF1->mvScaleFactors.data(), pKF2->mvScaleFactors.data(), (int)pKF2->mvScaleFactors.size(),
                                  Q1.valid.data(), Q1.u.data(), Q1.v.data(), Q1.level.data(), Q1.desc.data(), Q2.valid.data(),
                                  Q2.u.data(), Q2.v.data(), Q2.level.data(), Q2.desc.data(), th, m12.data(), &nFound),
              "orbm_search_by_sim3");
        for (int i1 = 0; i1 < N1; i1++) if (m12[i1] >= 0) vpMatches12[i1] = vpMapPoints2[m12[i1]];   // :1658-1671
        return nFound;
    }

    // Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th, const bool bRight) (src/ORBmatcher.cc:1148-1338),
    // monocular.  The window search has no state, so the device returns (best keypoint, distance) for every point that
    // passes the geometric tests; the map mutations run afterwards on the host in the reference's order, with the
    // mutable pre-checks (isBad / IsInKeyFrame, which Replace / AddObservation of earlier points can change) re-evaluated
    // at that moment exactly as the reference's loop does.
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0, const bool bRight = false) {
        if (bRight) throw std::runtime_error("ORBmatcher::Fuse: bRight (stereo fisheye) is outside the monocular drop-in");
        const auto Tcw = pKF->GetPose();
        const auto Ow = pKF->GetCameraCenter();
        const int nMPs = (int)vpMapPoints.size();
        Queries Q(nMPs);
        for (int i = 0; i < nMPs; i++) {                                       // :1192-1243 (state-independent part)
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP) continue;
            const auto p3Dw = pMP->GetWorldPos();
            const auto p3Dc = Tcw * p3Dw;
            if (p3Dc(2) < 0.0f) continue;
            const auto uv = pKF->mpCamera->project(p3Dc);
            if (!pKF->IsInImage(uv(0), uv(1))) continue;
            const auto PO = p3Dw - Ow;
            const float dist3D = PO.norm();
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            const auto Pn = pMP->GetNormal();
            if (PO.dot(Pn) < 0.5 * dist3D) continue;
            Q.set(i, uv(0), uv(1), pMP->PredictScale(dist3D, pKF), pMP->GetDescriptor());
        }
        bindFrame(*pKF, a_);
        std::vector<int> bestIdx(nMPs, -1), bestDist(nMPs, 256);
        check(orbm_fuse_candidates(h_, a_, pKF->mvScaleFactors.data(), pKF->mvInvLevelSigma2.data(), (int)pKF->mvScaleFactors.size(), nMPs,
                                   Q.valid.data(), Q.u.data(), Q.v.data(), Q.level.data(), Q.desc.data(), th, 1, bestIdx.data(),
                                   bestDist.data()), "orbm_fuse_candidates");
        int nFused = 0;
        for (int i = 0; i < nMPs; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;     // :1180-1196, at loop time
            if (!Q.valid[i] || bestIdx[i] < 0 || bestDist[i] > TH_LOW) continue;
            MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[i]);                  // :1312-1331
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, bestIdx[i]);
                pKF->AddMapPoint(pMP, bestIdx[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // Fuse(KeyFrame *pKF, Sophus::Sim3f &Scw, const vector<MapPoint*> &vpPoints, float th, vector<MapPoint*> &vpReplacePoint)
    // (src/ORBmatcher.cc:1340-1455).
    template <class KeyFrameT, class Sim3T, class MapPointT>
    int Fuse(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint) {
        typedef typename std::decay<decltype(pKF->GetPose())>::type SE3T;
        const SE3T Tcw(Scw.rotationMatrix(), Scw.translation() / Scw.scale());
        const auto Ow = Tcw.inverse().translation();
        const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
        const int nPoints = (int)vpPoints.size();
        Queries Q(nPoints);
        for (int iMP = 0; iMP < nPoints; iMP++) {                              // :1360-1403
            MapPointT* pMP = vpPoints[iMP];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;          // nothing in this loop changes either
            float u, v, dist3D;
            if (!sim3Visible(pKF, Tcw, Ow, pMP, u, v, dist3D)) continue;
            Q.set(iMP, u, v, pMP->PredictScale(dist3D, pKF), pMP->GetDescriptor());
        }
        bindFrame(*pKF, a_);
        std::vector<int> bestIdx(nPoints, -1), bestDist(nPoints, 256);
        check(orbm_fuse_candidates(h_, a_, pKF->mvScaleFactors.data(), nullptr, (int)pKF->mvScaleFactors.size(), nPoints, Q.valid.data(),
                                   Q.u.data(), Q.v.data(), Q.level.data(), Q.desc.data(), th, 0, bestIdx.data(), bestDist.data()),
              "orbm_fuse_candidates");
        int nFused = 0;
        for (int iMP = 0; iMP < nPoints; iMP++) {                              // :1433-1449, in order
            if (!Q.valid[iMP] || bestIdx[iMP] < 0 || bestDist[iMP] > TH_LOW) continue;
            MapPointT* pMP = vpPoints[iMP];
            MapPointT* pMPinKF = pKF->GetMapPoint(bestIdx[iMP]);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, bestIdx[iMP]);
                pKF->AddMapPoint(pMP, bestIdx[iMP]);
            }
            nFused++;
        }
        return nFused;
    }

    orbm_handle handle() const { return h_; }

protected:
    // src/ORBmatcher.cc:215-221 and :2012-2053 run inside the CUDA searches; kept for subclasses that call them
    float RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998f ? 2.5f : 4.0f; }

    float mfNNratio;
    bool mbCheckOrientation;

private:
    // flattened per-query inputs of the projection searches
    struct Queries {
        std::vector<uint8_t> valid, desc;
        std::vector<float> u, v;
        std::vector<int> level;
        explicit Queries(int n) : valid(n, 0), desc((size_t)(n > 0 ? n : 1) * 32), u(n, 0.f), v(n, 0.f), level(n, 0) {}
        template <class MatT> void set(int i, float uu, float vv, int lvl, const MatT& d) {
            valid[i] = 1; u[i] = uu; v[i] = vv; level[i] = lvl;
            std::memcpy(&desc[(size_t)i * 32], d.template ptr<uint8_t>(), 32);
        }
    };
    // DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned int>>) as the CSR orbm_featvec wants
    struct FeatVecCSR {
        std::vector<int> node, off, idx;
        orbm_featvec view;
        template <class FeatVecT> explicit FeatVecCSR(const FeatVecT& fv) {
            off.push_back(0);
            for (const auto& kv : fv) {
                node.push_back((int)kv.first);
                for (const auto i : kv.second) idx.push_back((int)i);
                off.push_back((int)idx.size());
            }
            view.n_nodes = (int)node.size(); view.node = node.data(); view.off = off.data(); view.idx = idx.data();
        }
    };
    // the visibility tests shared by the Sim3 projection searches and the Sim3 Fuse (:458-497, :1368-1397)
    template <class KeyFrameT, class SE3T, class Vec3T, class MapPointT>
    static bool sim3Visible(KeyFrameT* pKF, const SE3T& Tcw, const Vec3T& Ow, MapPointT* pMP, float& u, float& v, float& dist) {
        const auto p3Dw = pMP->GetWorldPos();
        const auto p3Dc = Tcw * p3Dw;
        if (p3Dc(2) < 0.0) return false;
        const auto uv = pKF->mpCamera->project(p3Dc);
        if (!pKF->IsInImage(uv(0), uv(1))) return false;
        const float maxDistance = pMP->GetMaxDistanceInvariance();
        const float minDistance = pMP->GetMinDistanceInvariance();
        const auto PO = p3Dw - Ow;
        dist = PO.norm();
        if (dist < minDistance || dist > maxDistance) return false;
        const auto Pn = pMP->GetNormal();
        if (PO.dot(Pn) < 0.5 * dist) return false;
        u = uv(0); v = uv(1);
        return true;
    }
    template <class KeyFrameT, class Sim3T, class MapPointT>
    int sim3Projection(KeyFrameT* pKF, Sim3T& Scw, const std::vector<MapPointT*>& vpPoints, const std::vector<MapPointT*>& vpMatched,
                       int th, float ratioHamming, std::vector<int>& kpMatch) {
        typedef typename std::decay<decltype(pKF->GetPose())>::type SE3T;
        const SE3T Tcw(Scw.rotationMatrix(), Scw.translation() / Scw.scale());
        const auto Ow = Tcw.inverse().translation();
        std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPointT*>(nullptr));
        const int nq = (int)vpPoints.size();
        Queries Q(nq);
        for (int iMP = 0; iMP < nq; iMP++) {
            MapPointT* pMP = vpPoints[iMP];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            float u, v, dist;
            if (!sim3Visible(pKF, Tcw, Ow, pMP, u, v, dist)) continue;
            Q.set(iMP, u, v, pMP->PredictScale(dist, pKF), pMP->GetDescriptor());
        }
        bindFrame(*pKF, a_);
        std::vector<uint8_t> matched(pKF->N, 0);
        for (int i = 0; i < pKF->N; i++) matched[i] = vpMatched[i] != nullptr;                            // :507-508
        kpMatch.assign(pKF->N, -1);
        int nmatches = 0;
        check(orbm_search_sim3(h_, a_, matched.data(), pKF->mvScaleFactors.data(), (int)pKF->mvScaleFactors.size(), nq, Q.valid.data(),
                               Q.u.data(), Q.v.data(), Q.level.data(), Q.desc.data(), th, ratioHamming, kpMatch.data(), &nmatches),
              "orbm_search_sim3");
        return nmatches;
    }
    // uploads mvKeysUn / mDescriptors / bounds of a Frame or KeyFrame and builds its grid on the device
    template <class FrameT> void bindFrame(const FrameT& F, orbm_frame& slot) {
        orbm_frame_view view;
        view.n = F.N;
        view.kps = reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn.data());
        view.desc = F.mDescriptors.template ptr<uint8_t>();
        view.min_x = F.mnMinX; view.min_y = F.mnMinY; view.max_x = F.mnMaxX; view.max_y = F.mnMaxY;
        view.on_device = 0;
        check(slot ? orbm_frame_update(h_, slot, &view) : orbm_frame_create(h_, &view, &slot), "orbm_frame_update");
    }
    void check(int rc, const char* what) {
        if (rc != ORBM_OK) throw std::runtime_error(std::string(what) + ": " + orbm_strerror(rc) + " " + orbm_last_error(h_));
    }
    struct PoolSet { orbm_handle h = nullptr; orbm_frame a = nullptr, b = nullptr; int device = 0; };
    struct Pool {
        std::vector<PoolSet> free;
        ~Pool() {
            for (PoolSet& s : free) {
                if (s.a) orbm_frame_destroy(s.h, s.a);
                if (s.b) orbm_frame_destroy(s.h, s.b);
                orbm_destroy(s.h);
            }
        }
    };
    static Pool& pool() { thread_local Pool p; return p; }
    orbm_handle h_ = nullptr;
    orbm_frame a_ = nullptr, b_ = nullptr;     // two device frame slots, reused across calls (and, through the pool, across objects)
    int device_ = 0;
};

}  // namespace MAM3SLAM
#endif
