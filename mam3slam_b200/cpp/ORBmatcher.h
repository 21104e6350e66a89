// Drop-in replacement of MAM3SLAM's ORBmatcher (reference include/ORBmatcher.h:36-103) on top of the C ABI
// (include/orbm.h).  Each method does on the host exactly what the reference does before and after its Hamming loop
// (projection, frustum / distance / viewing-angle tests, PredictScale — all MapPoint / pose dependent) and hands the
// loop itself (candidate gather through the frame grid, DescriptorDistance, best / second best, thresholds, ratio,
// occupancy, rotation histogram) to the CUDA library.  The methods are templates over the reference's own Frame /
// KeyFrame / MapPoint types so that this header has no dependency of its own; in a MAM3SLAM tree include it instead of
// include/ORBmatcher.h after Frame.h / KeyFrame.h / MapPoint.h.  INTEGRATION.md lists the remaining methods' mapping.
#ifndef MAM3SLAM_B200_ORBMATCHER_H
#define MAM3SLAM_B200_ORBMATCHER_H

#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/orbm.h"

namespace MAM3SLAM {

class ORBmatcher {
public:
    static const int TH_LOW = ORBM_TH_LOW;
    static const int TH_HIGH = ORBM_TH_HIGH;
    static const int HISTO_LENGTH = ORBM_HISTO_LENGTH;

    ORBmatcher(float nnratio = 0.6, bool checkOri = true, int device = 0) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {
        const int rc = orbm_create(device, &h_);
        if (rc != ORBM_OK) throw std::runtime_error(std::string("orbm_create: ") + orbm_strerror(rc));   // no CPU fallback
    }
    ~ORBmatcher() {
        if (cur_) orbm_frame_destroy(h_, cur_);
        orbm_destroy(h_);
    }
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
        std::vector<uint8_t> valid(N, 0), qdesc((size_t)N * 32);
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
            oct[i] = LastFrame.mvKeys[i].octave;
            ang[i] = LastFrame.mvKeysUn[i].angle;
            const auto d = pMP->GetDescriptor();
            std::memcpy(&qdesc[(size_t)i * 32], d.template ptr<uint8_t>(), 32);
        }
        bindFrame(CurrentFrame);
        std::vector<uint8_t> occ(CurrentFrame.N, 0);
        for (int i = 0; i < CurrentFrame.N; i++)                               // :1745-1747
            occ[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
        std::vector<int> kpMatch(CurrentFrame.N, -1);
        int nmatches = 0;
        check(orbm_search_frame(h_, cur_, occ.data(), CurrentFrame.mvScaleFactors.data(), (int)CurrentFrame.mvScaleFactors.size(), N,
                                valid.data(), u.data(), v.data(), oct.data(), ang.data(), qdesc.data(), th, TH_HIGH,
                                mbCheckOrientation, kpMatch.data(), &nmatches), "orbm_search_frame");
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
        std::vector<uint8_t> valid(nq, 0), qdesc((size_t)nq * 32);
        std::vector<float> px(nq, 0.f), py(nq, 0.f), vc(nq, 0.f);
        std::vector<int> lvl(nq, 0);
        for (int i = 0; i < nq; i++) {                                        // :51-63
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP->mbTrackInView) continue;
            if (bFarPoints && pMP->mTrackDepth > thFarPoints) continue;
            if (pMP->isBad()) continue;
            valid[i] = 1; px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; lvl[i] = pMP->mnTrackScaleLevel; vc[i] = pMP->mTrackViewCos;
            const auto d = pMP->GetDescriptor();
            std::memcpy(&qdesc[(size_t)i * 32], d.template ptr<uint8_t>(), 32);
        }
        bindFrame(F);
        std::vector<uint8_t> occ(F.N, 0);
        for (int i = 0; i < F.N; i++) occ[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;   // :92-94
        std::vector<int> kpMatch(F.N, -1);
        int nmatches = 0;
        check(orbm_search_local_map(h_, cur_, occ.data(), F.mvScaleFactors.data(), (int)F.mvScaleFactors.size(), nq, valid.data(),
                                    px.data(), py.data(), lvl.data(), vc.data(), qdesc.data(), th, mfNNratio, kpMatch.data(),
                                    &nmatches), "orbm_search_local_map");
        for (int i = 0; i < F.N; i++) if (kpMatch[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[kpMatch[i]];
        return nmatches;
    }

    orbm_handle handle() const { return h_; }

protected:
    float mfNNratio;
    bool mbCheckOrientation;

private:
    // uploads mvKeysUn / mDescriptors / bounds of a Frame or KeyFrame and builds its grid on the device
    template <class FrameT> void bindFrame(const FrameT& F) {
        orbm_frame_view view;
        view.n = F.N;
        view.kps = reinterpret_cast<const orbx_keypoint*>(F.mvKeysUn.data());
        view.desc = F.mDescriptors.template ptr<uint8_t>();
        view.min_x = F.mnMinX; view.min_y = F.mnMinY; view.max_x = F.mnMaxX; view.max_y = F.mnMaxY;
        view.on_device = 0;
        check(cur_ ? orbm_frame_update(h_, cur_, &view) : orbm_frame_create(h_, &view, &cur_), "orbm_frame_update");
    }
    void check(int rc, const char* what) {
        if (rc != ORBM_OK) throw std::runtime_error(std::string(what) + ": " + orbm_strerror(rc) + " " + orbm_last_error(h_));
    }
    orbm_handle h_ = nullptr;
    orbm_frame cur_ = nullptr;
};

}  // namespace MAM3SLAM
#endif
