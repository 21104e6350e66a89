// Drop-in replacement of MAM3SLAM's ORBextractor (reference include/ORBextractor.h:43-109) on top of the C ABI
// (include/orbx.h, libmam3slam_b200.so).  Same namespace, constructor, operator() signature, getters and public
// mvImagePyramid, so Frame::ExtractORB (src/Frame.cc:418-425) and Tracking (src/Tracking.cc:600-606) compile unchanged.
// Header-only; needs <opencv2/core.hpp> (or, in this repo's tests, the minimal cv:: shim in tests/native/cv_shim.h).
#ifndef MAM3SLAM_B200_ORBEXTRACTOR_H
#define MAM3SLAM_B200_ORBEXTRACTOR_H

#include <cassert>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/orbx.h"

namespace MAM3SLAM {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_, int device = 0)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_), minThFAST(minThFAST_) {
        orbx_config cfg{nfeatures_, scaleFactor_, nlevels_, iniThFAST_, minThFAST_, device, 1};
        const int rc = orbx_create(&cfg, &h_);
        if (rc != ORBX_OK) throw std::runtime_error(std::string("orbx_create: ") + orbx_strerror(rc));   // no CPU fallback
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
        orbx_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                        mnFeaturesPerLevel.data());
        mvImagePyramid.resize(nlevels);
    }
    ~ORBextractor() { orbx_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // src/ORBextractor.cc:1086-1168.  Mask is ignored like in the reference (include/ORBextractor.h:56).
    int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        if (_image.empty()) return -1;
        cv::Mat image = _image.getMat();
        assert(image.type() == CV_8UC1);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
        const int cap = orbx_max_keypoints(h_);
        kp_.resize(cap); desc_.resize((size_t)cap * 32);
        int n = 0, mono = 0;
        const int rc = orbx_extract(h_, image.data, image.cols, image.rows, (int)image.step, vLappingArea[0], vLappingArea[1],
                                    kp_.data(), desc_.data(), cap, &n, &mono);
        if (rc != ORBX_OK) throw std::runtime_error(std::string("orbx_extract: ") + orbx_strerror(rc) + " " + orbx_last_error(h_));
        _keypoints = std::vector<cv::KeyPoint>(n);
        if (n == 0) { _descriptors.release(); pyramidStale_ = true; return mono; }
        std::memcpy((void*)_keypoints.data(), kp_.data(), (size_t)n * sizeof(orbx_keypoint));
        _descriptors.create(n, 32, CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < n; ++i) std::memcpy(descriptors.ptr(i), &desc_[(size_t)i * 32], 32);
        pyramidStale_ = true;
        return mono;
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // The reference fills mvImagePyramid on every call (src/ORBextractor.cc:1170-1195) but only stereo matching reads it
    // (src/Frame.cc:818).  The pyramid stays in HBM; call this before reading mvImagePyramid to download the levels as
    // ROIs of 19-px-padded buffers, exactly the Mats the reference exposes.
    void DownloadImagePyramid() {
        if (!pyramidStale_) return;
        for (int l = 0; l < nlevels; ++l) {
            int w = 0, hgt = 0;
            if (orbx_level_size(h_, l, &w, &hgt) != ORBX_OK) return;
            cv::Mat padded(hgt + 38, w + 38, CV_8UC1);
            orbx_download_level(h_, 0, l, 1, padded.data, (int)padded.step);
            mvImagePyramid[l] = padded(cv::Rect(19, 19, w, hgt));
        }
        pyramidStale_ = false;
    }
    std::vector<cv::Mat> mvImagePyramid;

    orbx_handle handle() const { return h_; }   // device-resident results for the matcher (orbx_results_device)

protected:
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

private:
    orbx_handle h_ = nullptr;
    std::vector<orbx_keypoint> kp_;
    std::vector<uint8_t> desc_;
    bool pyramidStale_ = true;
};

}  // namespace MAM3SLAM
#endif
