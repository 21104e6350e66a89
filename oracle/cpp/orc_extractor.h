// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_primitives.h).  CPU restatement of the reference's
// ORBextractor (src/ORBextractor.cc, include/ORBextractor.h).  It keeps the reference's container choices
// (std::list for the quadtree, std::sort for the "largest node first" phase) so libstdc++'s list order and
// introsort tie behaviour are inherited rather than modelled.
//
// Parity: PINNED to the reference's own source.  oracle/_ref/liborb_ref.so is /root/reference/src/ORBextractor.cc compiled
// unmodified against the header shim oracle/ref_shim/ (recipe: oracle/Makefile `ref`), and tests/test_oracle_vs_ref.py asserts
// that this restatement equals it bit for bit (keypoints, order, responses, angles, descriptors, monoIndex, pyramid levels,
// per-level kept lists).  The pixel primitives both share are pinned to cv2 4.13.0 (tests/test_oracle_vs_cv2.py, tests/golden/).
#pragma once
#include "orc_primitives.h"
#include <list>
#include <utility>

namespace orc {

struct KeyPoint {          // the fields of cv::KeyPoint the reference fills (class_id stays -1)
    float x, y, size, angle, response;
    int octave;
};

class Extractor {
public:
    static const int kPatch = 31, kHalfPatch = 15, kEdge = 19;

    Extractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);

    // ORBextractor::operator() (src/ORBextractor.cc:1086-1168).  Returns monoIndex, or -1 on empty input.
    int extract(const uint8_t* img, int w, int h, int stride, int lap0, int lap1,
                std::vector<KeyPoint>& kps, std::vector<uint8_t>& desc);

    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;                       // the reference stores the float argument in a double member
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> featuresPerLevel, umax;
    std::vector<int> pattern;                 // 512 points x (x,y)

    // stage outputs of the last extract(), kept for stage-by-stage parity tests
    struct Level {
        int w = 0, h = 0, pstride = 0;
        std::vector<uint8_t> padded;          // (w+38) x (h+38), mvImagePyramid's backing buffer
        std::vector<uint8_t> blurred;         // w x h, the per-level workingMat
        std::vector<KeyPoint> cand;           // vToDistributeKeys (coords relative to minBorder)
        std::vector<KeyPoint> kept;           // after DistributeOctTree + border offset + angle
        const uint8_t* roi() const { return padded.data() + (size_t)kEdge * pstride + kEdge; }
    };
    std::vector<Level> levels;
    long sortTieCount = 0;                    // number of equal-key neighbours seen by compareNodes sorts (diagnostic)

    void computePyramid(const uint8_t* img, int w, int h, int stride);
    void computeKeyPointsOctTree();
    std::vector<KeyPoint> distributeOctTree(const std::vector<KeyPoint>& keys, int minX, int maxX, int minY, int maxY, int N);
    float icAngle(const Level& L, float px, float py) const;
    float icAngleAt(const uint8_t* center, int step) const;
    void orbDescriptor(const KeyPoint& kp, const uint8_t* img, int step, uint8_t* out) const;
};

}  // namespace orc
