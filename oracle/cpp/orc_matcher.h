// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_primitives.h).  CPU restatement of the reference's ORBmatcher searches
// (src/ORBmatcher.cc) and of the candidate-gather helpers they call (src/Frame.cc:385-416,657-735,
// src/KeyFrame.cc:704-748) over PLAIN ARRAYS: the reference works on a pointer graph (Frame / KeyFrame / MapPoint);
// here every MapPoint-dependent predicate (isBad, Observations()>0, mbTrackInView, projection into the image...)
// arrives as a precomputed per-element flag or number, exactly what the C++ adapter of the product passes to the
// C ABI (include/orbm.h).  Each function keeps the loop structure, scan order, comparison operators and float
// expressions of the function it restates, so it can be checked against the reference line by line.
//
// Parity: UNPINNED by the reference (no tests / golden vectors for this path).  DescriptorDistance
// is pinned against numpy popcount; the search functions are pinned only by being literal restatements.
#pragma once
#include <cstdint>
#include <vector>

namespace orc {

struct KP {              // binary layout of cv::KeyPoint
    float x, y, size, angle, response;
    int octave, class_id;
};

constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:35-37
constexpr int GRID_COLS = 64, GRID_ROWS = 48;                      // include/Frame.h:44-45

// DescriptorDistance (src/ORBmatcher.cc:2058-2074): SWAR popcount over 8 x 32 bits
int descriptor_distance(const uint8_t* a, const uint8_t* b);

// A Frame / KeyFrame as the searches see it: undistorted keypoints, descriptors, image bounds and the 64x48 grid.
struct FrameView {
    int n = 0;
    const KP* kps = nullptr;
    const uint8_t* desc = nullptr;           // n x 32
    float minX = 0, minY = 0, maxX = 0, maxY = 0;
    float invW = 0, invH = 0;                // mfGridElementWidthInv / HeightInv
    std::vector<std::vector<int>> grid;      // [ix*GRID_ROWS+iy] -> keypoint indices in insertion order

    void build(int n_, const KP* k, const uint8_t* d, float minX_, float minY_, float maxX_, float maxY_);
    // Frame::GetFeaturesInArea (src/Frame.cc:657-723); KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:704-748) is the
    // same walk with minLevel = maxLevel = -1
    void features_in_area(float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) const;
};

// ComputeThreeMaxima (src/ORBmatcher.cc:2012-2053) on bin sizes
void three_maxima(const int* sizes, int L, int& ind1, int& ind2, int& ind3);
// rotation-histogram bin of a match (e.g. src/ORBmatcher.cc:1784-1791)
int rot_bin(float angle1, float angle2);

// M1  SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) (src/ORBmatcher.cc:43-213), mono branch.
//     valid[q] = mbTrackInView && !far && !isBad;  occupied[i] = mvpMapPoints[i] && Observations()>0 (in/out: a new
//     assignment occupies the slot iff qocc[q] = the query's map point has Observations() > 0; qocc == nullptr: always).
//     kpMatch[i] = LAST query assigned to keypoint i; the return value counts assignments, as the reference's nmatches++ does.
int search_local_map(const FrameView& F, uint8_t* occupied, const float* scaleFactors, int nq, const uint8_t* valid,
                     const float* projX, const float* projY, const int* level, const float* viewCos,
                     const uint8_t* qdesc, float th, float nnratio, int* kpMatch, const uint8_t* qocc = nullptr);

// M2 / M3  SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) (:1676-1887, mono) and
//     SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:1889-2010): window th*scale[octave],
//     levels [octave-1, octave+1], accept best <= thDist, claim, rotation histogram.
int search_frame(const FrameView& F, uint8_t* occupied, const float* scaleFactors, int nq, const uint8_t* valid,
                 const float* u, const float* v, const int* octave, const float* qangle, const uint8_t* qdesc,
                 float th, int thDist, bool checkOri, int* kpMatch, const uint8_t* qocc = nullptr);

// M7  SearchByProjection(KeyFrame*, Sim3, points, matched, th, ratioHamming) (:427-532 and :534-646)
int search_sim3(const FrameView& KF, uint8_t* matched, const float* scaleFactors, int nq, const uint8_t* valid,
                const float* u, const float* v, const int* level, const uint8_t* qdesc, int th, float ratioHamming,
                int* kpMatch);

// M5  SearchForInitialization (:648-763).  prevMatched: n1 x 2 floats, updated like vbPrevMatched.
int search_initialization(const FrameView& F1, const FrameView& F2, float* prevMatched, int windowSize, float nnratio,
                          bool checkOri, int* matches12);

// DBoW2::FeatureVector as CSR: sorted node ids, offsets (nn+1), feature indices
struct FeatVec { int nn; const int* node; const int* off; const int* idx; };

// M4a SearchByBoW(KeyFrame*, Frame&, matches) (:223-425), mono.  elig1[i] = KF keypoint i has a good MapPoint.
//     matchF[j] = KF keypoint matched to frame keypoint j (vpMapPointMatches), -1 = NULL.
int search_bow_kf_frame(const FrameView& KF, const FeatVec& fvKF, const uint8_t* elig1, const FrameView& F,
                        const FeatVec& fvF, float nnratio, bool checkOri, int* matchF);
// M4b SearchByBoW(KeyFrame*, KeyFrame*, matches12) (:765-905).  matches12[i1] = i2.
int search_bow_kf_kf(const FrameView& K1, const FeatVec& fv1, const uint8_t* elig1, const FrameView& K2,
                     const FeatVec& fv2, const uint8_t* elig2, float nnratio, bool checkOri, int* matches12);

// M6  SearchForTriangulation (:907-1146), mono / Pinhole.  free1/free2 = keypoint has no MapPoint; ep = epipole in
//     image 2; F12 row-major as Pinhole::epipolarConstrain computes it (src/CameraModels/Pinhole.cpp:107-129);
//     sigma2 = mvLevelSigma2 of KF2.
int search_triangulation(const FrameView& K1, const FeatVec& fv1, const uint8_t* free1, const FrameView& K2,
                         const FeatVec& fv2, const uint8_t* free2, const float* ep, const float* scaleFactors2,
                         const float* F12, const float* sigma2, bool coarse, bool checkOri, int* matches12);

// M9  candidate part of Fuse (:1148-1338 mono, :1340-1455): per map point the best keypoint in the window.
void fuse_candidates(const FrameView& KF, const float* scaleFactors, const float* invSigma2, int nq,
                     const uint8_t* valid, const float* u, const float* v, const int* level, const uint8_t* qdesc,
                     float th, bool chi2Gate, int* bestIdx, int* bestDist);

// M8  SearchBySim3 (:1457-1674): bidirectional projection search + mutual agreement.  Query i of direction 1->2 is the
//     map point of KF1's keypoint i (descriptor row i of `qdesc1`), projected into KF2 at (u12, v12) with predicted
//     level level12; valid1[i] = it survives the reference's pre-checks (:1500-1531).  matches12[i1] = i2 or -1.
//     qdesc1 / qdesc2 = MapPoint::GetDescriptor() of each slot (n1 x 32, n2 x 32).
int search_by_sim3(const FrameView& K1, const FrameView& K2, const float* scaleFactors1, const float* scaleFactors2,
                   const uint8_t* valid1, const float* u12, const float* v12, const int* level12, const uint8_t* qdesc1,
                   const uint8_t* valid2, const float* u21, const float* v21, const int* level21, const uint8_t* qdesc2, float th,
                   int* matches12);

// 8(f) rank 4: Frame::ComputeStereoMatches (src/Frame.cc:811-981).  pyrL / pyrR = mvImagePyramid of the left / right
// extractor as plain planes (nlevels pointers + row strides), colsR = widths of the right levels.  Writes mvuRight /
// mvDepth (N floats each, -1 = no match); returns the number of stereo matches kept.
int stereo_matches(int N, const KP* keysL, const uint8_t* descL, int Nr, const KP* keysR, const uint8_t* descR, int nlevels,
                   const uint8_t* const* pyrL, const int* strideL, const uint8_t* const* pyrR, const int* strideR,
                   const int* colsR, const float* scaleFactors, const float* invScaleFactors, float mb, float mbf,
                   float* mvuRight, float* mvDepth);

// M11 brute-force k-NN (config 5): per query the k smallest distances, ties -> lowest row
void knn(const uint8_t* q, int nq, const uint8_t* db, long ndb, int k, uint16_t* dist, int32_t* idx, int threads);

// 8(f) rank 2: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:329-396): index of the descriptor with the
// least median distance to the others
int distinctive_descriptor(const uint8_t* desc, int n);

// 8(f) rank 1 (last part): Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:747-780,782-812).  Both call
// cv::undistortPoints(mat, mat, K, distCoef, Mat(), K) — OpenCV (external dependency, find_package(OpenCV 4.4),
// CMakeLists.txt:38; calib3d/src/undistort.dispatch.cpp, cvUndistortPointsInternal) with the default criteria
// (5 iterations, no epsilon test), R = I, P = K, no tilt terms.  All arithmetic is double on float inputs; the result is
// rounded to float.  Pinned bit-exact against cv2.undistortPoints (tests/golden/cv2_undistort.npz).
// K = {fx, fy, cx, cy}; dist = k1 k2 p1 p2 [k3 [k4 k5 k6 [s1 s2 s3 s4]]] (ndist in {4,5,8,12}); xy: n x 2 floats.
void undistort_points(const float* xy, int n, const float* K, const float* dist, int ndist, float* out);

// 8(f) rank 3: DBoW2 TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)
// (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1198 and the per-feature descent :1218-1260; BowVector.cpp:35-88;
// FeatureVector.cpp:32-46), called by Frame::ComputeBoW with levelsup = 4 (src/Frame.cc:738-745).
// The vocabulary tree is given as plain arrays: children of node p are children[childOff[p] .. childOff[p+1]) in the
// order of m_nodes[p].children; a node with no children is a leaf (word).  weighting: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY
// (BowVector.h:39-45); norm: 0 none, 1 L1, 2 L2 (what m_scoring_object->mustNormalize yields, ScoringObject.h:74-89).
struct VocabView {
    int nnodes, L;
    const int* childOff; const int* children; const uint8_t* desc; const int* wordId; const double* weight;
    int weighting, norm;
};
// per feature: word id, weight and the node `levelsup` levels above the leaves
void bow_descend(const VocabView& V, const uint8_t* feature, int levelsup, int* wordId, double* weight, int* nodeId);
// whole transform; outputs are flattened maps in key order.  bowWord/bowValue need n entries, fvNode n, fvOff n+1, fvIdx n.
void bow_transform(const VocabView& V, const uint8_t* desc, int n, int levelsup, int* featWord, int* featNode,
                   int* nBow, int* bowWord, double* bowValue, int* nFv, int* fvNode, int* fvOff, int* fvIdx);

}  // namespace orc
