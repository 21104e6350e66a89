/* orbm.h — C ABI of the B200-native ORB Hamming matcher (libmam3slam_b200.so).
 *
 * Drop-in boundary for MAM3SLAM's ORBmatcher (include/ORBmatcher.h:36-103, src/ORBmatcher.cc).  The reference has no
 * FFI layer: the boundary is the C++ class, whose methods take Frame / KeyFrame / MapPoint pointer graphs.  The C++
 * adapter (mam3slam_b200/cpp/ORBmatcher.h) flattens those graphs into the plain arrays below — every
 * MapPoint-dependent predicate (isBad, Observations()>0, mbTrackInView, in-image projection ...) becomes a per-element
 * flag or number computed on the host exactly where the reference computes it — and calls these entry points; all
 * Hamming work (candidate gather through the 64x48 frame grid or the BoW node lists, DescriptorDistance, best /
 * second-best selection in the reference's scan order, TH_LOW / TH_HIGH / NN-ratio acceptance, once-only slot
 * occupancy, rotation-histogram filtering) runs in CUDA kernels.  Results are bit-exact with the reference's loops.
 *
 * Conventions: opaque handles, int status (0 = ORBM_OK, negative = error), no C++/OpenCV types, caller-owned HOST
 * buffers unless stated, one CUDA stream per handle, one thread per handle at a time.  No CPU fallback.
 * Keypoints use cv::KeyPoint's 28-byte layout (orbx_keypoint) so std::vector<cv::KeyPoint>::data() can be passed as is.
 */
#ifndef ORBM_H_
#define ORBM_H_
#include <stdint.h>
#include <stddef.h>
#include "orbx.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ORBM_OK 0
#define ORBM_ERR_INVALID (-1)
#define ORBM_ERR_CUDA (-2)
#define ORBM_ERR_CAPACITY (-3)
#define ORBM_ERR_UNSUPPORTED (-4)
#define ORBM_ERR_NO_DEVICE (-5)

#define ORBM_TH_HIGH 100      /* ORBmatcher::TH_HIGH  (src/ORBmatcher.cc:35) */
#define ORBM_TH_LOW 50        /* ORBmatcher::TH_LOW   (:36) */
#define ORBM_HISTO_LENGTH 30  /* ORBmatcher::HISTO_LENGTH (:37) */

typedef struct orbm_matcher* orbm_handle;
typedef struct orbm_frame_s* orbm_frame;

const char* orbm_strerror(int status);
const char* orbm_last_error(orbm_handle h);
int orbm_create(int device, orbm_handle* out);
void orbm_destroy(orbm_handle h);
long long orbm_launch_count(orbm_handle h);
void* orbm_stream(orbm_handle h);

/* replaces ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:2058-2074): host-side, same SWAR arithmetic */
int orbm_distance(const uint8_t* a, const uint8_t* b);

/* ---- frames: device mirror of what the searches read from a Frame / KeyFrame ----
 * n undistorted keypoints (mvKeysUn), n x 32 descriptor bytes (mDescriptors), image bounds (mnMinX.. mnMaxY).  The 64x48
 * grid of Frame::AssignFeaturesToGrid (src/Frame.cc:385-416; PosInGrid :725-735) is built on the device.
 * on_device != 0: kps / desc are DEVICE pointers (e.g. from orbx_results_device), nothing crosses PCIe. */
typedef struct orbm_frame_view {
    int n;
    const orbx_keypoint* kps;
    const uint8_t* desc;
    float min_x, min_y, max_x, max_y;
    int on_device;
} orbm_frame_view;
int orbm_frame_create(orbm_handle h, const orbm_frame_view* view, orbm_frame* out);
int orbm_frame_update(orbm_handle h, orbm_frame f, const orbm_frame_view* view);   /* reuse the slot for a new frame */
void orbm_frame_destroy(orbm_handle h, orbm_frame f);
/* the grid as CSR (cell = ix*48+iy, off has 64*48+1 ints, idx has n ints): parity tap for AssignFeaturesToGrid */
int orbm_frame_grid(orbm_handle h, orbm_frame f, int* off, int* idx);
/* replaces Frame::GetFeaturesInArea (src/Frame.cc:657-723) and KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:704-748,
 * min_level = max_level = -1): indices in the reference's order.  Parity tap; the searches gather on the device. */
int orbm_features_in_area(orbm_handle h, orbm_frame f, float x, float y, float r, int min_level, int max_level,
                          int* out, int cap, int* n_out);

/* ---- projection-guided searches (queries = map points already projected by the caller) ----
 * Common arguments: nq queries; valid[q] = the query survives the reference's pre-checks; qdesc = nq x 32 bytes
 * (MapPoint::GetDescriptor); scale_factors = mvScaleFactors (nlevels floats); kp_match[n] (out) = index of the query
 * assigned to keypoint i or -1 (the reference writes F.mvpMapPoints[i] = pMP); *nmatches = the reference's return
 * value (it counts assignments).  occupied[n] (in/out) = slot already holds a map point that blocks matching.  Sequential
 * semantics of the reference loop: an accepted match is written into the slot, and it blocks the slot for later queries
 * iff q_occupies[q] != 0, i.e. the query's MapPoint has Observations() > 0 (src/ORBmatcher.cc:92-94, 1745-1747); a
 * later query may otherwise take the same keypoint and overwrite it (kp_match = the LAST one).  q_occupies = NULL: every
 * match occupies (the KeyFrame variant :1955, and every map point a mono system tracks). */

/* replaces SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints) (src/ORBmatcher.cc:43-213),
 * mono branch.  level = mnTrackScaleLevel, view_cos = mTrackViewCos (RadiusByViewingCos :215-221). */
int orbm_search_local_map(orbm_handle h, orbm_frame f, uint8_t* occupied, const float* scale_factors, int nlevels,
                          int nq, const uint8_t* valid, const float* proj_x, const float* proj_y, const int* level,
                          const float* view_cos, const uint8_t* qdesc, float th, float nnratio, int* kp_match,
                          int* nmatches, const uint8_t* q_occupies);
/* replaces SearchByProjection(Frame& Cur, const Frame& Last, th, bMono=true) (:1676-1887) and
 * SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:1889-2010): octave = last octave / predicted
 * level, qangle = angle of the source keypoint, th_dist = TH_HIGH or ORBdist. */
int orbm_search_frame(orbm_handle h, orbm_frame f, uint8_t* occupied, const float* scale_factors, int nlevels, int nq,
                      const uint8_t* valid, const float* u, const float* v, const int* octave, const float* qangle,
                      const uint8_t* qdesc, float th, int th_dist, int check_orientation, int* kp_match,
                      int* nmatches, const uint8_t* q_occupies);
/* replaces SearchByProjection(KeyFrame*, Sim3, points, matched, th, ratioHamming) (:427-532) and its pointKFs
 * overload (:534-646): matched[n] in/out = vpMatched[i] != NULL. */
int orbm_search_sim3(orbm_handle h, orbm_frame kf, uint8_t* matched, const float* scale_factors, int nlevels, int nq,
                     const uint8_t* valid, const float* u, const float* v, const int* level, const uint8_t* qdesc,
                     int th, float ratio_hamming, int* kp_match, int* nmatches);
/* replaces SearchForInitialization (:648-763): prev_matched = n1 x 2 floats in/out (vbPrevMatched), matches12[n1]. */
int orbm_search_initialization(orbm_handle h, orbm_frame f1, orbm_frame f2, float* prev_matched, int window_size,
                               float nnratio, int check_orientation, int* matches12, int* nmatches);
/* candidate part of Fuse(KeyFrame*, vector<MapPoint*>, th) (:1148-1338, chi2_gate = 1, needs inv_sigma2 =
 * mvInvLevelSigma2) and Fuse(KeyFrame*, Sim3, points, th, replace) (:1340-1455, chi2_gate = 0): best keypoint and
 * distance per map point; the map mutations (Replace / AddObservation) stay on the host, applied in order. */
int orbm_fuse_candidates(orbm_handle h, orbm_frame kf, const float* scale_factors, const float* inv_sigma2, int nlevels,
                         int nq, const uint8_t* valid, const float* u, const float* v, const int* level,
                         const uint8_t* qdesc, float th, int chi2_gate, int* best_idx, int* best_dist);

/* replaces SearchBySim3(KF1, KF2, vpMatches12, S12, th) (:1457-1674; no caller in the reference tree, kept for the class
 * surface).  Slot i of a keyframe stands for its map point: valid1[i] = the point survives the pre-checks (:1500-1531: present,
 * not already matched, not bad, positive depth, in image, inside the scale-invariance range), (u12, v12) = its projection
 * into KF2, level12 = PredictScale, qdesc1 = MapPoint::GetDescriptor rows (n1 x 32); likewise 2 -> 1.  Each direction keeps
 * the best keypoint (levels [l-1, l], first minimum) when its distance <= TH_HIGH; matches12[i1] = i2 where both
 * directions agree, else -1; *nfound = the reference's return value. */
int orbm_search_by_sim3(orbm_handle h, orbm_frame kf1, orbm_frame kf2, const float* scale_factors1,
                        const float* scale_factors2, int nlevels, const uint8_t* valid1, const float* u12, const float* v12,
                        const int* level12, const uint8_t* qdesc1, const uint8_t* valid2, const float* u21, const float* v21,
                        const int* level21, const uint8_t* qdesc2, float th, int* matches12, int* nfound);

/* Device-resident tracking step (BASELINE config[1]: per-frame extraction + SearchByProjection on one GPU): for every
 * frame i >= 1 of a batch of extractor results that never left HBM (d_kps / d_desc / d_counts = the arrays behind
 * orbx_results_device: frame f at d_kps + f*frame_stride, its keypoint count at d_counts[f*count_stride]) runs
 * SearchByProjection(Cur = frame i, Last = frame i-1, th, bMono = true) (src/ORBmatcher.cc:1676-1887) with every keypoint
 * of the last frame standing for a map point projected at its own position.  The frame grids are built on the device.
 * d_kp_match: nframes x frame_stride ints (row i = the reference's CurrentFrame.mvpMapPoints as indices into frame i-1,
 * -1 = none; row 0 all -1), d_nmatches: nframes ints.  Asynchronous on `stream` (NULL = the handle's stream). */
int orbm_track_batch_device(orbm_handle h, int nframes, const orbx_keypoint* d_kps, const uint8_t* d_desc, const int* d_counts,
                            int frame_stride, int count_stride, const float* bounds4, const float* scale_factors, int nlevels,
                            float th, int th_dist, int check_orientation, int* d_kp_match, int* d_nmatches, void* stream);

/* ---- BoW-guided searches: DBoW2::FeatureVector (std::map<NodeId, vector<unsigned>>) flattened to CSR ---- */
typedef struct orbm_featvec {
    int n_nodes;
    const int* node;     /* ascending node ids */
    const int* off;      /* n_nodes + 1 */
    const int* idx;      /* feature indices, insertion order */
} orbm_featvec;
/* replaces SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches) (:223-425), mono.  elig1[i] = KF keypoint i holds a good
 * map point.  match_f[nF] = KF keypoint matched to frame keypoint j, or -1. */
int orbm_search_bow_kf_frame(orbm_handle h, orbm_frame kf, const orbm_featvec* fv_kf, const uint8_t* elig1,
                             orbm_frame f, const orbm_featvec* fv_f, float nnratio, int check_orientation,
                             int* match_f, int* nmatches);
/* replaces SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12) (:765-905): matches12[n1] = keypoint of KF2 or -1 */
int orbm_search_bow_kf_kf(orbm_handle h, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* elig1, orbm_frame kf2,
                          const orbm_featvec* fv2, const uint8_t* elig2, float nnratio, int check_orientation,
                          int* matches12, int* nmatches);
/* replaces SearchForTriangulation (:907-1146), mono + Pinhole: free1/free2 = keypoint has no map point; ep = epipole
 * of KF1 in KF2 (2 floats); F12 = 9 floats row-major, computed by the caller as Pinhole::epipolarConstrain does
 * (src/CameraModels/Pinhole.cpp:107-112); sigma2 = KF2's mvLevelSigma2; coarse = bCoarse. */
int orbm_search_triangulation(orbm_handle h, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* free1,
                              orbm_frame kf2, const orbm_featvec* fv2, const uint8_t* free2, const float* ep,
                              const float* scale_factors2, const float* F12, const float* sigma2, int nlevels,
                              int coarse, int check_orientation, int* matches12, int* nmatches);
/* SearchForTriangulation (:907-1146) for cameras whose epipolar predicate stays with the caller: the reference calls the virtual
 * pCamera1->epipolarConstrain (:1072), and the only settings it ships are KannalaBrandt8, whose predicate triangulates through
 * an SVD (src/CameraModels/KannalaBrandt8.cpp:216-220, 306-380).  For every keypoint i1 of KF1 this returns the KF2 candidates
 * that survive all tests the reference makes BEFORE that predicate (same vocabulary node, no map point, DescriptorDistance <=
 * TH_LOW, outside the epipole disc :1026-1032), in the reference's scan order: cand_idx2 / cand_dist are n1 x cap ints,
 * cand_count[i1] the true number (ORBM_ERR_CAPACITY if any exceeds cap: call again with a larger cap).  The caller then runs
 * the reference's inner loop over them with its own camera object (mam3slam_b200/cpp/ORBmatcher.h does). */
int orbm_triangulation_candidates(orbm_handle h, orbm_frame kf1, const orbm_featvec* fv1, const uint8_t* free1,
                                  orbm_frame kf2, const orbm_featvec* fv2, const uint8_t* free2, const float* ep,
                                  const float* scale_factors2, int nlevels, int cap, int* cand_idx2, int* cand_dist,
                                  int* cand_count);

/* ---- brute-force k-NN Hamming search (BASELINE config 5: cross-map search at scale) ----
 * For each of nq queries the k smallest distances over ndb rows, ties to the lowest row (a strict '<' scan).
 * The database is uploaded once and stays in HBM; row_offset is added to every returned index (global row ids of a
 * shard).  dist: nq x k uint16, idx: nq x k int32 (HOST buffers). */
typedef struct orbm_db_s* orbm_db;
int orbm_db_create(orbm_handle h, const uint8_t* rows, long long ndb, int rows_on_device, long long row_offset,
                   orbm_db* out);
void orbm_db_destroy(orbm_handle h, orbm_db db);
int orbm_knn(orbm_handle h, orbm_db db, const uint8_t* q, int nq, int k, uint16_t* dist, int32_t* idx);
/* device-resident variant: d_q, d_dist, d_idx are device pointers; asynchronous on `stream` (NULL = handle stream) */
int orbm_knn_device(orbm_handle h, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx,
                    void* stream);
/* Allocation-free variant of orbm_knn_device.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same result bit for bit; the call itself performs no cudaMalloc / cudaFree / allocating
 * library call.  Its scratch (per-chunk partial top-k lists) is a file-scope device array of ORBM_KNN_STATIC_MAX_ENTRIES
 * 64-bit keys (8 MiB) that every GPU loading the library holds until the process exits.  The call needs
 * chunks x nq x k entries (chunks ~ 4 x SM count / ceil(nq / 256): ~74 on a B200 at nq = 2000); larger requests return
 * ORBM_ERR_CAPACITY with the numbers in orbm_last_error().  The database must be device-resident
 * (orbm_db_create(..., rows_on_device = 1, ...), created beforehand).  The first call on a device resolves the array's address
 * (cudaGetSymbolAddress): make that one call outside stream capture.  All calls on a device SHARE the array, which keeps the
 * previous call's contents (nothing needs to start at zero: every entry read is written first by the same call): calls
 * that overlap on the GPU - different streams, concurrent graph replays - race and must be serialised by the caller. */
#define ORBM_KNN_STATIC_MAX_ENTRIES (1u << 20)
int orbm_knn_device_static(orbm_handle h, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist,
                           int32_t* d_idx, void* stream);
/* test aid (no counterpart in the reference): number of allocating CUDA calls (cudaMalloc, cudaMallocHost, stream creation) this
 * library's matcher side has made since it was loaded.  The C++ adapter's pooled ORBmatcher objects and
 * orbm_knn_device_static must leave it unchanged in steady state (tests/test_cpp_adapter_gpu.py, tests/test_zz_no_alloc.py). */
long long orbm_debug_alloc_count(void);
/* merge of per-shard top-k lists (nshards x nq x k, e.g. after an all-gather): same tie rule on global row ids */
int orbm_knn_merge(const uint16_t* dist, const int32_t* idx, int nshards, int nq, int k, uint16_t* out_dist,
                   int32_t* out_idx);

/* the same merge on the device, for lists that an all-gather left in HBM: shard s has its nq x k uint16 distances at
 * d_dist + s * shard_stride_bytes and its nq x k int32 global row ids at d_idx + s * shard_stride_bytes (both device
 * pointers; one buffer per rank holding ids and distances can be gathered with a single collective).  Asynchronous on
 * `stream` (NULL = the handle's stream). */
int orbm_knn_merge_device(orbm_handle h, const void* d_dist, const void* d_idx, size_t shard_stride_bytes, int nshards, int nq,
                          int k, uint16_t* d_out_dist, int32_t* d_out_idx, void* stream);

/* The same search on the tensor cores (A/B beside orbm_knn_device, which stays the default; no counterpart in the reference):
 * Hamming(q, r) = popc(q) + popc(r) - 2 <q, r>, the dot products computed exactly by tcgen05.mma kind::i8 on bits expanded to
 * bytes on chip (mam3slam_b200/csrc/orbm_kernels_tc.cuh).  Same arguments, same result bit for bit. */
int orbm_knn_device_tc(orbm_handle h, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist, int32_t* d_idx,
                       void* stream);

/* ---- sharded k-NN over peer memory (map-sharded search of SURVEY.md 8e; the reference is one CPU process with one address
 * space and has no counterpart) ------------------------------------------------------------------------------------------
 * One process per GPU; rank r holds rows [row_offset, row_offset + n) of the database (orbm_db) and a replica of the queries.
 * orbm_knn_sharded = local scan + all-gather + merge, with the exchange done BY the kernels: the scan kernel stores every
 * finished query block of merged (distance, global row) keys straight into every rank's exchange buffer over NVLink peer
 * stores and bumps that rank's arrival counter; a second kernel waits for all counters and merges.  No NCCL call, no host
 * synchronisation.  The result equals the unsharded search bit for bit (same (distance, row) rule).
 *
 * Setup: every rank creates its buffer, the 64-byte CUDA IPC handles are exchanged by whatever the host has (the Python mirror
 * uses torch.distributed.all_gather_object), every rank opens every other rank's handle.  orbm_xchg_open_local links two
 * buffers of ONE process instead (several GPUs driven by one process, or two handles on one GPU in the tests).
 * Collective semantics: every rank must call orbm_knn_sharded the same number of times with the same nq and k.  A rank that
 * never arrives does not hang the GPU: the merge gives up after ORBM_XCHG_TIMEOUT_MS (default 4000), writes "no neighbour"
 * (-1 / 65535) and raises orbm_xchg_status(). */
typedef struct orbm_xchg_s* orbm_xchg;
#define ORBM_XCHG_HANDLE_BYTES 64
int orbm_xchg_create(orbm_handle h, int rank, int world, int max_nq, int max_k, orbm_xchg* out);
int orbm_xchg_ipc_handle(orbm_xchg x, void* handle64);
int orbm_xchg_open(orbm_xchg x, int peer_rank, const void* handle64);
int orbm_xchg_open_local(orbm_xchg x, int peer_rank, orbm_xchg peer);
int orbm_xchg_status(orbm_xchg x);            /* 0 = fine, 1 = a merge timed out waiting for a rank */
void orbm_xchg_destroy(orbm_xchg x);
int orbm_knn_sharded(orbm_handle h, orbm_xchg x, orbm_db db, const uint8_t* d_q, int nq, int k, uint16_t* d_dist,
                     int32_t* d_idx, void* stream);

/* replaces the selection in MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:366-396) for a batch of map
 * points: descriptors of point p are rows [off[p], off[p+1]); best[p] = row (relative) with the least median distance */
int orbm_distinctive_descriptors(orbm_handle h, const uint8_t* desc, const int* off, int npoints, int* best);

/* ---- Frame-side steps next to the searches ---- */

/* replaces cv::undistortPoints(mat, mat, K, mDistCoef, Mat(), mK) in Frame::UndistortKeyPoints (src/Frame.cc:747-780) and
 * Frame::ComputeImageBounds (src/Frame.cc:782-812).  K4 = {fx, fy, cx, cy}; dist = k1 k2 p1 p2 [k3 [k4 k5 k6 [s1..s4]]]
 * (ndist 4, 5, 8 or 12); xy / out: n x 2 floats (HOST; may alias).  Double arithmetic, result bit-identical to OpenCV's.
 * The caller keeps the reference's shortcut: when dist[0] == 0 the keypoints are used as they are. */
int orbm_undistort_points(orbm_handle h, const float* xy, int n, const float* K4, const float* dist, int ndist,
                          float* out);
/* device-resident batch: frame f holds d_counts[f * count_stride] keypoints at d_kps + f * frame_stride (the layout
 * orbx_extract_batch_device writes); d_kps_un receives the same records with pt undistorted (mvKeysUn).  Asynchronous. */
int orbm_undistort_keypoints_device(orbm_handle h, int nframes, const orbx_keypoint* d_kps, const int* d_counts,
                                    int frame_stride, int count_stride, const float* K4, const float* dist, int ndist,
                                    orbx_keypoint* d_kps_un, void* stream);

/* DBoW2 vocabulary tree as plain arrays (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h m_nodes): node 0 is the root; the
 * children of node p are children[child_off[p] .. child_off[p+1]) in the order of m_nodes[p].children; a node without
 * children is a word with id word_id[p] and weight weight[p].  weighting: 0 TF_IDF, 1 TF, 2 IDF, 3 BINARY
 * (BowVector.h:39-45); norm: 0 none, 1 L1, 2 L2 (m_scoring_object->mustNormalize; ORBvoc is TF_IDF + L1). */
typedef struct orbm_vocab_s* orbm_vocab;
typedef struct {
    int nnodes, levels;
    const int* child_off;      /* nnodes + 1 */
    const int* children;       /* child_off[nnodes] node ids */
    const uint8_t* node_desc;  /* nnodes x 32 (row 0 unused) */
    const int* word_id;        /* nnodes, read at leaves */
    const double* weight;      /* nnodes, read at leaves */
    int weighting, norm;
} orbm_vocab_view;
int orbm_vocab_create(orbm_handle h, const orbm_vocab_view* view, orbm_vocab* out);   /* validates and uploads the tree */
void orbm_vocab_destroy(orbm_handle h, orbm_vocab v);
/* replaces mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4) (src/Frame.cc:738-745;
 * TemplatedVocabulary.h:1127-1198).  desc: n x 32 (HOST).  feat_word[i] = word of feature i, or -(word+1) when the
 * word is stopped (weight 0) and the feature is left out of both vectors; feat_node[i] = its node `levelsup` levels above
 * the leaves (either may be NULL).  BowVector: *n_bow (word, value) pairs in ascending word order (capacity n).
 * FeatureVector as the CSR the SearchByBoW entry points take: *n_fv ascending node ids, fv_off (n + 1 ints),
 * fv_idx (n ints, ascending feature index inside a node). */
int orbm_bow_transform(orbm_handle h, orbm_vocab v, const uint8_t* desc, int n, int levelsup, int* feat_word,
                       int* feat_node, int* n_bow, int* bow_word, double* bow_value, int* n_fv, int* fv_node,
                       int* fv_off, int* fv_idx);
/* device-resident batch of the descent only (per-feature word / node ids), for descriptors left in HBM by the
 * extractor; same frame layout as orbm_track_batch_device.  Asynchronous on `stream`. */
int orbm_bow_transform_device(orbm_handle h, orbm_vocab v, int nframes, const uint8_t* d_desc, const int* d_counts,
                              int frame_stride, int count_stride, int levelsup, int* d_feat_word, int* d_feat_node,
                              void* stream);

/* Hamming evaluations performed by the last search call (for matches/s reporting); the track variant waits for the
 * device and reads the counter of the last orbm_track_batch_device */
long long orbm_track_evaluations(orbm_handle h);
long long orbm_last_evaluations(orbm_handle h);

#ifdef __cplusplus
}
#endif
#endif /* ORBM_H_ */
