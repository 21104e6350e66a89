/* orbx.h — C ABI of the B200-native ORB extractor (libmam3slam_b200.so).
 *
 * This is the drop-in boundary for MAM3SLAM's ORBextractor.  The reference has no FFI layer for this path: the
 * boundary is the C++ class itself (include/ORBextractor.h:43-109, linked into libMAM3SLAM.so,
 * CMakeLists.txt:76-112).  The adapter class in mam3slam_b200/cpp/ORBextractor.h keeps that class surface and
 * forwards to the entry points below; INTEGRATION.md shows the binding a maintainer adds.
 *
 * Conventions: opaque handle, int status (0 = ORBX_OK, negative = error, see orbx_strerror), no exceptions and no
 * C++/OpenCV/torch types across the boundary, caller-owned output buffers, one CUDA stream per handle, a handle
 * is used by one thread at a time (the reference's extractor object is owned by one Tracking thread,
 * include/Tracking.h:276-277).  There is no CPU fallback: every entry point that computes needs a CUDA device.
 */
#ifndef ORBX_H_
#define ORBX_H_
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_ERR_INVALID (-1)      /* bad argument */
#define ORBX_ERR_CUDA (-2)         /* CUDA runtime error; orbx_last_error() has the text */
#define ORBX_ERR_CAPACITY (-3)     /* caller's output buffer too small (n_out tells the need) */
#define ORBX_ERR_UNSUPPORTED (-4)  /* image too small for the 35-px FAST grid, too many features for on-chip state */
#define ORBX_ERR_NO_DEVICE (-5)    /* no CUDA device: the product path has no CPU fallback */

typedef struct orbx_extractor* orbx_handle;

/* Constructor arguments of ORBextractor (include/ORBextractor.h:49-50, src/ORBextractor.cc:409-469) plus
 * device placement.  max_batch = frames processed per pipeline pass (frames in flight on the GPU). */
typedef struct orbx_config {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
    int device;        /* CUDA device ordinal (agent i -> GPU i mod G) */
    int max_batch;     /* >= 1 */
} orbx_config;

/* Binary layout of cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id): 28 bytes, so the C++
 * adapter can copy straight into std::vector<cv::KeyPoint>.  class_id is always -1 (src/ORBextractor.cc never
 * sets it). */
typedef struct orbx_keypoint {
    float x, y;
    float size;
    float angle;       /* degrees [0,360), cv::fastAtan2 polynomial */
    float response;    /* FAST score */
    int32_t octave;
    int32_t class_id;
} orbx_keypoint;

int orbx_device_count(void);
/* PCI bus id of a device ("0000:1b:00.0"), for hosts that bind an agent's threads and pinned staging buffers to the NUMA
 * node its GPU hangs off (mam3slam_b200/numa.py; the reference runs on the CPU and has no counterpart).  out: >= 16 bytes */
int orbx_device_pci_bus_id(int device, char* out, int len);
/* Page-lock / release a host range the caller keeps frames or output arrays in (cudaHostRegister / cudaHostUnregister without
 * CUDA headers on the caller's side).  orbx_extract / orbx_extract_batch use pinned pointers in place; pageable ones are copied
 * through the handle's staging buffer first (a cv::Mat from cv::imread, src/Agent.cc:92-154, is pageable: the reference has no
 * counterpart because it never leaves the host). */
int orbx_pin_host(void* ptr, size_t bytes);
int orbx_unpin_host(void* ptr);
const char* orbx_strerror(int status);
const char* orbx_last_error(orbx_handle h);

/* replaces ORBextractor::ORBextractor (src/ORBextractor.cc:409-469) */
int orbx_create(const orbx_config* cfg, orbx_handle* out);
void orbx_destroy(orbx_handle h);

/* replaces the getters (include/ORBextractor.h:61-81); arrays of nlevels entries, any pointer may be NULL */
int orbx_get_tables(orbx_handle h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                    int* features_per_level);
/* upper bound of keypoints one frame can yield (quota per level + the 4-way split overshoot, SURVEY.md A9 vii) */
int orbx_max_keypoints(orbx_handle h);

/* replaces ORBextractor::operator() (src/ORBextractor.cc:1086-1168) for one HOST image (8-bit gray, `stride`
 * bytes per row).  lap0/lap1 = vLappingArea[0..1].  Writes *n_out keypoints + n_out x 32 descriptor bytes in the
 * reference's output order, *mono_index_out = the reference's return value.  Returns ORBX_OK, or -1
 * (ORBX_ERR_INVALID) for an empty image like the reference (:1090). */
int orbx_extract(orbx_handle h, const uint8_t* img, int width, int height, int stride, int lap0, int lap1,
                 orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out, int* mono_index_out);

/* Throughput mode: nimg HOST images of identical shape (frames in flight).  kps/desc hold
 * nimg x cap entries; n_out / mono_index_out hold nimg ints. */
int orbx_extract_batch(orbx_handle h, const uint8_t* const* imgs, int nimg, int width, int height, int stride,
                       int lap0, int lap1, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                       int* mono_index_out);

/* Device-resident mode: images already in HBM (image i at d_imgs + i*image_bytes, `stride` bytes per row,
 * nimg <= max_batch).  Runs asynchronously on the handle's stream (or on `stream` if non-NULL, a cudaStream_t);
 * results stay on the device until fetched. */
int orbx_extract_device(orbx_handle h, const uint8_t* d_imgs, size_t image_bytes, int nimg, int width, int height,
                        int stride, int lap0, int lap1, void* stream);
int orbx_synchronize(orbx_handle h);
/* per-frame results of the last pass: device pointers (valid until the next pass), counts read back */
int orbx_results_device(orbx_handle h, int frame, const orbx_keypoint** d_kps, const uint32_t** d_desc,
                        int* n_out, int* mono_index_out);
/* the batch arrays themselves (device): frame f's keypoints at *d_kps + f*(*frame_stride), descriptor rows at *d_desc +
 * f*(*frame_stride)*32, its {n, monoIndex} at (*d_counts)[2f .. 2f+1].  Feed these to orbm_track_batch_device. */
int orbx_results_batch_device(orbx_handle h, const orbx_keypoint** d_kps, const uint8_t** d_desc, const int** d_counts,
                              int* frame_stride);
int orbx_fetch_results(orbx_handle h, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                       int* mono_index_out);

/* replaces Frame::ComputeStereoMatches (src/Frame.cc:811-981; SURVEY.md 8f rank 4 — the product is monocular, the stereo
 * constructor is kept for completeness): `left` / `right` are the two extractor objects of Frame's stereo constructor
 * (src/Frame.cc:122-125), each holding the single frame of its last orbx_extract (keypoints, descriptors and
 * mvImagePyramid stay device-resident).  mb = baseline in metres, mbf = baseline * fx.  u_right / depth (HOST, cap floats
 * each) receive mvuRight / mvDepth for the *n_left left keypoints (-1 = no stereo match); *n_kept = matches kept after the
 * median-SAD outlier filter. */
int orbx_stereo_matches(orbx_handle left, orbx_handle right, float mb, float mbf, float* u_right, float* depth, int cap,
                        int* n_left, int* n_kept);

/* replaces the public mvImagePyramid (include/ORBextractor.h:83): level `level` of frame `frame` of the last
 * pass, with (padded != 0) or without the 19-px BORDER_REFLECT_101 ring.  dst_stride in bytes. */
int orbx_level_size(orbx_handle h, int level, int* width, int* height);
int orbx_download_level(orbx_handle h, int frame, int level, int padded, uint8_t* dst, int dst_stride);

/* ---- stage taps for parity tests (not needed by a caller of the reference API) ---- */
int orbx_download_blurred(orbx_handle h, int frame, int level, uint8_t* dst, int dst_stride);
/* FAST candidates before culling, reference order; out = n x 3 ints (x, y relative to minBorder, response) */
int orbx_download_candidates(orbx_handle h, int frame, int level, int* out, int cap, int* n_out);
/* keypoints kept by the quadtree, list order; out = n x 3 ints (x, y relative to minBorder, response) */
int orbx_download_level_keypoints(orbx_handle h, int frame, int level, int* out, int cap, int* n_out);
/* device-side emulation of libstdc++ std::sort with compareNodes (src/ORBextractor.cc:538-553): perm = order of
 * the input indices after sorting (size, ul_x) pairs.  mode 0 = full sort by one thread, 1 = heap-sort fallback only,
 * 2 = the warp-parallel form the quadtree kernel runs (n <= 65535) */
int orbx_debug_sort_nodes(orbx_handle h, const int* size, const int* ul_x, int n, int* perm, int mode);
/* standalone DistributeOctTree (src/ORBextractor.cc:555-779) on device: keys n x 3 ints (x, y, response) */
int orbx_debug_distribute(orbx_handle h, const int* keys, int n, int min_x, int max_x, int min_y, int max_y,
                          int n_features, int* out, int cap, int* n_out);
/* the same with the choice of form: use_pyramid = 1 tries the count-pyramid form first (*used_pyramid = 1 when it completed,
 * 0 when a node deeper than the pyramid made the partitioning form take over); both forms must give the same keys */
int orbx_debug_distribute2(orbx_handle h, const int* keys, int n, int min_x, int max_x, int min_y, int max_y,
                           int n_features, int use_pyramid, int* out, int cap, int* n_out, int* used_pyramid);

/* bookkeeping for bench.py: kernels launched so far by this handle, and the handle's cudaStream_t */
long long orbx_launch_count(orbx_handle h);
void* orbx_stream(orbx_handle h);
/* per-stage device time (ms) of the last pass measured with CUDA events when enabled:
 * [0]=pyramid [1]=fast [2]=octree [3]=blur [4]=finalize [5]=orient+brief [6]=total.  Returns ORBX_OK. */
int orbx_set_stage_timing(orbx_handle h, int enable);
int orbx_stage_times(orbx_handle h, float* ms7);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
