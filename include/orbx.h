/* orbx.h — C ABI of the B200-native ORB extractor (liborbx.so).
 *
 * Drop-in boundary for ORB_SLAM::ORBextractor (reference include/ORBextractor.h:32-77, src/ORBextractor.cc:457-822),
 * whose single call site is the Frame constructor, (*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors)
 * (reference src/Frame.cc:59).  Plain pointers and sizes only; no C++/torch/OpenCV types cross this boundary.
 * There is NO CPU fallback behind these entry points: every one of them runs hand-written sm_100a CUDA kernels and
 * returns ORBX_ERR_CUDA when no usable device is present.
 *
 * Error convention: 0 on success, negative ORBX_ERR_* otherwise, message via orbx_last_error().  The reference
 * asserts instead (src/ORBextractor.cc:82,725); the empty-image early return (:721-722) maps to "*n_out = 0, return 0".
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* == cv::KeyPoint memory layout (28 bytes): what vector<cv::KeyPoint>& keypoints receives, ORBextractor.h:43-45.
 * x,y     level coordinates * mvScaleFactor[octave]                         (src/ORBextractor.cc:769-775)
 * size    (float)(int)(31 * mvScaleFactor[octave])                           (:675,692)
 * angle   intensity-centroid orientation, degrees in [0,360)                 (:124-151)
 * response FAST-9 score, or the Harris response when score_type == 0         (:607-620)
 * octave  pyramid level, class_id = -1                                                                        */
typedef struct orbx_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

enum { ORBX_HARRIS_SCORE = 0, ORBX_FAST_SCORE = 1 }; /* ORBextractor.h:36 */

/* OpenCV-version dependent arithmetic (the reference does not pin its OpenCV, CMakeLists.txt:19):
 * blur_mode   0 = OpenCV 2.4.x sepFilter2D 8-bit path, taps 18,34,49,55,49,34,18 (default: the reference's era)
 *             1 = OpenCV >= 3.4.2 GaussianBlurFixedPoint, taps 18,34,48,56,48,34,18
 * retain_mode 0 = KeyPointsFilter::retainBest with nth_element(begin, begin+n, end)    (OpenCV 2.4.x, default)
 *             1 = nth_element(begin, begin+n-1, end)                                   (OpenCV >= 3.4)          */
enum { ORBX_BLUR_CV24 = 0, ORBX_BLUR_CV4_FIXED = 1 };
enum { ORBX_RETAIN_CV24 = 0, ORBX_RETAIN_CV4 = 1 };

typedef struct orbx_params {
  int32_t nfeatures;   /* ORBextractor.nFeatures   (Data/Settings.yaml:29, ctor arg 1, ORBextractor.h:38) */
  float scale_factor;  /* ORBextractor.scaleFactor (Settings.yaml:32) */
  int32_t nlevels;     /* ORBextractor.nLevels     (Settings.yaml:35) */
  int32_t score_type;  /* ORBextractor.nScoreType  (Settings.yaml:41): 0 Harris, 1 FAST */
  int32_t fast_th;     /* ORBextractor.fastTh      (Settings.yaml:38) */
  int32_t max_width, max_height; /* largest frame this handle will see (device buffers are sized once) */
  int32_t max_batch;   /* frames processed per kernel wave; larger batches are chunked internally */
  int32_t device;      /* CUDA device ordinal */
  int32_t blur_mode, retain_mode;
} orbx_params;

typedef struct orbx_handle orbx_handle;

enum {
  ORBX_OK = 0,
  ORBX_ERR_ARG = -1,         /* bad argument (null pointer, non-positive size, image larger than max_*) */
  ORBX_ERR_CUDA = -2,        /* CUDA runtime error or no device */
  ORBX_ERR_UNSUPPORTED = -3, /* configuration on which the reference itself divides by zero / asserts */
  ORBX_ERR_OVERFLOW = -4,    /* an internal fixed-capacity list overflowed (never silently truncated) */
  ORBX_ERR_IO = -5           /* settings file unreadable / key missing */
};

/* Fills the reference defaults: 1000, 1.2f, 8, FAST_SCORE, 20 (ORBextractor.h:38), 640x480, batch 256, device 0. */
int orbx_default_params(orbx_params* p);
/* Reads the five ORBextractor.* keys of a Settings.yaml (reference src/Tracking.cc:105-109). Other fields keep
 * their current values. */
int orbx_params_from_yaml(const char* settings_path, orbx_params* p);

int orbx_create(const orbx_params* p, orbx_handle** out);
void orbx_destroy(orbx_handle* h);
int orbx_get_levels(const orbx_handle* h);         /* ORBextractor::GetLevels,      ORBextractor.h:47 */
float orbx_get_scale_factor(const orbx_handle* h); /* ORBextractor::GetScaleFactor, ORBextractor.h:50 */
const char* orbx_last_error(const orbx_handle* h); /* h == NULL: error of the last failed orbx_create on this thread */

/* ORBextractor::operator()(image, mask, keypoints, descriptors)  (src/ORBextractor.cc:718-779), HOST buffers.
 * img: 8-bit single channel, `stride` bytes between rows.  kps: capacity nfeatures.  desc: capacity nfeatures*32,
 * row i belongs to kps[i].  The mask argument of the reference is accepted and ignored by its detection code
 * (cellMask is dead, :601-603), so it has no counterpart here.  Not re-entrant per handle (like the reference,
 * whose pyramid is a member, ORBextractor.h:74-75). */
int orbx_extract(orbx_handle* h, const uint8_t* img, int width, int height, size_t stride, orbx_keypoint* kps,
                 uint8_t* desc, int* n_out);

/* Batch of independent frames, HOST buffers: frame f starts at imgs + f*frame_stride; its outputs are
 * kps + f*nfeatures, desc + f*nfeatures*32, n_out[f].  H2D / kernels / D2H of successive chunks overlap. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs, int nframes, int width, int height, size_t stride,
                       size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int* n_out);

/* Same with DEVICE buffers (inputs already resident in HBM, outputs left in HBM), asynchronous on `stream`
 * (a cudaStream_t; NULL = the handle's own stream).  nframes may exceed max_batch. */
int orbx_extract_batch_device(orbx_handle* h, const uint8_t* d_imgs, int nframes, int width, int height,
                              size_t stride, size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc,
                              int* d_n_out, void* stream);
/* Waits for the handle's streams and reports deferred device-side errors (list overflow). */
int orbx_synchronize(orbx_handle* h);

/* Pinned (page-locked) host memory for the buffers handed to orbx_extract_batch, for hosts that do not link a CUDA
 * runtime of their own: copies from / to pinned memory are asynchronous and run at the link's speed.  write_combined != 0
 * asks for write-combined pages (fast for the DMA engine to read, very slow for the CPU to read back: input frames only).
 * No counterpart in the reference. */
int orbx_host_alloc(size_t bytes, int write_combined, void** out);
void orbx_host_free(void* p);

/* ---- measurement hooks (bench.py) ---------------------------------------------------------------------------- */
enum { ORBX_STAGE_PYRAMID = 0, ORBX_STAGE_FAST, ORBX_STAGE_SELECT, ORBX_STAGE_BLUR, ORBX_STAGE_DESCRIBE,
       ORBX_STAGE_COUNT };
/* When enabled, every chunk records CUDA events between stages on the launching stream. */
int orbx_profile_enable(orbx_handle* h, int on);
/* Accumulated per-stage milliseconds and launch counts since the last reset; resets the accumulators. */
int orbx_profile_read(orbx_handle* h, float ms[ORBX_STAGE_COUNT], int64_t launches[ORBX_STAGE_COUNT]);
/* Total kernels launched by this handle since creation. */
int64_t orbx_kernel_launches(const orbx_handle* h);

/* ---- parity introspection (tests): state of frame `frame` of the LAST chunk processed ------------------------- */
int orbx_debug_level_info(orbx_handle* h, int level, int32_t out8[8]); /* w,h,cols,rows,cellW,cellH,quota,nDesired */
/* which: 0 = unblurred level, 1 = blurred level; out: w*h bytes, contiguous */
int orbx_debug_level_image(orbx_handle* h, int frame, int level, int which, uint8_t* out);
/* per-cell FAST candidates before retention (level coords): count returned through *n (may exceed cap) */
int orbx_debug_cell_candidates(orbx_handle* h, int frame, int level, int cell, int32_t* xs, int32_t* ys,
                               float* resp, int cap, int* n);
int orbx_debug_level_cells(orbx_handle* h, int frame, int level, int32_t* n_total, int32_t* n_retain);
/* KeyPointsFilter::retainBest(list, keep[i]) + resize (src/ORBextractor.cc:683-685, 697-701) through the device port of
 * libstdc++'s std::nth_element, on nlists caller-made response lists (list i = resp[start[i] .. start[i+1])).
 * order[k] = original position, inside its list, of the record now at place k; the first min(keep, size) places of a
 * list are what the reference keeps.  batch: 0 = the scans of select_kernel, 1 = those of select_smem_kernel. */
int orbx_debug_retain_best(orbx_handle* h, const float* resp, const int32_t* start, const int32_t* keep, int nlists,
                           int retain_mode, int batch, int32_t* order);

/* ---- synthetic frames (bench utility; byte-identical CPU twin in oracle/synth.cc) ------------------------------ */
/* Frame f gets seed seed0+f.  Device buffer, asynchronous on `stream`. */
int orbx_synth_frames_device(uint8_t* d_out, int nframes, int width, int height, size_t stride, size_t frame_stride,
                             uint64_t seed0, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
