/*
 * rpe.h — C ABI of the B200-native two-view relative-pose engine (librpe_b200.so).
 *
 * This is the drop-in boundary for the reference's hot path
 *   ORB -> BFMatcher(Hamming) -> findEssentialMat(5-pt RANSAC) -> recoverPose -> Euler
 * i.e. the body of PoseEstimator.estimate() (reference: src/core/pose_estimator.py:487-569)
 * and the four OpenCV calls it makes (:85-91/:108, :131/:144-151, :522-527, :533), plus
 * rotation_to_euler (reference: src/utils/geometry.py:14-27).
 *
 * Plain pointers and sizes only; no exceptions cross the ABI; every function returns
 * RPE_OK (0) or a negative rpe_error.  One handle per GPU; a handle is not thread-safe.
 * There is NO CPU fallback: rpe_create fails with RPE_ERR_NO_DEVICE when no sm_100 GPU
 * is usable.
 */
#ifndef RPE_H_
#define RPE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RPE_ABI_VERSION 1
#define RPE_MAX_LEVELS 16
#define RPE_MAX_DIM 4095 /* keypoint coordinates are packed in 12 bits */

typedef enum rpe_error {
  RPE_OK = 0,
  RPE_ERR_INVALID = -1,   /* bad argument / config */
  RPE_ERR_NO_DEVICE = -2, /* no CUDA device, or not sm_100 */
  RPE_ERR_CUDA = -3,      /* CUDA runtime error (see rpe_last_error) */
  RPE_ERR_CAPACITY = -4   /* batch / image larger than the handle was created for */
} rpe_error;

/* Per-pair status, the batched form of the reference's RuntimeErrors
 * (reference: src/core/pose_estimator.py:508-509, :514-515, :529-530). */
typedef enum rpe_status {
  RPE_STATUS_OK = 0,
  RPE_STATUS_NO_DESCRIPTORS = 1,       /* "Could not compute descriptors for one of the images." */
  RPE_STATUS_INSUFFICIENT_MATCHES = 2, /* "Insufficient matches: n (minimum 5 required)" */
  RPE_STATUS_E_FAILED = 3              /* "Could not estimate Essential matrix." (also M == 5) */
} rpe_status;
#define RPE_STATUS_CODE(s) ((s) & 0xff)
#define RPE_STATUS_FLAG_OVERFLOW 0x100 /* a candidate / tie list hit its capacity (documented limit) */

enum { RPE_MATCH_CROSSCHECK = 0, RPE_MATCH_RATIO = 1 };
enum { RPE_EULER_YUP = 0, RPE_EULER_ZYX = 1 };
/* The reference's two colour -> gray paths (they differ by +-1 on about a third of the pixels):
 * cv2.imread + cv2.cvtColor(BGR2GRAY) (reference: src/utils/image_loader.py:23-28) and
 * cv2.imread(path, IMREAD_GRAYSCALE) of a colour PNG (reference: src/run_single_pair.py:31-32). */
enum { RPE_GRAY_CVTCOLOR = 0, RPE_GRAY_IMREAD_PNG = 1 };

/* Mirrors the constructor arguments and hard-coded constants of the reference
 * (reference: src/core/pose_estimator.py:19-32 ctor, :85-91 ORB, :131 matcher, :150 cut,
 *  :522-527 RANSAC).  rpe_default_config() fills the reference's values. */
typedef struct rpe_config {
  int32_t max_width;        /* workspace sizing: largest image the handle accepts */
  int32_t max_height;
  int32_t max_pairs;        /* largest batch (pairs) per call */
  int32_t nfeatures;        /* 4000 */
  int32_t nlevels;          /* 12 */
  float scale_factor;       /* 1.1f (a C float, as cv2.ORB_create takes it) */
  int32_t fast_threshold;   /* 15 */
  int32_t match_mode;       /* RPE_MATCH_CROSSCHECK (reference) | RPE_MATCH_RATIO (configs 3-5) */
  double ratio;             /* 0.75 (a Python float in the reference-style ratio test) */
  int32_t max_matches;      /* 500; <= 0 means "None" (no cut) */
  double ransac_prob;       /* 0.999 */
  double ransac_threshold;  /* 1.0 px */
  int32_t ransac_max_iters; /* 1000 (cv2 default) */
  int32_t euler_convention; /* RPE_EULER_YUP | RPE_EULER_ZYX */
} rpe_config;

typedef struct rpe_handle rpe_handle;

int rpe_abi_version(void);
void rpe_default_config(rpe_config* cfg);
const char* rpe_last_error(void);

/* replaces PoseEstimator.__init__ (reference: src/core/pose_estimator.py:19-69) */
int rpe_create(const rpe_config* cfg, int device, rpe_handle** out);
void rpe_destroy(rpe_handle* h);
/* run all work of this handle on an existing CUDA stream (cudaStream_t); NULL = own stream */
int rpe_set_stream(rpe_handle* h, void* cuda_stream);
/* Scheduling of one handle: the dense kernels (ORB, matcher) run in order on the handle's stream; image
 * uploads run ahead on an internal copy stream, and the latency-bound geometry tail (RANSAC, recoverPose,
 * result D2H) on an internal high-priority stream, so they overlap the dense kernels of the next chunk or
 * of the next submitted batch.  rpe_set_pipeline: n_streams = 1: a batch is one chunk (right when batches
 * are submitted back to back: they overlap each other); n_streams >= 2: a batch is cut into chunks of
 * `chunk_pairs` pairs, so that inside ONE blocking call the upload of chunk c+1 and the tail of chunk c
 * hide under the dense kernels.  Default: 32 pairs, chunked.  Results do not depend on these settings. */
int rpe_set_pipeline(rpe_handle* h, int32_t chunk_pairs, int32_t n_streams);
/* pixel format of the images given to the batch / sequence calls (host or device): channels = 1 (default):
 * H x W uint8 gray; channels = 3: H x W x 3 interleaved BGR as cv2.imread returns it, converted to gray on the
 * device with the arithmetic of `gray_mode` (RPE_GRAY_*) — replaces load_image's cvtColor
 * (reference: src/utils/image_loader.py:28).  `pitch` is then the byte pitch of a BGR row (>= 3 W). */
int rpe_set_input_format(rpe_handle* h, int32_t channels, int32_t gray_mode);
/* capacities derived from the config: max keypoints per image, max matches per pair */
int rpe_get_capacity(const rpe_handle* h, int32_t* kp_cap, int32_t* match_cap);

/* replaces PoseEstimator.estimate over a batch (reference: src/core/pose_estimator.py:487-569).
 * img1[i] / img2[i]: H x W uint8 gray, row pitch `pitch` bytes, HOST memory (copied H2D
 * inside the call).  Outputs (host, caller-owned): R[B*9] row-major, t[B*3], euler[B*3] =
 * (yaw, pitch, roll) degrees, status[B]; info[B*4] (optional, may be NULL) =
 * {num_matches, recoverPose count, RANSAC inliers, RANSAC iterations}.  Blocks until the
 * results are in host memory. */
int rpe_estimate_batch(rpe_handle* h, const uint8_t* const* img1, const uint8_t* const* img2,
                       int32_t B, int32_t H, int32_t W, size_t pitch, const double K[9],
                       double* R, double* t, double* euler, int32_t* status, int32_t* info);

/* asynchronous form of rpe_estimate_batch: rpe_submit_batch enqueues the uploads and kernels and
 * returns without waiting; rpe_wait_batch blocks until the results of the submitted batch are in the
 * output pointers given at submit time.  The images and the output arrays must stay valid until
 * rpe_wait_batch returns.  TWO batches may be pending per handle (a third submit, and every other call on
 * the handle, fails with RPE_ERR_INVALID meanwhile); rpe_wait_batch waits for the OLDEST one.  Submitting
 * batch n+1 before waiting for batch n keeps the GPU full: upload and ORB / matcher kernels of n+1 run
 * under the RANSAC / recoverPose tail of n (see bench.py). */
int rpe_submit_batch(rpe_handle* h, const uint8_t* const* img1, const uint8_t* const* img2,
                     int32_t B, int32_t H, int32_t W, size_t pitch, const double K[9],
                     double* R, double* t, double* euler, int32_t* status, int32_t* info);
int rpe_submit_batch_device(rpe_handle* h, const uint8_t* d_images, size_t img_stride,
                            int32_t B, int32_t H, int32_t W, size_t pitch, const double K[9],
                            double* R, double* t, double* euler, int32_t* status, int32_t* info);
int rpe_wait_batch(rpe_handle* h);

/* same, inputs already resident in device memory: d_images holds 2*B images, image 2*i is
 * view 1 of pair i and 2*i+1 is view 2, each H rows of `pitch` bytes, consecutive images
 * `img_stride` bytes apart.  Results are copied to the host pointers as above.
 * Gray frames whose layout already is the pyramid's (pitch == (W + 15) / 16 * 16, base and img_stride multiples
 * of 16) are read IN PLACE as pyramid level 0 (no copy); any other layout is copied first.  Either way the frames
 * must stay valid and unchanged until the call (or, for the submit forms, rpe_wait_batch) returns. */
int rpe_estimate_batch_device(rpe_handle* h, const uint8_t* d_images, size_t img_stride,
                              int32_t B, int32_t H, int32_t W, size_t pitch, const double K[9],
                              double* R, double* t, double* euler, int32_t* status, int32_t* info);

/* Consecutive-frame form of the batch call (the reference's sequence driver,
 * src/core/batch_processor.py:71-92, pairs frame k with frame k+1 and runs ORB on BOTH frames of every
 * pair, src/core/pose_estimator.py:505-506): F frames -> F-1 pairs, pair p = (frames[p], frames[p+1]),
 * ORB once per frame (frame-level feature cache).  Outputs hold F-1 entries and are bit-identical to
 * rpe_estimate_batch on the same F-1 pairs.  F <= max_pairs + 1.  To chain calls over a longer
 * stream, start the next call at the previous call's last frame.  submit/wait as above. */
int rpe_estimate_sequence(rpe_handle* h, const uint8_t* const* frames, int32_t F, int32_t H, int32_t W,
                          size_t pitch, const double K[9], double* R, double* t, double* euler,
                          int32_t* status, int32_t* info);
int rpe_submit_sequence(rpe_handle* h, const uint8_t* const* frames, int32_t F, int32_t H, int32_t W,
                        size_t pitch, const double K[9], double* R, double* t, double* euler,
                        int32_t* status, int32_t* info);
int rpe_estimate_sequence_device(rpe_handle* h, const uint8_t* d_frames, size_t img_stride, int32_t F,
                                 int32_t H, int32_t W, size_t pitch, const double K[9], double* R,
                                 double* t, double* euler, int32_t* status, int32_t* info);
int rpe_submit_sequence_device(rpe_handle* h, const uint8_t* d_frames, size_t img_stride, int32_t F,
                               int32_t H, int32_t W, size_t pitch, const double K[9], double* R,
                               double* t, double* euler, int32_t* status, int32_t* info);

/* ---- stage entry points (parity harness; same kernels as the batch path) ------------- */

/* replaces ORB.detectAndCompute(image, None) (reference: src/core/pose_estimator.py:98-109).
 * Keypoints come in canonical (octave, y_level, x_level) order.  Arrays must hold kp_cap
 * entries (desc: kp_cap*32 bytes).  Any output pointer may be NULL. */
int rpe_orb_detect(rpe_handle* h, const uint8_t* img, int32_t H, int32_t W, size_t pitch,
                   float* x, float* y, float* size, float* angle, float* response,
                   int32_t* octave, uint8_t* desc, int32_t* n_out, int32_t* flags_out);

/* replaces cv2.cvtColor(img, COLOR_BGR2GRAY) / the decoder-side conversion of IMREAD_GRAYSCALE (see
 * RPE_GRAY_*): bgr = H x W x 3 host image, gray = H x W tight host output (stage entry for parity). */
int rpe_bgr_to_gray(rpe_handle* h, const uint8_t* bgr, int32_t H, int32_t W, size_t pitch, int32_t gray_mode,
                    uint8_t* gray);

/* replaces _match_descriptors (reference: src/core/pose_estimator.py:133-153): matcher in the
 * handle's match_mode, stable sort by distance, max_matches cut.  matches: m_cap*3 int32
 * rows (queryIdx, trainIdx, distance). */
int rpe_match(rpe_handle* h, const uint8_t* desc1, int32_t n1, const uint8_t* desc2, int32_t n2,
              int32_t* matches, int32_t* m_out);

/* replaces cv2.findEssentialMat(pts1, pts2, K, RANSAC, prob, thr, maxIters)
 * (reference: src/core/pose_estimator.py:522-527).  pts: m x 2 float32.  E: 9 doubles,
 * mask: m bytes {0,1}; info[4] = {ok, best inlier count, iterations run, hypotheses solved}. */
int rpe_find_essential(rpe_handle* h, const float* pts1, const float* pts2, int32_t m,
                       const double K[9], double* E, uint8_t* mask, int32_t* info);

/* per-hypothesis trace of the last rpe_find_essential call: for hypothesis i < n_hyp,
 * subset[i*5..], nmodels[i], counts[i*10..] (inlier count per model), models[i*90..] */
int rpe_ransac_trace(rpe_handle* h, int32_t max_hyp, int32_t* subset, int32_t* nmodels,
                     int32_t* counts, double* models, int32_t* n_hyp);

/* replaces cv2.recoverPose(E, pts1, pts2, K) (reference: src/core/pose_estimator.py:533)
 * and rotation_to_euler (reference: src/utils/geometry.py:14-27).  mask: m bytes {0,255}. */
int rpe_recover_pose(rpe_handle* h, const double E[9], const float* pts1, const float* pts2,
                     int32_t m, const double K[9], double* R, double* t, double* euler,
                     int32_t* count, uint8_t* mask);

/* the geometry half of the path over a batch of pairs whose matched points are already known: replaces
 * cv2.findEssentialMat(pts1, pts2, K, RANSAC, prob, thr, maxIters) + cv2.recoverPose(E, pts1, pts2, K) +
 * rotation_to_euler (reference: src/core/pose_estimator.py:518-533, src/utils/geometry.py:14-27) for B pairs in one
 * launch sequence.  pts1 / pts2: B x cap x 2 float32 (pair p at element offset p*cap*2), counts[B] = matches of
 * pair p (<= cap <= match capacity).  Outputs as rpe_estimate_batch; a pair with fewer than 5 matches gets
 * RPE_STATUS_INSUFFICIENT_MATCHES.  Blocking.  rpe_get_stage (E, masks) then refers to this call. */
int rpe_estimate_from_matches(rpe_handle* h, const float* pts1, const float* pts2, const int32_t* counts,
                              int32_t B, int32_t cap, const double K[9], double* R, double* t, double* euler,
                              int32_t* status, int32_t* info);

/* replaces PoseEvaluator.evaluate_sequence + compute_summary_statistics (reference:
 * src/core/pose_evaluator.py:32-134, :136-157) with rotation_error / translation_direction_error /
 * euler_to_rotation (reference: src/utils/geometry.py:128-174, :82-125, :208-236), on the device.
 * Host inputs, n frames: est_R[n*9] (world rotations as BatchProcessor composes them), est_t[n*3],
 * est_euler[n*3] and gt_euler[n*3] = (yaw, pitch, roll) degrees, gt_pos[n*3] = (x, y, z).
 * errors[n*5] = roll, pitch, yaw (wrapped to [0, 180]), rotation, translation-direction error in degrees (the
 * last against gt_pos[i] - gt_pos[i-1]; 0 for i = 0); summary[5*5] = per metric, in that order: mean, std
 * (population), median, max, min.  Either output may be NULL. */
int rpe_evaluate_sequence(rpe_handle* h, int32_t n, const double* est_R, const double* est_t,
                          const double* est_euler, const double* gt_euler, const double* gt_pos,
                          int32_t euler_convention, double* errors, double* summary);

/* the summary half alone (compute_summary_statistics, reference: src/core/pose_evaluator.py:136-157) over a
 * host errors[n*5] table laid out as above */
int rpe_summarize_errors(rpe_handle* h, int32_t n, const double* errors, double* summary);

/* NO COUNTERPART IN THE REFERENCE — added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * rpe_evaluate_sequence without any allocation and without a handle: same kernels, same arithmetic, but the
 * per-frame error table (the scratch of that call) lives in a file-scope static __device__ array of
 * RPE_NOALLOC_MAX_FRAMES x 5 doubles (640 KiB, held by every GPU that loads the library until the process exits).
 * Runs on the CURRENT device; ALL pointers are DEVICE pointers owned by the caller: inputs laid out as for
 * rpe_evaluate_sequence, d_summary[25] as its `summary`.  The work is enqueued on `cuda_stream` (cudaStream_t; NULL =
 * the default stream) and the call returns without synchronising; it calls no cudaMalloc / cudaFree / allocating
 * library routine.  n > RPE_NOALLOC_MAX_FRAMES -> RPE_ERR_CAPACITY (rpe_last_error names the limit); host pointers
 * -> RPE_ERR_INVALID.  Every call on a device shares the one array (it keeps the last call's table; nothing in it
 * needs clearing, rows beyond n are never read): calls that overlap on the GPU — different streams without a
 * dependency, or two threads — race; serialise them on one stream or with events.  The first call on a device
 * resolves the array's address (cudaGetSymbolAddress): make it outside a stream capture. */
#define RPE_NOALLOC_MAX_FRAMES 16384
int rpe_evaluate_sequence_noalloc(int32_t n, const double* d_est_R, const double* d_est_t, const double* d_est_euler,
                                  const double* d_gt_euler, const double* d_gt_pos, int32_t euler_convention,
                                  double* d_summary, void* cuda_stream);

/* intermediate results of the LAST batch call, for stage-wise parity:
 * copies up to *n bytes to `out` and stores the byte count in *n.  `view` selects frame pair + view of a
 * sequence call (image slot pair + view) or view 1 / 2 of a pair-batch call (slot 2*pair + view). */
enum {
  RPE_STAGE_PYRAMID_LEVEL = 0, /* arg = level; tight w x h u8 */
  RPE_STAGE_BLURRED_LEVEL = 1,
  RPE_STAGE_KP_X = 2,          /* float[n]  */
  RPE_STAGE_KP_Y = 3,
  RPE_STAGE_KP_SIZE = 4,
  RPE_STAGE_KP_ANGLE = 5,
  RPE_STAGE_KP_RESPONSE = 6,
  RPE_STAGE_KP_OCTAVE = 7,     /* int32[n] */
  RPE_STAGE_DESCRIPTORS = 8,   /* uint8[n*32] */
  RPE_STAGE_MATCHES = 9,       /* int32[m*3] (q, t, dist); view ignored */
  RPE_STAGE_PTS1 = 10,         /* float[m*2] */
  RPE_STAGE_PTS2 = 11,
  RPE_STAGE_E = 12,            /* double[9] */
  RPE_STAGE_RANSAC_MASK = 13,  /* uint8[m] {0,1} */
  RPE_STAGE_POSE_MASK = 14     /* uint8[m] {0,255} */
};
int rpe_get_stage(rpe_handle* h, int32_t pair, int32_t view, int32_t stage, int32_t arg,
                  void* out, size_t* n);

/* per-stage device time of the last batch call in ms (CUDA events on the handle's stream);
 * enabled with rpe_set_profiling(h, 1).  Order: see RPE_T_* */
enum {
  RPE_T_H2D = 0, RPE_T_PYRAMID, RPE_T_FAST, RPE_T_SELECT, RPE_T_BLUR, RPE_T_DESCRIBE,
  RPE_T_MATCH, RPE_T_MATCH_SORT, RPE_T_RANSAC, RPE_T_POSE, RPE_T_D2H, RPE_T_COUNT
};
int rpe_set_profiling(rpe_handle* h, int32_t on);
int rpe_get_stage_times(rpe_handle* h, float* ms /* RPE_T_COUNT */, int32_t* launches /* RPE_T_COUNT */);

#ifdef __cplusplus
}
#endif
#endif /* RPE_H_ */
