// rpe_device.cuh — device-side data layout shared by the kernels and the host engine.
//
// HBM layout (per handle; NI = 2 * max_pairs image slots, NP = max_pairs pair slots):
//   pyr / blur   u8  [NI][pyr_stride]      all pyramid levels of one image back to back, each
//                                          level 256-B aligned, row pitch = align16(w)
//   cand_key     u32 [NI][cand_stride]     FAST+NMS candidates per level (score<<24 | y<<12 | x)
//   cand_cnt     u32 [NI][L],  hist u32 [NI][L][256] (FAST-score histogram per level)
//   sel_key/resp u32/f32 [NI][cand_stride] survivors of the first retainBest + Harris response
//   lvl_key/resp u32/f32 [NI][kp_stride]   final per-level keypoints in (y, x) order; lvl_cnt [NI][L]
//   kp_*         f32/i32 [NI][kp_stride]   final keypoints, canonical (octave, y, x) order; kp_n [NI]
//   desc         u8  [NI][kp_stride][32]
//   fwd1/fwd2/bwd1 u32 [NP][kp_stride]     packed (dist<<20 | index) nearest / second nearest
//   m_q/m_t/m_d  i32 [NP][mcap], pts1/pts2 f32 [NP][mcap][2], npts f64 [NP][mcap][4], m_n, pstat [NP]
//   E f64 [NP][9], rmask/pmask u8 [NP][mcap], rinfo i32 [NP][4], R/t/euler f64, status i32 [NP]
// Everything from m_q down exists TWICE (set = batch sequence number & 1): the geometry tail of batch n
// (RANSAC, recoverPose, D2H; its own stream) runs under the ORB / matcher kernels of batch n + 1.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/rpe.h"

// Checked build (-DRPE_CHECKED=1; python relative-pose-estimation_b200/build.py with RPE_LIB_VARIANT=checked
// RPE_EXTRA_DEFS=-DRPE_CHECKED=1): device-side assertions on every hand-computed shared-memory / global index of the
// kernels.  compute-sanitizer is closed on the GPU pool (profiles/r02_compute_sanitizer_closed.txt), so the GPU test
// suite is run once against this variant instead; a violated assertion traps the kernel and fails the test.
#if defined(RPE_CHECKED) && defined(__CUDACC__)
#include <cassert>
#define RPE_ASSERT(c) assert(c)
#else
#define RPE_ASSERT(c) ((void)0)
#endif

namespace rpe {

struct LevelGeom {
  int w, h, pitch;
  int quota;
  float scale;
  uint32_t off;            // byte offset inside one image's pyramid slab
  int cand_off, cand_cap;  // slice of the per-image candidate arrays
  int kp_off, kp_cap;      // slice of the per-image per-level keypoint arrays
  int xtab, ytab;          // tap-table offsets for producing this level from the previous one
  int tile0, tiles_x, tiles_y;     // 64 x 32 tiles (blur)
  int ftile0, ftiles_x, ftiles_y;  // kFastTW x kFastTH tiles restricted to the rows/cols FAST needs
};

struct OrbGeom {
  int nlevels;
  int fast_threshold;
  int total_tiles;
  int total_ftiles;
  int total_bitems;        // blur work items (one warp each): 128 columns x two kBlurRows-row strips
  int cand_stride;
  int kp_stride;
  int sort_cap;            // >= max kp_cap (select kernel: kept-keypoint buffer)
  int sel_row_cap;         // rows of the tallest level (select kernel: per-row tables of the (y, x) ordering)
  int resize_span;         // most source rows any 8-row output strip needs (k_resize_tma stages them all)
  int resize_xspan;        // most source bytes (from the aligned start) any 128-column output group taps (k_resize_tma stages them)
  int resize_rows_strict;  // the first vertical tap row is strictly increasing over the output rows of every level (k_resize_tma's walk)
  int resize_strip_ok;     // every aligned group of 4 output columns taps <= 8 consecutive source bytes
  size_t pyr_stride;
  LevelGeom lv[RPE_MAX_LEVELS];
};

constexpr int kTileW = 64;
constexpr int kTileH = 32;
constexpr int kFastTW = 120;
constexpr int kFastTH = 14;     // tile rows; + 2 = 16 score rows = two bands of 8 (k_fast: one warp per tile)
// (strip height, 512 images 1280x720: 28 / 32 / 35 / 42 / 49 rows 2.08 / 2.20 / 2.06 / 2.30 / 2.17 ms: whole groups of 7 rows
// waste nothing, taller strips cost occupancy)
constexpr int kBlurRows = 35;   // blur work item = 128 columns x two strips of kBlurRows rows (one warp)
// Layout of a BLURRED level (k_blur writes it, k_describe gathers from it): vertical strips of kBlurStripW columns,
// each strip all h rows of 16 bytes, strip after strip.  A 128-byte line is then 8 rows x 16 columns instead of 128
// columns of one row, and the 32 rotated pattern points of one gather instruction fall into ~6 lines instead of ~11
// (tools/gather_bench.cu: 0.197 -> 0.138 ms for 320 000 keypoints; surface / texture reads of a block-linear array:
// 0.277 ms).  Level size is unchanged: ceil(w / 16) * 16 * h = pitch * h.
constexpr int kBlurStripLog = 4, kBlurStripW = 1 << kBlurStripLog;   // (8-column strips: k_describe the same, k_blur 6 % slower)
__host__ __device__ inline size_t blur_strip_offset(int x, int y, int h) {
  return (size_t)(x >> kBlurStripLog) * ((size_t)h * kBlurStripW) + (size_t)y * kBlurStripW + (size_t)(x & (kBlurStripW - 1));
}
constexpr int kFastOrigin = 24;  // FAST tiles start at x = y = 24 (a multiple of 4): only scores in [30, size-30) can matter

// Where the pyramid levels of image slot `img` live: level l >= 1 inside the handle's pyramid slab; level 0 there too
// (copied in by the upload), or - when the caller's device buffer already has the pyramid's row pitch and 16-byte
// alignment - in the caller's buffer itself (lvl0 != nullptr: no copy at all).
struct PyrView {
  uint8_t* pyr;
  const uint8_t* lvl0;
  size_t lvl0_stride;
#if defined(__CUDACC__)
  __device__ __forceinline__ const uint8_t* level(const OrbGeom& g, int img, int l) const {
    return (l == 0 && lvl0 != nullptr) ? lvl0 + (size_t)img * lvl0_stride : pyr + (size_t)img * g.pyr_stride + g.lv[l].off;
  }
#endif
};

struct OrbBuffers {
  uint8_t* pyr;
  const uint8_t* lvl0;     // see PyrView
  size_t lvl0_stride;
  uint8_t* blur;
  uint32_t* cand_key;
  uint32_t* cand_cnt;
  uint32_t* hist;
  uint32_t* sel_key;
  float* sel_resp;
  uint32_t* lvl_key;
  float* lvl_resp;
  uint32_t* lvl_cnt;
  uint32_t* img_flags;
  const int2* taps;
  const uint32_t* ftab;   // FAST tile table: level<<24 | tile_y<<12 | tile_x
  const uint32_t* btab;   // blur work items: level<<24 | strip<<12 | segment
  float* kp_x; float* kp_y; float* kp_size; float* kp_angle; float* kp_resp;
  int32_t* kp_oct;
  uint8_t* desc;
  int32_t* kp_n;
};

struct MatchBuffers {
  uint32_t* fwd1; uint32_t* fwd2; uint32_t* bwd1;
  int32_t* m_q; int32_t* m_t; int32_t* m_d; int32_t* m_n;
  float* pts1; float* pts2;
  int32_t* pstat;   // [NP] per pair: bit 0 = an image without descriptors, RPE_STATUS_FLAG_OVERFLOW = list overflow
};

// per-pair RANSAC state between the kernels of a super-round (kernels_geom.cu)
struct RansacState {
  unsigned long long rng;   // OpenCV's MWC generator, advanced exactly as cv2's loop advances it
  int best, niters, it0, done, have_best, nh, solved, overflow;   // nh = hypotheses of the super-round in flight
  double best_model[9];
};
// a model that can win OpenCV's sequential update: a prefix maximum of its chunk
struct RansacCand {
  int h, r, count, pad;     // hypothesis index inside the super-round, model index, inlier count
  double E[9];
};

struct GeomBuffers {
  double* npts;     // [NP][mcap][4] normalised (x1, y1, x2, y2)
  double* E;        // [NP][9]
  uint8_t* rmask;   // [NP][mcap]
  int32_t* rinfo;   // [NP][4] {ok, best count, iterations, hypotheses solved}
  double* R; double* t; double* euler;
  int32_t* pose_cnt;
  uint8_t* pmask;
  int32_t* status;
  RansacState* rstate;      // [NP]
  int32_t* rsubsets;        // [NP][kMaxChunks * kChunk][5]
  RansacCand* rcands;       // [NP][kMaxChunks][kCandCap]
  int32_t* rncand;          // [NP][kMaxChunks]
  double* rmodels;          // [NP][kMaxChunks * kChunk][10][9]  models of the super-round in flight
  int32_t* rnm;             // [NP][kMaxChunks * kChunk]          models per hypothesis
  // optional per-hypothesis trace of pair 0 (stage entry point only)
  int32_t* tr_subset; int32_t* tr_nmodels; int32_t* tr_counts; double* tr_models; int32_t tr_cap;
};

struct RansacParams {
  double prob;
  double threshold;   // pixels
  int max_iters;
  int euler_convention;
};

// packed nearest-neighbour key
__host__ __device__ inline uint32_t pack_nn(uint32_t dist, uint32_t idx) { return (dist << 20) | idx; }

// launchers (defined in the kernel .cu files)
// BGR (interleaved, 4-byte aligned rows) -> level 0 of n_img pyramid slabs; mode: RPE_GRAY_*
void launch_bgr2gray(const uint8_t* bgr, size_t src_stride, int src_pitch, uint8_t* pyr, size_t pyr_stride,
                     int dst_pitch, int w, int h, int n_img, int mode, cudaStream_t s, int* launches);
void launch_pyramid(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches);
void launch_fast(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches);
void launch_select(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches);
void launch_blur(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches);
void launch_describe(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches);
// pstep: image slots of pair p are pstep*p and pstep*p + 1 (2 = pair batch, 1 = consecutive frames)
void launch_match(const OrbBuffers& ob, const MatchBuffers& mb, int kp_stride, int nfeatures, int n_pairs, int pstep, int mode,
                  cudaStream_t s, int* launches);
void launch_build_matches(const OrbBuffers& ob, const MatchBuffers& mb, int kp_stride, int mcap,
                          int sort_cap, int n_pairs, int pstep, int mode, double ratio, int max_matches,
                          cudaStream_t s, int* launches);
void launch_normalize(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, const double* dK,
                      cudaStream_t s, int* launches);
void launch_ransac(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, const double* dK,
                   RansacParams rp, cudaStream_t s, int* launches);
void launch_pose(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, RansacParams rp,
                 const int32_t* pstat, cudaStream_t s, int* launches);
// sequence evaluation: per-frame errors [n][5] + summary [5][5], all device pointers
void launch_evaluate(int n, const double* est_R, const double* est_t, const double* est_euler, const double* gt_euler,
                     const double* gt_pos, int convention, double* errors, double* summary, cudaStream_t s);
void launch_eval_summary(int n, const double* errors, double* summary, cudaStream_t s);
int report_error(int code, const char* msg);   // sets rpe_last_error() of the calling thread, returns code (engine.cu)
size_t select_smem_bytes(int sort_cap, int row_cap);   // dynamic shared memory of k_select
int setup_orb_attributes(int select_smem);
// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the FUNCTION (per device), not of a handle: a second handle
// with a smaller configuration must not lower it under the first.  Raises the limit of `func` on the current device to
// at least `bytes` (never lowers it); returns 0 on success.  (engine.cu)
int raise_dynamic_smem_limit(const void* func, int bytes);
int setup_match_attributes(int sort_cap);
int setup_geom_attributes();
constexpr int kRansacSuperRoundCap = 8 * 64;   // kMaxChunks * kChunk (kernels_geom.cu)
constexpr int kRansacCandPerPair = 8 * 32;     // kMaxChunks * kCandCap

}  // namespace rpe
