// kernels_eval.cu — on-device evaluation of an estimated pose sequence against ground truth.
// Replaces PoseEvaluator.evaluate_sequence + compute_summary_statistics
// (reference: src/core/pose_evaluator.py:32-134, :136-157, :187-198) and rotation_error /
// translation_direction_error / euler_to_rotation (reference: src/utils/geometry.py:128-174, :82-125, :208-236)
// for long streams (BASELINE config 5: 100 000 pairs), where the reference's per-frame Python loop is the cost.
//
//   k_eval_errors  : one thread per frame, FP64: wrapped |angle| errors, geodesic rotation error, translation
//                    direction error against the ground-truth position delta (0 for the first frame).
//   k_eval_summary : one CTA per metric: mean, population std, median (exact: 64-bit radix select on the
//                    orderable bit pattern, 8 passes of 8 bits), max, min.  A NaN anywhere in a metric makes all
//                    five of its statistics NaN, as numpy does.
#include <cstdio>
#include <mutex>

#include "rpe_device.cuh"

namespace rpe {

constexpr double kDegPerRad = 57.29577951308232;     // 180 / pi, as np.rad2deg
constexpr double kRadPerDeg = 0.017453292519943295;  // pi / 180, as np.deg2rad / math.radians

// ((e + 180) % 360) - 180 with Python's float modulo (sign of the divisor), then abs (pose_evaluator.py:197-198)
__device__ __forceinline__ double wrap_abs(double e) {
  double m = fmod(e + 180.0, 360.0);
  if (m < 0.0) m += 360.0;
  return fabs(m - 180.0);
}
__device__ __forceinline__ double clip1(double v) { return v < -1.0 ? -1.0 : (v > 1.0 ? 1.0 : v); }   // NaN passes through

// geometry.py:82-125 (yup: Ry(yaw) Rx(pitch) Rz(roll)) and :208-236 (zyx: Rz(yaw) Ry(pitch) Rx(roll))
__device__ void euler_to_rotation_dev(double yaw_deg, double pitch_deg, double roll_deg, int convention, double R[9]) {
  double sy, cy, sp, cp, sr, cr;
  sincos(yaw_deg * kRadPerDeg, &sy, &cy);
  sincos(pitch_deg * kRadPerDeg, &sp, &cp);
  sincos(roll_deg * kRadPerDeg, &sr, &cr);
  if (convention == RPE_EULER_ZYX) {
    R[0] = cy * cp; R[1] = cy * sp * sr - sy * cr; R[2] = cy * sp * cr + sy * sr;
    R[3] = sy * cp; R[4] = sy * sp * sr + cy * cr; R[5] = sy * sp * cr - cy * sr;
    R[6] = -sp;     R[7] = cp * sr;                R[8] = cp * cr;
  } else {
    // (Ry Rx) Rz with Ry Rx = [[cy, sy sp, sy cp], [0, cp, -sp], [-sy, cy sp, cy cp]]
    const double a01 = sy * sp, a02 = sy * cp, a21 = cy * sp, a22 = cy * cp;
    R[0] = cy * cr + a01 * sr;  R[1] = -cy * sr + a01 * cr; R[2] = a02;
    R[3] = cp * sr;             R[4] = cp * cr;             R[5] = -sp;
    R[6] = -sy * cr + a21 * sr; R[7] = sy * sr + a21 * cr;  R[8] = a22;
  }
}

// errors[i*5 + {0..4}] = roll, pitch, yaw, rotation, translation direction error of frame i (degrees)
__global__ void __launch_bounds__(256)
k_eval_errors(int n, const double* __restrict__ est_R, const double* __restrict__ est_t,
              const double* __restrict__ est_euler, const double* __restrict__ gt_euler,
              const double* __restrict__ gt_pos, int convention, double* __restrict__ errors) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double gy = gt_euler[3 * i], gp = gt_euler[3 * i + 1], gr = gt_euler[3 * i + 2];
  errors[5 * i + 0] = wrap_abs(fabs(est_euler[3 * i + 2] - gr));
  errors[5 * i + 1] = wrap_abs(fabs(est_euler[3 * i + 1] - gp));
  errors[5 * i + 2] = wrap_abs(fabs(est_euler[3 * i + 0] - gy));
  // rotation_error: acos(clip((trace(R_est R_gt^T) - 1) / 2)); trace(A B^T) = sum_ij A_ij B_ij
  double Rg[9];
  euler_to_rotation_dev(gy, gp, gr, convention, Rg);
  double tr = 0.0;
#pragma unroll
  for (int k = 0; k < 9; ++k) tr += est_R[9 * i + k] * Rg[k];
  errors[5 * i + 3] = acos(clip1((tr - 1.0) / 2.0)) * kDegPerRad;
  // translation_direction_error(est_t[i], gt_pos[i] - gt_pos[i-1]); first frame: 0.0 (pose_evaluator.py:112-118)
  double te = 0.0;
  if (i > 0) {
    const double e0 = est_t[3 * i], e1 = est_t[3 * i + 1], e2 = est_t[3 * i + 2];
    const double d0 = gt_pos[3 * i] - gt_pos[3 * i - 3], d1 = gt_pos[3 * i + 1] - gt_pos[3 * i - 2],
                 d2 = gt_pos[3 * i + 2] - gt_pos[3 * i - 1];
    const double ne = sqrt(e0 * e0 + e1 * e1 + e2 * e2), nd = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
    const double dot = (e0 / ne) * (d0 / nd) + (e1 / ne) * (d1 / nd) + (e2 / ne) * (d2 / nd);   // 0/0 -> NaN as numpy
    te = acos(clip1(dot)) * kDegPerRad;
  }
  errors[5 * i + 4] = te;
}

constexpr int kSumThreads = 1024;

__device__ __forceinline__ unsigned long long f64_orderable(double v) {
  const unsigned long long u = (unsigned long long)__double_as_longlong(v);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_from_orderable(unsigned long long k) {
  const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)u);
}

// block-wide reduction of (sum, min, max, nan flag); result valid in every thread
__device__ void block_reduce(double& s, double& mn, double& mx, int& nan, double* sh_s, double* sh_mn, double* sh_mx, int* sh_nan) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    nan |= __shfl_xor_sync(0xffffffffu, nan, o);
  }
  __syncthreads();
  if (lane == 0) { sh_s[warp] = s; sh_mn[warp] = mn; sh_mx[warp] = mx; sh_nan[warp] = nan; }
  __syncthreads();
  s = 0.0; mn = INFINITY; mx = -INFINITY; nan = 0;
  for (int w = 0; w < kSumThreads / 32; ++w) { s += sh_s[w]; mn = fmin(mn, sh_mn[w]); mx = fmax(mx, sh_mx[w]); nan |= sh_nan[w]; }
}

// k-th smallest (0-based) of the metric column by radix select on the orderable keys
__device__ double radix_select(const double* __restrict__ col, int n, int k, uint32_t* s_hist, unsigned long long* s_prefix, int* s_k) {
  if (threadIdx.x == 0) { *s_prefix = 0ull; *s_k = k; }
  for (int pass = 7; pass >= 0; --pass) {
    const int shift = pass * 8;
    const unsigned long long mask = pass == 7 ? 0ull : (~0ull << (shift + 8));
    if (threadIdx.x < 256) s_hist[threadIdx.x] = 0;
    __syncthreads();
    const unsigned long long prefix = *s_prefix;
    for (int i = threadIdx.x; i < n; i += kSumThreads) {
      const unsigned long long key = f64_orderable(col[5 * (size_t)i]);
      if ((key & mask) == prefix) atomicAdd(&s_hist[(key >> shift) & 0xff], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int kk = *s_k;
      int b = 0;
      for (; b < 255; ++b) {
        if (kk < (int)s_hist[b]) break;
        kk -= (int)s_hist[b];
      }
      *s_k = kk;
      *s_prefix = prefix | ((unsigned long long)b << shift);
    }
    __syncthreads();
  }
  return f64_from_orderable(*s_prefix);
}

// summary[m*5 + {0..4}] = mean, std, median, max, min of metric m (compute_summary_statistics order)
__global__ void __launch_bounds__(kSumThreads)
k_eval_summary(int n, const double* __restrict__ errors, double* __restrict__ summary) {
  __shared__ double sh_s[kSumThreads / 32], sh_mn[kSumThreads / 32], sh_mx[kSumThreads / 32];
  __shared__ int sh_nan[kSumThreads / 32];
  __shared__ uint32_t s_hist[256];
  __shared__ unsigned long long s_prefix;
  __shared__ int s_k;
  const int m = blockIdx.x;
  const double* col = errors + m;
  double s = 0.0, mn = INFINITY, mx = -INFINITY;
  int nan = 0;
  for (int i = threadIdx.x; i < n; i += kSumThreads) {
    const double v = col[5 * (size_t)i];
    s += v; mn = fmin(mn, v); mx = fmax(mx, v); nan |= (v != v);
  }
  block_reduce(s, mn, mx, nan, sh_s, sh_mn, sh_mx, sh_nan);
  const double mean = s / (double)n;
  double q = 0.0, d0 = 0.0, d1 = 0.0;
  int dn = 0;
  for (int i = threadIdx.x; i < n; i += kSumThreads) {
    const double d = col[5 * (size_t)i] - mean;
    q += d * d;
  }
  block_reduce(q, d0, d1, dn, sh_s, sh_mn, sh_mx, sh_nan);
  double med = 0.0;
  if (!nan) {
    const double lo = radix_select(col, n, (n - 1) / 2, s_hist, &s_prefix, &s_k);
    __syncthreads();
    const double hi = (n & 1) ? lo : radix_select(col, n, n / 2, s_hist, &s_prefix, &s_k);
    med = (n & 1) ? lo : 0.5 * (lo + hi);   // np.median: mean of the two middle values
  }
  if (threadIdx.x == 0) {
    const double qnan = __longlong_as_double(0x7ff8000000000000ll);
    summary[5 * m + 0] = mean;
    summary[5 * m + 1] = sqrt(q / (double)n);
    summary[5 * m + 2] = nan ? qnan : med;
    summary[5 * m + 3] = nan ? qnan : mx;
    summary[5 * m + 4] = nan ? qnan : mn;
  }
}

void launch_evaluate(int n, const double* est_R, const double* est_t, const double* est_euler, const double* gt_euler,
                     const double* gt_pos, int convention, double* errors, double* summary, cudaStream_t s) {
  k_eval_errors<<<(n + 255) / 256, 256, 0, s>>>(n, est_R, est_t, est_euler, gt_euler, gt_pos, convention, errors);
  k_eval_summary<<<5, kSumThreads, 0, s>>>(n, errors, summary);
}

void launch_eval_summary(int n, const double* errors, double* summary, cudaStream_t s) {
  k_eval_summary<<<5, kSumThreads, 0, s>>>(n, errors, summary);
}

}  // namespace rpe

// ---------------------------------------------------------------------------------------------------------------
// rpe_evaluate_sequence_noalloc: the evaluation above for callers that may not allocate (include/rpe.h).
// The per-frame error table [n][5] — the scratch rpe_evaluate_sequence keeps in a cudaMalloc'ed, grown-on-demand
// buffer — lives here in a file-scope device array instead.  Every GPU that loads the library holds these
// RPE_NOALLOC_MAX_FRAMES * 5 * 8 bytes (640 KiB) until the process exits.
//
// `g_noalloc_errors` below is, in HOST code, only the host-side stand-in of the symbol: its device address is asked
// for once per device (cudaGetSymbolAddress with that device current) and that pointer is what the kernels get.
// All calls on one device share the array; it keeps the previous call's table.  Nothing in it has to start at
// zero: k_eval_errors writes rows 0 .. n-1 before k_eval_summary (next on the same stream) reads exactly those
// rows, so rows >= n left by a longer earlier call are never read.  Two calls that overlap on the GPU (different
// streams, no dependency) race on the array.
static __device__ __align__(256) double g_noalloc_errors[(size_t)RPE_NOALLOC_MAX_FRAMES * 5];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
double* g_noalloc_ptr[kMaxDevices] = {};   // device address of g_noalloc_errors on device d (nullptr: not asked yet)

bool is_device_pointer(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
}  // namespace

extern "C" int rpe_evaluate_sequence_noalloc(int32_t n, const double* d_est_R, const double* d_est_t,
                                             const double* d_est_euler, const double* d_gt_euler,
                                             const double* d_gt_pos, int32_t euler_convention, double* d_summary,
                                             void* cuda_stream) {
  char msg[256];
  if (n < 1) return rpe::report_error(RPE_ERR_INVALID, "rpe_evaluate_sequence_noalloc: n must be >= 1");
  if (n > RPE_NOALLOC_MAX_FRAMES) {
    snprintf(msg, sizeof msg, "rpe_evaluate_sequence_noalloc: n = %d exceeds RPE_NOALLOC_MAX_FRAMES = %d (the static "
             "scratch is not grown; use rpe_evaluate_sequence for longer sequences)", (int)n, (int)RPE_NOALLOC_MAX_FRAMES);
    return rpe::report_error(RPE_ERR_CAPACITY, msg);
  }
  if (euler_convention != RPE_EULER_YUP && euler_convention != RPE_EULER_ZYX)
    return rpe::report_error(RPE_ERR_INVALID, "rpe_evaluate_sequence_noalloc: bad euler convention");
  const void* ptrs[6] = {d_est_R, d_est_t, d_est_euler, d_gt_euler, d_gt_pos, d_summary};
  for (const void* p : ptrs)
    if (!p || !is_device_pointer(p))
      return rpe::report_error(RPE_ERR_INVALID, "rpe_evaluate_sequence_noalloc: inputs and d_summary must be device pointers");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    snprintf(msg, sizeof msg, "rpe_evaluate_sequence_noalloc: cudaGetDevice -> %s (device %d)", cudaGetErrorString(e), dev);
    return rpe::report_error(RPE_ERR_NO_DEVICE, msg);
  }
  double* scratch;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {                 // set-up, once per device, that device current
      double* p = nullptr;
      e = cudaGetSymbolAddress((void**)&p, g_noalloc_errors);
      if (e != cudaSuccess || !p) {
        snprintf(msg, sizeof msg, "rpe_evaluate_sequence_noalloc: cudaGetSymbolAddress(g_noalloc_errors) on device %d -> %s",
                 dev, cudaGetErrorString(e));
        return rpe::report_error(RPE_ERR_CUDA, msg);
      }
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  rpe::launch_evaluate(n, d_est_R, d_est_t, d_est_euler, d_gt_euler, d_gt_pos, euler_convention, scratch, d_summary,
                       (cudaStream_t)cuda_stream);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(msg, sizeof msg, "rpe_evaluate_sequence_noalloc: launch -> %s", cudaGetErrorString(e));
    return rpe::report_error(RPE_ERR_CUDA, msg);
  }
  return RPE_OK;
}
