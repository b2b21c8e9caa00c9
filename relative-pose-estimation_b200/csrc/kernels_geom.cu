// kernels_geom.cu — five-point essential-matrix RANSAC and recoverPose on sm_100a (FP64).
// Replaces cv2.findEssentialMat(pts1, pts2, K, RANSAC, prob, thr, maxIters) and
// cv2.recoverPose(E, pts1, pts2, K) as the reference calls them
// (src/core/pose_estimator.py:522-527, :533) plus rotation_to_euler (src/utils/geometry.py:14-27).
//
// k_ransac: one CTA per pair.  OpenCV's loop is sequential (strict ">" best update, adaptive
// iteration count); here hypotheses are generated in chunks of kChunk from the SAME RNG stream,
// solved one per thread, scored one model per warp (Sampson error over all M points), and the
// sequential update rule is then replayed over the chunk's per-model inlier counts, so the CTA
// stops at exactly the iteration where OpenCV would have stopped and holds the model it held.
#include "rpe_device.cuh"
#include "geom_math.cuh"

namespace rpe {

__global__ void __launch_bounds__(256)
k_normalize(const float* __restrict__ pts1, const float* __restrict__ pts2, const int32_t* __restrict__ m_n,
            int mcap, const double* __restrict__ K, double* __restrict__ npts) {
  const int pair = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m_n[pair]) return;
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  const size_t o = (size_t)pair * mcap + i;
  double4 v;
  v.x = ((double)pts1[2 * o] - cx) / fx;
  v.y = ((double)pts1[2 * o + 1] - cy) / fy;
  v.z = ((double)pts2[2 * o] - cx) / fx;
  v.w = ((double)pts2[2 * o + 1] - cy) / fy;
  reinterpret_cast<double4*>(npts)[o] = v;
}

void launch_normalize(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, const double* dK,
                      cudaStream_t s, int* launches) {
  dim3 grid((mcap + 255) / 256, n_pairs);
  k_normalize<<<grid, 256, 0, s>>>(mb.pts1, mb.pts2, mb.m_n, mcap, dK, gb.npts);
  ++*launches;
}

// Hypotheses per round.  The five-point solver keeps 2.5 KB of per-thread scratch in local memory (the 10 x 20
// elimination matrix), interleaved by lane: 16 solver threads in the lower half-warp touch 40 KB, which stays in L1
// next to the dense kernels sharing the SM; 24 and more spill out of it (B200, 128 easy pairs per step: 0.38 / 0.44 /
// 1.08 / 1.11 / 1.24 / 3.2 ms for rounds of 8 / 16 / 24 / 32 / 64 / 128), and small rounds solve fewer hypotheses
// past the adaptive stop.  A round costs one solver latency (~0.3 ms) whatever its size, though, so a pair that still
// needs many iterations (low inlier ratio: the reference's simulator frames run to the 1000-iteration cap) takes
// rounds of kChunkMax instead: 58 such pairs 22 ms -> 10 ms.  The result does not depend on the round sizes (same
// RNG stream, exactly the drawn subsets consumed, sequential update rule replayed per round).
#ifndef RPE_RANSAC_CHUNK
#define RPE_RANSAC_CHUNK 16
#endif
constexpr int kChunkSmall = RPE_RANSAC_CHUNK;   // first round, and while few iterations remain
constexpr int kChunkMax = 64;                   // later rounds while more than kChunkSwitch iterations remain
#ifndef RPE_RANSAC_SWITCH
#define RPE_RANSAC_SWITCH 48
#endif
constexpr int kChunkSwitch = RPE_RANSAC_SWITCH;
constexpr int kChunk = kChunkMax;               // array sizes
static_assert(kChunkSmall <= kChunkMax, "round sizes");
#ifndef RPE_RANSAC_THREADS
#define RPE_RANSAC_THREADS 128
#endif
constexpr int kRansacThreads = RPE_RANSAC_THREADS;
static_assert(kRansacThreads >= kChunk && kRansacThreads % 32 == 0, "one solver thread per hypothesis of a round");

struct RansacShared {
  double models[kChunk][10][9];
  int nmodels[kChunk];
  int counts[kChunk][10];
  int subset[kChunk][5];
  double best_model[9];
  unsigned long long rng_state;
  int best, niters, it0, done, have_best, nh;
};

__global__ void __launch_bounds__(kRansacThreads)
k_ransac(const int32_t* __restrict__ m_n, int mcap, const double* __restrict__ npts,
         const double* __restrict__ K, RansacParams rp, double* __restrict__ E_out,
         uint8_t* __restrict__ rmask, int32_t* __restrict__ rinfo, int32_t* tr_subset, int32_t* tr_nmodels,
         int32_t* tr_counts, double* tr_models, int tr_cap) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  RansacShared& S = *reinterpret_cast<RansacShared*>(s_raw);
  const int pair = blockIdx.x, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int M = m_n[pair];
  const double4* P = reinterpret_cast<const double4*>(npts) + (size_t)pair * mcap;
  const bool trace = (tr_subset != nullptr) && pair == 0;
  if (M < 5) {   // run() returns false; the host maps this to INSUFFICIENT_MATCHES / E_FAILED
    if (tid < 9) E_out[pair * 9 + tid] = 0.0;
    if (tid < 4) rinfo[pair * 4 + tid] = 0;
    return;
  }
  const double thr = rp.threshold / ((K[0] + K[4]) * 0.5);
  const float t2 = (float)(thr * thr);
  if (tid == 0) {
    S.rng_state = 0xFFFFFFFFFFFFFFFFull;
    S.best = 0; S.niters = rp.max_iters; S.it0 = 0; S.done = 0; S.have_best = 0;
  }
  __syncthreads();
  int solved = 0;
  while (true) {
    // (a) subsets for this chunk, sequentially from the shared RNG stream
    if (tid == 0) {
      const int remaining = S.niters - S.it0;
      // the first round is always small: it is what brings niters down from max_iters for an easy pair
      const int nh_round = min((S.it0 > 0 && remaining > kChunkSwitch) ? kChunkMax : kChunkSmall, remaining);
      S.nh = nh_round;                              // hypotheses that can still matter
      CvRng rng{S.rng_state};
      for (int h = 0; h < nh_round; ++h) {
        if (M == 5) { for (int k = 0; k < 5; ++k) S.subset[h][k] = k; }
        else draw_subset5(rng, M, S.subset[h]);
      }
      S.rng_state = rng.state;
    }
    __syncthreads();
    const int it0 = S.it0;
    const int nh = S.nh;
    // (b) minimal solver, one hypothesis per thread
    const int hyp = tid;
    if (hyp < nh) {
      double x1[10], x2[10];
      for (int k = 0; k < 5; ++k) {
        const double4 p = P[S.subset[hyp][k]];
        x1[2 * k] = p.x; x1[2 * k + 1] = p.y; x2[2 * k] = p.z; x2[2 * k + 1] = p.w;
      }
      S.nmodels[hyp] = five_point_solve(x1, x2, &S.models[hyp][0][0]);
    }
    __syncthreads();
    solved += nh;
    // (c) Sampson scoring, one model per warp
    for (int mi = warp; mi < nh * 10; mi += kRansacThreads / 32) {
      const int h = mi / 10, r = mi % 10;
      if (r >= S.nmodels[h]) continue;
      const double* E = S.models[h][r];
      int cnt = 0;
      for (int i = lane; i < M; i += 32) {
        const double4 p = P[i];
        cnt += (sampson_err_f32(E, p.x, p.y, p.z, p.w) <= t2) ? 1 : 0;
      }
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0) S.counts[h][r] = cnt;
    }
    __syncthreads();
    // (d) replay OpenCV's sequential update rule over the chunk
    if (tid == 0) {
      int h = 0;
      for (; h < nh; ++h) {
        if (it0 + h >= S.niters) break;
        for (int r = 0; r < S.nmodels[h]; ++r) {
          const int c = S.counts[h][r];
          if (c > max(S.best, 4)) {
            S.best = c;
            S.have_best = 1;
            for (int k = 0; k < 9; ++k) S.best_model[k] = S.models[h][r][k];
            S.niters = ransac_update_iters(rp.prob, (double)(M - c) / M, 5, S.niters);
          }
        }
      }
      S.it0 = it0 + nh;
      if (S.it0 >= S.niters || M == 5) S.done = 1;
    }
    if (trace) {
      for (int i = tid; i < nh; i += kRansacThreads) {
        const int g = it0 + i;
        if (g < tr_cap) {
          tr_nmodels[g] = S.nmodels[i];
          for (int k = 0; k < 5; ++k) tr_subset[g * 5 + k] = S.subset[i][k];
          for (int r = 0; r < 10; ++r) {
            tr_counts[g * 10 + r] = r < S.nmodels[i] ? S.counts[i][r] : -1;
            for (int k = 0; k < 9; ++k) tr_models[(size_t)g * 90 + r * 9 + k] = r < S.nmodels[i] ? S.models[i][r][k] : 0.0;
          }
        }
      }
    }
    __syncthreads();
    if (S.done) break;
  }
  // inlier mask of the held model
  const int ok = S.have_best && M > 5;   // M == 5: cv2 returns stacked roots -> E_FAILED
  for (int i = tid; i < M; i += kRansacThreads) {
    uint8_t v = 0;
    if (S.have_best) {
      const double4 p = P[i];
      v = sampson_err_f32(S.best_model, p.x, p.y, p.z, p.w) <= t2;
    }
    rmask[(size_t)pair * mcap + i] = v;
  }
  if (tid < 9) E_out[pair * 9 + tid] = S.have_best ? S.best_model[tid] : 0.0;
  if (tid == 0) {
    rinfo[pair * 4 + 0] = ok;
    rinfo[pair * 4 + 1] = S.best;
    rinfo[pair * 4 + 2] = min(S.it0, S.niters);
    rinfo[pair * 4 + 3] = solved;
  }
}

void launch_ransac(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, const double* dK,
                   RansacParams rp, cudaStream_t s, int* launches) {
  k_ransac<<<n_pairs, kRansacThreads, sizeof(RansacShared), s>>>(mb.m_n, mcap, gb.npts, dK, rp, gb.E, gb.rmask,
                                                                 gb.rinfo, gb.tr_subset, gb.tr_nmodels,
                                                                 gb.tr_counts, gb.tr_models, gb.tr_cap);
  ++*launches;
}

// ------------------------------------------------------------------ K12/K13: recoverPose + Euler
constexpr int kPoseThreads = 128;

__global__ void __launch_bounds__(kPoseThreads)
k_pose(const int32_t* __restrict__ m_n, int mcap, const double* __restrict__ npts, const double* __restrict__ E_in,
       const int32_t* __restrict__ rinfo, RansacParams rp, const int32_t* __restrict__ pstat,
       double* __restrict__ R_out, double* __restrict__ t_out,
       double* __restrict__ euler_out, int32_t* __restrict__ pose_cnt, uint8_t* __restrict__ pmask,
       int32_t* __restrict__ status) {
  __shared__ double sR[2][9];
  __shared__ double st[3];
  __shared__ int s_good[4];
  __shared__ int s_sel;
  const int pair = blockIdx.x, tid = threadIdx.x;
  const int M = m_n[pair];
  // status, mirroring the reference's RuntimeErrors
  int stat = RPE_STATUS_OK;
  const int ps = pstat != nullptr ? pstat[pair] : 0;     // written by k_build_matches
  if (ps & 1) stat = RPE_STATUS_NO_DESCRIPTORS;
  else if (M < 5) stat = RPE_STATUS_INSUFFICIENT_MATCHES;
  else if (rinfo != nullptr && rinfo[pair * 4] == 0) stat = RPE_STATUS_E_FAILED;
  const uint32_t fl = (uint32_t)ps & RPE_STATUS_FLAG_OVERFLOW;
  if (stat != RPE_STATUS_OK) {
    if (tid < 9) R_out[pair * 9 + tid] = 0.0;
    if (tid < 3) { t_out[pair * 3 + tid] = 0.0; euler_out[pair * 3 + tid] = 0.0; }
    if (tid == 0) { pose_cnt[pair] = 0; status[pair] = stat | (fl ? RPE_STATUS_FLAG_OVERFLOW : 0); }
    return;
  }
  if (tid == 0) {
    decompose_essential(E_in + pair * 9, sR[0], sR[1], st);
    s_good[0] = s_good[1] = s_good[2] = s_good[3] = 0;
  }
  __syncthreads();
  const double4* P = reinterpret_cast<const double4*>(npts) + (size_t)pair * mcap;
  uint8_t* pm = pmask + (size_t)pair * mcap;
  const double tp[3] = {st[0], st[1], st[2]}, tn[3] = {-st[0], -st[1], -st[2]};
  int local[4] = {0, 0, 0, 0};
  for (int i = tid; i < M; i += kPoseThreads) {
    const double4 p = P[i];
    uint8_t bits = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const bool ok = triangulate_cheirality(sR[c & 1], (c & 2) ? tn : tp, p.x, p.y, p.z, p.w, 50.0);
      bits |= (uint8_t)ok << c;
      local[c] += ok;
    }
    pm[i] = bits;
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    int v = local[c];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) atomicAdd(&s_good[c], v);
  }
  __syncthreads();
  if (tid == 0) {
    const int g1 = s_good[0], g2 = s_good[1], g3 = s_good[2], g4 = s_good[3];
    int sel;
    if (g1 >= g2 && g1 >= g3 && g1 >= g4) sel = 0;
    else if (g2 >= g1 && g2 >= g3 && g2 >= g4) sel = 1;
    else if (g3 >= g1 && g3 >= g2 && g3 >= g4) sel = 2;
    else sel = 3;
    s_sel = sel;
    const double* R = sR[sel & 1];
    const double sg = (sel & 2) ? -1.0 : 1.0;
    for (int k = 0; k < 9; ++k) R_out[pair * 9 + k] = R[k];
    for (int k = 0; k < 3; ++k) t_out[pair * 3 + k] = sg * st[k];
    rotation_to_euler(R, rp.euler_convention, euler_out + pair * 3);
    pose_cnt[pair] = s_good[sel];
    status[pair] = RPE_STATUS_OK | (fl ? RPE_STATUS_FLAG_OVERFLOW : 0);
  }
  __syncthreads();
  const int sel = s_sel;
  for (int i = tid; i < M; i += kPoseThreads) pm[i] = ((pm[i] >> sel) & 1) ? 255 : 0;
}

void launch_pose(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, RansacParams rp,
                 const int32_t* pstat, cudaStream_t s, int* launches) {
  k_pose<<<n_pairs, kPoseThreads, 0, s>>>(mb.m_n, mcap, gb.npts, gb.E, gb.rinfo, rp, pstat, gb.R, gb.t,
                                          gb.euler, gb.pose_cnt, gb.pmask, gb.status);
  ++*launches;
}

int setup_geom_attributes() {
  return (int)cudaFuncSetAttribute(k_ransac, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)sizeof(RansacShared));
}

}  // namespace rpe
