// kernels_geom.cu — five-point essential-matrix RANSAC and recoverPose on sm_100a (FP64).
// Replaces cv2.findEssentialMat(pts1, pts2, K, RANSAC, prob, thr, maxIters) and
// cv2.recoverPose(E, pts1, pts2, K) as the reference calls them
// (src/core/pose_estimator.py:522-527, :533) plus rotation_to_euler (src/utils/geometry.py:14-27).
//
// RANSAC: hypotheses come from the SAME RNG stream as OpenCV's, are solved one per thread and scored one model per
// warp (Sampson error over all M points) by many CTAs per pair, and OpenCV's sequential update rule is then
// replayed over their inlier counts, so every pair stops at exactly the iteration where OpenCV would have stopped
// and holds the model it held (see k_ransac_* below).
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

// RANSAC over a batch, hypotheses in parallel ACROSS CTAs.  OpenCV's loop is sequential (strict ">" best update,
// adaptive iteration count); here it advances in super-rounds:
//   k_ransac_begin   thread per pair: state (cv2's RNG seed, niters = maxIters), subsets of the first, small round
//   k_ransac_solve   CTA per (chunk of kChunk hypotheses, pair): minimal solver, one hypothesis per thread -> models
//   k_ransac_score   CTA per (chunk, pair): Sampson scoring one model per warp, then the chunk's PREFIX MAXIMA (the
//                    models whose inlier count beats everything before them in the chunk and the best held at the
//                    start of the super-round): only those can ever win the sequential update, so only they
//                    (<= kCandCap, with their E) are handed to the replay
//   k_ransac_replay  thread per pair: cv2's sequential rule replayed over the chunks' prefix maxima in hypothesis
//                    order (stop where cv2 stops: it0 + h >= niters), then the next super-round's subsets from the
//                    pair's RNG stream
//   k_ransac_finish  CTA per pair: inlier mask of the held model, E, rinfo
// [solve, score, replay] is enqueued a fixed number of times (enough for maxIters); CTAs of finished pairs, and chunks past
// the iterations still needed, return at once.  Same RNG stream, exactly the drawn subsets consumed in order, same
// update rule => the same stop iteration and the same held model as the one-CTA-per-pair loop of round 1, whatever
// the chunking.  What changes is latency: a pair that needs 1000 iterations (the reference's simulator frames) takes
// 1 + ceil(984 / (chunks * 64)) solver passes instead of 16, with its chunks on different SMs (the five-point solver
// keeps 2.5 KB of per-thread scratch in local memory: 64 solver threads per SM stay in L1, more spill out of it:
// B200 round 1, rounds of 16 / 64 / 128 hypotheses on one SM: 0.44 / 1.24 / 3.2 ms per pass).
constexpr int kChunk = 64;                      // hypotheses per solve CTA
constexpr int kFirstRound = 16;                 // the first round is small: it brings niters down for an easy pair
constexpr int kMaxChunks = 8;                   // chunks per pair and super-round
constexpr int kCandCap = 32;                    // prefix maxima kept per chunk
constexpr int kScoreThreads = 256;              // k_ransac_score / k_ransac_finish
constexpr int kRansacThreads = kScoreThreads;

__global__ void __launch_bounds__(128)
k_ransac_begin(const int32_t* __restrict__ m_n, int n_pairs, RansacParams rp, RansacState* __restrict__ state,
               int32_t* __restrict__ subsets, int sr_cap) {
  const int pair = blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= n_pairs) return;
  const int M = m_n[pair];
  RansacState st;
  st.rng = 0xFFFFFFFFFFFFFFFFull;
  st.best = 0; st.niters = rp.max_iters; st.it0 = 0; st.have_best = 0; st.solved = 0; st.overflow = 0;
  st.done = M < 5;                              // run() returns false; the host maps this to INSUFFICIENT_MATCHES / E_FAILED
  st.nh = st.done ? 0 : min(kFirstRound, st.niters);
  for (int k = 0; k < 9; ++k) st.best_model[k] = 0.0;
  int32_t* sub = subsets + (size_t)pair * sr_cap * 5;
  CvRng rng{st.rng};
  for (int h = 0; h < st.nh; ++h) {
    if (M == 5) { for (int k = 0; k < 5; ++k) sub[h * 5 + k] = k; }
    else draw_subset5(rng, M, sub + h * 5);
  }
  st.rng = rng.state;
  state[pair] = st;
}

// One thread = one hypothesis; 64 threads per CTA and nothing in shared memory, so that four CTAs (eight solver
// warps) share an SM: the solver is a long dependent FP64 chain (Durand-Kerner), bound by latency, not by the pipe.
// Models go to HBM (<= 10 x 9 doubles per hypothesis) for k_ransac_score.
__global__ void __launch_bounds__(kChunk)
k_ransac_solve(const int32_t* __restrict__ m_n, int mcap, const double* __restrict__ npts,
               const RansacState* __restrict__ state, const int32_t* __restrict__ subsets, int sr_cap,
               double* __restrict__ gmodels, int32_t* __restrict__ gnm) {
  const int pair = blockIdx.y, chunk = blockIdx.x, tid = threadIdx.x;
  const RansacState& st = state[pair];
  if (st.done) return;
  const int h0 = chunk * kChunk;
  if (h0 + tid >= st.nh) return;
  const int M = m_n[pair];
  RPE_ASSERT(chunk < kMaxChunks && st.nh <= sr_cap && M >= 5 && M <= mcap);
  const double4* P = reinterpret_cast<const double4*>(npts) + (size_t)pair * mcap;
  const int32_t* sub = subsets + ((size_t)pair * sr_cap + h0 + tid) * 5;
  double x1[10], x2[10];
  for (int k = 0; k < 5; ++k) {
    const int idx = sub[k];
    RPE_ASSERT(idx >= 0 && idx < M);
    const double4 p = P[idx];
    x1[2 * k] = p.x; x1[2 * k + 1] = p.y; x2[2 * k] = p.z; x2[2 * k + 1] = p.w;
  }
  const size_t slot = ((size_t)pair * kMaxChunks + chunk) * kChunk + tid;
  const int nm = five_point_solve(x1, x2, gmodels + slot * 90);
  RPE_ASSERT(nm >= 0 && nm <= 10);
  gnm[slot] = nm;
}

// Sampson scoring of a chunk's models, one model per warp (eight warps, few registers: the FP64 pipe is the bound
// here and many warps keep it busy), then the chunk's prefix maxima.
__global__ void __launch_bounds__(kScoreThreads)
k_ransac_score(const int32_t* __restrict__ m_n, int mcap, const double* __restrict__ npts, const double* __restrict__ K,
               RansacParams rp, const RansacState* __restrict__ state, const int32_t* __restrict__ subsets, int sr_cap,
               const double* __restrict__ gmodels, const int32_t* __restrict__ gnm, RansacCand* __restrict__ cands,
               int32_t* __restrict__ n_cand, int32_t* tr_subset, int32_t* tr_nmodels, int32_t* tr_counts,
               double* tr_models, int tr_cap) {
  __shared__ int s_nm[kChunk];
  __shared__ int s_counts[kChunk][10];
  const int pair = blockIdx.y, chunk = blockIdx.x, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const RansacState& st = state[pair];
  if (st.done) return;
  const int h0 = chunk * kChunk;
  if (h0 >= st.nh) return;
  const int nh = min(kChunk, st.nh - h0);
  const int it0 = st.it0, best0 = st.best;
  const int M = m_n[pair];
  const double4* P = reinterpret_cast<const double4*>(npts) + (size_t)pair * mcap;
  const double thr = rp.threshold / ((K[0] + K[4]) * 0.5);
  const SampsonGate gate = sampson_gate((float)(thr * thr));
  const size_t slot0 = ((size_t)pair * kMaxChunks + chunk) * kChunk;
  if (tid < nh) s_nm[tid] = gnm[slot0 + tid];
  __syncthreads();
  for (int mi = warp; mi < nh * 10; mi += kScoreThreads / 32) {
    const int h = mi / 10, r = mi % 10;
    if (r >= s_nm[h]) continue;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = gmodels[(slot0 + h) * 90 + r * 9 + k];
    int cnt = 0;
    for (int i = lane; i < M; i += 32) {
      const double4 p = P[i];
      cnt += sampson_is_inlier(E, p.x, p.y, p.z, p.w, gate) ? 1 : 0;
    }
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0) s_counts[h][r] = cnt;
  }
  __syncthreads();
  // prefix maxima of the chunk, in (hypothesis, model) order, above what the pair held when the super-round began
  if (tid == 0) {
    RansacCand* out = cands + ((size_t)pair * kMaxChunks + chunk) * kCandCap;
    int run = max(best0, 4), n = 0;
    for (int h = 0; h < nh; ++h)
      for (int r = 0; r < s_nm[h]; ++r) {
        const int c = s_counts[h][r];
        if (c > run) {
          run = c;
          if (n < kCandCap) {
            out[n].h = h0 + h; out[n].r = r; out[n].count = c; out[n].pad = 0;
            for (int k = 0; k < 9; ++k) out[n].E[k] = gmodels[(slot0 + h) * 90 + r * 9 + k];
          }
          ++n;
        }
      }
    n_cand[pair * kMaxChunks + chunk] = n;
  }
  if (tr_subset != nullptr && pair == 0) {
    const int32_t* sub = subsets + ((size_t)pair * sr_cap + h0) * 5;
    for (int i = tid; i < nh; i += kScoreThreads) {
      const int g = it0 + h0 + i;
      if (g < tr_cap) {
        tr_nmodels[g] = s_nm[i];
        for (int k = 0; k < 5; ++k) tr_subset[g * 5 + k] = sub[i * 5 + k];
        for (int r = 0; r < 10; ++r) {
          tr_counts[g * 10 + r] = r < s_nm[i] ? s_counts[i][r] : -1;
          for (int k = 0; k < 9; ++k) tr_models[(size_t)g * 90 + r * 9 + k] = r < s_nm[i] ? gmodels[(slot0 + i) * 90 + r * 9 + k] : 0.0;
        }
      }
    }
  }
}

__global__ void __launch_bounds__(128)
k_ransac_replay(const int32_t* __restrict__ m_n, int n_pairs, RansacParams rp, RansacState* __restrict__ state,
                int32_t* __restrict__ subsets, int sr_cap, int chunks, const RansacCand* __restrict__ cands,
                const int32_t* __restrict__ n_cand) {
  const int pair = blockIdx.x * blockDim.x + threadIdx.x;
  if (pair >= n_pairs) return;
  RansacState st = state[pair];
  if (st.done) return;
  const int M = m_n[pair];
  // OpenCV's sequential update rule over the super-round: `for h: if (it0 + h >= niters) break; for every model of h:
  // if (count > max(best, 4)) take it and shrink niters`.  A model that is not a prefix maximum of its chunk cannot
  // exceed the best held when it is visited, so visiting the prefix maxima in order is the same loop.
  bool stop = false;
  int last_h = -1;
  const int used_chunks = (st.nh + kChunk - 1) / kChunk;
  for (int c = 0; c < used_chunks && !stop; ++c) {
    const RansacCand* in = cands + ((size_t)pair * kMaxChunks + c) * kCandCap;
    const int n = n_cand[pair * kMaxChunks + c];
    if (n > kCandCap) st.overflow = 1;
    for (int i = 0; i < min(n, kCandCap); ++i) {
      const int h = in[i].h;
      RPE_ASSERT(h >= c * kChunk && h < min((c + 1) * kChunk, st.nh) && in[i].count <= M && h >= last_h);
      if (h != last_h && st.it0 + h >= st.niters) { stop = true; break; }   // (models of one hypothesis are all visited)
      last_h = h;
      const int cnt = in[i].count;
      if (cnt > max(st.best, 4)) {
        st.best = cnt;
        st.have_best = 1;
        for (int k = 0; k < 9; ++k) st.best_model[k] = in[i].E[k];
        st.niters = ransac_update_iters(rp.prob, (double)(M - cnt) / M, 5, st.niters);
      }
    }
  }
  st.it0 += st.nh;
  st.solved += st.nh;
  if (st.it0 >= st.niters || M == 5) {
    st.done = 1;
    st.nh = 0;
  } else {
    // subsets of the next super-round, sequentially from the pair's RNG stream
    st.nh = min(chunks * kChunk, st.niters - st.it0);
    int32_t* sub = subsets + (size_t)pair * sr_cap * 5;
    CvRng rng{st.rng};
    for (int h = 0; h < st.nh; ++h) draw_subset5(rng, M, sub + h * 5);
    st.rng = rng.state;
  }
  state[pair] = st;
}

__global__ void __launch_bounds__(kRansacThreads)
k_ransac_finish(const int32_t* __restrict__ m_n, int mcap, const double* __restrict__ npts, const double* __restrict__ K,
                RansacParams rp, const RansacState* __restrict__ state, double* __restrict__ E_out,
                uint8_t* __restrict__ rmask, int32_t* __restrict__ rinfo) {
  __shared__ double sE[9];
  const int pair = blockIdx.x, tid = threadIdx.x;
  const RansacState& st = state[pair];
  const int M = m_n[pair];
  if (M < 5) {
    if (tid < 9) E_out[pair * 9 + tid] = 0.0;
    if (tid < 4) rinfo[pair * 4 + tid] = 0;
    return;
  }
  if (tid < 9) sE[tid] = st.best_model[tid];
  __syncthreads();
  const double4* P = reinterpret_cast<const double4*>(npts) + (size_t)pair * mcap;
  const double thr = rp.threshold / ((K[0] + K[4]) * 0.5);
  const SampsonGate gate = sampson_gate((float)(thr * thr));
  for (int i = tid; i < M; i += kRansacThreads) {
    uint8_t v = 0;
    if (st.have_best) {
      const double4 p = P[i];
      v = sampson_is_inlier(sE, p.x, p.y, p.z, p.w, gate);
    }
    rmask[(size_t)pair * mcap + i] = v;
  }
  if (tid < 9) E_out[pair * 9 + tid] = st.have_best ? sE[tid] : 0.0;
  if (tid == 0) {
    rinfo[pair * 4 + 0] = st.have_best && M > 5;   // M == 5: cv2 returns stacked roots -> E_FAILED
    rinfo[pair * 4 + 1] = st.best;
    rinfo[pair * 4 + 2] = min(st.it0, st.niters);
    rinfo[pair * 4 + 3] = st.solved;
  }
}

int ransac_chunks(int n_pairs, int max_iters) {
  // enough chunks to put ~2 CTAs on every SM when the batch is small, and to keep the number of solver passes that
  // maxIters can need at about a dozen; never more than kMaxChunks
  const int fill = (2 * 148 + n_pairs - 1) / max(n_pairs, 1);
  const int passes = (max_iters + 12 * kChunk - 1) / (12 * kChunk);
  return max(1, min(kMaxChunks, max(fill, passes)));
}

void launch_ransac(const MatchBuffers& mb, const GeomBuffers& gb, int mcap, int n_pairs, const double* dK,
                   RansacParams rp, cudaStream_t s, int* launches) {
  const int chunks = ransac_chunks(n_pairs, rp.max_iters);
  const int sr_cap = kMaxChunks * kChunk;
  const int tb = (n_pairs + 127) / 128;
  k_ransac_begin<<<tb, 128, 0, s>>>(mb.m_n, n_pairs, rp, gb.rstate, gb.rsubsets, sr_cap);
  ++*launches;
  const int rounds = 1 + (max(rp.max_iters - kFirstRound, 0) + chunks * kChunk - 1) / (chunks * kChunk);
  for (int r = 0; r < rounds; ++r) {
    dim3 grid(r == 0 ? 1 : chunks, n_pairs);
    k_ransac_solve<<<grid, kChunk, 0, s>>>(mb.m_n, mcap, gb.npts, gb.rstate, gb.rsubsets, sr_cap, gb.rmodels, gb.rnm);
    k_ransac_score<<<grid, kScoreThreads, 0, s>>>(mb.m_n, mcap, gb.npts, dK, rp, gb.rstate, gb.rsubsets, sr_cap, gb.rmodels,
                                                  gb.rnm, gb.rcands, gb.rncand, gb.tr_subset, gb.tr_nmodels, gb.tr_counts,
                                                  gb.tr_models, gb.tr_cap);
    k_ransac_replay<<<tb, 128, 0, s>>>(mb.m_n, n_pairs, rp, gb.rstate, gb.rsubsets, sr_cap, chunks, gb.rcands, gb.rncand);
    *launches += 3;
  }
  k_ransac_finish<<<n_pairs, kRansacThreads, 0, s>>>(mb.m_n, mcap, gb.npts, dK, rp, gb.rstate, gb.E, gb.rmask, gb.rinfo);
  ++*launches;
}

// ------------------------------------------------------------------ K12/K13: recoverPose + Euler
constexpr int kPoseThreads = 512;        // 128 points x 4 candidates per pass (launch bound caps the registers at 128)

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
  // one thread per (point, candidate): the four candidates of a point sit in four adjacent lanes (a 4 x 4 Jacobi SVD in
  // FP64 per thread is a long dependent chain; with all four candidates of four points in one thread, as first written,
  // recoverPose was a quarter of the latency of a single-pair call)
  const int c = tid & 3;
  const double tc[3] = {(c & 2) ? -st[0] : st[0], (c & 2) ? -st[1] : st[1], (c & 2) ? -st[2] : st[2]};
  int local = 0;
  for (int i0 = 0; i0 < M; i0 += kPoseThreads / 4) {        // warp-uniform trip count
    const int i = i0 + (tid >> 2);
    bool ok = false;
    if (i < M) {
      const double4 p = P[i];
      ok = triangulate_cheirality(sR[c & 1], tc, p.x, p.y, p.z, p.w, 50.0);
    }
    local += ok;
    uint32_t bits = (uint32_t)ok << c;
    bits |= __shfl_xor_sync(0xffffffffu, bits, 1);
    bits |= __shfl_xor_sync(0xffffffffu, bits, 2);
    if (c == 0 && i < M) pm[i] = (uint8_t)bits;
  }
  // per-candidate counts: lanes with equal (lane & 3) hold the same candidate
  for (int o = 16; o >= 4; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((tid & 31) < 4) atomicAdd(&s_good[c], local);
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

int setup_geom_attributes() { return 0; }

}  // namespace rpe
