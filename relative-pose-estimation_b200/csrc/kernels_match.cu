// kernels_match.cu — brute-force Hamming matcher on sm_100a.  Replaces
// cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match + the reference's stable sort and
// max_matches cut (src/core/pose_estimator.py:131, :144-151) and, for BASELINE configs 3-5,
// knnMatch(k=2) + ratio test.
//
// k_hamming: bound by the XU pipe (POPC, 16 lanes/clk/SM) and the ALU pipe (LOP3, 64 lanes/clk/SM) together,
// not by HBM.  Operands are linearly pre-combined (match_math.cuh: hamming_mix) so that three carry-save adders
// cost one XOR (sum) + one LOP3 (carry) each; one train descriptor in four takes the 11 LOP3 + 5 POPC form, the others
// the 13 LOP3 + 4 POPC form, which balances the two pipes; the key arithmetic runs on the FMA pipe (IMAD).  Each
// thread keeps kQPT query descriptors (8 x u32 each) in registers, train descriptors stream through a
// shared-memory tile read with 128-bit broadcast loads; nearest / second nearest are tracked as packed
// (dist << 20 | index) keys, which reproduces OpenCV's tie rule (lowest index) for free.
// Measured alternatives (B200, 128 pairs x 5000 x 5000, round 1): plain 8 XOR + 8 POPC 4.14 ms; three / four
// adders 4.25 / 3.96 ms; + FMA-pipe keys and paired top-2 update 3.95 / 3.84 ms; pre-combined operands with four
// adders 3.60, alternating 3.58 (this kernel), three adders 3.92 ms.
#include "rpe_device.cuh"
#include "match_math.cuh"

namespace rpe {

constexpr int kMatchThreads = 128;
// (queries per thread, ratio mode, this kernel form, B200, 256 pairs x 5000 x 5000: 2 / 3 / 4 -> 6.73 / 7.03 / 6.91 ms at
// 56 / 72 / 80 registers: fewer shared-memory reads per distance do not pay for the lost occupancy)
constexpr int kQptTop2 = 2;                   // queries per thread, ratio mode
constexpr int kQptCross = 2;                  // queries per thread, cross-check mode (the column-minimum code is written for two)
constexpr int kTrainTile = 128;               // train descriptors per shared-memory tile

__device__ __forceinline__ uint32_t umin3(uint32_t a, uint32_t b, uint32_t c) {
  return __vimin3_u32(a, b, c);   // VIMNMX3 on sm_100
}

// kTop2 (ratio mode, knnMatch k = 2): nearest and second nearest train descriptor of every query, view 1 -> view 2.
// !kTop2 (cross-check mode): every distance D[q][t] is computed ONCE and feeds both directions: the row minimum
// (nearest train descriptor of query q, key dist<<20|t) stays in the thread's registers as above; the column
// minimum (nearest query of train descriptor t, key dist<<20|q) is the minimum over the warp's 64 queries - one
// REDUX.MIN per warp and train descriptor - kept by lane t % 32, folded into a shared-memory minimum per tile slot
// (one atomicMin per lane and 32 train descriptors) and flushed to bwd1 with one global atomicMin per CTA and train
// descriptor.  Keys carry the index in the low bits, so ties go to the lowest index in both directions, as cv2's.
template <bool kTop2, int kQPT>
__global__ void __launch_bounds__(kMatchThreads)
k_hamming(const uint8_t* __restrict__ desc, const int32_t* __restrict__ kp_n, int kp_stride, int pstep, uint32_t m20,
          int nsplit, uint32_t* __restrict__ fwd1, uint32_t* __restrict__ fwd2, uint32_t* __restrict__ bwd1) {
  __shared__ uint4 s_train[kTrainTile * 2];
  __shared__ uint32_t s_col[kTrainTile];
  constexpr int kQPB = kMatchThreads * kQPT;
  static_assert(kTop2 || kQPT == 2, "cross-check path handles two queries per thread");
  const int pair = blockIdx.y;
  // image slots of the two views: pstep = 2 (pair batch: 2p, 2p+1) or 1 (consecutive frames: p, p+1)
  const int qi = pstep * pair, ti = pstep * pair + 1;
  const int nq = kp_n[qi], nt_all = kp_n[ti];
  // blockIdx.x = query block * nsplit + split: each CTA scans one slice of the train set (finer work units: the
  // last wave of a launch stays full); slices are merged with atomicMin on the packed keys, see below
  const int split = nsplit > 1 ? blockIdx.x % nsplit : 0;
  const int q0 = (nsplit > 1 ? blockIdx.x / nsplit : blockIdx.x) * kQPB;
  if (q0 >= nq) return;
  const int slice = nsplit > 1 ? (((nt_all + nsplit - 1) / nsplit + 1) & ~1) : nt_all;
  const int tbeg = split * slice, nt = min(nt_all, tbeg + slice);
  if (tbeg >= nt && nsplit > 1) return;
  const uint4* dq = reinterpret_cast<const uint4*>(desc + (size_t)qi * kp_stride * 32);
  const uint4* dt = reinterpret_cast<const uint4*>(desc + (size_t)ti * kp_stride * 32);
  const int tid = threadIdx.x, lane = tid & 31;

  uint32_t q[kQPT][8];
  uint32_t best1[kQPT], best2[kQPT];
  int qidx[kQPT];
#pragma unroll
  for (int k = 0; k < kQPT; ++k) {
    // queries past the end repeat the last one: same keys, so they change neither a row nor a column minimum
    const int qq = min(q0 + tid + k * kMatchThreads, nq - 1);
    qidx[k] = qq;
    const uint4 a = dq[2 * qq], b = dq[2 * qq + 1];
    q[k][0] = a.x; q[k][1] = a.y; q[k][2] = a.z; q[k][3] = a.w;
    q[k][4] = b.x; q[k][5] = b.y; q[k][6] = b.z; q[k][7] = b.w;
    hamming_mix(q[k]);
    best1[k] = 0xffffffffu;
    best2[k] = 0xffffffffu;
  }
  static_assert(kTrainTile == kMatchThreads, "one staged train descriptor per thread");
  for (int t0 = tbeg; t0 < nt; t0 += kTrainTile) {
    const int tn = min(kTrainTile, nt - t0);
    __syncthreads();
    if (!kTop2 && t0 > tbeg && tid < kTrainTile) {           // flush the previous tile's column minima
      const uint32_t v = s_col[tid];
      RPE_ASSERT(t0 - kTrainTile + tid >= 0 && t0 - kTrainTile + tid < nt_all && (v == 0xffffffffu || (int)(v & 0xfffffu) < nq));
      if (v != 0xffffffffu) atomicMin(&bwd1[(size_t)pair * kp_stride + t0 - kTrainTile + tid], v);
    }
    if (!kTop2) s_col[tid] = 0xffffffffu;
    if (tid < tn) {
      const uint4 a = dt[2 * (t0 + tid)], b = dt[2 * (t0 + tid) + 1];
      uint32_t d[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
      hamming_mix(d);
      s_train[2 * tid] = make_uint4(d[0], d[1], d[2], d[3]);
      s_train[2 * tid + 1] = make_uint4(d[4], d[5], d[6], d[7]);
    }
    __syncthreads();
    if (kTop2) {
      // pairs of train descriptors: with m <= M the two new keys, the two smallest of {best1, best2, m, M}
      // are min(best1, m) and min(max(best1, m), best2, M): five min / max per two distances instead of six
      int j = 0;
      // groups of four train descriptors: the first takes the 11 LOP3 + 5 POPC form, the other three the 13 LOP3 + 4
      // POPC form (B200, 256 pairs x 5000 x 5000: every other one 6.73 ms, one in four 6.61 ms, none 6.62 ms)
      for (; j + 3 < tn; j += 4) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int jj = j + 2 * h;
          const uint4 a0 = s_train[2 * jj], b0 = s_train[2 * jj + 1], a1 = s_train[2 * jj + 2], b1 = s_train[2 * jj + 3];
#pragma unroll
          for (int k = 0; k < kQPT; ++k) {
            const uint32_t k0 = h == 0
                ? hamming256_mixed_key<false>(q[k], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + jj), m20)
                : hamming256_mixed_key<true>(q[k], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + jj), m20);
            const uint32_t k1 = hamming256_mixed_key<true>(q[k], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + jj + 1), m20);
            const uint32_t m = min(k0, k1), M = max(k0, k1);
            best2[k] = umin3(max(best1[k], m), best2[k], M);
            best1[k] = min(best1[k], m);
          }
        }
      }
      for (; j + 1 < tn; j += 2) {
        const uint4 a0 = s_train[2 * j], b0 = s_train[2 * j + 1], a1 = s_train[2 * j + 2], b1 = s_train[2 * j + 3];
#pragma unroll
        for (int k = 0; k < kQPT; ++k) {
          const uint32_t k0 = hamming256_mixed_key<false>(q[k], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + j), m20);
          const uint32_t k1 = hamming256_mixed_key<true>(q[k], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + j + 1), m20);
          const uint32_t m = min(k0, k1), M = max(k0, k1);
          best2[k] = umin3(max(best1[k], m), best2[k], M);
          best1[k] = min(best1[k], m);
        }
      }
      if (j < tn) {
        const uint4 a = s_train[2 * j], b = s_train[2 * j + 1];
#pragma unroll
        for (int k = 0; k < kQPT; ++k) {
          const uint32_t key = hamming256_mixed_key<false>(q[k], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, (uint32_t)(t0 + j), m20);
          best2[k] = min(best2[k], max(best1[k], key));
          best1[k] = min(best1[k], key);
        }
      }
    } else {
      // column key of (q, t) = row key + (q - t): the distance bits are shared
      const int dq0 = qidx[0] - t0, dq1 = qidx[1] - t0;
      for (int jb = 0; jb < tn; jb += 32) {
        uint32_t mine = 0xffffffffu;                         // column minimum of train descriptor t0 + jb + lane
        const int je = min(jb + 32, tn);
        int j = jb;
#pragma unroll 2
        for (; j + 1 < je; j += 2) {
          const uint4 a0 = s_train[2 * j], b0 = s_train[2 * j + 1], a1 = s_train[2 * j + 2], b1 = s_train[2 * j + 3];
          const uint32_t r00 = hamming256_mixed_key<false>(q[0], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + j), m20);
          const uint32_t r10 = hamming256_mixed_key<false>(q[1], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + j), m20);
          const uint32_t r01 = hamming256_mixed_key<true>(q[0], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + j + 1), m20);
          const uint32_t r11 = hamming256_mixed_key<true>(q[1], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + j + 1), m20);
          best1[0] = umin3(best1[0], r00, r01);
          best1[1] = umin3(best1[1], r10, r11);
          const uint32_t c0 = __reduce_min_sync(0xffffffffu, min(r00 + (uint32_t)(dq0 - j), r10 + (uint32_t)(dq1 - j)));
          const uint32_t c1 = __reduce_min_sync(0xffffffffu, min(r01 + (uint32_t)(dq0 - j - 1), r11 + (uint32_t)(dq1 - j - 1)));
          if (lane == ((j - jb) & 31)) mine = c0;
          if (lane == ((j + 1 - jb) & 31)) mine = c1;
        }
        if (j < je) {
          const uint4 a = s_train[2 * j], b = s_train[2 * j + 1];
          const uint32_t r0 = hamming256_mixed_key<false>(q[0], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, (uint32_t)(t0 + j), m20);
          const uint32_t r1 = hamming256_mixed_key<false>(q[1], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, (uint32_t)(t0 + j), m20);
          best1[0] = min(best1[0], r0);
          best1[1] = min(best1[1], r1);
          const uint32_t c = __reduce_min_sync(0xffffffffu, min(r0 + (uint32_t)(dq0 - j), r1 + (uint32_t)(dq1 - j)));
          if (lane == ((j - jb) & 31)) mine = c;
        }
        RPE_ASSERT(jb + 31 < kTrainTile || jb + lane >= tn || jb + lane < kTrainTile);
        if (jb + lane < tn) atomicMin(&s_col[jb + lane], mine);
      }
    }
  }
  if (!kTop2) {                                              // column minima of the last tile
    __syncthreads();
    const int t_last = tbeg + ((nt - tbeg - 1) / kTrainTile) * kTrainTile;
    const uint32_t v = s_col[tid];
    if (t_last + tid < nt && v != 0xffffffffu) atomicMin(&bwd1[(size_t)pair * kp_stride + t_last + tid], v);
  }
  uint32_t* o1 = fwd1 + (size_t)pair * kp_stride;
#pragma unroll
  for (int k = 0; k < kQPT; ++k) {
    const int qq = q0 + tid + k * kMatchThreads;
    RPE_ASSERT(nq <= kp_stride && nt_all <= kp_stride && (best1[k] == 0xffffffffu || (int)(best1[k] & 0xfffffu) < nt_all));
    if (qq < nq) {
      if (nsplit > 1) {
        // outputs preset to 0xffffffff.  Every key except the overall minimum loses one atomicMin on o1 (as
        // the new value or as the displaced old one) and is then offered to the second-nearest slot
        const uint32_t old = atomicMin(&o1[qq], best1[k]);
        if (kTop2) {
          uint32_t* o2 = fwd2 + (size_t)pair * kp_stride + qq;
          const uint32_t loser = max(old, best1[k]);
          if (loser != 0xffffffffu) atomicMin(o2, loser);
          if (best2[k] != 0xffffffffu) atomicMin(o2, best2[k]);
        }
      } else {
        o1[qq] = best1[k];
        if (kTop2) fwd2[(size_t)pair * kp_stride + qq] = best2[k];
      }
    }
  }
}

// train-set slices per query block.  Finer work units keep every SM busy to the end of the launch (B200, 128 pairs x
// 5000 x 5000: 3.58 / 3.41 / 3.38 ms with 1 / 4 / 8 slices); a slice keeps >= ~5 shared-memory tiles of work so
// that loading the queries and the three atomics per query stay negligible
static int pick_nsplit(int nfeatures) { return max(1, min(32, (nfeatures + 5 * kTrainTile / 2) / (5 * kTrainTile))); }

void launch_match(const OrbBuffers& ob, const MatchBuffers& mb, int kp_stride, int nfeatures, int n_pairs, int pstep,
                  int mode, cudaStream_t s, int* launches) {
  const bool top2 = mode == RPE_MATCH_RATIO;
  const int qpb = kMatchThreads * (top2 ? kQptTop2 : kQptCross);
  const int qblocks = (kp_stride + qpb - 1) / qpb;
  const int nsplit = pick_nsplit(min(nfeatures, kp_stride));
  const size_t bytes = (size_t)n_pairs * kp_stride * sizeof(uint32_t);
  if (nsplit > 1) cudaMemsetAsync(mb.fwd1, 0xff, bytes, s);
  if (top2 && nsplit > 1) cudaMemsetAsync(mb.fwd2, 0xff, bytes, s);
  if (!top2) cudaMemsetAsync(mb.bwd1, 0xff, bytes, s);       // column minima always arrive through atomicMin
  dim3 grid(qblocks * nsplit, n_pairs);
  if (top2) k_hamming<true, kQptTop2><<<grid, kMatchThreads, 0, s>>>(ob.desc, ob.kp_n, kp_stride, pstep, 1u << 20, nsplit, mb.fwd1, mb.fwd2, mb.bwd1);
  else k_hamming<false, kQptCross><<<grid, kMatchThreads, 0, s>>>(ob.desc, ob.kp_n, kp_stride, pstep, 1u << 20, nsplit, mb.fwd1, mb.fwd2, mb.bwd1);
  ++*launches;
}

// ------------------------------------------------------------------ K9: filter + stable sort + cut + gather
// One CTA per pair.  Valid matches (cross-check or ratio test) are ranked by the unique key
// (dist << 20 | queryIdx), which is exactly the order of Python's stable sort by distance of a
// query-ascending list; the first max_matches are kept: everything below the cut distance,
// plus the first few (in query order) at the cut distance.
constexpr int kBuildThreads = 512;

__device__ __forceinline__ uint32_t match_key(int q, int mode, double ratio, int n1, int n2,
                                              const uint32_t* f1, const uint32_t* f2, const uint32_t* b1) {
  // returns packed (dist << 20 | q) or 0xffffffff when query q yields no match
  if (q >= n1 || n2 <= 0) return 0xffffffffu;
  const uint32_t k1 = f1[q];
  const uint32_t d = k1 >> 20, t = k1 & 0xfffffu;
  if (mode == RPE_MATCH_RATIO) {
    if (n2 < 2) return 0xffffffffu;
    const uint32_t d2 = f2[q] >> 20;
    if (!((double)d < ratio * (double)d2)) return 0xffffffffu;
  } else {
    if ((b1[t] & 0xfffffu) != (uint32_t)q) return 0xffffffffu;
  }
  return pack_nn(d, (uint32_t)q);
}

__global__ void __launch_bounds__(kBuildThreads)
k_build_matches(const float* __restrict__ kp_x, const float* __restrict__ kp_y, const int32_t* __restrict__ kp_n,
                int kp_stride, int pstep, const uint32_t* __restrict__ fwd1, const uint32_t* __restrict__ fwd2,
                const uint32_t* __restrict__ bwd1, int mode, double ratio, int max_matches, int mcap,
                int sort_cap, int32_t* __restrict__ m_q, int32_t* __restrict__ m_t, int32_t* __restrict__ m_d,
                int32_t* __restrict__ m_n, float* __restrict__ pts1, float* __restrict__ pts2,
                const uint32_t* __restrict__ img_flags, int32_t* __restrict__ pstat) {
  extern __shared__ uint32_t s_keys[];          // sort_cap entries
  __shared__ uint32_t s_hist[258];
  __shared__ uint32_t s_warp[kBuildThreads / 32];
  __shared__ uint32_t s_cut, s_take, s_n, s_running;
  const int pair = blockIdx.x, tid = threadIdx.x;
  const int v1 = pstep * pair, v2 = v1 + 1;   // image slots of the two views
  const int n1 = kp_n[v1], n2 = kp_n[v2];
  const uint32_t* f1 = fwd1 + (size_t)pair * kp_stride;
  const uint32_t* f2 = fwd2 + (size_t)pair * kp_stride;
  const uint32_t* b1 = bwd1 + (size_t)pair * kp_stride;
  const uint32_t limit = (uint32_t)min(max_matches > 0 ? max_matches : mcap, mcap);

  for (int i = tid; i < 258; i += kBuildThreads) s_hist[i] = 0;
  if (tid == 0) { s_n = 0; s_running = 0; }
  for (int i = tid; i < sort_cap; i += kBuildThreads) s_keys[i] = 0xffffffffu;
  __syncthreads();
  for (int q = tid; q < n1; q += kBuildThreads) {
    const uint32_t k = match_key(q, mode, ratio, n1, n2, f1, f2, b1);
    if (k != 0xffffffffu) atomicAdd(&s_hist[k >> 20], 1u);
  }
  __syncthreads();
  if (tid == 0) {
    uint32_t cum = 0, cut = 257, take = 0;
    for (uint32_t d = 0; d <= 256; ++d) {
      if (cum + s_hist[d] >= limit) { cut = d; take = limit - cum; break; }
      cum += s_hist[d];
    }
    s_cut = cut;     // distances < cut are all kept; `take` more at distance == cut, in query order
    s_take = take;
  }
  __syncthreads();
  const uint32_t cut = s_cut, take = s_take;
  // ordered pass over queries (chunks of blockDim) for the rank among distance == cut
  for (int base = 0; base < n1; base += kBuildThreads) {
    const int q = base + tid;
    const uint32_t k = match_key(q, mode, ratio, n1, n2, f1, f2, b1);
    const uint32_t d = k >> 20;
    const bool valid = k != 0xffffffffu;
    const bool at_cut = valid && d == cut;
    const uint32_t bal = __ballot_sync(0xffffffffu, at_cut);
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    uint32_t before = s_running;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    const uint32_t rank = before + __popc(bal & ((1u << lane) - 1u));
    const bool keep = valid && (d < cut || (at_cut && rank < take));
    if (keep) {
      const uint32_t j = atomicAdd(&s_n, 1u);
      RPE_ASSERT(j < (uint32_t)sort_cap);      // limit <= mcap <= sort_cap, and at most `limit` keys are kept
      if (j < (uint32_t)sort_cap) s_keys[j] = k;
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t tot = 0;
      for (int w = 0; w < kBuildThreads / 32; ++w) tot += s_warp[w];
      s_running += tot;
    }
    __syncthreads();
  }
  const uint32_t n = min(s_n, (uint32_t)sort_cap);
  uint32_t m = 1;
  while (m < n) m <<= 1;
  for (uint32_t k = 2; k <= m; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t i = tid; i < m; i += kBuildThreads) {
        const uint32_t ixj = i ^ j;
        if (ixj > i) {
          const uint32_t a = s_keys[i], b = s_keys[ixj];
          const bool up = (i & k) == 0;
          if ((a > b) == up) { s_keys[i] = b; s_keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
  const float* x1 = kp_x + (size_t)v1 * kp_stride;
  const float* y1 = kp_y + (size_t)v1 * kp_stride;
  const float* x2 = kp_x + (size_t)v2 * kp_stride;
  const float* y2 = kp_y + (size_t)v2 * kp_stride;
  for (uint32_t i = tid; i < n; i += kBuildThreads) {
    const uint32_t k = s_keys[i];
    const int q = k & 0xfffffu, d = k >> 20;
    const int t = f1[q] & 0xfffffu;
    RPE_ASSERT(q < n1 && t < n2 && i < (uint32_t)mcap);
    const size_t o = (size_t)pair * mcap + i;
    m_q[o] = q; m_t[o] = t; m_d[o] = d;
    pts1[2 * o] = x1[q]; pts1[2 * o + 1] = y1[q];
    pts2[2 * o] = x2[t]; pts2[2 * o + 1] = y2[t];
  }
  if (tid == 0) {
    m_n[pair] = (int32_t)n;
    // what k_pose needs to know about the two images, so that the geometry tail never touches the ORB
    // buffers (the next batch's ORB may already be overwriting them): bit 0 = an image without
    // descriptors, bit 8 = a candidate / tie list overflowed
    const uint32_t fl = img_flags != nullptr ? (img_flags[v1] | img_flags[v2]) : 0u;
    pstat[pair] = ((n1 == 0 || n2 == 0) ? 1 : 0) | (fl ? RPE_STATUS_FLAG_OVERFLOW : 0);
  }
}

void launch_build_matches(const OrbBuffers& ob, const MatchBuffers& mb, int kp_stride, int mcap, int sort_cap,
                          int n_pairs, int pstep, int mode, double ratio, int max_matches, cudaStream_t s,
                          int* launches) {
  k_build_matches<<<n_pairs, kBuildThreads, (size_t)sort_cap * 4, s>>>(
      ob.kp_x, ob.kp_y, ob.kp_n, kp_stride, pstep, mb.fwd1, mb.fwd2, mb.bwd1, mode, ratio, max_matches, mcap,
      sort_cap,
      mb.m_q, mb.m_t, mb.m_d, mb.m_n, mb.pts1, mb.pts2, ob.img_flags, mb.pstat);
  ++*launches;
}

int setup_match_attributes(int sort_cap_match) {
  return raise_dynamic_smem_limit((const void*)k_build_matches, sort_cap_match * 4);
}

}  // namespace rpe
