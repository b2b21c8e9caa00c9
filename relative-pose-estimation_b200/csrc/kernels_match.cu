// kernels_match.cu — brute-force Hamming matcher on sm_100a.  Replaces
// cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match + the reference's stable sort and
// max_matches cut (src/core/pose_estimator.py:131, :144-151) and, for BASELINE configs 3-5,
// knnMatch(k=2) + ratio test.
//
// k_hamming: bound by the XU pipe (POPC, 16 lanes/clk/SM) and the ALU pipe (LOP3, 64 lanes/clk/SM) together,
// not by HBM; carry-save adders on the ALU pipe (match_math.cuh) cut the POPCs per 256-bit distance from 8
// to 5 or 4, the key arithmetic runs on the FMA pipe (IMAD), and the default variant alternates the 5- and
// 4-POPC forms so that both pipes sit near 18/64 clk per distance.  Each thread keeps kQPT query
// descriptors (8 x u32 each) in registers, train descriptors stream through a shared-memory
// tile read with 128-bit broadcast loads; nearest / second nearest are tracked as packed
// (dist << 20 | index) keys, which reproduces OpenCV's tie rule (lowest index) for free.
#include <cstdlib>
#include "rpe_device.cuh"
#include "match_math.cuh"

namespace rpe {

constexpr int kMatchThreads = 128;
constexpr int kQPT = 2;                       // queries per thread
constexpr int kQPB = kMatchThreads * kQPT;    // queries per block
constexpr int kTrainTile = 128;               // train descriptors per shared-memory tile

__device__ __forceinline__ uint32_t umin3(uint32_t a, uint32_t b, uint32_t c) {
  return __vimin3_u32(a, b, c);   // VIMNMX3 on sm_100
}

// kFma: keys built with multiply-adds (FMA pipe) and two train descriptors folded per top-2 update;
// kMix (with kFma): the even train descriptor of each pair uses three carry-save adders + five POPCs, the odd
// one four adders + four POPCs, which balances the ALU pipe (LOP3) against the XU pipe (POPC)
// kLin (with kFma): operands in the linearly pre-combined layout of match_math.cuh (hamming_mix): queries are
// rewritten in registers, train descriptors when a thread stages them in shared memory (one descriptor per
// thread); 13 LOP3 + 4 POPC per distance
template <bool kTop2, bool kCsa4, bool kFma, bool kMix = false, bool kLin = false>
__global__ void __launch_bounds__(kMatchThreads)
k_hamming(const uint8_t* __restrict__ desc, const int32_t* __restrict__ kp_n, int kp_stride, int pstep, uint32_t m20,
          int nsplit, uint32_t* __restrict__ fwd1, uint32_t* __restrict__ fwd2, uint32_t* __restrict__ bwd1) {
  __shared__ uint4 s_train[kTrainTile * 2];
  const int pair = blockIdx.y;
  const int dir = blockIdx.z;                  // 0: view1 -> view2, 1: view2 -> view1
  // image slots of the two views: pstep = 2 (pair batch: 2p, 2p+1) or 1 (consecutive frames: p, p+1)
  const int qi = pstep * pair + dir, ti = pstep * pair + 1 - dir;
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
  const int tid = threadIdx.x;

  uint32_t q[kQPT][8];
  uint32_t best1[kQPT], best2[kQPT];
#pragma unroll
  for (int k = 0; k < kQPT; ++k) {
    const int qq = min(q0 + tid + k * kMatchThreads, nq - 1);
    const uint4 a = dq[2 * qq], b = dq[2 * qq + 1];
    q[k][0] = a.x; q[k][1] = a.y; q[k][2] = a.z; q[k][3] = a.w;
    q[k][4] = b.x; q[k][5] = b.y; q[k][6] = b.z; q[k][7] = b.w;
    if (kLin) hamming_mix(q[k]);
    best1[k] = 0xffffffffu;
    best2[k] = 0xffffffffu;
  }
  static_assert(kTrainTile == kMatchThreads, "one staged train descriptor per thread");
  for (int t0 = tbeg; t0 < nt; t0 += kTrainTile) {
    const int tn = min(kTrainTile, nt - t0);
    __syncthreads();
    if (kLin) {
      if (tid < tn) {
        const uint4 a = dt[2 * (t0 + tid)], b = dt[2 * (t0 + tid) + 1];
        uint32_t d[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        hamming_mix(d);
        s_train[2 * tid] = make_uint4(d[0], d[1], d[2], d[3]);
        s_train[2 * tid + 1] = make_uint4(d[4], d[5], d[6], d[7]);
      }
    } else {
      for (int i = tid; i < tn * 2; i += kMatchThreads) s_train[i] = dt[2 * t0 + i];
    }
    __syncthreads();
    if (kFma) {
      // pairs of train descriptors: with m <= M the two new keys, the two smallest of {best1, best2, m, M}
      // are min(best1, m) and min(max(best1, m), best2, M): five min / max per two distances instead of six
      int j = 0;
#pragma unroll 2
      for (; j + 1 < tn; j += 2) {
        const uint4 a0 = s_train[2 * j], b0 = s_train[2 * j + 1], a1 = s_train[2 * j + 2], b1 = s_train[2 * j + 3];
#pragma unroll
        for (int k = 0; k < kQPT; ++k) {
          const uint32_t k0 = kLin ? hamming256_mixed_key<kCsa4 && !kMix>(q[k], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + j), m20)
                                   : hamming256_key<kCsa4 && !kMix>(q[k], a0.x, a0.y, a0.z, a0.w, b0.x, b0.y, b0.z, b0.w, (uint32_t)(t0 + j), m20);
          const uint32_t k1 = kLin ? hamming256_mixed_key<kCsa4 || kMix>(q[k], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + j + 1), m20)
                                   : hamming256_key<kCsa4 || kMix>(q[k], a1.x, a1.y, a1.z, a1.w, b1.x, b1.y, b1.z, b1.w, (uint32_t)(t0 + j + 1), m20);
          if (kTop2) {
            const uint32_t m = min(k0, k1), M = max(k0, k1);
            best2[k] = umin3(max(best1[k], m), best2[k], M);
            best1[k] = min(best1[k], m);
          } else {
            best1[k] = umin3(best1[k], k0, k1);
          }
        }
      }
      if (j < tn) {
        const uint4 a = s_train[2 * j], b = s_train[2 * j + 1];
#pragma unroll
        for (int k = 0; k < kQPT; ++k) {
          const uint32_t key = kLin ? hamming256_mixed_key<kCsa4>(q[k], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, (uint32_t)(t0 + j), m20)
                                    : hamming256_key<kCsa4>(q[k], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, (uint32_t)(t0 + j), m20);
          if (kTop2) best2[k] = min(best2[k], max(best1[k], key));
          best1[k] = min(best1[k], key);
        }
      }
    } else {
#pragma unroll 4
      for (int j = 0; j < tn; ++j) {
        const uint4 a = s_train[2 * j], b = s_train[2 * j + 1];
#pragma unroll
        for (int k = 0; k < kQPT; ++k) {
          const uint32_t d = hamming256<kCsa4>(q[k], a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w);
          const uint32_t key = pack_nn(d, (uint32_t)(t0 + j));
          if (kTop2) best2[k] = min(best2[k], max(best1[k], key));
          best1[k] = min(best1[k], key);
        }
      }
    }
  }
  uint32_t* o1 = (dir == 0 ? fwd1 : bwd1) + (size_t)pair * kp_stride;
#pragma unroll
  for (int k = 0; k < kQPT; ++k) {
    const int qq = q0 + tid + k * kMatchThreads;
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
static int pick_nsplit(int nfeatures) {
  static const int forced = getenv("RPE_MATCH_SPLIT") ? atoi(getenv("RPE_MATCH_SPLIT")) : 0;   // tuning knob
  if (forced > 0) return forced;
  return max(1, min(32, (nfeatures + 5 * kTrainTile / 2) / (5 * kTrainTile)));
}

void launch_match(const OrbBuffers& ob, const MatchBuffers& mb, int kp_stride, int nfeatures, int n_pairs, int pstep,
                  int mode, cudaStream_t s, int* launches) {
  const int qblocks = (kp_stride + kQPB - 1) / kQPB;
  static const int variant = getenv("RPE_MATCH_VARIANT") ? atoi(getenv("RPE_MATCH_VARIANT")) : 6;   // tuning knob (B200: 4.14 / 4.25 / 3.96 / 3.95 / 3.84 / 3.60 / 3.58 / 3.92 ms for 0..7)
  const bool top2 = mode == RPE_MATCH_RATIO;

  // 0: CSA3, 1: CSA4, 2: CSA3 + FMA-pipe keys + paired update, 3: CSA4 + FMA-pipe keys + paired update,
  // 4: as 2 / 3 with CSA3 and CSA4 alternating
  // 5: pre-combined operands, CSA4 (13 LOP3 + 4 POPC); 6: pre-combined, CSA3 / CSA4 alternating; 7: pre-combined, CSA3
  auto k = top2 ? (variant == 5 ? k_hamming<true, true, true, false, true> : variant == 6 ? k_hamming<true, false, true, true, true>
                   : variant == 7 ? k_hamming<true, false, true, false, true> : variant == 1 ? k_hamming<true, true, false> : variant == 2 ? k_hamming<true, false, true>
                   : variant == 3 ? k_hamming<true, true, true> : variant == 4 ? k_hamming<true, false, true, true>
                   : k_hamming<true, false, false>)
                : (variant == 5 ? k_hamming<false, true, true, false, true> : variant == 6 ? k_hamming<false, false, true, true, true>
                   : variant == 7 ? k_hamming<false, false, true, false, true> : variant == 1 ? k_hamming<false, true, false> : variant == 2 ? k_hamming<false, false, true>
                   : variant == 3 ? k_hamming<false, true, true> : variant == 4 ? k_hamming<false, false, true, true>
                   : k_hamming<false, false, false>);
  const int ndir = top2 ? 1 : 2;
  const int nsplit = pick_nsplit(min(nfeatures, kp_stride));
  if (nsplit > 1) {
    const size_t bytes = (size_t)n_pairs * kp_stride * sizeof(uint32_t);
    cudaMemsetAsync(mb.fwd1, 0xff, bytes, s);
    cudaMemsetAsync(top2 ? mb.fwd2 : mb.bwd1, 0xff, bytes, s);
  }
  dim3 grid(qblocks * nsplit, n_pairs, ndir);
  k<<<grid, kMatchThreads, 0, s>>>(ob.desc, ob.kp_n, kp_stride, pstep, 1u << 20, nsplit, mb.fwd1, mb.fwd2, mb.bwd1);
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
  return (int)cudaFuncSetAttribute(k_build_matches, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   sort_cap_match * 4);
}

}  // namespace rpe
