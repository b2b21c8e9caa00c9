// kernels_orb.cu — ORB stage on sm_100a: pyramid, FAST-9 + NMS, two-step retainBest with
// Harris scoring, 7x7 blur, orientation + steered rBRIEF.  Replaces
// cv2.ORB.detectAndCompute as the reference calls it (src/core/pose_estimator.py:85-91, :108).
// Every launch covers the whole batch (grid.y / grid.z = image slot); the dense kernels
// (FAST, blur) cover all pyramid levels in one launch through a flat tile table.
#include "rpe_device.cuh"
#include "orb_math.cuh"
#include "tma.cuh"

namespace rpe {

// ------------------------------------------------------------------ K0: BGR -> gray ingest
// HBM-bound: 3 bytes in, 1 byte out per pixel.  One thread = 4 pixels = three aligned word loads
// (12 bytes of interleaved BGR) and one word store into level 0 of the image's pyramid slab.
__global__ void __launch_bounds__(256)
k_bgr2gray(const uint8_t* __restrict__ bgr, size_t src_stride, int src_pitch, uint8_t* __restrict__ pyr,
           size_t pyr_stride, int dst_pitch, int w, int h, int mode) {
  const int x4 = (blockIdx.x * 64 + threadIdx.x) * 4;
  const int y = blockIdx.y * 4 + threadIdx.y;
  if (y >= h || x4 >= w) return;
  const uint32_t* p = reinterpret_cast<const uint32_t*>(bgr + blockIdx.z * src_stride + (size_t)y * src_pitch + 3 * x4);
  const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);   // B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
  const uint32_t o0 = gray_from_bgr(w0 & 0xff, (w0 >> 8) & 0xff, (w0 >> 16) & 0xff, mode);
  const uint32_t o1 = gray_from_bgr(w0 >> 24, w1 & 0xff, (w1 >> 8) & 0xff, mode);
  const uint32_t o2 = gray_from_bgr((w1 >> 16) & 0xff, w1 >> 24, w2 & 0xff, mode);
  const uint32_t o3 = gray_from_bgr((w2 >> 8) & 0xff, (w2 >> 16) & 0xff, w2 >> 24, mode);
  // columns >= w of the last quad fall into the row padding (pitch = align16(w)): harmless
  *reinterpret_cast<uint32_t*>(pyr + blockIdx.z * pyr_stride + (size_t)y * dst_pitch + x4) =
      o0 | (o1 << 8) | (o2 << 16) | (o3 << 24);
}

void launch_bgr2gray(const uint8_t* bgr, size_t src_stride, int src_pitch, uint8_t* pyr, size_t pyr_stride,
                     int dst_pitch, int w, int h, int n_img, int mode, cudaStream_t s, int* launches) {
  dim3 grid((w + 255) / 256, (h + 3) / 4, n_img), block(64, 4);
  k_bgr2gray<<<grid, block, 0, s>>>(bgr, src_stride, src_pitch, pyr, pyr_stride, dst_pitch, w, h, mode);
  ++*launches;
}

// ------------------------------------------------------------------ K1: pyramid level l from l-1
__global__ void __launch_bounds__(256)
k_resize(const __grid_constant__ OrbGeom g, const uint8_t* __restrict__ src0, size_t src_stride,
         uint8_t* __restrict__ pyr, const int2* __restrict__ taps, int level) {
  const LevelGeom& d = g.lv[level];
  const LevelGeom& s = g.lv[level - 1];
  const int img = blockIdx.z;
  const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int y = blockIdx.y * 8 + threadIdx.y;
  if (y >= d.h || x4 >= d.pitch) return;
  const uint8_t* src = src0 + (size_t)img * src_stride;      // source level of this image (level 0 may be the caller's buffer)
  uint8_t* dst = pyr + (size_t)img * g.pyr_stride + d.off;
  const int2 ty = taps[d.ytab + y];
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = min(x4 + k, d.w - 1);
    const int2 tx = taps[d.xtab + x];
    out |= (uint32_t)resize_px(src, s.pitch, s.w, s.h, tx.x, tx.y, ty.x, ty.y) << (8 * k);
  }
  *reinterpret_cast<uint32_t*>(dst + (size_t)y * d.pitch + x4) = out;
}

// Fast path for scale factors <= 4/3 (ORB's 1.1 / 1.2): one thread = 4 adjacent output columns x a
// strip of kResizeRows output rows.  The 4 x 2 horizontal taps of the thread lie within 8
// consecutive source bytes, so one source row costs three aligned word loads, two funnel shifts,
// and per output column one PRMT (pick the two neighbours) + one DP2A (c0 * a + c1 * b in 8.8).
// Row results are kept in registers while the strip walks down: consecutive output rows share
// a source row, so ~1.1 horizontal passes per output row instead of 2.  Same integer
// arithmetic as resize_px (orb_math.cuh), bit for bit.
constexpr int kResizeRows = 8;

__global__ void __launch_bounds__(256)
k_resize_strip(const __grid_constant__ OrbGeom g, const uint8_t* __restrict__ src0, size_t src_stride,
               uint8_t* __restrict__ pyr, const int2* __restrict__ taps, int level) {
  const LevelGeom& d = g.lv[level];
  const LevelGeom& s = g.lv[level - 1];
  const int img = blockIdx.z;
  const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
  const int y_first = (blockIdx.y * 8 + threadIdx.y) * kResizeRows;
  if (y_first >= d.h || x4 >= d.pitch) return;
  const uint8_t* src = src0 + (size_t)img * src_stride;      // source level of this image (level 0 may be the caller's buffer)
  uint8_t* dst = pyr + (size_t)img * g.pyr_stride + d.off;
  // horizontal taps of the four columns
  uint32_t sel[4], wgt[4];
  int xbase;
  {
    const int2 t0 = taps[d.xtab + min(x4, d.w - 1)];
    xbase = t0.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int2 t = taps[d.xtab + min(x4 + k, d.w - 1)];
      const uint32_t r = (uint32_t)(t.x - xbase);          // 0 .. 6 (checked on the host)
      sel[k] = r | ((r + 1u) << 4) | 0x8800u;               // bytes r, r+1 of the window; upper bytes unused
      wgt[k] = (256u - (uint32_t)t.y) | ((uint32_t)t.y << 16);   // {c0, c1} as two u16
    }
  }
  const int abase = xbase & ~3, sh8 = (xbase & 3) * 8;
  const int sp = s.pitch, last_row = s.h - 1;
  // the three words of a source row are requested one row ahead of their use, and the vertical tap of the next
  // output row one row ahead too: the strip is otherwise bound by load latency (ncu stall_long_sb 65 %)
  auto fetch = [&](int row, uint32_t* r) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (size_t)min(row, last_row) * sp + abase);
    r[0] = __ldg(p); r[1] = __ldg(p + 1); r[2] = __ldg(p + 2);
  };
  auto hpass = [&](const uint32_t* r, uint32_t* h) {
    const uint32_t lo = __funnelshift_r(r[0], r[1], sh8), hi = __funnelshift_r(r[1], r[2], sh8);
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __dp2a_lo(wgt[k], __byte_perm(lo, hi, sel[k]), 0u);
  };
  const int y_end = min(y_first + kResizeRows, d.h);
  int2 ty = __ldg(&taps[d.ytab + y_first]);
  int ra = ty.x;
  uint32_t hA[4], hB[4], nx[3];
  {
    uint32_t r0[3], r1[3];
    fetch(ra, r0); fetch(ra + 1, r1); fetch(ra + 2, nx);
    hpass(r0, hA); hpass(r1, hB);
  }
  for (int y = y_first; y < y_end; ++y) {
    const int2 ty_next = __ldg(&taps[d.ytab + min(y + 1, d.h - 1)]);
    // 16.16 result in bits 16 .. 23; no clamp needed: (256 * 256 * 255 + 32768) >> 16 == 255
    const uint32_t cy1 = (uint32_t)ty.y, cy0 = 256u - cy1;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = cy0 * hA[k] + (cy1 * hB[k] + 32768u);
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * d.pitch + x4) =
        __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
    // advance the two-row window to the next output row's source rows (by one row, sometimes two)
    if (y + 1 < y_end) {
      while (ra < ty_next.x) {
#pragma unroll
        for (int k = 0; k < 4; ++k) hA[k] = hB[k];
        hpass(nx, hB);
        ++ra;
        fetch(ra + 2, nx);
      }
    }
    ty = ty_next;
  }
}


constexpr int kResizeSrcRows = 11;          // source rows of an 8-row output strip at scale <= ~1.25

// Same strip, source rows staged by the TMA engine: one WARP = 128 output columns x kResizeRows output rows; its
// <= kResizeSrcRows source rows (<= kResizeTmaW bytes each, from the 16-byte-aligned column at or left of lane 0's
// first tap) arrive by bulk copies on the warp's own mbarrier; the horizontal pass reads its three words of a row at
// the lane's own offset.  No CTA-wide barrier.  History (B200, 256 pairs, 11 launches): k_resize_strip on every level
// 2.20 ms; rows staged per thread with 33 LDG + 33 STS and two horizontal passes per output row 1.82 ms (bound by just
// that: l1tex 78 %, LSU pipe 56 %, ncu; 4-byte cp.async instead of LDG + STS: no better); TMA rows 1.68 ms; + one
// horizontal pass per staged row (40 registers instead of 60) 1.46 ms.
constexpr int kResizeTmaW = 192;
struct __align__(16) ResizeWarpSmem {
  uint8_t rows[kResizeSrcRows][kResizeTmaW];
  uint64_t bar;
  uint64_t pad;
};

__global__ void __launch_bounds__(256)
k_resize_tma(const __grid_constant__ OrbGeom g, const uint8_t* __restrict__ src0, size_t src_stride,
             uint8_t* __restrict__ pyr, const int2* __restrict__ taps, int level) {
  __shared__ ResizeWarpSmem s_rz[8];
  const LevelGeom& d = g.lv[level];
  const LevelGeom& s = g.lv[level - 1];
  const int img = blockIdx.z, lane = threadIdx.x, warp = threadIdx.y;
  const int x4 = (blockIdx.x * 32 + lane) * 4;
  const int y_first = (blockIdx.y * 8 + warp) * kResizeRows;
  if (y_first >= d.h) return;                                  // the whole warp
  ResizeWarpSmem& S = s_rz[warp];
  const uint8_t* src = src0 + (size_t)img * src_stride;
  uint8_t* dst = pyr + (size_t)img * g.pyr_stride + d.off;
  const int y_end = min(y_first + kResizeRows, d.h);
  int2 ty = __ldg(&taps[d.ytab + y_first]);
  const int ra0 = ty.x;
  const int n_src = __ldg(&taps[d.ytab + y_end - 1]).x + 2 - ra0;      // <= kResizeSrcRows (host-checked)
  const int sp = s.pitch, last_row = s.h - 1;
  const int astart = __ldg(&taps[d.xtab + min(blockIdx.x * 128, d.w - 1)]).x & ~15;   // lane 0's first tap, aligned down
  const uint32_t row_bytes = (uint32_t)min(kResizeTmaW, sp - astart);
  if (lane == 0) {
    mbar_init(&S.bar, 1);
    mbar_arrive_expect_tx(&S.bar, row_bytes * (uint32_t)n_src);
  }
  __syncwarp();
  RPE_ASSERT(n_src <= kResizeSrcRows && row_bytes % 16 == 0 && row_bytes > 0);
  if (lane < n_src) bulk_copy_g2s(S.rows[lane], src + (size_t)min(ra0 + lane, last_row) * sp + astart, row_bytes, &S.bar);
  // horizontal / vertical taps while the rows are in flight
  const int xbase = __ldg(&taps[d.xtab + min(x4, d.w - 1)]).x;
  const int aoff = (xbase & ~3) - astart, sh8 = (xbase & 3) * 8;
  uint32_t sel[4], wgt[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int2 t = __ldg(&taps[d.xtab + min(x4 + k, d.w - 1)]);
    const uint32_t r = (uint32_t)(t.x - xbase);          // 0 .. 6 (checked on the host)
    sel[k] = r | ((r + 1u) << 4) | 0x8800u;
    wgt[k] = (256u - (uint32_t)t.y) | ((uint32_t)t.y << 16);
  }
  mbar_wait(&S.bar, 0);
  if (x4 >= d.pitch) return;
  // (bytes past row_bytes - at the right edge of the source level - are only ever multiplied into columns >= d.w)
  RPE_ASSERT(aoff >= 0 && aoff + 12 <= kResizeTmaW);
  auto hpass = [&](int i, uint32_t* h) {
    const uint32_t* r = reinterpret_cast<const uint32_t*>(S.rows[i] + aoff);
    const uint32_t r0 = r[0], r1 = r[1], r2 = r[2];
    const uint32_t lo = __funnelshift_r(r0, r1, sh8), hi = __funnelshift_r(r1, r2, sh8);
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = __dp2a_lo(wgt[k], __byte_perm(lo, hi, sel[k]), 0u);
  };
  // Every staged row takes ONE horizontal pass (not two per output row: 16 for <= 11 rows).  The walk is
  // static over the staged rows; the first tap row of the output rows is strictly increasing (host-checked:
  // resize_rows_strict), so staged row i closes at most one output row, the one whose first tap is i - 1.  All
  // decisions are warp-uniform.
  uint8_t* out = dst + (size_t)y_first * d.pitch + x4;
  uint32_t hrow[2][4];
  hpass(0, hrow[0]);
  int y = y_first;                                           // next output row
#pragma unroll
  for (int i = 1; i < kResizeSrcRows; ++i) {
    if (i < n_src) {
      hpass(i, hrow[i & 1]);
      if (y < y_end && ty.x - ra0 == i - 1) {
        const uint32_t cy1 = (uint32_t)ty.y, cy0 = 256u - cy1;
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = cy0 * hrow[(i - 1) & 1][k] + (cy1 * hrow[i & 1][k] + 32768u);
        *reinterpret_cast<uint32_t*>(out) = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
        out += d.pitch;
        ++y;
        ty = __ldg(&taps[d.ytab + min(y, d.h - 1)]);
      }
    }
  }
  RPE_ASSERT(y == y_end);
}

static PyrView pyr_view(const OrbBuffers& b) { return PyrView{b.pyr, b.lvl0, b.lvl0_stride}; }

void launch_pyramid(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches) {
  for (int l = 1; l < g.nlevels; ++l) {
    const LevelGeom& d = g.lv[l];
    // the source level: inside the pyramid slab, or - level 0 of device-resident frames used in place - the caller's buffer
    const bool ext = l == 1 && b.lvl0 != nullptr;
    const uint8_t* src0 = ext ? b.lvl0 : b.pyr + g.lv[l - 1].off;
    const size_t sst = ext ? b.lvl0_stride : g.pyr_stride;
    if (g.resize_strip_ok && g.resize_span <= kResizeSrcRows && g.resize_xspan <= kResizeTmaW && g.resize_rows_strict) {
      dim3 grid((d.pitch / 4 + 31) / 32, (d.h + 8 * kResizeRows - 1) / (8 * kResizeRows), n_img), block(32, 8);
      k_resize_tma<<<grid, block, 0, s>>>(g, src0, sst, b.pyr, b.taps, l);
    } else if (g.resize_strip_ok) {
      dim3 grid((d.pitch / 4 + 31) / 32, (d.h + 8 * kResizeRows - 1) / (8 * kResizeRows), n_img), block(32, 8);
      k_resize_strip<<<grid, block, 0, s>>>(g, src0, sst, b.pyr, b.taps, l);
    } else {
      dim3 grid((d.pitch / 4 + 31) / 32, (d.h + 7) / 8, n_img), block(32, 8);
      k_resize<<<grid, block, 0, s>>>(g, src0, sst, b.pyr, b.taps, l);
    }
    ++*launches;
  }
}

// ------------------------------------------------------------------ K2: FAST-9 + 3x3 NMS + border
// One WARP = one 120x14 tile of one level of one image (flat tile table: ftab).  Warps are independent: own slice
// of shared memory (6.3 KB: 32 warps per SM), own mbarrier, no CTA-wide barrier.  The scored region (tile + 1 px
// for the NMS neighbours) is 16 rows x 32 aligned pixel quads: lane = quad column, the warp walks down the rows.
//  (1) TMA: one bulk copy per image row of the tile (+halo) onto the warp's mbarrier (UBLKCP), issued by the lanes;
//      the score tile is cleared while the rows are in flight;
//  (2) packed pre-test, four pixels per lane and row (fast_compass4_unpacked): the column word of an image row
//      serves as D, C and U of three score rows (7-row register window).  The pass bits of the lane's 16 x 4 pixels
//      collect in two registers (one per band of 8 rows);
//  (3) compaction per CANDIDATE, not per pixel: one warp scan over the lanes' candidate counts, then every lane
//      expands its own set bits into the warp's queue (round 1 spent 133 instructions per thread and tile on
//      per-pixel predicated stores: 9.5 per pixel; ncu);
//  (4) one lane per queued pixel: FAST-9 corner test + score in one packed min/max network (fast9_score) -> score
//      tile; the corners are compacted in place at the head of the queue (ballot);
//  (5) one lane per corner: strict 3x3 NMS on the score tile, 31-px border filter; the survivors' keys
//      (score<<24 | y<<12 | x) collect in shared memory (over the image rows, no longer needed) and go to the
//      level's candidate list with ONE global atomicAdd per tile (a returning atomic per corner left the warp
//      waiting on L2 round trips: 30 % of the stall samples of the first version of this kernel), + histogram.
// A tile with more candidates than the queue holds (dense noise) is processed in half-bands of 4 rows and its NMS
// scans the score tile instead of a corner list.
// Measured on B200, 512 images 1280x720, 12 levels (profiles/): one CTA of 8 warps per 120x30 tile, per-pixel
// predicated queue stores (round 1): 3.78 ms at 44.6 thread-instructions per pixel.  One warp per tile, round 2:
// 120x30 tiles (12 warps / SM) with a returning global atomic per corner 5.67 ms; 120x14 tiles (32 warps / SM), one
// slot reservation per tile 3.83 ms; + three-input minima in the score network, queue codes decoded by the consumer
// 3.55 ms; + pass bits rotated across lanes, straight-line NMS 3.32 ms.  Tile height with everything else equal:
// 14 / 22 / 30 rows (32 / 24 / 20 warps per SM) 3.32 / 3.43 / 3.86 ms: occupancy outweighs the smaller halo.
constexpr int kFastIP = 144;                 // staged bytes per image row (16-B aligned start; columns ib-3 .. ib+130 used)
constexpr int kFastIH = kFastTH + 8;         // image rows y0-4 .. y0+TH+3
constexpr int kFastSP = 128;                 // score tile pitch: x0-4 .. x0+123
constexpr int kFastSH = kFastTH + 2;         // score rows y0-1 .. y0+TH
constexpr int kFastIPW = kFastIP / 4;
constexpr int kFastBands = kFastSH / 8;
static_assert(kFastSH % 8 == 0 && kFastTW + 8 == kFastSP, "score rows in bands of 8, 32 quads per row");
constexpr int kFastWarps = 4;                // independent warps (tiles) per CTA
constexpr int kFastRot = 5;                  // row sr's pass bits are held by lane (quad column + kFastRot * sr) % 32
constexpr int kFastQueue = 512;              // queue entries per warp; a half-band of 4 rows x 128 px always fits

// unpack a column word into its two u16x2 halves (volatile: keep the order of the window updates as written)
__device__ __forceinline__ void fast_unpack_keep(uint32_t w, uint32_t& lo, uint32_t& hi) {
  asm volatile("prmt.b32 %0, %1, 0, 0x4140;" : "=r"(lo) : "r"(w));
  asm volatile("prmt.b32 %0, %1, 0, 0x4342;" : "=r"(hi) : "r"(w));
}

struct __align__(16) FastWarpSmem {
  uint8_t img[kFastIH * kFastIP];
  uint8_t score[kFastSH * kFastSP];
  uint16_t queue[kFastQueue];
  uint64_t bar;
  uint32_t n_keep, pad;
};
static_assert(sizeof(FastWarpSmem) % 16 == 0, "warp slices stay 16-byte aligned");
static_assert(kFastIH * kFastIP >= 4 * ((kFastTH + 1) / 2) * ((kFastTW + 1) / 2), "kept keys (NMS maxima: at most one pixel in four) fit over the image rows");

// (8 CTAs per SM, 61 registers; asking for 9 CTAs, 56 registers, changes nothing: 3.32 / 3.31 ms)
__global__ void __launch_bounds__(kFastWarps * 32, 8)
k_fast(const __grid_constant__ OrbGeom g, const PyrView pv, const uint32_t* __restrict__ ftab,
       uint32_t* __restrict__ cand_key, uint32_t* __restrict__ cand_cnt, uint32_t* __restrict__ hist) {
  extern __shared__ __align__(128) uint8_t s_fast[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int item = blockIdx.x * kFastWarps + warp, img = blockIdx.y;
  if (item >= g.total_ftiles) return;                       // warps are independent: no CTA-wide barrier below
  FastWarpSmem& S = *reinterpret_cast<FastWarpSmem*>(s_fast + warp * sizeof(FastWarpSmem));
  const uint32_t te = __ldg(ftab + item);                   // level<<24 | ty<<12 | tx
  const int level = te >> 24;
  const LevelGeom& L = g.lv[level];
  const int x0 = kFastOrigin + (int)(te & 0xfff) * kFastTW, y0 = kFastOrigin + (int)((te >> 12) & 0xfff) * kFastTH;
  const int astart = (x0 - 8) & ~15;
  const int pitch = L.pitch, lw = L.w, lh = L.h;
  const uint8_t* src = pv.level(g, img, level) + (size_t)(y0 - 4) * pitch + astart;

  // (1) rows y0-4 .. y0+TH+3, columns astart .. astart+143, clipped to the level buffer.  Bytes outside the level
  //     are never read for a pixel that can become a candidate
  const uint32_t row_bytes = (uint32_t)min(kFastIP, pitch - astart);
  const int n_rows = min(kFastIH, lh - (y0 - 4));
  if (lane == 0) {
    S.n_keep = 0;
    mbar_init(&S.bar, 1);
    mbar_arrive_expect_tx(&S.bar, row_bytes * (uint32_t)n_rows);
  }
  __syncwarp();
  RPE_ASSERT(level < g.nlevels && n_rows > 0 && row_bytes > 0 && row_bytes % 16 == 0 && y0 - 4 >= 0 && astart >= 0 &&
             astart + (int)row_bytes <= pitch && (x0 - 4 - astart == 4 || x0 - 4 - astart == 12));
  for (int r = lane; r < n_rows; r += 32) bulk_copy_g2s(S.img + r * kFastIP, src + (size_t)r * pitch, row_bytes, &S.bar);
#pragma unroll
  for (int i = 0; i < kFastSH * kFastSP / (16 * 32); ++i) reinterpret_cast<uint4*>(S.score)[i * 32 + lane] = make_uint4(0, 0, 0, 0);
  const int thr = g.fast_threshold;
  const uint32_t t4 = (uint32_t)min(thr, 127) * 0x01010101u;
  // scores are only ever needed for x in [30, w-30), y in [30, h-30)
  const int xlo = max(30, x0 - 1), xhi = min(lw - 30, x0 + kFastTW + 1);
  const int ylo = max(30, y0 - 1), yhi = min(lh - 30, y0 + kFastTH + 1);
  const int ib = x0 - 4 - astart;                            // staged byte column of score column 0 (4 or 12)
  const int xq = x0 - 4 + 4 * lane;                          // first pixel of this lane's quad
  const int lo = min(max(xlo - xq, 0), 4), hi = min(max(xhi - xq, 0), 4);
  const uint32_t vmask = (uint32_t)(0x80808080ull << (8 * lo)) & (uint32_t)(0x80808080ull >> (8 * (4 - hi)));
  const uint32_t* wimg = reinterpret_cast<const uint32_t*>(S.img) + (ib >> 2) + lane;   // this lane's word of image row 0
  mbar_wait(&S.bar, 0);

  // (2) score row sr = image rows sr (U), sr + 3 (C), sr + 6 (D); window slot of image row r = r % 7.  Straight
  //     line: rows outside [ylo, yhi) (first / last tile row of a level only) are computed and masked out
  uint32_t wl[7], wh[7];
#pragma unroll
  for (int r = 0; r < 6; ++r) fast_unpack_keep(wimg[r * kFastIPW], wl[r], wh[r]);
  const uint32_t kt = fast_compass_kt(t4);
  const int sr_lo = ylo - (y0 - 1), sr_n = yhi - ylo;       // valid score rows: sr_lo <= sr < sr_lo + sr_n
  uint32_t m[kFastBands];                                    // bit 7 - k of byte j of m[b] = score row 8 b + k, pixel j of the quad
#pragma unroll
  for (int b = 0; b < kFastBands; ++b) m[b] = 0u;
#pragma unroll
  for (int sr = 0; sr < kFastSH; ++sr) {
    fast_unpack_keep(wimg[(sr + 6) * kFastIPW], wl[(sr + 6) % 7], wh[(sr + 6) % 7]);
    const uint32_t* row = wimg + (sr + 3) * kFastIPW;
    const uint32_t rmask = (uint32_t)(sr - sr_lo) < (uint32_t)sr_n ? vmask : 0u;
    uint32_t cw = fast_compass4_unpacked(wl[(sr + 3) % 7], wh[(sr + 3) % 7], wl[sr % 7], wh[sr % 7],
                                         wl[(sr + 6) % 7], wh[(sr + 6) % 7], row[-1], row[0], row[1], kt) & rmask;
    // candidates cluster (the worst lane of a band held ~22 of the band's ~75): the pass bits of row sr go to lane
    // (lane + 5 sr) % 32, so that a cluster spreads over many lanes and the lanes' expansion loops below, which run
    // to the worst lane's count, stay short.  The consumer undoes the rotation when it decodes a queue entry
    cw = __shfl_sync(0xffffffffu, cw, (lane - kFastRot * sr) & 31);
    m[sr >> 3] |= cw >> (sr & 7);
  }

  const uint32_t lt = (1u << lane) - 1u;
  // queue the set bits of the bands in mm[]; returns the warp's total (<= kFastQueue by the caller's choice of mm)
  auto push = [&](const uint32_t* mm) -> int {
    int c = 0;
#pragma unroll
    for (int b = 0; b < kFastBands; ++b) c += __popc(mm[b]);
    int inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    // queue entry = band << 10 | lane << 5 | q with q = clz of the bit: the lanes' loops are divergent (they run to
    // the worst lane's count), so they only find the bit and store this code; (row, column) are decoded by the
    // consumer, where every lane is busy.  Bit 31 - q = pixel 3 - (q >> 3) of a quad, row 8 band + (q & 7); the
    // quad's column is (lane - kFastRot * row) % 32
    uint16_t* out = S.queue + (inc - c);
#pragma unroll
    for (int b = 0; b < kFastBands; ++b) {
      uint32_t bits = mm[b];
      const uint32_t code0 = (uint32_t)(b << 10) | (uint32_t)(lane << 5);
      while (bits) {
        const int q = __clz(bits);
        bits &= ~(0x80000000u >> q);
        RPE_ASSERT(out >= S.queue && out < S.queue + kFastQueue);
        *out++ = (uint16_t)(code0 + q);
      }
    }
    __syncwarp();
    return __shfl_sync(0xffffffffu, inc, 31);
  };
  // (4) corner test + score of the queued pixels; corners are compacted in place (they never outrun the reads:
  //     n_corner <= entries consumed, and a round's entries are in registers before its ballot)
  auto score = [&](int total) -> int {
    int n_corner = 0;
    for (int base = 0; base < total; base += 32) {
      const int i = base + lane;
      int e = 0, sc = 0;
      if (i < total) {
        const int code = S.queue[i];
        const int row = ((code >> 10) << 3) | (code & 7);
        e = (row << 7) | (4 * (((code >> 5) - kFastRot * row) & 31) + 3 - ((code >> 3) & 3));          // row << 7 | column
        RPE_ASSERT(e < kFastSH * kFastSP && ((e >> 7) + 6) * kFastIP + ib + (e & 127) + 3 < kFastIH * kFastIP &&
                   (e >> 7) + 6 < n_rows);
        sc = fast9_score(&S.img[((e >> 7) + 3) * kFastIP + ib + (e & 127)], kFastIP, thr);
      }
      const uint32_t bal = __ballot_sync(0xffffffffu, sc > 0);
      if (sc > 0) {
        S.score[e] = (uint8_t)sc;
        S.queue[n_corner + __popc(bal & lt)] = (uint16_t)e;
      }
      n_corner += __popc(bal);
    }
    __syncwarp();
    return n_corner;
  };
  // (5) NMS + border filter for the corner at score-tile offset e; survivors' keys -> shared memory
  uint32_t* keep_keys = reinterpret_cast<uint32_t*>(S.img);  // (all scores are in the score tile by now)
  auto nms = [&](int e) {
    const int sr = e >> 7, c = e & 127;
    const int x = x0 - 4 + c, y = y0 - 1 + sr;
    // inside the tile proper (not its 1-px NMS rim) and at least 31 px from the level border; straight-line
    const bool inside = (unsigned)(sr - 1) < (unsigned)kFastTH && (unsigned)(c - 4) < (unsigned)kFastTW &&
                        (unsigned)(x - 31) < (unsigned)(lw - 62) && (unsigned)(y - 31) < (unsigned)(lh - 62);
    if (!inside) return;
    RPE_ASSERT(e - kFastSP - 1 >= 0 && e + kFastSP + 1 < kFastSH * kFastSP);
    const uint8_t* p = &S.score[e];
    const int s = p[0];
    const int hi = max(max(max(p[-kFastSP - 1], p[-kFastSP]), max(p[-kFastSP + 1], p[-1])),
                       max(max(p[1], p[kFastSP - 1]), max(p[kFastSP], p[kFastSP + 1])));
    if (s > hi) {                                            // strict maximum of its 3 x 3 neighbourhood
      const uint32_t slot = atomicAdd(&S.n_keep, 1u);
      RPE_ASSERT(4 * (slot + 1) <= (uint32_t)(kFastIH * kFastIP));
      keep_keys[slot] = ((uint32_t)s << 24) | ((uint32_t)y << 12) | (uint32_t)x;
    }
  };

  int all = 0;
#pragma unroll
  for (int b = 0; b < kFastBands; ++b) all += __popc(m[b]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) all += __shfl_xor_sync(0xffffffffu, all, o);
  if (all == 0) return;
  if (all <= kFastQueue) {
    const int n_corner = score(push(m));
    for (int i = lane; i < n_corner; i += 32) nms(S.queue[i]);
  } else {
    // dense tile: half-bands of 4 rows (at most 4 x 128 = kFastQueue pixels), then NMS over the score tile
#pragma unroll 1
    for (int hb = 0; hb < 2 * kFastBands; ++hb) {
      uint32_t one[kFastBands];
#pragma unroll
      for (int b = 0; b < kFastBands; ++b) one[b] = (hb >> 1) == b ? (m[b] & ((hb & 1) ? 0x0f0f0f0fu : 0xf0f0f0f0u)) : 0u;
      score(push(one));
    }
    for (int sr = 1; sr <= kFastTH; ++sr) {
      const uint32_t w = reinterpret_cast<const uint32_t*>(S.score)[sr * (kFastSP / 4) + lane];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if ((w >> (8 * j)) & 0xffu) nms(sr * kFastSP + 4 * lane + j);
    }
  }
  __syncwarp();
  // one slot reservation per tile, then coalesced key stores + the score histogram
  const uint32_t n_keep = S.n_keep;
  if (n_keep == 0) return;
  uint32_t* cnt = cand_cnt + img * g.nlevels + level;
  uint32_t* hst = hist + (size_t)(img * g.nlevels + level) * 256;
  uint32_t* out_key = cand_key + (size_t)img * g.cand_stride + L.cand_off;
  const uint32_t cap = (uint32_t)L.cand_cap;
  uint32_t base = 0;
  if (lane == 0) base = atomicAdd(cnt, n_keep);
  base = __shfl_sync(0xffffffffu, base, 0);
  for (uint32_t i = lane; i < n_keep; i += 32) {
    const uint32_t key = keep_keys[i];
    RPE_ASSERT((key & 0xfff) < (uint32_t)lw && ((key >> 12) & 0xfff) < (uint32_t)lh && (key >> 24) > 0);
    if (base + i < cap) out_key[base + i] = key;
    atomicAdd(&hst[key >> 24], 1u);
  }
}

void launch_fast(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches) {
  if (g.total_ftiles == 0) return;
  dim3 grid((g.total_ftiles + kFastWarps - 1) / kFastWarps, n_img);
  k_fast<<<grid, kFastWarps * 32, kFastWarps * sizeof(FastWarpSmem), s>>>(g, pyr_view(b), b.ftab, b.cand_key, b.cand_cnt, b.hist);
  ++*launches;
}

// ------------------------------------------------------------------ K3/K4: retainBest x2 + Harris
// One CTA per (level, image).  (1) FAST-score threshold for 2*quota from the histogram (ties
// kept), (2) Harris response of the survivors, (3) exact quota-th largest response by 4-pass
// radix select (ties kept), (4) the kept keypoints put in (y, x) order by rows -> canonical order.
__device__ __forceinline__ uint32_t f32_orderable(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int kSelThreads = 512;

// harris_response (orb_math.cuh) with the 9 x 9 neighbourhood fetched as 27 aligned words requested up front (one
// memory round trip per candidate instead of one per row of the three-row sliding window) and the same integer sums.
// c = the candidate pixel in a level whose rows start 16-byte aligned; x = its column.
__device__ __forceinline__ float harris_response_words(const uint8_t* c, int pitch, int x) {
  const int sh8 = ((x - 4) & 3) * 8;
  const uint8_t* p0 = c - 4 - ((x - 4) & 3) - 4 * pitch;      // aligned word holding column x - 4, row -4
  uint32_t w[9][3];
#pragma unroll
  for (int r = 0; r < 9; ++r) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(p0 + r * pitch);
    w[r][0] = __ldg(p); w[r][1] = __ldg(p + 1); w[r][2] = __ldg(p + 2);
  }
  int a = 0, b = 0, cc = 0;
  int r0[9], r1[9], r2[9];
  auto unpack = [&](int r, int* o) {
    const uint32_t lo = __funnelshift_r(w[r][0], w[r][1], sh8), mid = __funnelshift_r(w[r][1], w[r][2], sh8);
    const uint32_t hi = w[r][2] >> sh8;
    o[0] = lo & 0xff; o[1] = (lo >> 8) & 0xff; o[2] = (lo >> 16) & 0xff; o[3] = lo >> 24;
    o[4] = mid & 0xff; o[5] = (mid >> 8) & 0xff; o[6] = (mid >> 16) & 0xff; o[7] = mid >> 24;
    o[8] = hi & 0xff;
  };
  unpack(0, r0);
  unpack(1, r1);
#pragma unroll
  for (int y = 2; y < 9; ++y) {
    unpack(y, r2);
#pragma unroll
    for (int i = 1; i < 8; ++i) {
      const int Ix = (r1[i + 1] - r1[i - 1]) * 2 + (r0[i + 1] - r0[i - 1]) + (r2[i + 1] - r2[i - 1]);
      const int Iy = (r2[i] - r0[i]) * 2 + (r2[i - 1] - r0[i - 1]) + (r2[i + 1] - r0[i + 1]);
      a += Ix * Ix;
      b += Iy * Iy;
      cc += Ix * Iy;
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) { r0[i] = r1[i]; r1[i] = r2[i]; }
  }
  return harris_from_sums(a, b, cc);
}

// Called by the 32 lanes of one warp: the largest bin b of a 256-bin histogram with sum(hist[b..255]) >= k, and
// k minus the count above b (b = 0 when the whole histogram holds fewer than k).  Every lane returns the result.
__device__ __forceinline__ void warp_select_bin(const uint32_t* hist, uint32_t k, uint32_t* bin, uint32_t* k_rest) {
  const int lane = threadIdx.x & 31;
  uint32_t loc[8], sum = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { loc[j] = hist[255 - 8 * lane - j]; sum += loc[j]; }   // lane 0 owns the top bins
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  const uint32_t exc = inc - sum, total = __shfl_sync(0xffffffffu, inc, 31);
  uint32_t b = 0, r = k - (total - loc[7]);      // fallback (lane 31: loc[7] = hist[0])
  const bool mine = total >= k && exc < k && inc >= k;
  if (mine) {
    uint32_t cum = exc;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (cum + loc[j] >= k) { b = 255 - 8 * lane - j; r = k - cum; break; }
      cum += loc[j];
    }
  }
  const uint32_t who = __ballot_sync(0xffffffffu, mine);
  const int src_lane = who ? (__ffs(who) - 1) : 31;
  *bin = __shfl_sync(0xffffffffu, b, src_lane);
  *k_rest = __shfl_sync(0xffffffffu, r, src_lane);
}

__global__ void __launch_bounds__(kSelThreads, 2)
k_select(const __grid_constant__ OrbGeom g, const PyrView pv,
         const uint32_t* __restrict__ cand_key, const uint32_t* __restrict__ cand_cnt,
         const uint32_t* __restrict__ hist, uint32_t* __restrict__ sel_key, float* __restrict__ sel_resp,
         uint32_t* __restrict__ lvl_key, float* __restrict__ lvl_resp, uint32_t* __restrict__ lvl_cnt,
         uint32_t* __restrict__ img_flags) {
  extern __shared__ __align__(16) unsigned long long s_sort[];   // g.sort_cap kept (key, response) entries, then the row tables of (4)
  __shared__ uint32_t s_hist[256];
  __shared__ uint32_t s_cnt, s_thr1, s_prefix, s_k, s_keep;
  const int level = blockIdx.x, img = blockIdx.y;
  const LevelGeom& L = g.lv[level];
  const int tid = threadIdx.x;
  const uint32_t raw = cand_cnt[img * g.nlevels + level];
  const uint32_t n = min(raw, (uint32_t)L.cand_cap);
  if (raw > (uint32_t)L.cand_cap && tid == 0) atomicOr(&img_flags[img], 1u);
  const uint32_t quota = (uint32_t)L.quota;
  const uint8_t* src = pv.level(g, img, level);
  const uint32_t* ck = cand_key + (size_t)img * g.cand_stride + L.cand_off;
  uint32_t* sk = sel_key + (size_t)img * g.cand_stride + L.cand_off;
  float* sr = sel_resp + (size_t)img * g.cand_stride + L.cand_off;

  // (1) threshold on the FAST score for retainBest(2 * quota): the histogram goes to shared memory and warp 0
  //     finds the cut (a one-thread scan of 256 global words cost more than the Harris pass)
  if (tid < 256) s_hist[tid] = hist[(size_t)(img * g.nlevels + level) * 256 + tid];
  if (tid == 0) { s_thr1 = 0; s_cnt = 0; }
  __syncthreads();
  if (tid < 32 && n > 2 * quota) {
    uint32_t rest;
    warp_select_bin(s_hist, 2 * quota, &s_thr1, &rest);
  }
  __syncthreads();
  const uint32_t thr1 = s_thr1;
  // (2) survivors are compacted first, so that every lane of a warp has a Harris response to compute
  //     (about one candidate in ten survives: computing it in place left 3 of 32 lanes busy)
  if (quota > 0) {
    for (uint32_t i = tid; i < n; i += kSelThreads) {
      const uint32_t key = ck[i];
      if ((key >> 24) >= thr1) sk[atomicAdd(&s_cnt, 1u)] = key & 0xffffffu;
    }
  }
  __syncthreads();
  const uint32_t n1 = s_cnt;
  for (uint32_t j = tid; j < n1; j += kSelThreads) {
    const uint32_t key = sk[j];
    const int x = key & 0xfff, y = (key >> 12) & 0xfff;
    sr[j] = harris_response_words(src + (size_t)y * L.pitch + x, L.pitch, x);
  }
  __syncthreads();
  // (3) quota-th largest response (exact), ties kept
  uint32_t thr_key = 0;
  if (n1 > quota) {
    if (tid == 0) { s_prefix = 0; s_k = quota; }
    for (int pass = 3; pass >= 0; --pass) {
      const int shift = pass * 8;
      const uint32_t mask = (pass == 3) ? 0u : (0xffffffffu << (shift + 8));
      if (tid < 256) s_hist[tid] = 0;
      __syncthreads();
      const uint32_t prefix = s_prefix;
      for (uint32_t i = tid; i < n1; i += kSelThreads) {
        const uint32_t k = f32_orderable(sr[i]);
        if ((k & mask) == (prefix & mask)) atomicAdd(&s_hist[(k >> shift) & 255u], 1u);
      }
      __syncthreads();
      if (tid < 32) {
        uint32_t b, k;
        warp_select_bin(s_hist, s_k, &b, &k);
        __syncwarp();
        if (tid == 0) {
          s_k = k;
          s_prefix = prefix | (b << shift);
        }
      }
      __syncthreads();
    }
    thr_key = s_prefix;
  }
  // (4) keep, put in (y, x) order, write.  The order is found by ROWS, not by a comparison sort: count the kept keys of
  //     every image row, scan the counts into row offsets, drop every key into its row's segment, and rank it inside the
  //     segment by x (rows hold one or two keypoints on average).  Five short passes and one block scan instead of the
  //     66 barrier-separated stages of a 2048-element bitonic sort (ncu: 40 % of this kernel's instructions, 25 % of its
  //     stall samples on the barriers).
  uint32_t* s_rcnt = reinterpret_cast<uint32_t*>(s_sort + g.sort_cap);      // [row_cap] kept keys per row
  uint32_t* s_roff = s_rcnt + g.sel_row_cap;                                 // [row_cap] segment start; segment END once filled
  uint16_t* s_tmp = reinterpret_cast<uint16_t*>(s_roff + g.sel_row_cap);     // [sort_cap] kept-list index per segment slot
  __shared__ uint32_t s_wsum[kSelThreads / 32];
  if (tid == 0) s_keep = 0;
  for (int r = tid; r < L.h; r += kSelThreads) s_rcnt[r] = 0;
  __syncthreads();
  for (uint32_t i = tid; i < n1; i += kSelThreads) {
    const float r = sr[i];
    if (f32_orderable(r) < thr_key) continue;
    const uint32_t j = atomicAdd(&s_keep, 1u);
    if (j < (uint32_t)g.sort_cap) {
      const uint32_t key = sk[i];
      s_sort[j] = ((unsigned long long)key << 32) | __float_as_uint(r);
      RPE_ASSERT((int)(key >> 12) < L.h);
      atomicAdd(&s_rcnt[key >> 12], 1u);
    }
  }
  __syncthreads();
  const uint32_t nk_raw = s_keep;
  const uint32_t nk = min(nk_raw, (uint32_t)L.kp_cap);
  const uint32_t n_sorted = min(nk_raw, (uint32_t)g.sort_cap);
  if (nk_raw > (uint32_t)L.kp_cap && tid == 0) atomicOr(&img_flags[img], 2u);
  {
    // exclusive scan of the row counts: consecutive rows per thread, warp scan, scan of the warp totals
    const int per = (L.h + kSelThreads - 1) / kSelThreads, r0 = tid * per;
    uint32_t sum = 0;
    for (int k = 0; k < per; ++k) sum += (r0 + k < L.h) ? s_rcnt[r0 + k] : 0u;
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if ((tid & 31) >= o) inc += v;
    }
    if ((tid & 31) == 31) s_wsum[tid >> 5] = inc;
    __syncthreads();
    uint32_t base = inc - sum;
    for (int w = 0; w < (tid >> 5); ++w) base += s_wsum[w];
    for (int k = 0; k < per; ++k)
      if (r0 + k < L.h) { s_roff[r0 + k] = base; base += s_rcnt[r0 + k]; }
  }
  __syncthreads();
  for (uint32_t i = tid; i < n_sorted; i += kSelThreads) {
    const uint32_t y = (uint32_t)(s_sort[i] >> 44);
    s_tmp[atomicAdd(&s_roff[y], 1u)] = (uint16_t)i;
  }
  __syncthreads();
  uint32_t* ok = lvl_key + (size_t)img * g.kp_stride + L.kp_off;
  float* orr = lvl_resp + (size_t)img * g.kp_stride + L.kp_off;
  for (uint32_t i = tid; i < n_sorted; i += kSelThreads) {
    const unsigned long long v = s_sort[i];
    const uint32_t key = (uint32_t)(v >> 32), y = key >> 12;
    const uint32_t cnt = s_rcnt[y], base = s_roff[y] - cnt;
    uint32_t rank = 0;
    for (uint32_t k = 0; k < cnt; ++k) rank += (uint32_t)(s_sort[s_tmp[base + k]] >> 32) < key;
    const uint32_t pos = base + rank;                         // keys are unique: a permutation of 0 .. n_sorted - 1
    if (pos < nk) { ok[pos] = key; orr[pos] = __uint_as_float((uint32_t)v); }
  }
  if (tid == 0) lvl_cnt[img * g.nlevels + level] = nk;
}

size_t select_smem_bytes(int sort_cap, int row_cap) { return (size_t)sort_cap * 10 + (size_t)row_cap * 8; }

void launch_select(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches) {
  dim3 grid(g.nlevels, n_img);
  k_select<<<grid, kSelThreads, select_smem_bytes(g.sort_cap, g.sel_row_cap), s>>>(g, pyr_view(b), b.cand_key, b.cand_cnt, b.hist, b.sel_key,
                                                             b.sel_resp, b.lvl_key, b.lvl_resp, b.lvl_cnt,
                                                             b.img_flags);
  ++*launches;
}

// ------------------------------------------------------------------ K6: 7x7 sigma-2 blur (f32 FMA separable)
// One WARP = one work item (btab): 128 columns x TWO row strips of kBlurRows rows that are walked in
// lockstep, so that every f32 operation is a packed f32x2 instruction (FFMA2 / FADD2 / FMUL2 of sm_100:
// lane 0 = strip A, lane 1 = strip B; each lane is an IEEE round-to-nearest f32 operation, so the results
// are those of the scalar blur_row7 / blur_col7_sum in orb_math.cuh bit for bit).
//  (1) TMA: the warp's lanes issue one bulk copy per input row (rows y-3 .. y+n+2, reflected at the top /
//      bottom of the level by choosing the SOURCE row; 16 bytes of halo left and right) onto the warp's own
//      mbarrier: no registers, no software prefetch (ncu on the register-prefetch version: 18 % of the
//      issued instructions were MOVs, stall_long_sb 25 %);
//  (2) edge warps mirror three columns (reflect-101) into the halo in shared memory, so the row loop has
//      no edge cases at all;
//  (3) per row: six LDS.32 (pixels x-4 .. x+7 of both strips), ten u8 -> f32 conversions with the 2^23
//      mantissa trick (PRMT + FADD2: no I2F), the row pass for four columns into a 7-row register window,
//      the column pass out of that window, round-to-nearest-even by the 1.5 * 2^23 trick, one STG.32 per
//      strip.  The loop is unrolled by 7, so the window needs no register moves.
constexpr int kBlurSP = 160;                      // staged bytes per row: pixels x0 - 16 .. x0 + 143
constexpr int kBlurSR = 6 + 7 * ((kBlurRows + 6) / 7);   // staged row slots per strip (loop runs in groups of 7)
constexpr int kBlurWarpBytes = 2 * kBlurSR * kBlurSP;

__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 u8x2_to_f32(uint32_t wa, uint32_t wb, uint32_t sel) {
  // bytes {b, 0, 0, 0x4b} = 2^23 + b exactly; subtracting 2^23 is exact
  return __fadd2_rn(f2(__uint_as_float(__byte_perm(wa, 0x4b000000u, sel)),
                       __uint_as_float(__byte_perm(wb, 0x4b000000u, sel))), f2(-8388608.0f, -8388608.0f));
}
__device__ __forceinline__ float2 blur_row7_x2(const float2* v) {
  float2 s = __fmul2_rn(f2(RPE_BLUR_K0, RPE_BLUR_K0), v[0]);
  s = __ffma2_rn(f2(RPE_BLUR_K1, RPE_BLUR_K1), v[1], s);
  s = __ffma2_rn(f2(RPE_BLUR_K2, RPE_BLUR_K2), v[2], s);
  s = __ffma2_rn(f2(RPE_BLUR_K3, RPE_BLUR_K3), v[3], s);
  s = __ffma2_rn(f2(RPE_BLUR_K2, RPE_BLUR_K2), v[4], s);
  s = __ffma2_rn(f2(RPE_BLUR_K1, RPE_BLUR_K1), v[5], s);
  s = __ffma2_rn(f2(RPE_BLUR_K0, RPE_BLUR_K0), v[6], s);
  return s;
}
__device__ __forceinline__ float2 blur_col7_x2(float2 r0, float2 r1, float2 r2, float2 r3, float2 r4, float2 r5,
                                               float2 r6) {
  float2 s = __fmul2_rn(f2(RPE_BLUR_K3, RPE_BLUR_K3), r3);
  s = __ffma2_rn(f2(RPE_BLUR_K2, RPE_BLUR_K2), __fadd2_rn(r2, r4), s);
  s = __ffma2_rn(f2(RPE_BLUR_K1, RPE_BLUR_K1), __fadd2_rn(r1, r5), s);
  s = __ffma2_rn(f2(RPE_BLUR_K0, RPE_BLUR_K0), __fadd2_rn(r0, r6), s);
  return s;
}

// row pass of one staged row of both strips for this lane's four columns
__device__ __forceinline__ void blur_rows_x2(const uint8_t* rowA, const uint8_t* rowB, float2* out) {
  const uint32_t a0 = *reinterpret_cast<const uint32_t*>(rowA), a1 = *reinterpret_cast<const uint32_t*>(rowA + 4),
                 a2 = *reinterpret_cast<const uint32_t*>(rowA + 8);
  const uint32_t b0 = *reinterpret_cast<const uint32_t*>(rowB), b1 = *reinterpret_cast<const uint32_t*>(rowB + 4),
                 b2 = *reinterpret_cast<const uint32_t*>(rowB + 8);
  float2 v[10];   // pixels x-3 .. x+6
  v[0] = u8x2_to_f32(a0, b0, 0x7441); v[1] = u8x2_to_f32(a0, b0, 0x7442); v[2] = u8x2_to_f32(a0, b0, 0x7443);
  v[3] = u8x2_to_f32(a1, b1, 0x7440); v[4] = u8x2_to_f32(a1, b1, 0x7441); v[5] = u8x2_to_f32(a1, b1, 0x7442);
  v[6] = u8x2_to_f32(a1, b1, 0x7443);
  v[7] = u8x2_to_f32(a2, b2, 0x7440); v[8] = u8x2_to_f32(a2, b2, 0x7441); v[9] = u8x2_to_f32(a2, b2, 0x7442);
#pragma unroll
  for (int c = 0; c < 4; ++c) out[c] = blur_row7_x2(v + c);
}

// (128 registers, 4 CTAs per SM.  Capping the registers for 5 / 6 CTAs per SM spills the 7-row window: 2.48 / 2.74 ms
// against 2.21 ms; B200, 512 images 1280x720.)
__global__ void __launch_bounds__(128, 4)
k_blur(const __grid_constant__ OrbGeom g, const PyrView pv, uint8_t* __restrict__ blur,
       const uint32_t* __restrict__ btab) {
  extern __shared__ __align__(128) uint8_t s_blur[];
  __shared__ __align__(8) uint64_t s_bar[4];
  const int img = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int item = blockIdx.x * 4 + warp;
  if (item >= g.total_bitems) return;                       // warps are independent: no CTA-wide barrier below
  const uint32_t e = __ldg(btab + item);                    // level<<24 | strip pair<<12 | segment
  const LevelGeom& L = g.lv[e >> 24];
  const int w = L.w, h = L.h, pitch = L.pitch;
  const int x0 = (int)(e & 0xfff) * 128;
  const int yA = (int)((e >> 12) & 0xfff) * (2 * kBlurRows), yB = yA + kBlurRows;
  const int nA = min(kBlurRows, h - yA), nB = min(kBlurRows, h - yB);   // nB may be <= 0: strip B idle
  const uint8_t* src = pv.level(g, img, (int)(e >> 24));
  uint8_t* sA = s_blur + warp * kBlurWarpBytes;
  uint8_t* sB = sA + kBlurSR * kBlurSP;
  uint64_t* bar = &s_bar[warp];

  // (1) stage rows: global columns [gx0, gx1) land at shared column gx0 - (x0 - 16)
  const int gx0 = max(x0 - 16, 0), gx1 = min(x0 + 144, pitch);
  const uint32_t rb = (uint32_t)(gx1 - gx0);
  const int rowsA = nA + 6, rowsB = nB > 0 ? nB + 6 : 0;
  if (lane == 0) {
    mbar_init(bar, 1);
    mbar_arrive_expect_tx(bar, rb * (uint32_t)(rowsA + rowsB));
  }
  __syncwarp();
  for (int r = lane; r < rowsA + rowsB; r += 32) {
    const bool isB = r >= rowsA;
    const int rr = isB ? r - rowsA : r;
    const int sy = min(max(reflect101((isB ? yB : yA) - 3 + rr, h), 0), h - 1);
    RPE_ASSERT(rr < kBlurSR && gx0 - (x0 - 16) >= 0 && gx0 - (x0 - 16) + (int)rb <= kBlurSP && rb % 16 == 0 && gx1 <= pitch);
    bulk_copy_g2s((isB ? sB : sA) + rr * kBlurSP + (gx0 - (x0 - 16)), src + (size_t)sy * pitch + gx0, rb, bar);
  }
  mbar_wait(bar, 0);
  // (2) reflect-101 columns for the warps at the left / right border of the level
  if (x0 == 0 || x0 + 128 + 3 > w) {
    for (int r = lane; r < rowsA + rowsB; r += 32) {
      uint8_t* row = (r >= rowsA ? sB + (r - rowsA) * kBlurSP : sA + r * kBlurSP) + 16 - x0;   // row[p] = pixel p
      if (x0 == 0) { row[-1] = row[1]; row[-2] = row[2]; row[-3] = row[3]; }
      if (x0 + 128 + 3 > w) {
#pragma unroll
        for (int k = 0; k < 3; ++k) row[w + k] = row[max(w - 2 - k, 0)];
      }
    }
    __syncwarp();
  }
  // (3) rows
  const int x = x0 + 4 * lane;
  const uint8_t* pA = sA + 12 + 4 * lane;                  // pixels x - 4 .. x + 7 of a staged row
  const uint8_t* pB = sB + 12 + 4 * lane;
  // the blurred level is stored in STRIPS (kBlurStripW columns x all rows, strip after strip; a 128-byte line = 8 rows
  // of one strip) for k_describe's rotated gathers: see blur_strip_offset
  uint8_t* dA = blur + (size_t)img * g.pyr_stride + L.off + blur_strip_offset(x, yA, h);
  uint8_t* dB = dA + (size_t)kBlurRows * kBlurStripW;
  const bool col_ok = x < w;
  float2 win[7][4];
#pragma unroll
  for (int i = 0; i < 6; ++i) blur_rows_x2(pA + i * kBlurSP, pB + i * kBlurSP, win[i]);
  for (int base = 0; base < nA; base += 7) {
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int o = base + i;                               // output row of the strip; its last input row is o + 6
      const int slot = (6 + i) % 7;
      blur_rows_x2(pA + (o + 6) * kBlurSP, pB + (o + 6) * kBlurSP, win[slot]);
      uint32_t oa[4], ob[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float2 sum = blur_col7_x2(win[(slot + 1) % 7][c], win[(slot + 2) % 7][c], win[(slot + 3) % 7][c],
                                        win[(slot + 4) % 7][c], win[(slot + 5) % 7][c], win[(slot + 6) % 7][c],
                                        win[slot][c]);
        // 0 <= sum <= 255: adding 1.5 * 2^23 leaves rint(sum) (ties to even) in the low byte
        const float2 q = __fadd2_rn(sum, f2(12582912.0f, 12582912.0f));
        oa[c] = __float_as_uint(q.x);
        ob[c] = __float_as_uint(q.y);
      }
      if (col_ok && o < nA) {
        const uint32_t lo = __byte_perm(oa[0], oa[1], 0x0040), hi = __byte_perm(oa[2], oa[3], 0x0040);
        *reinterpret_cast<uint32_t*>(dA + o * kBlurStripW) = __byte_perm(lo, hi, 0x5410);
      }
      if (col_ok && o < nB) {
        const uint32_t lo = __byte_perm(ob[0], ob[1], 0x0040), hi = __byte_perm(ob[2], ob[3], 0x0040);
        *reinterpret_cast<uint32_t*>(dB + o * kBlurStripW) = __byte_perm(lo, hi, 0x5410);
      }
    }
  }
}

void launch_blur(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches) {
  if (g.total_bitems == 0) return;
  dim3 grid((g.total_bitems + 3) / 4, n_img);
  k_blur<<<grid, 128, 4 * kBlurWarpBytes, s>>>(g, pyr_view(b), b.blur, b.btab);
  ++*launches;
}

// ------------------------------------------------------------------ K5/K7: orientation + rBRIEF + keypoint fields
// One CTA (4 warps) = 32 consecutive keypoints of one image; one WARP per keypoint for the two
// gather-heavy parts, so that every load instruction of a warp stays inside one patch:
//  (0) threads 0..31: locate keypoint (level, index in level), write x / y / size / response / octave;
//  (1) warp per keypoint, lane = patch column u = lane - 15: integer moments m10, m01 over the
//      31 rows of the circular patch (31 coalesced byte loads), warp-reduced, fastAtan2;
//  (2) threads 0..31: steering coefficients (double cos / sin), one keypoint per thread;
//  (3) warp per keypoint, lane = descriptor byte: its 8 tests (pattern slice held in registers for
//      the whole CTA), 16 gathers from the blurred level, one coalesced 32-byte store.
constexpr int kDescKp = 32;   // keypoints per CTA

// orientation weights: [alignment sh = (x - 15) & 3][patch row 0..30][word 0..9] -> {s8x4 column offsets u (0 outside
// the disc |u| <= umax[|v|]), u8x4 disc mask}; byte b of word w is column u = 4 w + b - 15 - sh.  Filled at setup.
static __device__ uint2 kOriTab_dev[4 * 31 * 10];

__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

static void build_ori_table(uint2* tab) {
  for (int sh = 0; sh < 4; ++sh)
    for (int r = 0; r < 31; ++r)
      for (int w = 0; w < 10; ++w) {
        const int v = r - 15, d = kUmax_host[v < 0 ? -v : v];
        uint32_t cu = 0, cm = 0;
        for (int b = 0; b < 4; ++b) {
          const int u = 4 * w + b - 15 - sh;
          if (u >= -d && u <= d) {
            cu |= (uint32_t)(uint8_t)(int8_t)u << (8 * b);
            cm |= 1u << (8 * b);
          }
        }
        tab[(sh * 31 + r) * 10 + w] = make_uint2(cu, cm);
      }
}

// brief_test (orb_math.cuh) against the strip layout of the blurred level.  Same non-fused rotation; cvRound by the
// 1.5 * 2^23 trick (the low mantissa bits of d + 1.5 * 2^23 are rint(d), ties to even, in two's complement: no F2I on
// the XU pipe), so X = cx + ix is one integer add on the float's bits (cxm = cx - kRintBits).
constexpr int kRintBits = 0x4B400000;             // bits of 12582912.0f = 1.5 * 2^23
__device__ __forceinline__ uint32_t brief_test_strips(const uint8_t* base, int strip_skip, int cxm, int cym, float a,
                                                      float b, uint32_t t) {
  const float xa = (float)(int8_t)(t & 0xff), ya = (float)(int8_t)((t >> 8) & 0xff);
  const float xb = (float)(int8_t)((t >> 16) & 0xff), yb = (float)(int8_t)(t >> 24);
  const int Xa = __float_as_int(f32_add(f32_sub(f32_mul(xa, a), f32_mul(ya, b)), 12582912.0f)) + cxm;
  const int Ya = __float_as_int(f32_add(f32_add(f32_mul(xa, b), f32_mul(ya, a)), 12582912.0f)) + cym;
  const int Xb = __float_as_int(f32_add(f32_sub(f32_mul(xb, a), f32_mul(yb, b)), 12582912.0f)) + cxm;
  const int Yb = __float_as_int(f32_add(f32_add(f32_mul(xb, b), f32_mul(yb, a)), 12582912.0f)) + cym;
  // (X >> 4) * strip_bytes + (X & 15) = X + (X >> 4) * (strip_bytes - 16): strip_skip = strip_bytes - kBlurStripW
  const int va = base[(Xa >> kBlurStripLog) * strip_skip + Xa + Ya * kBlurStripW];
  const int vb = base[(Xb >> kBlurStripLog) * strip_skip + Xb + Yb * kBlurStripW];
  return (uint32_t)(va < vb);
}

// (8 CTAs per SM = 64 registers.  Measured, 512 images x 5000 keypoints: no bound given to the compiler, 124 registers,
// 2.45 ms; 8 CTAs 1.87 ms; 9 CTAs, 56 registers, 1.89 ms; 10 / 12 / 16 CTAs spill: 4.30 / 2.39 / 2.89 ms.)
__global__ void __launch_bounds__(128, 8)
k_describe(const __grid_constant__ OrbGeom g, const PyrView pv, const uint8_t* __restrict__ blur,
           const uint32_t* __restrict__ lvl_key, const float* __restrict__ lvl_resp,
           const uint32_t* __restrict__ lvl_cnt, float* __restrict__ kp_x, float* __restrict__ kp_y,
           float* __restrict__ kp_size, float* __restrict__ kp_angle, float* __restrict__ kp_resp,
           int32_t* __restrict__ kp_oct, uint8_t* __restrict__ desc, int32_t* __restrict__ kp_n) {
  __shared__ int s_level[kDescKp], s_x[kDescKp], s_y[kDescKp], s_cx[kDescKp], s_cy[kDescKp];
  __shared__ float s_a[kDescKp], s_b[kDescKp];
  __shared__ int s_total;
  const int img = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int i0 = blockIdx.x * kDescKp;
  // (0) locate
  if (tid < kDescKp) {
    const int i = i0 + tid;
    int level = -1, base = 0, total = 0;
    for (int l = 0; l < g.nlevels; ++l) {
      const int c = (int)lvl_cnt[img * g.nlevels + l];
      if (level < 0 && i < total + c) { level = l; base = total; }
      total += c;
    }
    if (i == 0) kp_n[img] = total;
    if (tid == 0) s_total = total;
    s_level[tid] = level;
    if (level >= 0) {
      const LevelGeom& L = g.lv[level];
      const int j = i - base;
      const uint32_t key = lvl_key[(size_t)img * g.kp_stride + L.kp_off + j];
      const int x = key & 0xfff, y = (key >> 12) & 0xfff;
      s_x[tid] = x; s_y[tid] = y;
      const float px = f32_mul((float)x, L.scale), py = f32_mul((float)y, L.scale);
      const size_t o = (size_t)img * g.kp_stride + i;
      kp_x[o] = px;
      kp_y[o] = py;
      kp_size[o] = f32_mul(31.0f, L.scale);
      kp_resp[o] = lvl_resp[(size_t)img * g.kp_stride + L.kp_off + j];
      kp_oct[o] = level;
      // descriptor centre exactly as computeOrbDescriptors: cvRound(pt * (1.f / scale))
      const float inv = f32_div(1.0f, L.scale);
      s_cx[tid] = f32_rint(f32_mul(px, inv));
      s_cy[tid] = f32_rint(f32_mul(py, inv));
    }
  }
  __syncthreads();
  const int nk = min(kDescKp, s_total - i0);
  if (nk <= 0) return;
  // this lane's 8 tests (32 bytes of the pattern table)
  uint32_t pat[8];
  {
    const uint4* pp = reinterpret_cast<const uint4*>(kBriefPattern_dev + 32 * lane);
    const uint4 p0 = pp[0], p1 = pp[1];
    pat[0] = p0.x; pat[1] = p0.y; pat[2] = p0.z; pat[3] = p0.w;
    pat[4] = p1.x; pat[5] = p1.y; pat[6] = p1.z; pat[7] = p1.w;
  }
  // (1) orientation: integer moments of the radius-15 disc.  A lane owns one aligned word of one patch
  //     row: 10 words cover columns x-15 .. x+15 at any alignment, 3 rows per warp instruction, 11 steps for the
  //     31 rows; the per-byte column weights u (zero outside the disc) and the 0/1 disc mask come from a table
  //     indexed by (alignment, row, word), so a step is one LDG.32, one table load, two DP4A and one IMAD
  //     (the byte-per-lane form spent 47 % of the kernel's instructions here).
  //     Two keypoints per iteration (both sets of loads in flight together); lane j keeps the sums of the warp's
  //     j-th keypoint, so that the angle (fastAtan2: two divisions, a polynomial) and the steering coefficients
  //     (double-precision sincos) of all eight keypoints of the warp are computed ONCE, in eight lanes side by side
  {
    const int rsub = lane / 10, w = lane - 10 * rsub;      // lanes 30, 31 idle
    int my_m10 = 0, my_m01 = 0;
    for (int k = warp, j = 0; k < nk; k += 8, j += 2) {
      const int kb = (k + 4 < nk) ? k + 4 : k;              // second keypoint (or the first again: result unused)
      const LevelGeom& La = g.lv[s_level[k]];
      const LevelGeom& Lb = g.lv[s_level[kb]];
      const int xA = s_x[k], xB = s_x[kb];
      const uint8_t* ca = pv.level(g, img, s_level[k]) + (size_t)(s_y[k] - 15 + rsub) * La.pitch + ((xA - 15) & ~3) + 4 * w;
      const uint8_t* cb = pv.level(g, img, s_level[kb]) + (size_t)(s_y[kb] - 15 + rsub) * Lb.pitch + ((xB - 15) & ~3) + 4 * w;
      const uint2* ta = kOriTab_dev + (((xA - 15) & 3) * 31 + rsub) * 10 + w;
      const uint2* tb = kOriTab_dev + (((xB - 15) & 3) * 31 + rsub) * 10 + w;
      int a10 = 0, a01 = 0, b10 = 0, b01 = 0;
      if (lane < 30) {
        uint32_t pa[11], pb[11];
#pragma unroll
        for (int it = 0; it < 11; ++it) {
          const bool ok = it < 10 || rsub == 0;             // row 3 it + rsub <= 30
          pa[it] = ok ? __ldg(reinterpret_cast<const uint32_t*>(ca + (size_t)(3 * it) * La.pitch)) : 0u;
          pb[it] = ok ? __ldg(reinterpret_cast<const uint32_t*>(cb + (size_t)(3 * it) * Lb.pitch)) : 0u;
        }
#pragma unroll
        for (int it = 0; it < 11; ++it) {
          const int r = 3 * it + rsub;
          if (it < 10 || rsub == 0) {
            const uint2 t = __ldg(ta + 30 * it), u = __ldg(tb + 30 * it);
            a10 = dp4a_u8s8(pa[it], t.x, a10);                     // u8 pixels x s8 column offsets
            a01 += (r - 15) * (int)__dp4a(pa[it], t.y, 0u);        // row offset x masked row sum
            b10 = dp4a_u8s8(pb[it], u.x, b10);
            b01 += (r - 15) * (int)__dp4a(pb[it], u.y, 0u);
          }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        a10 += __shfl_xor_sync(0xffffffffu, a10, o);
        a01 += __shfl_xor_sync(0xffffffffu, a01, o);
        b10 += __shfl_xor_sync(0xffffffffu, b10, o);
        b01 += __shfl_xor_sync(0xffffffffu, b01, o);
      }
      if (lane == j) { my_m10 = a10; my_m01 = a01; }
      if (lane == j + 1) { my_m10 = b10; my_m01 = b01; }
    }
    // (2) angle + steering coefficients of the warp's keypoints, one per lane
    if (lane < kDescKp / 4) {
      const int k = warp + 4 * lane;
      if (k < nk) {
        const float ang = fast_atan2_deg((float)my_m01, (float)my_m10);
        kp_angle[(size_t)img * g.kp_stride + i0 + k] = ang;
        brief_steer(ang, &s_a[k], &s_b[k]);
      }
    }
  }
  // a warp only ever touches its own keypoints (k = warp, warp + 4, ...): warp-level synchronisation is enough, and
  // the four warps of the CTA drift through the phases independently
  __syncwarp();
  // (3) descriptor: lanes = descriptor bytes, 8 steered tests each, gathered straight from the blurred level.
  //     (Staging the keypoint's 37 x 37 window in shared memory first - 13 row-contiguous word loads per lane, then
  //     LDS gathers - was measured and lost: 2.02 ms against 1.87 ms for 512 images x 5000 keypoints; the extra
  //     staging instructions and the lower occupancy, 80 registers, cost more than the scattered sectors.)
  //     Two / four keypoints of a warp in flight together (loop unrolled): 1.92 / 1.95 ms.
  for (int k = warp; k < nk; k += 4) {
    const LevelGeom& L = g.lv[s_level[k]];
    const uint8_t* base = blur + (size_t)img * g.pyr_stride + L.off;
    RPE_ASSERT(s_cx[k] >= 19 && s_cx[k] < L.w - 19 && s_cy[k] >= 19 && s_cy[k] < L.h - 19);   // pattern radius 18.4
    const float a = s_a[k], b = s_b[k];
    const int strip_skip = (L.h - 1) * kBlurStripW;
    const int cxm = s_cx[k] - kRintBits, cym = s_cy[k] - kRintBits;
    uint32_t byte = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) byte |= brief_test_strips(base, strip_skip, cxm, cym, a, b, pat[j]) << j;
    desc[((size_t)img * g.kp_stride + i0 + k) * 32 + lane] = (uint8_t)byte;
  }
}

void launch_describe(const OrbGeom& g, const OrbBuffers& b, int n_img, cudaStream_t s, int* launches) {
  dim3 grid((g.kp_stride + kDescKp - 1) / kDescKp, n_img);
  k_describe<<<grid, 128, 0, s>>>(g, pyr_view(b), b.blur, b.lvl_key, b.lvl_resp, b.lvl_cnt, b.kp_x, b.kp_y, b.kp_size,
                                  b.kp_angle, b.kp_resp, b.kp_oct, b.desc, b.kp_n);
  ++*launches;
}

int setup_orb_attributes(int select_smem) {
  static uint2 ori[4 * 31 * 10];
  build_ori_table(ori);
  if (cudaMemcpyToSymbol(kOriTab_dev, ori, sizeof ori) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(k_blur, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * kBlurWarpBytes) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(k_fast, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kFastWarps * sizeof(FastWarpSmem))) != cudaSuccess) return 1;
  if (cudaFuncSetAttribute(k_fast, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess) return 1;
  return raise_dynamic_smem_limit((const void*)k_select, select_smem);
}

}  // namespace rpe
