// orb_math.cuh — per-pixel / per-keypoint arithmetic of the ORB stage, bit-exact with what
// cv2.ORB.detectAndCompute computes (reference call site: src/core/pose_estimator.py:85-91,
// :108; arithmetic spec: SURVEY.md Appendix A.1-A.8).  All functions take a pointer to the
// centre pixel and a row pitch so they work on global memory and shared-memory tiles alike.
#pragma once
#include "rpe_hd.cuh"

namespace rpe {

// ------------------------------------------------------------------ ingest: BGR -> gray
// The two gray paths of the reference differ and are matched separately:
//   mode 0 = cv2.cvtColor(BGR2GRAY) after cv2.imread (src/utils/image_loader.py:23-28):
//            (3735 B + 19235 G + 9798 R + 2^14) >> 15;
//   mode 1 = cv2.imread(path, IMREAD_GRAYSCALE) of a colour PNG (src/run_single_pair.py:31-32), where
//            libpng converts with truncated 15-bit weights: (9797 R + 19234 G + 3737 B) >> 15.
// Both probed bit-exact against cv2 4.13 (tests/test_oracle_gray.py).
RPE_HD uint32_t gray_from_bgr(uint32_t b, uint32_t g, uint32_t r, int mode) {
  return mode == 0 ? (3735u * b + 19235u * g + 9798u * r + 16384u) >> 15
                   : (3737u * b + 19234u * g + 9797u * r) >> 15;
}

// ------------------------------------------------------------------ A.1 pyramid (resize)
// One INTER_LINEAR_EXACT output pixel from host-built per-axis tap tables: ofs = left/top
// source index, c1 = 8.8 weight of the right/bottom neighbour (c0 = 256 - c1).
RPE_HD uint8_t resize_px(const uint8_t* src, int pitch, int sw, int sh, int x0, int cx1, int y0,
                         int cy1) {
  const int x1 = imin(x0 + 1, sw - 1), y1 = imin(y0 + 1, sh - 1);
  const uint32_t cx0 = 256u - (uint32_t)cx1, cy0 = 256u - (uint32_t)cy1;
  const uint8_t* r0 = src + (size_t)y0 * pitch;
  const uint8_t* r1 = src + (size_t)y1 * pitch;
  const uint32_t h0 = cx0 * r0[x0] + (uint32_t)cx1 * r0[x1];  // 8.8
  const uint32_t h1 = cx0 * r1[x0] + (uint32_t)cx1 * r1[x1];
  const uint32_t v = cy0 * h0 + (uint32_t)cy1 * h1;           // 16.16
  const uint32_t o = (v + 32768u) >> 16;
  return (uint8_t)(o > 255u ? 255u : o);
}

// ------------------------------------------------------------------ A.2 FAST-9/16
// Exact corner test: 9 contiguous ring pixels all brighter than v + t or all darker than v - t.
RPE_HD bool fast9_is_corner(const uint8_t* c, int pitch, int t) {
  const int v = c[0];
  int d[16];
#define RPE_LD(k, dx, dy) d[k] = v - (int)c[(dy) * pitch + (dx)]
  RPE_LD(0, 0, 3);  RPE_LD(1, 1, 3);   RPE_LD(2, 2, 2);    RPE_LD(3, 3, 1);
  RPE_LD(4, 3, 0);  RPE_LD(5, 3, -1);  RPE_LD(6, 2, -2);   RPE_LD(7, 1, -3);
  RPE_LD(8, 0, -3); RPE_LD(9, -1, -3); RPE_LD(10, -2, -2); RPE_LD(11, -3, -1);
  RPE_LD(12, -3, 0); RPE_LD(13, -3, 1); RPE_LD(14, -2, 2); RPE_LD(15, -1, 3);
#undef RPE_LD
  uint32_t mp = 0, mn = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    mp |= (uint32_t)(d[k] > t) << k;
    mn |= (uint32_t)(d[k] < -t) << k;
  }
  // 9 contiguous set bits on the 16-ring
  uint32_t xp = mp | (mp << 16), xn = mn | (mn << 16);
  uint32_t ap = xp & (xp >> 1); ap &= ap >> 2; ap &= ap >> 4; ap &= xp >> 8;
  uint32_t an = xn & (xn >> 1); an &= an >> 2; an &= an >> 4; an &= xn >> 8;
  return ((ap | an) & 0xffffu) != 0;
}

// Score of a pixel KNOWN to be a corner: max(a, b) - 1 with a = max over the 16 arcs of min(d),
// b = max over arcs of min(-d) = -(min over arcs of max(d)).  Always >= threshold.
RPE_HD int fast9_corner_score(const uint8_t* c, int pitch) {
  const int v = c[0];
  int d[16];
#define RPE_LD(k, dx, dy) d[k] = v - (int)c[(dy) * pitch + (dx)]
  RPE_LD(0, 0, 3);  RPE_LD(1, 1, 3);   RPE_LD(2, 2, 2);    RPE_LD(3, 3, 1);
  RPE_LD(4, 3, 0);  RPE_LD(5, 3, -1);  RPE_LD(6, 2, -2);   RPE_LD(7, 1, -3);
  RPE_LD(8, 0, -3); RPE_LD(9, -1, -3); RPE_LD(10, -2, -2); RPE_LD(11, -3, -1);
  RPE_LD(12, -3, 0); RPE_LD(13, -3, 1); RPE_LD(14, -2, 2); RPE_LD(15, -1, 3);
#undef RPE_LD
  int lo2[16], hi2[16], lo4[16], hi4[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    lo2[k] = imin(d[k], d[(k + 1) & 15]);
    hi2[k] = imax(d[k], d[(k + 1) & 15]);
  }
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    lo4[k] = imin(lo2[k], lo2[(k + 2) & 15]);
    hi4[k] = imax(hi2[k], hi2[(k + 2) & 15]);
  }
  int a = -1000, bmin = 1000;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int lo9 = imin(imin(lo4[k], lo4[(k + 4) & 15]), d[(k + 8) & 15]);
    const int hi9 = imax(imax(hi4[k], hi4[(k + 4) & 15]), d[(k + 8) & 15]);
    a = imax(a, lo9);
    bmin = imin(bmin, hi9);
  }
  return imax(a, -bmin) - 1;
}

// returns 0 for a non-corner, else the corner score (scalar reference form; tests only)
RPE_HD int fast9_score_scalar(const uint8_t* c, int pitch, int t) {
  return fast9_is_corner(c, pitch, t) ? fast9_corner_score(c, pitch) : 0;
}

// Packed pre-test on four horizontally adjacent pixels (bytes of C = pixels x .. x+3; Lw / Rw
// = the words left / right of C on the same row; U / D = the words at x .. x+3 three rows up /
// down).  Returns 0x80 in byte j iff pixel j passes both opposite-pair tests on the compass
// points of the ring, (0,+-3) and (+-3,0), for one polarity — a necessary condition for a
// 9-arc, since an arc of 9 contains at least one point of every opposite pair:
//   bright: min(max(N,S), max(E,W)) > v + t      dark: max(min(N,S), min(E,W)) < v - t.
// Two pixels per register as u16 lanes (VIMNMX.U16x2 is native on sm_100; byte-wise min / max /
// compare are not): the two comparisons become one 32-bit add / subtract each whose lane bit 15
// is the answer (lanes stay within [0x7e00, 0x80ff]: no carry or borrow crosses lanes).
// t4 = the threshold replicated in the four bytes (threshold < 128).
RPE_HD uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(a, b, sel);
#else
  const uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
  return r;
#endif
}
RPE_HD uint32_t umax16x2(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __vmaxu2(a, b);
#else
  const uint32_t lo = (a & 0xffff) > (b & 0xffff) ? (a & 0xffff) : (b & 0xffff);
  const uint32_t hi = (a >> 16) > (b >> 16) ? (a >> 16) : (b >> 16);
  return lo | (hi << 16);
#endif
}
RPE_HD uint32_t umin16x2(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
  return __vminu2(a, b);
#else
  const uint32_t lo = (a & 0xffff) < (b & 0xffff) ? (a & 0xffff) : (b & 0xffff);
  const uint32_t hi = (a >> 16) < (b >> 16) ? (a >> 16) : (b >> 16);
  return lo | (hi << 16);
#endif
}
RPE_HD uint32_t umin3_16x2(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  return __vimin3_u16x2(a, b, c);
#else
  return umin16x2(umin16x2(a, b), c);
#endif
}
RPE_HD uint32_t umax3_16x2(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  return __vimax3_u16x2(a, b, c);
#else
  return umax16x2(umax16x2(a, b), c);
#endif
}

// Corner test + score in one pass, both polarities at once: lane 0 of P[k] = 256 + d_k, lane 1 =
// 256 - d_k (d_k = v - p_k; one IMAD per ring pixel: p * 0xffff + base).  With
//   a = max over the 16 arcs of min(d over 9 contiguous k),  b = the same for -d,
// the pixel is a corner iff max(a, b) > t and its score is max(a, b) - 1.
// Returns 0 for a non-corner, else the score (>= t >= 1).
RPE_HD int fast9_score(const uint8_t* c, int pitch, int t) {
  const uint32_t v = c[0];
  const uint32_t base = (256u + v) | ((256u - v) << 16);
  uint32_t P[16];
#define RPE_LD(k, dx, dy) P[k] = (uint32_t)c[(dy) * pitch + (dx)] * 0xffffu + base
  RPE_LD(0, 0, 3);  RPE_LD(1, 1, 3);   RPE_LD(2, 2, 2);    RPE_LD(3, 3, 1);
  RPE_LD(4, 3, 0);  RPE_LD(5, 3, -1);  RPE_LD(6, 2, -2);   RPE_LD(7, 1, -3);
  RPE_LD(8, 0, -3); RPE_LD(9, -1, -3); RPE_LD(10, -2, -2); RPE_LD(11, -3, -1);
  RPE_LD(12, -3, 0); RPE_LD(13, -3, 1); RPE_LD(14, -2, 2); RPE_LD(15, -1, 3);
#undef RPE_LD
  // min over the 9 contiguous ring positions k .. k+8 = min3 of three windows of three (VIMNMX3.U16x2): 16 + 16
  // three-input minima and 8 three-input maxima, 40 instructions for both polarities
  uint32_t m3[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) m3[k] = umin3_16x2(P[k], P[(k + 1) & 15], P[(k + 2) & 15]);
  uint32_t best = 0;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const uint32_t m9a = umin3_16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
    const uint32_t m9b = umin3_16x2(m3[k + 1], m3[(k + 4) & 15], m3[(k + 7) & 15]);
    best = umax3_16x2(best, m9a, m9b);
  }
  const int a = (int)(best & 0xffffu) - 256, b = (int)(best >> 16) - 256;
  const int s = imax(a, b);
  return s > t ? s - 1 : 0;
}

#if defined(__CUDACC__)
static __device__ __constant__ uint32_t kFastImad[2] = {1u, 0xffffffffu};
#endif
// bit 15 of each lane = the two-pixel pre-test; kt = 0x7fff7fff - (t | t << 16)
RPE_HD uint32_t fast_compass2(uint32_t c, uint32_t n, uint32_t s, uint32_t e, uint32_t w, uint32_t kt) {
  const uint32_t bright = umin16x2(umax16x2(n, s), umax16x2(e, w));
  const uint32_t dark = umax16x2(umin16x2(n, s), umin16x2(e, w));
#if defined(__CUDA_ARCH__)
  // the four additions as IMADs with multipliers the assembler cannot see through (+1 / -1 from constant memory): they
  // issue on the FMA pipe, which k_fast leaves idle, instead of the saturated ALU pipe (k_fast 3.32 -> 3.28 ms)
  uint32_t a, b;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a) : "r"(c), "r"(kFastImad[1]), "r"(kt));        // kt - c
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(a) : "r"(bright), "r"(kFastImad[0]), "r"(a));    // + bright
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b) : "r"(c), "r"(kFastImad[0]), "r"(kt));        // kt + c
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(b) : "r"(dark), "r"(kFastImad[1]), "r"(b));      // - dark
  return a | b;
#else
  return (bright + (kt - c)) | ((kt + c) - dark);
#endif
}
// The same test with the column words already unpacked (k_fast keeps the unpacked words of the rows above / below in
// a register window while it walks down a tile: the word of a row serves as D, C and U of three output rows).
// clo / chi = bytes {0, 1} / {2, 3} of the centre word as u16 lanes (fast_unpack_lo / _hi), same for U and D.
RPE_HD uint32_t fast_unpack_lo(uint32_t w) { return prmt(w, 0, 0x4140); }
RPE_HD uint32_t fast_unpack_hi(uint32_t w) { return prmt(w, 0, 0x4342); }
RPE_HD uint32_t fast_compass_kt(uint32_t t4) { return 0x7fff7fffu - (t4 & 0x00ff00ffu); }
RPE_HD uint32_t fast_compass4_unpacked(uint32_t clo, uint32_t chi, uint32_t ulo, uint32_t uhi, uint32_t dlo, uint32_t dhi,
                                       uint32_t Lw, uint32_t C, uint32_t Rw, uint32_t kt) {
  const uint32_t E = (C >> 24) | (Rw << 8);                // pixels x+3 .. x+6
  const uint32_t W = (Lw >> 8) | (C << 24);                // pixels x-3 .. x
  const uint32_t lo = fast_compass2(clo, ulo, dlo, fast_unpack_lo(E), fast_unpack_lo(W), kt);
  const uint32_t hi = fast_compass2(chi, uhi, dhi, fast_unpack_hi(E), fast_unpack_hi(W), kt);
  return prmt(lo, hi, 0x7531) & 0x80808080u;
}
RPE_HD uint32_t fast_compass4(uint32_t Lw, uint32_t C, uint32_t Rw, uint32_t U, uint32_t D, uint32_t t4) {
  return fast_compass4_unpacked(fast_unpack_lo(C), fast_unpack_hi(C), fast_unpack_lo(U), fast_unpack_hi(U),
                                fast_unpack_lo(D), fast_unpack_hi(D), Lw, C, Rw, fast_compass_kt(t4));
}

// ------------------------------------------------------------------ A.5 Harris response
// response from the three integer sums over the 7x7 block (non-fused f32, A.5)
RPE_HD float harris_from_sums(int a, int b, int cc) {
  const float s = f32_div(1.0f, f32_mul(28.0f, 255.0f));   // 1.f/(4*7*255.f)
  const float s4 = f32_mul(f32_mul(f32_mul(s, s), s), s);
  const float fa = (float)a, fb = (float)b, fc = (float)cc;
  const float apb = f32_add(fa, fb);
  const float t1 = f32_mul(fa, fb);
  const float t2 = f32_mul(fc, fc);
  const float t3 = f32_mul(f32_mul(0.04f, apb), apb);
  return f32_mul(f32_sub(f32_sub(t1, t2), t3), s4);
}

RPE_HD float harris_response(const uint8_t* c, int pitch) {
  int a = 0, b = 0, cc = 0;
  // 9x9 neighbourhood, rows -4..4; three-row sliding window
  int r0[9], r1[9], r2[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    r0[i] = c[-4 * pitch + (i - 4)];
    r1[i] = c[-3 * pitch + (i - 4)];
  }
#pragma unroll
  for (int y = -3; y <= 3; ++y) {
#pragma unroll
    for (int i = 0; i < 9; ++i) r2[i] = c[(y + 1) * pitch + (i - 4)];
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

// ------------------------------------------------------------------ A.6 orientation
RPE_DEFINE_TABLE(int8_t, kUmax, 16, {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3})

RPE_HD float fast_atan2_deg(float y, float x) {
  // cv::fastAtan2 scalar path: 7th-order odd polynomial, non-fused f32
  const float k = 57.29577951308232f;  // (float)(180/CV_PI)
  const float p1 = f32_mul(0.9997878412794807f, k), p3 = f32_mul(-0.3258083974640975f, k);
  const float p5 = f32_mul(0.1555786518463281f, k), p7 = f32_mul(-0.04432655554792128f, k);
  const float eps = 2.220446049250313e-16f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = f32_div(ay, f32_add(ax, eps));
    c2 = f32_mul(c, c);
    a = f32_mul(f32_add(f32_mul(f32_add(f32_mul(f32_add(f32_mul(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = f32_div(ax, f32_add(ay, eps));
    c2 = f32_mul(c, c);
    a = f32_sub(90.0f, f32_mul(f32_add(f32_mul(f32_add(f32_mul(f32_add(f32_mul(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = f32_sub(180.0f, a);
  if (y < 0) a = f32_sub(360.0f, a);
  return a;
}

RPE_HD float ic_angle(const uint8_t* c, int pitch) {
  int m01 = 0, m10 = 0;
  for (int u = -15; u <= 15; ++u) m10 += u * (int)c[u];
  for (int v = 1; v <= 15; ++v) {
    int vsum = 0;
    const int d = RPE_TBL(kUmax)[v];
    const uint8_t* pp = c + v * pitch;
    const uint8_t* pm = c - v * pitch;
    for (int u = -d; u <= d; ++u) {
      const int vp = pp[u], vm = pm[u];
      vsum += vp - vm;
      m10 += u * (vp + vm);
    }
    m01 += v * vsum;
  }
  return fast_atan2_deg((float)m01, (float)m10);
}

// ------------------------------------------------------------------ A.8 blur taps
// f32 bit patterns of getGaussianKernel(7, 2): k[0..3] (symmetric)
#define RPE_BLUR_K0 0.0701593235135078430f /* 0x3d8fafb1 */
#define RPE_BLUR_K1 0.1310748755931854248f /* 0x3e06387e */
#define RPE_BLUR_K2 0.1907128244638442993f /* 0x3e434a39 */
#define RPE_BLUR_K3 0.2161059379577636719f /* 0x3e5d4ae0 */

// row pass on 7 consecutive u8: s = k0*x0; s = fma(k_j, x_j, s) left to right
RPE_HD float blur_row7(float x0, float x1, float x2, float x3, float x4, float x5, float x6) {
  float s = f32_mul(RPE_BLUR_K0, x0);
  s = f32_fma(RPE_BLUR_K1, x1, s);
  s = f32_fma(RPE_BLUR_K2, x2, s);
  s = f32_fma(RPE_BLUR_K3, x3, s);
  s = f32_fma(RPE_BLUR_K2, x4, s);
  s = f32_fma(RPE_BLUR_K1, x5, s);
  s = f32_fma(RPE_BLUR_K0, x6, s);
  return s;
}
// column pass on 7 row-pass results r[-3..3]: symmetric pairing, then round + saturate
RPE_HD float blur_col7_sum(float r0, float r1, float r2, float r3, float r4, float r5, float r6) {
  float s = f32_mul(RPE_BLUR_K3, r3);
  s = f32_fma(RPE_BLUR_K2, f32_add(r2, r4), s);
  s = f32_fma(RPE_BLUR_K1, f32_add(r1, r5), s);
  s = f32_fma(RPE_BLUR_K0, f32_add(r0, r6), s);
  return s;
}
RPE_HD uint8_t blur_col7(float r0, float r1, float r2, float r3, float r4, float r5, float r6) {
  int v = f32_rint(blur_col7_sum(r0, r1, r2, r3, r4, r5, r6));
  return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}
RPE_HD int reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}

// ------------------------------------------------------------------ A.8 steered rBRIEF
static const int8_t kBriefPattern_host[1024] = {
#include "orb_pattern.inc"
};
#if defined(__CUDACC__)
static __device__ __constant__ int8_t kBriefPattern_dev[1024] = {
#include "orb_pattern.inc"
};
#endif

// steering coefficients of a keypoint: a = (float)cos(theta), b = (float)sin(theta) with theta the f32
// angle in radians and cos / sin evaluated in double, as computeOrbDescriptors does
RPE_HD void brief_steer(float angle_deg, float* a, float* b) {
  const float theta = f32_mul(angle_deg, 0.017453292519943295f);  // (float)(CV_PI/180)
#if defined(__CUDA_ARCH__)
  double sn, cs;                       // one range reduction for both (same polynomials as sin() / cos())
  sincos((double)theta, &sn, &cs);
  *a = (float)cs;
  *b = (float)sn;
#else
  *a = (float)cos((double)theta);
  *b = (float)sin((double)theta);
#endif
}

// one test; t = (xa, ya, xb, yb) as four int8 packed little-endian; c = blurred level at the keypoint centre
RPE_HD uint32_t brief_test(const uint8_t* c, int pitch, float a, float b, uint32_t t) {
  const float xa = (float)(int8_t)(t & 0xff), ya = (float)(int8_t)((t >> 8) & 0xff);
  const float xb = (float)(int8_t)((t >> 16) & 0xff), yb = (float)(int8_t)(t >> 24);
  const int ixa = f32_rint(f32_sub(f32_mul(xa, a), f32_mul(ya, b)));
  const int iya = f32_rint(f32_add(f32_mul(xa, b), f32_mul(ya, a)));
  const int ixb = f32_rint(f32_sub(f32_mul(xb, a), f32_mul(yb, b)));
  const int iyb = f32_rint(f32_add(f32_mul(xb, b), f32_mul(yb, a)));
  const int va = c[iya * pitch + ixa], vb = c[iyb * pitch + ixb];
  return (uint32_t)(va < vb);
}

// one descriptor byte = 8 tests; p = 8 x (xa, ya, xb, yb)
RPE_HD uint32_t brief_byte(const uint8_t* c, int pitch, float a, float b, const int8_t* p) {
  uint32_t bits = 0;
  for (int j = 0; j < 8; ++j, p += 4) {
    const uint32_t t = (uint32_t)(uint8_t)p[0] | ((uint32_t)(uint8_t)p[1] << 8) | ((uint32_t)(uint8_t)p[2] << 16) |
                       ((uint32_t)(uint8_t)p[3] << 24);
    bits |= brief_test(c, pitch, a, b, t) << j;
  }
  return bits;
}

// c = blurred level at the keypoint centre; angle in degrees; out = 8 x u32 (32 bytes, LSB first)
RPE_HD void brief256(const uint8_t* c, int pitch, float angle_deg, uint32_t out[8]) {
  float a, b;
  brief_steer(angle_deg, &a, &b);
  const int8_t* P = RPE_TBL(kBriefPattern);
  for (int w = 0; w < 8; ++w) {
    uint32_t bits = 0;
    for (int j = 0; j < 4; ++j) bits |= brief_byte(c, pitch, a, b, P + 32 * (4 * w + j)) << (8 * j);
    out[w] = bits;
  }
}

}  // namespace rpe
