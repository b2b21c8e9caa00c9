// hostcheck.cpp — TEST-ONLY host instantiation of the __host__ __device__ arithmetic in
// relative-pose-estimation_b200/csrc/*_math.cuh, so the exact per-element math of the CUDA
// kernels can be diffed against the oracle on a machine without a GPU.  Never loaded by the
// product (librpe_b200.so has no CPU path).  Build: g++ -O2 -ffp-contract=off -shared -fPIC.
#include <cstring>
#include "orb_math.cuh"
#include "geom_math.cuh"
#include "match_math.cuh"
using namespace rpe;

extern "C" {

// all-pairs Hamming distances with the kernel's carry-save form (mode 0: 3 adders, 1: 4 adders, 2: plain)
void hc_hamming(const uint32_t* q, int nq, const uint32_t* t, int nt, int mode, int* out) {
  for (int i = 0; i < nq; ++i)
    for (int j = 0; j < nt; ++j) {
      const uint32_t* a = q + 8 * i;
      const uint32_t* b = t + 8 * j;
      if (mode == 5 || mode == 6) {   // linearly pre-combined operands (hamming_mix), 4- / 3-adder forms
        uint32_t am[8], bm[8];
        memcpy(am, a, 32); memcpy(bm, b, 32);
        hamming_mix(am); hamming_mix(bm);
        out[i * nt + j] = mode == 5 ? (int)(hamming256_mixed_key<true>(am, bm[0], bm[1], bm[2], bm[3], bm[4], bm[5], bm[6], bm[7], 0u, 1u << 20) >> 20)
                                    : (int)(hamming256_mixed_key<false>(am, bm[0], bm[1], bm[2], bm[3], bm[4], bm[5], bm[6], bm[7], 0u, 1u << 20) >> 20);
        continue;
      }
      out[i * nt + j] = mode == 0   ? (int)hamming256<false>(a, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7])
                        : mode == 1 ? (int)hamming256<true>(a, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7])
                        : mode == 3 ? (int)(hamming256_key<false>(a, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], 0u, 1u << 20) >> 20)
                        : mode == 4 ? (int)(hamming256_key<true>(a, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], 0u, 1u << 20) >> 20)
                                    : (int)hamming256_plain(a, b);
    }
}

void hc_gray(const uint8_t* bgr, int n, int mode, uint8_t* out) {
  for (int i = 0; i < n; ++i) out[i] = (uint8_t)gray_from_bgr(bgr[3 * i], bgr[3 * i + 1], bgr[3 * i + 2], mode);
}
void hc_resize(const uint8_t* src, int sw, int sh, uint8_t* dst, int dw, int dh, const int* xo,
               const int* xc, const int* yo, const int* yc) {
  for (int y = 0; y < dh; ++y)
    for (int x = 0; x < dw; ++x) dst[y * dw + x] = resize_px(src, sw, sw, sh, xo[x], xc[x], yo[y], yc[y]);
}
void hc_fast_scores(const uint8_t* img, int w, int h, int t, int* out) {
  memset(out, 0, sizeof(int) * w * h);
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) out[y * w + x] = fast9_score(img + y * w + x, w, t);
}
// packed pre-test on every aligned quad of rows 3..h-4 / cols 4..w-8: out[y*w+x] = 1 if it passes
// scalar reference form (exact 9-arc test, then min / max network) for the same pixels
void hc_fast_scores_scalar(const uint8_t* img, int w, int h, int t, int* out) {
  memset(out, 0, sizeof(int) * w * h);
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) out[y * w + x] = fast9_score_scalar(img + y * w + x, w, t);
}
void hc_fast_compass(const uint8_t* img, int w, int h, int t, uint8_t* out) {
  memset(out, 0, (size_t)w * h);
  const uint32_t t4 = (uint32_t)t * 0x01010101u;
  for (int y = 3; y < h - 3; ++y)
    for (int x = 4; x + 8 <= w; x += 4) {
      uint32_t Lw, C, Rw, U, D;
      memcpy(&Lw, img + y * w + x - 4, 4); memcpy(&C, img + y * w + x, 4); memcpy(&Rw, img + y * w + x + 4, 4);
      memcpy(&U, img + (y - 3) * w + x, 4); memcpy(&D, img + (y + 3) * w + x, 4);
      const uint32_t c = fast_compass4(Lw, C, Rw, U, D, t4);
      for (int j = 0; j < 4; ++j) out[y * w + x + j] = (c >> (8 * j + 7)) & 1;
    }
}
void hc_harris(const uint8_t* img, int w, const int* xs, const int* ys, int n, float* out) {
  for (int i = 0; i < n; ++i) out[i] = harris_response(img + ys[i] * w + xs[i], w);
}
void hc_angle(const uint8_t* img, int w, const int* xs, const int* ys, int n, float* out) {
  for (int i = 0; i < n; ++i) out[i] = ic_angle(img + ys[i] * w + xs[i], w);
}
void hc_atan2(const float* y, const float* x, int n, float* out) {
  for (int i = 0; i < n; ++i) out[i] = fast_atan2_deg(y[i], x[i]);
}
void hc_blur(const uint8_t* img, int w, int h, uint8_t* out) {
  float* rows = new float[(size_t)w * h];
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      float v[7];
      for (int k = 0; k < 7; ++k) v[k] = (float)img[y * w + reflect101(x + k - 3, w)];
      rows[y * w + x] = blur_row7(v[0], v[1], v[2], v[3], v[4], v[5], v[6]);
    }
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      float v[7];
      for (int k = 0; k < 7; ++k) v[k] = rows[reflect101(y + k - 3, h) * w + x];
      out[y * w + x] = blur_col7(v[0], v[1], v[2], v[3], v[4], v[5], v[6]);
    }
  delete[] rows;
}
void hc_brief(const uint8_t* blurred, int w, const int* cx, const int* cy, const float* angle, int n,
              uint8_t* desc) {
  for (int i = 0; i < n; ++i) {
    uint32_t o[8];
    brief256(blurred + cy[i] * w + cx[i], w, angle[i], o);
    memcpy(desc + 32 * i, o, 32);
  }
}
void hc_subsets(int count, int n, int* out) {
  CvRng rng{0xFFFFFFFFFFFFFFFFull};
  for (int i = 0; i < n; ++i) draw_subset5(rng, count, out + 5 * i);
}
int hc_update_iters(double p, double ep, int mp, int mx) { return ransac_update_iters(p, ep, mp, mx); }
int hc_five_point(const double* x1, const double* x2, double* E) { return five_point_solve(x1, x2, E); }
void hc_sampson(const double* E, const double* x1, const double* x2, int n, float* err) {
  for (int i = 0; i < n; ++i) err[i] = sampson_err_f32(E, x1[2 * i], x1[2 * i + 1], x2[2 * i], x2[2 * i + 1]);
}
void hc_sampson_inlier(const double* E, const double* x1, const double* x2, int n, float t2, uint8_t* ok) {
  const SampsonGate g = sampson_gate(t2);
  for (int i = 0; i < n; ++i) ok[i] = sampson_is_inlier(E, x1[2 * i], x1[2 * i + 1], x2[2 * i], x2[2 * i + 1], g);
}
void hc_decompose(const double* E, double* R1, double* R2, double* t) { decompose_essential(E, R1, R2, t); }
void hc_cheirality(const double* R, const double* t, const double* x1, const double* x2, int n, double dist,
                   uint8_t* ok) {
  for (int i = 0; i < n; ++i)
    ok[i] = triangulate_cheirality(R, t, x1[2 * i], x1[2 * i + 1], x2[2 * i], x2[2 * i + 1], dist);
}
void hc_euler(const double* R, int conv, double* ypr) { rotation_to_euler(R, conv, ypr); }
int hc_dk_iters(void) { return g_dk_iters_last; }
void hc_solve_poly10(const double* c, double* roots_re_im) {
  Cplx r[10];
  solve_poly10(c, r, 300);
  for (int i = 0; i < 10; ++i) { roots_re_im[2 * i] = r[i].re; roots_re_im[2 * i + 1] = r[i].im; }
}
}
