// geom_math.cuh — FP64 geometry of the hot path: OpenCV's RANSAC driver pieces (MWC RNG,
// subset draw, adaptive iteration count), the five-point essential-matrix solver (Nister /
// Stewenius-Li formulation: null space -> 10 cubic constraints -> 10x20 elimination ->
// degree-10 polynomial -> Durand-Kerner roots), Sampson error, essential-matrix decomposition,
// DLT triangulation + cheirality, and rotation -> Euler.
// Replaces cv2.findEssentialMat / cv2.recoverPose as the reference calls them
// (src/core/pose_estimator.py:522-527, :533) and rotation_to_euler (src/utils/geometry.py:14-27);
// arithmetic spec: SURVEY.md A.10, A.11.
#pragma once
#include "rpe_hd.cuh"

namespace rpe {

// ------------------------------------------------------------------ RANSAC driver pieces
struct CvRng {
  uint64_t state;
  RPE_HD uint32_t next() {
    state = (uint64_t)(uint32_t)state * 4164903690ull + (state >> 32);
    return (uint32_t)state;
  }
  RPE_HD int uniform(int n) { return (int)(next() % (uint32_t)n); }
};

// RANSACPointSetRegistrator::getSubset with modelPoints = 5
RPE_HD void draw_subset5(CvRng& rng, int count, int idx[5]) {
  for (int i = 0; i < 5; ++i) {
    int v;
    bool dup;
    do {
      v = rng.uniform(count);
      dup = false;
      for (int j = 0; j < i; ++j) dup = dup || (idx[j] == v);
    } while (dup);
    idx[i] = v;
  }
}

// cv::RANSACUpdateNumIters
RPE_HD int ransac_update_iters(double p, double ep, int model_points, int max_iters) {
  const double DBL_MIN_ = 2.2250738585072014e-308;
  p = fmax(p, 0.0); p = fmin(p, 1.0);
  ep = fmax(ep, 0.0); ep = fmin(ep, 1.0);
  double num = fmax(1.0 - p, DBL_MIN_);
  double denom = 1.0 - pow(1.0 - ep, (double)model_points);
  if (denom < DBL_MIN_) return 0;
  num = log(num);
  denom = log(denom);
  return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : f64_rint(num / denom);
}

// EMEstimatorCallback::computeError for one correspondence (f64 Sampson, stored as f32).  Every product and sum is
// rounded on its own, in cv2's order (Matx row times vector left to right, then the dot product, then a + b + c + d):
// given the same E and points the error is cv2's to the last bit, so `err <= threshold` decides as cv2 does.
RPE_HD float sampson_err_f32(const double* E, double x1, double y1, double x2, double y2) {
  const double Ex0 = f64_add(f64_add(f64_mul(E[0], x1), f64_mul(E[1], y1)), E[2]);
  const double Ex1 = f64_add(f64_add(f64_mul(E[3], x1), f64_mul(E[4], y1)), E[5]);
  const double Ex2 = f64_add(f64_add(f64_mul(E[6], x1), f64_mul(E[7], y1)), E[8]);
  const double Et0 = f64_add(f64_add(f64_mul(E[0], x2), f64_mul(E[3], y2)), E[6]);
  const double Et1 = f64_add(f64_add(f64_mul(E[1], x2), f64_mul(E[4], y2)), E[7]);
  const double x2tEx1 = f64_add(f64_add(f64_mul(x2, Ex0), f64_mul(y2, Ex1)), Ex2);
  const double den = f64_add(f64_add(f64_add(f64_mul(Ex0, Ex0), f64_mul(Ex1, Ex1)), f64_mul(Et0, Et0)), f64_mul(Et1, Et1));
  return (float)(f64_mul(x2tEx1, x2tEx1) / den);
}

// The inlier decision `(float)(num2 / den) <= t2` of the RANSAC loop WITHOUT the FP64 division (a third of the
// scoring instructions; ncu).  (float)q <= t2 holds iff q lies below the midpoint `mid` between t2 and the next
// float (or on it when the tie rounds down to t2), and q = fl(num2 / den) lies on the same side of mid as the exact
// quotient whenever that is further than a few ulps from mid: then sign(num2 - mid * den) decides.  The rare pair
// within 1e-13 of the boundary (and den == 0) takes the division, so the decision is the reference's in every case.
struct SampsonGate {
  float t2;
  double mid;
};
RPE_HD SampsonGate sampson_gate(float t2) {
  SampsonGate g;
  g.t2 = t2;
  union { float f; uint32_t u; } a, b;
  a.f = t2;
  b.u = a.u + 1u;                                   // next float above t2 (t2 > 0, finite)
  g.mid = ((double)a.f + (double)b.f) * 0.5;        // exact in double
  return g;
}
RPE_HD bool sampson_is_inlier(const double* E, double x1, double y1, double x2, double y2, const SampsonGate& g) {
  const double Ex0 = f64_add(f64_add(f64_mul(E[0], x1), f64_mul(E[1], y1)), E[2]);
  const double Ex1 = f64_add(f64_add(f64_mul(E[3], x1), f64_mul(E[4], y1)), E[5]);
  const double Ex2 = f64_add(f64_add(f64_mul(E[6], x1), f64_mul(E[7], y1)), E[8]);
  const double Et0 = f64_add(f64_add(f64_mul(E[0], x2), f64_mul(E[3], y2)), E[6]);
  const double Et1 = f64_add(f64_add(f64_mul(E[1], x2), f64_mul(E[4], y2)), E[7]);
  const double x2tEx1 = f64_add(f64_add(f64_mul(x2, Ex0), f64_mul(y2, Ex1)), Ex2);
  const double den = f64_add(f64_add(f64_add(f64_mul(Ex0, Ex0), f64_mul(Ex1, Ex1)), f64_mul(Et0, Et0)), f64_mul(Et1, Et1));
  const double num2 = f64_mul(x2tEx1, x2tEx1);
  const double md = g.mid * den;
  const double r = num2 - md;
  if (fabs(r) > 1e-13 * md) return r < 0.0;         // (false for NaN / inf operands as well: they take the division)
  return (float)(num2 / den) <= g.t2;
}

// ------------------------------------------------------------------ five-point solver
// monomial bookkeeping: linear {x,y,z,1}; quadratic (10); cubic (20) in the column order
// [x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1]
RPE_DEFINE_TABLE(int8_t, kQ2, 16, {0, 1, 2, 3, 1, 4, 5, 6, 2, 5, 7, 8, 3, 6, 8, 9})
RPE_DEFINE_TABLE(int8_t, kC3, 40, {0, 2, 4, 5, 2, 3, 8, 9, 4, 8, 10, 11, 5, 9, 11, 12, 3, 1, 6, 7,
                                   8, 6, 13, 14, 9, 7, 14, 15, 10, 13, 16, 17, 11, 14, 17, 18,
                                   12, 15, 18, 19})

// out(10) += s * a(4) * b(4)
RPE_HD void poly_mul11(const double* a, const double* b, double s, double* out) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) out[RPE_TBL(kQ2)[4 * i + j]] += s * a[i] * b[j];
}
// out(20) += s * q(10) * l(4)
RPE_HD void poly_mul21(const double* q, const double* l, double s, double* out) {
  for (int i = 0; i < 10; ++i)
    for (int j = 0; j < 4; ++j) out[RPE_TBL(kC3)[4 * i + j]] += s * q[i] * l[j];
}

struct Cplx {
  double re, im;
};
RPE_HD Cplx cmul(Cplx a, Cplx b) { return Cplx{a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
RPE_HD Cplx cdiv(Cplx a, Cplx b) {
  const double t = 1.0 / (b.re * b.re + b.im * b.im);
  return Cplx{(a.re * b.re + a.im * b.im) * t, (a.im * b.re - a.re * b.im) * t};
}

// cv::solvePoly (Durand-Kerner with in-place updates, start points 1,(1+i),(1+i)^2,...),
// degree 10, coeffs c[i]*x^i.  OpenCV iterates until the largest update is exactly 0 or 300
// iterations; in double precision the updates of ill-conditioned roots stall at rounding-noise
// level, so that loop usually runs all 300 rounds (ncu: 96 % of k_ransac).  Here the SAME
// Durand-Kerner sequence runs only until every root's update is below 1e-7 of its magnitude
// (root labels / order are fixed long before that), then each root is polished independently
// by Newton steps on the polynomial itself (quadratic convergence to the rounding floor).
#if !defined(__CUDA_ARCH__)
static int g_dk_iters_last = 0;   // TEST-ONLY instrumentation of the host instantiation
#endif
RPE_HDN void solve_poly10(const double* c_in, Cplx* roots_out, int max_iters) {
  // coefficients and roots live in registers: every loop over them is fully unrolled (with the roots behind the
  // caller's pointer the inner loop was 20 local-memory loads per root and iteration; ncu, round 2)
  constexpr int n = 10;
  double c[n + 1];
#pragma unroll
  for (int i = 0; i <= n; ++i) c[i] = c_in[i];
  Cplx roots[n];
  Cplx p{1.0, 0.0};
  const Cplx r{1.0, 1.0};
#pragma unroll
  for (int i = 0; i < n; ++i) {
    roots[i] = p;
    p = cmul(p, r);
  }
  int iter = 0;
  for (; iter < max_iters; ++iter) {
    bool settled = true;
#pragma unroll
    for (int i = 0; i < n; ++i) {
      p = roots[i];
      Cplx num{c[n], 0.0}, den{c[n], 0.0};
#pragma unroll
      for (int j = 0; j < n; ++j) {
        num = cmul(num, p);
        num.re += c[n - j - 1];
        if (j != i) {
          const Cplx d{p.re - roots[j].re, p.im - roots[j].im};
          if (d.re != 0.0 || d.im != 0.0) den = cmul(den, d);
        }
      }
      num = cdiv(num, den);
      roots[i].re = p.re - num.re;
      roots[i].im = p.im - num.im;
      const double upd = fabs(num.re) + fabs(num.im), mag = fabs(p.re) + fabs(p.im);
      if (!(upd <= 1e-7 * (1.0 + mag))) settled = false;
      if (upd != upd) iter = max_iters;             // NaN: give up (as the reference loop does)
    }
    if (settled) break;
  }
#if !defined(__CUDA_ARCH__)
  g_dk_iters_last = iter;
#endif
  // Newton polish: z <- z - f(z) / f'(z), both by Horner
#pragma unroll
  for (int i = 0; i < n; ++i) {
    Cplx z = roots[i];
    for (int k = 0; k < 8; ++k) {
      Cplx f{c[n], 0.0}, df{0.0, 0.0};
#pragma unroll
      for (int j = n - 1; j >= 0; --j) {
        df = cmul(df, z);
        df.re += f.re; df.im += f.im;
        f = cmul(f, z);
        f.re += c[j];
      }
      if (df.re == 0.0 && df.im == 0.0) break;
      const Cplx dz = cdiv(f, df);
      const double upd = fabs(dz.re) + fabs(dz.im), mag = fabs(z.re) + fabs(z.im);
      if (!(upd <= 1e-3 * (1.0 + mag))) break;      // not in the quadratic basin (or NaN): keep the DK value
      z.re -= dz.re; z.im -= dz.im;
      if (upd <= 4e-16 * mag) break;
    }
    roots_out[i] = z;
  }
}

// All real essential matrices through five normalised correspondences (x2' E x1 = 0), each
// with Frobenius norm 1.  x1, x2: 5 x (x, y).  Eout: up to 10 x 9 (row-major).  Returns count.
RPE_HDN int five_point_solve(const double* x1, const double* x2, double* Eout) {
  // --- 0. static sample (x2 == x1 for all five points: the camera did not move, e.g. the last 5 pairs of the
  // reference's simulator sequence).  Every skew-symmetric matrix then satisfies the constraints, the polynomial
  // system is rank deficient and its numerical solution is noise; cv2 4.13 (LAPACK build) returns
  // E = [[0,-1,-1],[1,0,-1],[1,1,0]] / sqrt(6) for ANY such sample, whatever the points and K
  // (tests/test_oracle_geometry.py pins this against live cv2), i.e. R = I exactly.  Same here.
  {
    bool is_static = true;
    for (int i = 0; i < 10; ++i) is_static = is_static && (x1[i] == x2[i]);
    if (is_static) {
      const double a = 0.40824829046386302;   // 1 / sqrt(6)
      Eout[0] = 0.0; Eout[1] = -a;  Eout[2] = -a;
      Eout[3] = a;   Eout[4] = 0.0; Eout[5] = -a;
      Eout[6] = a;   Eout[7] = a;   Eout[8] = 0.0;
      return 1;
    }
  }
  // --- 1. null space of the 5x9 epipolar system, in the SAME basis cv2 uses.  The essential-matrix parametrisation
  // E = x N0 + y N1 + z N2 + N3 depends on this basis, and so do (a) the conditioning of near-degenerate samples and
  // (b) the ORDER in which the roots of the degree-10 polynomial in z come out of the root finder, i.e. the order of
  // the models of a hypothesis, which decides RANSAC ties.  cv2 (also when built with LAPACK: its HAL hands matrices
  // this small to OpenCV's own one-sided Jacobi SVD) does not return a LAPACK basis: JacobiSVDImpl_ computes the five
  // right singular vectors and FILLS the remaining four rows of Vt by Gram-Schmidt: row i = a fixed pseudo-random
  // vector of +-1/9 (cv::RNG seeded 0x12345678 on every call, bit 8 of each draw), orthogonalised twice against all
  // previous rows with an L1 rescale after every projection, then normalised.  Only the ROW SPACE of the system
  // enters (projections onto a subspace do not depend on its basis), so any orthonormal basis of it will do: rows
  // 0..4 of the Householder Q of A = L Q (dgelq2 / dlarfg reflectors).  Verified against cv2.SVDecomp: 1e-15; and
  // against cv2's five-point solver: same models in the same order (tests/test_kernel_math_host.py).
  double Rw[5][9], tau[5];
  for (int i = 0; i < 5; ++i) {
    const double a = x1[2 * i], b = x1[2 * i + 1], c = x2[2 * i], d = x2[2 * i + 1];
    Rw[i][0] = c * a; Rw[i][1] = c * b; Rw[i][2] = c;
    Rw[i][3] = d * a; Rw[i][4] = d * b; Rw[i][5] = d;
    Rw[i][6] = a;     Rw[i][7] = b;     Rw[i][8] = 1.0;
  }
  for (int i = 0; i < 5; ++i) {
    const double alpha = Rw[i][i];
    double xn = 0;
    for (int k = i + 1; k < 9; ++k) xn += Rw[i][k] * Rw[i][k];
    if (!(xn > 0)) { tau[i] = 0; continue; }            // H = I
    const double nrm = sqrt(alpha * alpha + xn);
    const double beta = alpha >= 0 ? -nrm : nrm;
    tau[i] = (beta - alpha) / beta;
    const double sc = 1.0 / (alpha - beta);
    for (int k = i + 1; k < 9; ++k) Rw[i][k] *= sc;       // v (v_i = 1 implied)
    for (int r = i + 1; r < 5; ++r) {                     // apply H_i to the remaining rows
      double w = Rw[r][i];
      for (int k = i + 1; k < 9; ++k) w += Rw[r][k] * Rw[i][k];
      w *= tau[i];
      Rw[r][i] -= w;
      for (int k = i + 1; k < 9; ++k) Rw[r][k] -= w * Rw[i][k];
    }
  }
  double Vt[9][9];                                        // rows 0..4: row space; rows 5..8: cv2's null-space basis
  for (int m = 0; m < 5; ++m) {                           // row m of Q = e_m' H5 H4 H3 H2 H1
    double v[9];
    for (int k = 0; k < 9; ++k) v[k] = (k == m) ? 1.0 : 0.0;
    for (int i = 4; i >= 0; --i) {
      if (tau[i] == 0) continue;
      double w = v[i];
      for (int k = i + 1; k < 9; ++k) w += v[k] * Rw[i][k];
      w *= tau[i];
      v[i] -= w;
      for (int k = i + 1; k < 9; ++k) v[k] -= w * Rw[i][k];
    }
    for (int k = 0; k < 9; ++k) Vt[m][k] = v[k];
  }
  for (int i = 5; i < 9; ++i) {
    // the 36 draws of cv::RNG(0x12345678): bit k of kSign[i - 5] set = +1/9, clear = -1/9
    const unsigned kSign[4] = {0x184u, 0x17du, 0x1b1u, 0x0e9u};
    double v[9];
    for (int k = 0; k < 9; ++k) v[k] = ((kSign[i - 5] >> k) & 1u) ? (1.0 / 9.0) : -(1.0 / 9.0);
    for (int iter = 0; iter < 2; ++iter)
      for (int j = 0; j < i; ++j) {
        double sd = 0;
        for (int k = 0; k < 9; ++k) sd += v[k] * Vt[j][k];
        double asum = 0;
        for (int k = 0; k < 9; ++k) {
          const double t = v[k] - sd * Vt[j][k];
          v[k] = t;
          asum += fabs(t);
        }
        asum = asum > 2.220446049250313e-16 * 10 * 100 ? 1.0 / asum : 0.0;
        for (int k = 0; k < 9; ++k) v[k] *= asum;
      }
    double nn = 0;
    for (int k = 0; k < 9; ++k) nn += v[k] * v[k];
    if (!(nn > 0)) return 0;                              // NaN / degenerate guard
    nn = 1.0 / sqrt(nn);
    for (int k = 0; k < 9; ++k) Vt[i][k] = v[k] * nn;
  }
  double N[4][9];
  for (int m = 0; m < 4; ++m)
    for (int k = 0; k < 9; ++k) N[m][k] = Vt[5 + m][k];
  // --- 2. constraint matrix A (10 x 20): det(E) = 0 and E E' E - tr(E E')/2 E = 0
  double e[9][4];
  for (int k = 0; k < 9; ++k)
    for (int m = 0; m < 4; ++m) e[k][m] = N[m][k];
  double A[10][20];
  for (int i = 0; i < 10; ++i)
    for (int j = 0; j < 20; ++j) A[i][j] = 0.0;
  {
    double q[10];
    // det = e0 (e4 e8 - e5 e7) - e1 (e3 e8 - e5 e6) + e2 (e3 e7 - e4 e6)
    for (int j = 0; j < 10; ++j) q[j] = 0;
    poly_mul11(e[4], e[8], 1.0, q); poly_mul11(e[5], e[7], -1.0, q);
    poly_mul21(q, e[0], 1.0, A[0]);
    for (int j = 0; j < 10; ++j) q[j] = 0;
    poly_mul11(e[3], e[8], 1.0, q); poly_mul11(e[5], e[6], -1.0, q);
    poly_mul21(q, e[1], -1.0, A[0]);
    for (int j = 0; j < 10; ++j) q[j] = 0;
    poly_mul11(e[3], e[7], 1.0, q); poly_mul11(e[4], e[6], -1.0, q);
    poly_mul21(q, e[2], 1.0, A[0]);
  }
  {
    double L[6][10];  // Lambda = E E' - tr/2 I, upper triangle: 00 01 02 11 12 22
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 10; ++j) L[i][j] = 0;
    int idx = 0;
    for (int r = 0; r < 3; ++r)
      for (int c = r; c < 3; ++c, ++idx)
        for (int k = 0; k < 3; ++k) poly_mul11(e[3 * r + k], e[3 * c + k], 1.0, L[idx]);
    double tr[10];
    for (int j = 0; j < 10; ++j) tr[j] = 0.5 * (L[0][j] + L[3][j] + L[5][j]);
    for (int j = 0; j < 10; ++j) { L[0][j] -= tr[j]; L[3][j] -= tr[j]; L[5][j] -= tr[j]; }
    const int sym[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c)
        for (int k = 0; k < 3; ++k) poly_mul21(L[sym[r][k]], e[3 * k + c], 1.0, A[1 + 3 * r + c]);
  }
  // --- 3. Gauss-Jordan on the first ten monomials (partial pivoting)
  for (int col = 0; col < 10; ++col) {
    int piv = col;
    double best = fabs(A[col][col]);
    for (int r = col + 1; r < 10; ++r)
      if (fabs(A[r][col]) > best) { best = fabs(A[r][col]); piv = r; }
    if (!(best > 1e-300)) return 0;
    if (piv != col)
      for (int j = col; j < 20; ++j) { const double t = A[col][j]; A[col][j] = A[piv][j]; A[piv][j] = t; }
    const double inv = 1.0 / A[col][col];
    for (int j = col; j < 20; ++j) A[col][j] *= inv;
    for (int r = 0; r < 10; ++r) {
      if (r == col) continue;
      const double f = A[r][col];
      if (f != 0.0)
        for (int j = col; j < 20; ++j) A[r][j] -= f * A[col][j];
    }
  }
  // --- 4. B(z): rows <x2z> - z<x2>, <y2z> - z<y2>, <xyz> - z<xy>; polys stored low-order first
  double P[3][3][5];  // P[row][col][power]; cols 0,1 cubic, col 2 quartic
  for (int i = 0; i < 3; ++i) {
    const double* r1 = &A[2 * i + 4][10];
    const double* r2 = &A[2 * i + 5][10];
    P[i][0][3] = -r2[0]; P[i][0][2] = r1[0] - r2[1]; P[i][0][1] = r1[1] - r2[2]; P[i][0][0] = r1[2]; P[i][0][4] = 0;
    P[i][1][3] = -r2[3]; P[i][1][2] = r1[3] - r2[4]; P[i][1][1] = r1[4] - r2[5]; P[i][1][0] = r1[5]; P[i][1][4] = 0;
    P[i][2][4] = -r2[6]; P[i][2][3] = r1[6] - r2[7]; P[i][2][2] = r1[7] - r2[8]; P[i][2][1] = r1[8] - r2[9];
    P[i][2][0] = r1[9];
  }
  // --- 5. det B(z) (degree 10): expansion along row 0
  double cf[11];
  for (int i = 0; i < 11; ++i) cf[i] = 0;
  for (int c0 = 0; c0 < 3; ++c0) {
    const int ca = (c0 + 1) % 3, cb = (c0 + 2) % 3;   // cyclic => sign +1
    double minor[9];
    for (int i = 0; i < 9; ++i) minor[i] = 0;
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 5; ++j)
        minor[i + j] += P[1][ca][i] * P[2][cb][j] - P[1][cb][i] * P[2][ca][j];
    for (int i = 0; i < 5; ++i)
      for (int j = 0; j < 9; ++j)
        if (i + j < 11) cf[i + j] += P[0][c0][i] * minor[j];
  }
  for (int i = 0; i < 11; ++i)
    if (!(fabs(cf[i]) < 1e300)) return 0;  // inf / nan guard
  // --- 6. roots, back-substitution
  Cplx roots[10];
  solve_poly10(cf, roots, 300);
  int count = 0;
  for (int i = 0; i < 10; ++i) {
    if (!(fabs(roots[i].im) <= 1e-10)) continue;
    const double z = roots[i].re;
    double Bz[3][3];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double s = P[r][c][4];
        for (int k = 3; k >= 0; --k) s = s * z + P[r][c][k];
        Bz[r][c] = s;
      }
    // null vector of the (rank-2) 3x3: best of the three row cross products
    double bv[3] = {0, 0, 0}, bn = -1;
    for (int a = 0; a < 3; ++a) {
      const int b = (a + 1) % 3;
      const double cx = Bz[a][1] * Bz[b][2] - Bz[a][2] * Bz[b][1];
      const double cy = Bz[a][2] * Bz[b][0] - Bz[a][0] * Bz[b][2];
      const double cz = Bz[a][0] * Bz[b][1] - Bz[a][1] * Bz[b][0];
      const double nn = cx * cx + cy * cy + cz * cz;
      if (nn > bn) { bn = nn; bv[0] = cx; bv[1] = cy; bv[2] = cz; }
    }
    if (!(bn > 0)) continue;
    const double inv = 1.0 / sqrt(bn);
    if (!(fabs(bv[2] * inv) >= 1e-10)) continue;
    const double x = bv[0] / bv[2], y = bv[1] / bv[2];
    double* E = Eout + 9 * count;
    double nn = 0;
    for (int k = 0; k < 9; ++k) {
      E[k] = x * N[0][k] + y * N[1][k] + z * N[2][k] + N[3][k];
      nn += E[k] * E[k];
    }
    nn = 1.0 / sqrt(nn);
    for (int k = 0; k < 9; ++k) E[k] *= nn;
    ++count;
  }
  return count;
}

// ------------------------------------------------------------------ small SVDs
// One-sided Jacobi on the columns of A (M x N, row-major in a[M*N]); V (N x N) accumulates the
// right rotations; on return the columns of A are orthogonal (A_in V = A_out).
template <int M, int N>
RPE_HD void jacobi_cols(double* a, double* v) {
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) v[i * N + j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 40; ++sweep) {
    bool changed = false;
    for (int p = 0; p < N - 1; ++p)
      for (int q = p + 1; q < N; ++q) {
        double al = 0, be = 0, ga = 0;
        for (int i = 0; i < M; ++i) {
          al += a[i * N + p] * a[i * N + p];
          be += a[i * N + q] * a[i * N + q];
          ga += a[i * N + p] * a[i * N + q];
        }
        if (!(fabs(ga) > 2.3e-16 * sqrt(al * be)) || ga == 0.0) continue;
        changed = true;
        const double zeta = (be - al) / (2.0 * ga);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
        for (int i = 0; i < M; ++i) {
          const double x = a[i * N + p], y = a[i * N + q];
          a[i * N + p] = c * x - s * y;
          a[i * N + q] = s * x + c * y;
        }
        for (int i = 0; i < N; ++i) {
          const double x = v[i * N + p], y = v[i * N + q];
          v[i * N + p] = c * x - s * y;
          v[i * N + q] = s * x + c * y;
        }
      }
    if (!changed) break;
  }
}

// decomposeEssentialMat: E -> R1, R2, t (unit).  Candidate order used downstream:
// (R1,t) (R2,t) (R1,-t) (R2,-t), as cv2.recoverPose.
RPE_HDN void decompose_essential(const double* E, double* R1, double* R2, double* t) {
  {
    // the static-pair matrix of five_point_solve step 0 has two equal singular values, so U and V are not unique;
    // cv2.recoverPose on it returns R = I, t = (1, -1, 1) / sqrt(3) (count 0): R1 = I makes candidate 0 win
    const double a0 = 0.40824829046386302;
    if (E[0] == 0.0 && E[4] == 0.0 && E[8] == 0.0 && E[1] == -a0 && E[2] == -a0 && E[5] == -a0 && E[3] == a0 &&
        E[6] == a0 && E[7] == a0) {
      const double q = 0.57735026918962576;      // 1 / sqrt(3)
      t[0] = q; t[1] = -q; t[2] = q;
      for (int i = 0; i < 9; ++i) R1[i] = (i % 4 == 0) ? 1.0 : 0.0;
      for (int i = 0; i < 3; ++i)                // R2 = 2 t t' - I: the half-turn about t (the twisted candidate)
        for (int j = 0; j < 3; ++j) R2[3 * i + j] = 2.0 * t[i] * t[j] - (i == j ? 1.0 : 0.0);
      return;
    }
  }
  double a[9], v[9];
  for (int i = 0; i < 9; ++i) a[i] = E[i];
  jacobi_cols<3, 3>(a, v);
  double s[3];
  for (int j = 0; j < 3; ++j) s[j] = a[j] * a[j] + a[3 + j] * a[3 + j] + a[6 + j] * a[6 + j];
  int o0 = 0, o1 = 1, o2 = 2;  // sort descending
  if (s[o0] < s[o1]) { int tt = o0; o0 = o1; o1 = tt; }
  if (s[o1] < s[o2]) { int tt = o1; o1 = o2; o2 = tt; }
  if (s[o0] < s[o1]) { int tt = o0; o0 = o1; o1 = tt; }
  double U[3][3], V[3][3];
  const double n0 = 1.0 / sqrt(s[o0]), n1 = 1.0 / sqrt(s[o1]);
  for (int i = 0; i < 3; ++i) {
    U[i][0] = a[3 * i + o0] * n0; U[i][1] = a[3 * i + o1] * n1;
    V[i][0] = v[3 * i + o0];      V[i][1] = v[3 * i + o1];
  }
  // third columns by cross product: det(U) = det(V) = +1 (cv2 flips whole matrices instead;
  // the candidate SET {R1,R2} x {t,-t} is identical)
  U[0][2] = U[1][0] * U[2][1] - U[2][0] * U[1][1];
  U[1][2] = U[2][0] * U[0][1] - U[0][0] * U[2][1];
  U[2][2] = U[0][0] * U[1][1] - U[1][0] * U[0][1];
  V[0][2] = V[1][0] * V[2][1] - V[2][0] * V[1][1];
  V[1][2] = V[2][0] * V[0][1] - V[0][0] * V[2][1];
  V[2][2] = V[0][0] * V[1][1] - V[1][0] * V[0][1];
  // U W = [-u1 u0 u2], U W' = [u1 -u0 u2]
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      R1[3 * i + j] = -U[i][1] * V[j][0] + U[i][0] * V[j][1] + U[i][2] * V[j][2];
      R2[3 * i + j] = U[i][1] * V[j][0] - U[i][0] * V[j][1] + U[i][2] * V[j][2];
    }
  t[0] = U[0][2]; t[1] = U[1][2]; t[2] = U[2][2];
}

// DLT triangulation with P0 = [I|0], P1 = [R|t] and the cv2.recoverPose cheirality test.
RPE_HDN bool triangulate_cheirality(const double* R, const double* t, double x1, double y1,
                                    double x2, double y2, double dist_thresh) {
  double a[16], v[16];
  a[0] = -1; a[1] = 0;  a[2] = x1; a[3] = 0;
  a[4] = 0;  a[5] = -1; a[6] = y1; a[7] = 0;
  a[8]  = x2 * R[6] - R[0]; a[9]  = x2 * R[7] - R[1]; a[10] = x2 * R[8] - R[2]; a[11] = x2 * t[2] - t[0];
  a[12] = y2 * R[6] - R[3]; a[13] = y2 * R[7] - R[4]; a[14] = y2 * R[8] - R[5]; a[15] = y2 * t[2] - t[1];
  jacobi_cols<4, 4>(a, v);
  int m = 0;
  double best = 1e300;
  for (int j = 0; j < 4; ++j) {
    const double s = a[j] * a[j] + a[4 + j] * a[4 + j] + a[8 + j] * a[8 + j] + a[12 + j] * a[12 + j];
    if (s < best) { best = s; m = j; }
  }
  const double Q0 = v[m], Q1 = v[4 + m], Q2 = v[8 + m], Q3 = v[12 + m];
  bool ok = (Q2 * Q3) > 0;
  const double X = Q0 / Q3, Y = Q1 / Q3, Z = Q2 / Q3;
  ok = ok && (Z < dist_thresh);
  const double z2 = R[6] * X + R[7] * Y + R[8] * Z + t[2];
  ok = ok && (z2 > 0) && (z2 < dist_thresh);
  return ok;
}

// ------------------------------------------------------------------ Euler (degrees)
// rotation_to_euler_yup / _zyx (reference: src/utils/geometry.py:48-80, 181-205): (yaw, pitch, roll)
RPE_HD void rotation_to_euler(const double* R, int convention, double* ypr) {
  const double r2d = 57.29577951308232;
  double yaw, pitch, roll;
  if (convention == 1) {
    const double sy = sqrt(R[0] * R[0] + R[3] * R[3]);
    if (!(sy < 1e-6)) {
      roll = atan2(R[7], R[8]); pitch = atan2(-R[6], sy); yaw = atan2(R[3], R[0]);
    } else {
      roll = atan2(-R[5], R[4]); pitch = atan2(-R[6], sy); yaw = 0.0;
    }
  } else {
    pitch = asin(R[7]);
    if (fabs(R[7]) > 0.9999) {
      roll = atan2(-R[5], R[4]); yaw = 0.0;
    } else {
      yaw = atan2(-R[6], R[0]); roll = atan2(R[3], R[4]);
    }
  }
  ypr[0] = yaw * r2d; ypr[1] = pitch * r2d; ypr[2] = roll * r2d;
}

}  // namespace rpe
