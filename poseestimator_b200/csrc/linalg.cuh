// Small dense linear algebra shared by rt.cu (GNC-TLS rotation) and refine.cu (Umeyama step of ICP, PCA axes).
#pragma once
#include <cuda_runtime.h>

namespace b2r {

__device__ __forceinline__ double det3(const double* M) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// R = V diag(1,1,det(U)det(V)) U^T from the SVD H = U S V^T (one-sided Jacobi on the columns of H).
// warm (optional, 9 doubles, row-major): an orthogonal start for V -- the V of a previous, similar H (successive GNC iterations);
// the sweeps then run on H * warm, whose columns are already nearly orthogonal, and converge in one or two instead of five.  The
// final V is written back.  First call: pass the identity.
__device__ inline void rotation_from_H(const double* H, double* R, double* warm = nullptr) {
  double a[3][3], v[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  if (warm) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        v[i][j] = warm[3 * i + j];
        a[i][j] = (H[3 * i] * warm[j] + H[3 * i + 1] * warm[3 + j]) + H[3 * i + 2] * warm[6 + j];
      }
  } else {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) a[i][j] = H[3 * i + j];
  }
  // Columns of (numerically) zero length -- H of two or three TIMs has rank <= 2 -- take no part in the sweeps: their inner products
  // with the others are pure rounding noise, the cosine `off` of noise against noise is O(1) and never falls, and the loop ran all 40
  // sweeps (336 k cycles on a 3-clique frame; 85 k with the test).  1e-26 * |H|_F^2 = (1e-13 |H|_F)^2.
  double fro2 = 0;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) fro2 += a[i][j] * a[i][j];
  const double null2 = 1e-26 * fro2;
  for (int sweep = 0; sweep < 40; ++sweep) {
    double off = 0;
#pragma unroll
    for (int pq = 0; pq < 3; ++pq) {
      const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
      const double alpha = a[0][p] * a[0][p] + a[1][p] * a[1][p] + a[2][p] * a[2][p];
      const double beta = a[0][q] * a[0][q] + a[1][q] * a[1][q] + a[2][q] * a[2][q];
      const double gamma = a[0][p] * a[0][q] + a[1][p] * a[1][q] + a[2][p] * a[2][q];
      if (gamma != 0.0 && alpha > null2 && beta > null2) {
        off = fmax(off, fabs(gamma) * rsqrt(alpha * beta + 1e-300));
        const double zeta = (beta - alpha) / (2 * gamma);
        const double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1 + zeta * zeta));
        const double c = rsqrt(1 + tt * tt), s = c * tt;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const double x = a[i][p], y = a[i][q];
          a[i][p] = c * x - s * y; a[i][q] = s * x + c * y;
          const double vx = v[i][p], vy = v[i][q];
          v[i][p] = c * vx - s * vy; v[i][q] = s * vx + c * vy;
        }
      }
    }
    // `off` is the largest column cosine seen BEFORE this sweep's rotations; cyclic Jacobi converges quadratically, so after a sweep that
    // started below 1e-10 the columns are orthogonal to ~1e-20.  (The former test, off < 1e-16, sat inside the rounding noise: most
    // calls ran all 40 sweeps, ~12 k cycles each -- the whole cost of a GNC iteration.)
    if (off < 1e-10) break;
  }
  if (warm) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) warm[3 * i + j] = v[i][j];
  }
  double sv[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) sv[j] = sqrt(a[0][j] * a[0][j] + a[1][j] * a[1][j] + a[2][j] * a[2][j]);
  // order columns by decreasing singular value (3-element sorting network on indices)
  int o0 = 0, o1 = 1, o2 = 2;
  if (sv[o0] < sv[o1]) { int t = o0; o0 = o1; o1 = t; }
  if (sv[o1] < sv[o2]) { int t = o1; o1 = o2; o2 = t; }
  if (sv[o0] < sv[o1]) { int t = o0; o0 = o1; o1 = t; }
  const int ord[3] = {o0, o1, o2};
  double U[9], V[9];
#pragma unroll
  for (int jj = 0; jj < 3; ++jj) {
    int j = ord[jj];
    double s = 0, c0 = 0, c1 = 0, c2 = 0, w0 = 0, w1 = 0, w2 = 0;
#pragma unroll
    for (int q = 0; q < 3; ++q)
      if (q == j) { s = sv[q]; c0 = a[0][q]; c1 = a[1][q]; c2 = a[2][q]; w0 = v[0][q]; w1 = v[1][q]; w2 = v[2][q]; }
    const double inv = s > 0 ? 1.0 / s : 0.0;
    U[jj] = c0 * inv; U[3 + jj] = c1 * inv; U[6 + jj] = c2 * inv;
    V[jj] = w0; V[3 + jj] = w1; V[6 + jj] = w2;
  }
  const double s0 = sv[o0], s1 = sv[o1], s2 = sv[o2];
  // complete U when H is (numerically) rank deficient, so that R stays a rotation
  if (!(s0 > 0)) { U[0] = 1; U[3] = 0; U[6] = 0; }
  if (!(s1 > 1e-12 * s0) || !(s0 > 0)) {
    const double ax = fabs(U[0]), ay = fabs(U[3]), az = fabs(U[6]);
    double e0 = 0, e1 = 0, e2 = 0;
    if (ax <= ay && ax <= az) e0 = 1; else if (ay <= az) e1 = 1; else e2 = 1;
    const double dd = e0 * U[0] + e1 * U[3] + e2 * U[6];
    double w0 = e0 - dd * U[0], w1 = e1 - dd * U[3], w2 = e2 - dd * U[6];
    const double nw = rsqrt(w0 * w0 + w1 * w1 + w2 * w2);
    U[1] = w0 * nw; U[4] = w1 * nw; U[7] = w2 * nw;
  }
  if (!(s2 > 1e-12 * s0) || !(s0 > 0)) {
    U[2] = U[3] * U[7] - U[6] * U[4];
    U[5] = U[6] * U[1] - U[0] * U[7];
    U[8] = U[0] * U[4] - U[3] * U[1];
  }
  if (det3(U) * det3(V) < 0) { V[2] = -V[2]; V[5] = -V[5]; V[8] = -V[8]; }
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) R[3 * r + c] = V[3 * r] * U[3 * c] + V[3 * r + 1] * U[3 * c + 1] + V[3 * r + 2] * U[3 * c + 2];
}

// The same rotation without an SVD when H is well conditioned: R = V U^T is the orthogonal polar factor of H^T, and Newton's iteration
// X <- (X + X^-T) / 2 from X0 = H^T / |H|_F converges to it quadratically (every singular value s -> (s + 1/s) / 2).  One iteration is a
// cofactor matrix and one division -- ~200 cycles of dependent latency against ~1500 for a Jacobi sweep of three rotations with their
// rsqrt / sqrt / div chains; the GNC loop calls this once per iteration from a single thread while 255 others wait.  Declines (returns
// false: the caller runs the Jacobi SVD) when det(H) <= 0 (the optimum is then V diag(1,1,-1) U^T, not the polar factor), when H is
// close to rank deficient, or when the iteration has not settled after 24 steps.  Agreement with the SVD path: ~1e-15.
__device__ inline bool rotation_from_H_polar(const double* H, double* R) {
  double f = 0.0;
#pragma unroll
  for (int i = 0; i < 9; ++i) f += H[i] * H[i];
  if (!(f > 0.0) || !(f < 1e300)) return false;
  const double inv = rsqrt(f);
  double X[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) X[3 * r + c] = H[3 * c + r] * inv;
  const double d0 = det3(X);
  if (!(d0 > 1e-5)) return false;   // |X|_F = 1: det <= 3^-1.5; 1e-5 keeps the smallest singular value above ~1e-4 of the largest
  // Determinant scaling (Higham): X0 * det^(-1/3) has singular values of geometric mean 1, which removes the slow first steps of the
  // unscaled iteration (s -> (s + 1/s)/2 from s ~ 0.05 needs five steps to reach 1; from the scaled start, three).
  const double gam = (double)(1.0f / cbrtf((float)d0));
#pragma unroll
  for (int i = 0; i < 9; ++i) X[i] *= gam;
  int exact_steps = 0;
  bool fast = true;   // far from the fixed point 1/det needs ~1e-12, not 0.5 ulp: float reciprocal + one Newton step instead of a division
  for (int it = 0; it < 24; ++it) {
    double C[9];
    C[0] = X[4] * X[8] - X[5] * X[7]; C[1] = X[5] * X[6] - X[3] * X[8]; C[2] = X[3] * X[7] - X[4] * X[6];
    C[3] = X[2] * X[7] - X[1] * X[8]; C[4] = X[0] * X[8] - X[2] * X[6]; C[5] = X[1] * X[6] - X[0] * X[7];
    C[6] = X[1] * X[5] - X[2] * X[4]; C[7] = X[2] * X[3] - X[0] * X[5]; C[8] = X[0] * X[4] - X[1] * X[3];
    const double det = X[0] * C[0] + X[1] * C[1] + X[2] * C[2];
    if (!(det > 0.0)) return false;
    double hd;
    if (fast) {
      const double r0 = (double)__frcp_rn((float)det);
      hd = 0.5 * (r0 * (2.0 - det * r0));
    } else {
      hd = 0.5 / det;
      ++exact_steps;
    }
    double dv[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const double xn = 0.5 * X[i] + hd * C[i];    // X^-T = cofactor matrix / det
      dv[i] = fabs(xn - X[i]);
      X[i] = xn;
    }
    // (a tree: the nine-deep fmax chain was a quarter of a step's dependent latency)
    const double diff = fmax(fmax(fmax(dv[0], dv[1]), fmax(dv[2], dv[3])), fmax(fmax(fmax(dv[4], dv[5]), fmax(dv[6], dv[7])), dv[8]));
    if (fast) {
      if (diff < 1e-4) fast = false;   // the remaining steps use the exact division
      continue;
    }
    // quadratic convergence: an exact step of size diff leaves an error of ~diff^2; two exact steps after diff < 1e-4 leave ~1e-32
    if (diff < 1e-9 || exact_steps >= 2) {
      if (!(diff < 1e-7)) return false;
#pragma unroll
      for (int i = 0; i < 9; ++i) R[i] = X[i];
      return true;
    }
  }
  return false;
}

// Cyclic Jacobi eigen-decomposition of a symmetric 3x3 (row-major C): eigenvalues w, eigenvectors = columns of Q (unsorted).
__device__ inline void eig3_sym(const double* C, double* w, double* Q) {
  double a[3][3], q[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = C[3 * i + j];
  for (int sweep = 0; sweep < 60; ++sweep) {
    const double off = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
    const double dia = fabs(a[0][0]) + fabs(a[1][1]) + fabs(a[2][2]);
    if (off <= 1e-300 || off <= 1e-17 * dia) break;
    for (int p = 0; p < 2; ++p) for (int r = p + 1; r < 3; ++r) {
      if (a[p][r] == 0.0) continue;
      const double theta = (a[r][r] - a[p][p]) / (2 * a[p][r]);
      const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(1 + theta * theta));
      const double c = 1.0 / sqrt(1 + t * t), s = c * t;
      for (int k = 0; k < 3; ++k) { const double x = a[k][p], y = a[k][r]; a[k][p] = c * x - s * y; a[k][r] = s * x + c * y; }
      for (int k = 0; k < 3; ++k) { const double x = a[p][k], y = a[r][k]; a[p][k] = c * x - s * y; a[r][k] = s * x + c * y; }
      for (int k = 0; k < 3; ++k) { const double x = q[k][p], y = q[k][r]; q[k][p] = c * x - s * y; q[k][r] = s * x + c * y; }
    }
  }
  for (int i = 0; i < 3; ++i) { w[i] = a[i][i]; for (int j = 0; j < 3; ++j) Q[3 * i + j] = q[i][j]; }
}

}  // namespace b2r
