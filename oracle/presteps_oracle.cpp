// =============================================================================
// oracle/presteps_oracle.cpp  --  TEST INFRASTRUCTURE ONLY.  NOT PRODUCT CODE.
//
// CPU restatement of the Open3D operations the reference runs before TEASER++
// (SURVEY.md Appendix B, rows a-1a .. a-2 of section 8):
//   back-projection        RealSenseClass.py:56-67, registration_utils.py:54-90, PoseEstimator.py:130-135
//   statistical outliers   RealSenseClass.py:68, registration_utils.py:93
//   voxel_down_sample      PoseEstimator.py:87,93
//   estimate_normals       HelpersRealtime.py:215-219
//   compute_fpfh_feature   HelpersRealtime.py:221-224
//   correspondences        HelpersRealtime.py:194-213 -> north-star: mutual nearest neighbour on FPFH
// (paths under /root/reference/src/pose_estimator/EstimHelpers/).
//
// Open3D (hint ==0.19.0, pyproject.toml:49) is NOT on this machine: this file restates its published algorithms
// from SURVEY.md Appendix B.  PARITY UNPINNED -- validated in tests/test_oracle_presteps.py against scipy.cKDTree,
// numpy.linalg.eigh and from-the-definition numpy restatements.
//
// Canonical choices where Open3D's own behaviour is unspecified (documented deviations):
//   * neighbour order: ascending (squared distance, index); Open3D's order on ties depends on KD-tree traversal;
//   * voxel output order: ascending voxel key (ix + nx*(iy + ny*iz)); Open3D's is unordered_map iteration order;
//     inside a voxel points are summed in ascending point index, as Open3D's sequential insert does;
//   * normal sign: as produced by the analytic eigen-solver below, optionally flipped towards / away from a viewpoint.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may load this library.
// Build: oracle/Makefile (-O3 -fopenmp -ffp-contract=off).
// =============================================================================
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <vector>
#include <omp.h>

// acos / cos / atan2 from +, -, *, /, sqrt, fma only (double-double): the SAME header the CUDA kernels include, so both sides
// bin pair features and solve the 3x3 eigen-problem on identical values; checked against mpmath and glibc in
// tests/test_oracle_exactmath.py.  Open3D itself calls the host libm, which these agree with to the last ulp except on the rare
// inputs where glibc is not correctly rounded.
#include "../include/b2r_exactmath.h"

namespace {

inline double sq3(double x, double y, double z) { return (x * x + y * y) + z * z; }

struct Nb { double d2; int idx; };
inline bool nb_less(const Nb& a, const Nb& b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.idx < b.idx); }

// B.4 / Open3D KDTreeFlann::SearchHybrid + SearchKNN semantics: the max_nn nearest points of the cloud (the query
// point itself included, d2 = 0), then -- when radius > 0 -- only those with d2 < radius^2.
// Brute force on purpose: obviously correct, and the canonical (d2, index) order is explicit.
int knn_one(const double* pts, int n, const double* q, int max_nn, double radius, std::vector<Nb>& heap, Nb* out) {
  heap.clear();
  const double r2 = radius * radius;
  for (int j = 0; j < n; ++j) {
    const double dx = pts[3 * j] - q[0], dy = pts[3 * j + 1] - q[1], dz = pts[3 * j + 2] - q[2];
    const double d2 = sq3(dx, dy, dz);
    if (radius > 0 && !(d2 < r2)) continue;
    Nb c{d2, j};
    if ((int)heap.size() < max_nn) {
      heap.push_back(c);
      std::push_heap(heap.begin(), heap.end(), nb_less);
    } else if (nb_less(c, heap.front())) {
      std::pop_heap(heap.begin(), heap.end(), nb_less);
      heap.back() = c;
      std::push_heap(heap.begin(), heap.end(), nb_less);
    }
  }
  std::sort_heap(heap.begin(), heap.end(), nb_less);
  const int k = (int)heap.size();
  for (int i = 0; i < k; ++i) out[i] = heap[i];
  return k;
}

void knn_all(const double* pts, int n, const double* queries, int nq, int max_nn, double radius, int* idx, double* d2, int* cnt) {
#pragma omp parallel
  {
    std::vector<Nb> heap;
    std::vector<Nb> out(max_nn);
#pragma omp for schedule(dynamic, 64)
    for (int i = 0; i < nq; ++i) {
      const int k = knn_one(pts, n, queries + 3 * i, max_nn, radius, heap, out.data());
      cnt[i] = k;
      for (int t = 0; t < max_nn; ++t) {
        idx[(size_t)i * max_nn + t] = t < k ? out[t].idx : -1;
        d2[(size_t)i * max_nn + t] = t < k ? out[t].d2 : 0.0;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------
// B.4 analytic eigenvector of the smallest eigenvalue of a symmetric 3x3 (Open3D FastEigen3x3, after Eberly,
// "A Robust Eigensolver for 3x3 Symmetric Matrices").  A is row-major, symmetric.
// ---------------------------------------------------------------------------------------------------------
inline void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}
inline double dot3(const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

void eigvec0(const double A[9], double ev, double out[3]) {
  const double r0[3] = {A[0] - ev, A[1], A[2]};
  const double r1[3] = {A[1], A[4] - ev, A[5]};
  const double r2[3] = {A[2], A[5], A[8] - ev};
  double c01[3], c02[3], c12[3];
  cross3(r0, r1, c01); cross3(r0, r2, c02); cross3(r1, r2, c12);
  const double d0 = dot3(c01, c01), d1 = dot3(c02, c02), d2 = dot3(c12, c12);
  double dmax = d0; int imax = 0;
  if (d1 > dmax) { dmax = d1; imax = 1; }
  if (d2 > dmax) { imax = 2; }
  const double* c = imax == 0 ? c01 : (imax == 1 ? c02 : c12);
  const double d = imax == 0 ? d0 : (imax == 1 ? d1 : d2);
  const double s = std::sqrt(d);
  out[0] = c[0] / s; out[1] = c[1] / s; out[2] = c[2] / s;
}

void eigvec1(const double A[9], const double e0[3], double ev1, double out[3]) {
  double U[3], V[3];
  if (std::fabs(e0[0]) > std::fabs(e0[1])) {
    const double inv = 1.0 / std::sqrt(e0[0] * e0[0] + e0[2] * e0[2]);
    U[0] = -e0[2] * inv; U[1] = 0; U[2] = e0[0] * inv;
  } else {
    const double inv = 1.0 / std::sqrt(e0[1] * e0[1] + e0[2] * e0[2]);
    U[0] = 0; U[1] = e0[2] * inv; U[2] = -e0[1] * inv;
  }
  cross3(e0, U, V);
  const double AU[3] = {(A[0] * U[0] + A[1] * U[1]) + A[2] * U[2], (A[1] * U[0] + A[4] * U[1]) + A[5] * U[2],
                        (A[2] * U[0] + A[5] * U[1]) + A[8] * U[2]};
  const double AV[3] = {(A[0] * V[0] + A[1] * V[1]) + A[2] * V[2], (A[1] * V[0] + A[4] * V[1]) + A[5] * V[2],
                        (A[2] * V[0] + A[5] * V[1]) + A[8] * V[2]};
  double m00 = dot3(U, AU) - ev1, m01 = dot3(U, AV), m11 = dot3(V, AV) - ev1;
  const double a00 = std::fabs(m00), a01 = std::fabs(m01), a11 = std::fabs(m11);
  if (a00 >= a11) {
    const double mx = std::max(a00, a01);
    if (mx > 0) {
      if (a00 >= a01) { m01 /= m00; m00 = 1 / std::sqrt(1 + m01 * m01); m01 *= m00; }
      else { m00 /= m01; m01 = 1 / std::sqrt(1 + m00 * m00); m00 *= m01; }
      for (int i = 0; i < 3; ++i) out[i] = m01 * U[i] - m00 * V[i];
    } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
  } else {
    const double mx = std::max(a11, a01);
    if (mx > 0) {
      if (a11 >= a01) { m01 /= m11; m11 = 1 / std::sqrt(1 + m01 * m01); m01 *= m11; }
      else { m11 /= m01; m01 = 1 / std::sqrt(1 + m11 * m11); m11 *= m01; }
      for (int i = 0; i < 3; ++i) out[i] = m11 * U[i] - m01 * V[i];
    } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
  }
}

void smallest_eigvec(const double C[9], double out[3]) {
  double A[9];
  double mx = C[0];
  for (int i = 1; i < 9; ++i) mx = std::max(mx, C[i]);
  if (mx == 0) { out[0] = out[1] = out[2] = 0; return; }
  for (int i = 0; i < 9; ++i) A[i] = C[i] / mx;
  const double norm = (A[1] * A[1] + A[2] * A[2]) + A[5] * A[5];
  if (norm > 0) {
    const double q = ((A[0] + A[4]) + A[8]) / 3.0;
    const double b00 = A[0] - q, b11 = A[4] - q, b22 = A[8] - q;
    const double p = std::sqrt((((b00 * b00 + b11 * b11) + b22 * b22) + norm * 2.0) / 6.0);
    const double c00 = b11 * b22 - A[5] * A[5];
    const double c01 = A[1] * b22 - A[5] * A[2];
    const double c02 = A[1] * A[5] - b11 * A[2];
    const double det = ((b00 * c00 - A[1] * c01) + A[2] * c02) / ((p * p) * p);
    double half = det * 0.5;
    half = std::min(std::max(half, -1.0), 1.0);
    const double angle = b2r_acos(half) / 3.0;
    const double two_thirds_pi = 2.09439510239319549;
    const double beta2 = b2r_cos(angle) * 2.0;
    const double beta0 = b2r_cos(angle + two_thirds_pi) * 2.0;
    const double beta1 = -(beta0 + beta2);
    const double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;  // e0 <= e1 <= e2
    double v0[3], v1[3], v2[3];
    if (half >= 0) {
      eigvec0(A, e2, v2);
      if (e2 < e0 && e2 < e1) { std::memcpy(out, v2, 24); return; }
      eigvec1(A, v2, e1, v1);
      if (e1 < e0 && e1 < e2) { std::memcpy(out, v1, 24); return; }
      cross3(v1, v2, out);
    } else {
      eigvec0(A, e0, v0);
      if (e0 < e1 && e0 < e2) { std::memcpy(out, v0, 24); return; }
      eigvec1(A, v0, e1, v1);
      if (e1 < e0 && e1 < e2) { std::memcpy(out, v1, 24); return; }
      cross3(v0, v1, out);
    }
  } else {  // diagonal
    if (C[0] < C[4] && C[0] < C[8]) { out[0] = 1; out[1] = 0; out[2] = 0; }
    else if (C[4] < C[0] && C[4] < C[8]) { out[0] = 0; out[1] = 1; out[2] = 0; }
    else { out[0] = 0; out[1] = 0; out[2] = 1; }
  }
}

}  // namespace

extern "C" {

// ---------------------------------------------------------------------------------------------------------
// B.1  depth image -> point cloud (row-major scan order inside the bbox [u0,u1) x [v0,v1)).
//   d32 = float(depth) * float(depth_scale)          RealSenseClass.py:57 (numpy float32 * python float -> float32)
//   mask == 0 -> 0                                   RealSenseClass.py:58 (mask may be NULL)
//   zmin/zmax (> 0): keep zmin < d32 < zmax          registration_utils.py:75
//   Open3D RGBDImage: d32 >= depth_trunc -> 0  (depth_trunc <= 0: no truncation); default 3.0 on the camera path
//   Open3D create_from_rgbd_image: d32 > 0: z = double(d32); x = (u - cx) * z / fx; y = (v - cy) * z / fy
// returns M; pts: M x 3 doubles, pix: M linear pixel ids (v * W + u).
int orcp_backproject(const uint16_t* depth, const uint8_t* mask, int H, int W, int u0, int v0, int u1, int v1, double fx,
                     double fy, double cx, double cy, double depth_scale, double depth_trunc, double zmin, double zmax,
                     double* pts, int* pix) {
  int m = 0;
  u0 = std::max(u0, 0); v0 = std::max(v0, 0); u1 = std::min(u1, W); v1 = std::min(v1, H);
  const float scale = (float)depth_scale;
  for (int v = v0; v < v1; ++v)
    for (int u = u0; u < u1; ++u) {
      float d = (float)depth[(size_t)v * W + u] * scale;
      if (mask && mask[(size_t)v * W + u] == 0) d = 0.f;
      if (zmin > 0 || zmax > 0) {
        const double dd = (double)d;  // numpy compares the float32 array with python floats in double
        if (!(dd > zmin && dd < zmax)) d = 0.f;
      }
      if (depth_trunc > 0 && d >= (float)depth_trunc) d = 0.f;
      if (d > 0.f) {
        const double z = (double)d;
        pts[3 * m] = (u - cx) * z / fx;
        pts[3 * m + 1] = (v - cy) * z / fy;
        pts[3 * m + 2] = z;
        if (pix) pix[m] = v * W + u;
        ++m;
      }
    }
  return m;
}

// hybrid / knn search for every query (radius <= 0: pure knn).  idx/d2: nq x max_nn (padding -1 / 0), cnt: nq.
void orcp_knn(const double* pts, int n, const double* queries, int nq, int max_nn, double radius, int* idx, double* d2,
              int* cnt) {
  knn_all(pts, n, queries, nq, max_nn, radius, idx, d2, cnt);
}

// B.2 remove_statistical_outlier(nb_neighbors, std_ratio): keep[i] in {0,1}; returns number kept.
int orcp_statistical_outlier(const double* pts, int n, int nb_neighbors, double std_ratio, uint8_t* keep, double* avg_out) {
  if (n == 0) return 0;
  const int k = std::min(nb_neighbors, n);
  std::vector<int> idx((size_t)n * k), cnt(n);
  std::vector<double> d2((size_t)n * k), avg(n);
  knn_all(pts, n, pts, n, k, 0.0, idx.data(), d2.data(), cnt.data());
  size_t valid = 0;
  for (int i = 0; i < n; ++i) {
    double mean = -1.0;
    if (cnt[i] > 0) {
      double s = 0;
      for (int t = 0; t < cnt[i]; ++t) s += std::sqrt(d2[(size_t)i * k + t]);  // sequential, ascending distance
      mean = s / (double)cnt[i];
      ++valid;
    }
    avg[i] = mean;
  }
  if (avg_out) std::copy(avg.begin(), avg.end(), avg_out);
  if (valid == 0) { std::fill(keep, keep + n, 0); return 0; }
  double sum = 0;
  for (int i = 0; i < n; ++i) if (avg[i] > 0) sum += avg[i];
  const double cloud_mean = sum / (double)valid;
  double sq = 0;
  for (int i = 0; i < n; ++i) sq += avg[i] > 0 ? (avg[i] - cloud_mean) * (avg[i] - cloud_mean) : 0.0;
  const double std_dev = std::sqrt(sq / (double)(valid - 1));
  const double thr = cloud_mean + std_ratio * std_dev;
  int kept = 0;
  for (int i = 0; i < n; ++i) { keep[i] = (avg[i] > 0 && avg[i] < thr) ? 1 : 0; kept += keep[i]; }
  return kept;
}

// B.3 voxel_down_sample(vs): returns V; out: V x 3 means in ascending voxel key order; vox_of_point (n) optional.
int orcp_voxel_down(const double* pts, int n, double vs, double* out, int* vox_of_point, int* counts_out) {
  if (n == 0) return 0;
  double mn[3] = {pts[0], pts[1], pts[2]}, mx[3] = {pts[0], pts[1], pts[2]};
  for (int i = 1; i < n; ++i)
    for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], pts[3 * i + a]); mx[a] = std::max(mx[a], pts[3 * i + a]); }
  double org[3];
  long long dim[3];
  for (int a = 0; a < 3; ++a) {
    org[a] = mn[a] - vs * 0.5;
    dim[a] = (long long)std::floor((mx[a] - org[a]) / vs) + 1;
  }
  std::vector<std::pair<long long, int>> keyed(n);
  for (int i = 0; i < n; ++i) {
    long long c[3];
    for (int a = 0; a < 3; ++a) c[a] = (long long)std::floor((pts[3 * i + a] - org[a]) / vs);
    keyed[i] = {c[0] + dim[0] * (c[1] + dim[1] * c[2]), i};
  }
  std::sort(keyed.begin(), keyed.end());  // (key, point index): ascending index inside a voxel
  int V = 0;
  for (int s = 0; s < n;) {
    int e = s;
    double acc[3] = {0, 0, 0};
    while (e < n && keyed[e].first == keyed[s].first) {
      const int i = keyed[e].second;
      acc[0] += pts[3 * i]; acc[1] += pts[3 * i + 1]; acc[2] += pts[3 * i + 2];
      if (vox_of_point) vox_of_point[i] = V;
      ++e;
    }
    const double c = (double)(e - s);
    out[3 * V] = acc[0] / c; out[3 * V + 1] = acc[1] / c; out[3 * V + 2] = acc[2] / c;
    if (counts_out) counts_out[V] = e - s;
    ++V;
    s = e;
  }
  return V;
}

// B.4 estimate_normals(KDTreeSearchParamHybrid(radius, max_nn)), fast_normal_computation=True.
// orient: 0 none, 1 towards `view`, 2 away from `view` (Open3D orient_normals_towards_camera_location and its mirror).
void orcp_normals(const double* pts, int n, double radius, int max_nn, int orient, const double* view, double* normals) {
  const int K = std::max(1, std::min(max_nn, n));
  std::vector<int> idx((size_t)n * K), cnt(n);
  std::vector<double> d2((size_t)n * K);
  knn_all(pts, n, pts, n, K, radius, idx.data(), d2.data(), cnt.data());
  for (int i = 0; i < n; ++i) {
    double nrm[3] = {0, 0, 1};
    if (cnt[i] >= 3) {
      double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int t = 0; t < cnt[i]; ++t) {  // cumulants in neighbour order
        const double* p = pts + 3 * idx[(size_t)i * K + t];
        c[0] += p[0]; c[1] += p[1]; c[2] += p[2];
        c[3] += p[0] * p[0]; c[4] += p[0] * p[1]; c[5] += p[0] * p[2];
        c[6] += p[1] * p[1]; c[7] += p[1] * p[2]; c[8] += p[2] * p[2];
      }
      const double m = (double)cnt[i];
      for (int j = 0; j < 9; ++j) c[j] /= m;
      double C[9];
      C[0] = c[3] - c[0] * c[0]; C[4] = c[6] - c[1] * c[1]; C[8] = c[8] - c[2] * c[2];
      C[1] = C[3] = c[4] - c[0] * c[1];
      C[2] = C[6] = c[5] - c[0] * c[2];
      C[5] = C[7] = c[7] - c[1] * c[2];
      smallest_eigvec(C, nrm);
      const double len = std::sqrt(sq3(nrm[0], nrm[1], nrm[2]));
      if (len == 0.0 || !(len == len)) { nrm[0] = 0; nrm[1] = 0; nrm[2] = 1; }
    }
    if (orient && view) {
      const double dx = view[0] - pts[3 * i], dy = view[1] - pts[3 * i + 1], dz = view[2] - pts[3 * i + 2];
      double dp = (nrm[0] * dx + nrm[1] * dy) + nrm[2] * dz;
      if (orient == 2) dp = -dp;
      if (dp < 0) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
    }
    normals[3 * i] = nrm[0]; normals[3 * i + 1] = nrm[1]; normals[3 * i + 2] = nrm[2];
  }
}

}  // extern "C"

namespace {
// B.5 pair features (f0 = atan2 term, f1 = v.n2, f2 = angle, f3 = distance)
inline void pair_features(const double* p1, const double* n1, const double* p2, const double* n2, double f[4]) {
  double dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  f[0] = f[1] = f[2] = 0;
  f[3] = std::sqrt(sq3(dp[0], dp[1], dp[2]));
  if (f[3] == 0.0) return;
  const double* a = n1;
  const double* b = n2;
  const double angle1 = dot3(n1, dp) / f[3];
  const double angle2 = dot3(n2, dp) / f[3];
  if (b2r_acos(std::fabs(angle1)) > b2r_acos(std::fabs(angle2))) {
    a = n2; b = n1;
    dp[0] = -dp[0]; dp[1] = -dp[1]; dp[2] = -dp[2];
    f[2] = -angle2;
  } else {
    f[2] = angle1;
  }
  double v[3];
  cross3(dp, a, v);
  const double vn = std::sqrt(sq3(v[0], v[1], v[2]));
  if (vn == 0.0) { f[0] = f[1] = f[2] = f[3] = 0; return; }
  v[0] /= vn; v[1] /= vn; v[2] /= vn;
  double w[3];
  cross3(a, v, w);
  f[1] = dot3(v, b);
  f[0] = b2r_atan2(dot3(w, b), dot3(a, b));
}
inline int bin11(double x) {
  int h = (int)std::floor(x);
  return h < 0 ? 0 : (h > 10 ? 10 : h);
}
}  // namespace

extern "C" {

// B.5 compute_fpfh_feature(KDTreeSearchParamHybrid(radius, max_nn)).  out: 33 x n, row-major (Open3D Feature.data),
// spfh_out (optional) likewise.
void orcp_fpfh(const double* pts, const double* normals, int n, double radius, int max_nn, double* out, double* spfh_out) {
  const int K = std::max(1, std::min(max_nn, n));
  std::vector<int> idx((size_t)n * K), cnt(n);
  std::vector<double> d2((size_t)n * K);
  knn_all(pts, n, pts, n, K, radius, idx.data(), d2.data(), cnt.data());
  std::vector<double> spfh((size_t)33 * n, 0.0);
  const double kPi = 3.14159265358979323846;
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; ++i) {
    if (cnt[i] > 1) {
      const double incr = 100.0 / (double)(cnt[i] - 1);
      for (int t = 1; t < cnt[i]; ++t) {
        const int j = idx[(size_t)i * K + t];
        double f[4];
        pair_features(pts + 3 * i, normals + 3 * i, pts + 3 * j, normals + 3 * j, f);
        spfh[(size_t)bin11(11 * (f[0] + kPi) / (2.0 * kPi)) * n + i] += incr;
        spfh[(size_t)(bin11(11 * (f[1] + 1.0) * 0.5) + 11) * n + i] += incr;
        spfh[(size_t)(bin11(11 * (f[2] + 1.0) * 0.5) + 22) * n + i] += incr;
      }
    }
  }
  std::fill(out, out + (size_t)33 * n, 0.0);
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < n; ++i) {
    if (cnt[i] > 1) {
      double sum[3] = {0, 0, 0};
      for (int t = 1; t < cnt[i]; ++t) {
        const double dist = d2[(size_t)i * K + t];
        if (dist == 0.0) continue;
        const int j = idx[(size_t)i * K + t];
        for (int b = 0; b < 33; ++b) {
          const double val = spfh[(size_t)b * n + j] / dist;
          sum[b / 11] += val;
          out[(size_t)b * n + i] += val;
        }
      }
      for (int g = 0; g < 3; ++g) if (sum[g] != 0.0) sum[g] = 100.0 / sum[g];
      for (int b = 0; b < 33; ++b) {
        out[(size_t)b * n + i] *= sum[b / 11];
        out[(size_t)b * n + i] += spfh[(size_t)b * n + i];
      }
    }
  }
  if (spfh_out) std::copy(spfh.begin(), spfh.end(), spfh_out);
}

// B.6 (north-star form): nearest neighbour in feature space both ways + mutual filter.
// fa: na x dim, fb: nb x dim (row-major, one descriptor per row).  Distances: sequential sum_j (a_j-b_j)^2 in fp64,
// lowest index wins ties.  nn01 (na), nn10 (nb); pairs: (i, nn01[i]) with nn10[nn01[i]] == i, ascending i; returns count.
int orcp_mutual_nn(const double* fa, int na, const double* fb, int nb, int dim, int* nn01, int* nn10, int* pairs) {
  auto nearest = [dim](const double* q, const double* base, int n) {
    int best = -1;
    double bd = std::numeric_limits<double>::infinity();
    for (int j = 0; j < n; ++j) {
      const double* r = base + (size_t)j * dim;
      double s = 0;
      for (int t = 0; t < dim; ++t) { const double d = q[t] - r[t]; s += d * d; }
      if (s < bd) { bd = s; best = j; }
    }
    return best;
  };
#pragma omp parallel for schedule(static)
  for (int i = 0; i < na; ++i) nn01[i] = nearest(fa + (size_t)i * dim, fb, nb);
#pragma omp parallel for schedule(static)
  for (int j = 0; j < nb; ++j) nn10[j] = nearest(fb + (size_t)j * dim, fa, na);
  int c = 0;
  for (int i = 0; i < na; ++i)
    if (nn01[i] >= 0 && nn10[nn01[i]] == i) {
      if (pairs) { pairs[2 * c] = i; pairs[2 * c + 1] = nn01[i]; }
      ++c;
    }
  return c;
}

void orcp_smallest_eigvec(const double* C9, double* out3) { smallest_eigvec(C9, out3); }
int orcp_num_threads() { return omp_get_max_threads(); }

// include/b2r_exactmath.h on arrays (tests/test_oracle_exactmath.py): which = 0 acos(a), 1 cos(a), 2 atan2(a, b)
void orcp_exactmath(int which, const double* a, const double* b, int n, double* out) {
  for (int i = 0; i < n; ++i) out[i] = which == 0 ? b2r_acos(a[i]) : which == 1 ? b2r_cos(a[i]) : b2r_atan2(a[i], b[i]);
}

}  // extern "C"
