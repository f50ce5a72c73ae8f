// =============================================================================
// oracle/teaser_oracle.cpp  --  TEST INFRASTRUCTURE ONLY.  NOT PRODUCT CODE.
//
// CPU restatement of the TEASER++ robust-registration algorithm that the
// reference calls through `teaserpp_python`:
//   /root/reference/src/pose_estimator/EstimHelpers/HelpersRealtime.py:158-172
//   /root/reference/src/pose_estimator/EstimHelpers/registration_utils.py:251-273
// The arithmetic lives in MIT-SPARK/TEASER-plusplus (+ PMC), an UN-VENDORED,
// UN-PINNED submodule (/root/reference/.gitmodules:1-3, ignored by .gitignore:30).
// No TEASER++/PMC/Eigen source exists on this machine, so this file restates
// the published algorithm (SURVEY.md Appendix A, items A.1-A.8 cited inline).
//
// PARITY UNPINNED: the reference ships no tests / golden vectors for this path
// and cannot be executed here; this oracle is validated by ground-truth
// recovery and by independent cross-checks (networkx max clique, numpy SVD,
// a numpy restatement of the edge test) in tests/.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.  The product path
// (poseestimator_b200/) never does.
//
// Structure follows upstream on purpose (materialised TIMs + index maps,
// adjacency lists, k-core -> heuristic -> exact branch-and-bound with OpenMP)
// so that timing it is a fair stand-in for the reference CPU path.
//
// Build: see oracle/Makefile  (-O3 -fopenmp -ffp-contract=off, no -march=native:
// upstream wheels are generic x86-64, so no FMA contraction -- A.4-3).
// =============================================================================
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <numeric>
#include <utility>
#include <vector>
#include <omp.h>

namespace {

using u64 = uint64_t;
using u32 = uint32_t;

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ----------------------------------------------------------------------------
// Parameters (A.1).  Plain C layout, mirrored by oracle/__init__.py (ctypes).
// ----------------------------------------------------------------------------
struct Params {
  double noise_bound;                 // 0.01
  double cbar2;                       // 1
  int estimate_scaling;               // upstream default 1; reference sets 0
  int rotation_estimation_algorithm;  // 0 GNC_TLS, 1 FGR
  double rotation_gnc_factor;         // 1.4
  int rotation_max_iterations;        // 100
  double rotation_cost_threshold;     // 1e-6
  int rotation_tim_graph;             // 0 CHAIN, 1 COMPLETE
  int inlier_selection_mode;          // 0 PMC_EXACT, 1 PMC_HEU, 2 KCORE_HEU, 3 NONE
  double kcore_heuristic_threshold;   // 0.5
  double max_clique_time_limit;       // 3000 s (not enforced by the oracle)
  int max_clique_num_threads;         // <=0 -> omp max
  double rotation_noise_bound;        // <=0 -> noise_bound (A.6-2 compounding is the caller's job)
  int materialize_tims;               // 1: like upstream; 0: stream the edge test (large N)
};

// (x^2 + y^2) + z^2 with separate mul / add  (A.4-3).
inline double sqn3(double x, double y, double z) { return (x * x + y * y) + z * z; }

// ----------------------------------------------------------------------------
// Dense bit graph (64-bit words).
// ----------------------------------------------------------------------------
struct BitGraph {
  int n = 0, W = 0;
  std::vector<u64> b;
  void init(int n_) { n = n_; W = (n + 63) / 64; b.assign((size_t)n * W, 0); }
  u64* row(int v) { return b.data() + (size_t)v * W; }
  const u64* row(int v) const { return b.data() + (size_t)v * W; }
  void set(int i, int j) { row(i)[j >> 6] |= (u64)1 << (j & 63); }
  bool has(int i, int j) const { return (row(i)[j >> 6] >> (j & 63)) & 1; }
};

inline int popc(const u64* p, int W) { int c = 0; for (int w = 0; w < W; ++w) c += __builtin_popcountll(p[w]); return c; }

// ----------------------------------------------------------------------------
// Solver state (what the getters read).
// ----------------------------------------------------------------------------
struct State {
  Params p;
  int n = 0;
  int64_t P = 0;                       // number of TIMs  N(N-1)/2
  std::vector<double> src_tims, dst_tims;   // 3 x P, column-major (materialised like upstream)
  std::vector<int> tims_map;                // 2 x P
  std::vector<uint8_t> scale_mask;          // P
  std::vector<std::vector<int>> adj_list;   // inlier graph (A.4-4)
  BitGraph G;
  int64_t n_edges = 0;
  std::vector<int> cores; int max_core = 0;
  std::vector<int> clique;                  // sorted, original ids
  int heuristic_lb = 0; int omega_exact = 0;
  double scale = 1.0, R[9] = {1,0,0,0,1,0,0,0,1}, t[3] = {0,0,0};
  int valid = 0;
  std::vector<double> rot_src, rot_dst;     // 3 x C pruned TIMs for rotation
  std::vector<int> rot_map;                 // 2 x C
  std::vector<uint8_t> rot_mask;            // C
  std::vector<uint8_t> trans_mask;          // |rotation inliers|
  std::vector<double> gnc_weights;
  int gnc_iters = 0; double gnc_cost = 0;
  double timing[8] = {0};                   // tims, scale, graph, clique, rot, trans, total
};

// ----------------------------------------------------------------------------
// A.4-1  computeTIMs : v_j - v_i for all i<j, TIM k = i*N - i(i+1)/2 + (j-i-1)
// ----------------------------------------------------------------------------
void compute_tims(const double* v, int n, std::vector<double>& tims, std::vector<int>* map) {
  const int64_t P = (int64_t)n * (n - 1) / 2;
  tims.resize((size_t)3 * P);
  if (map) map->resize((size_t)2 * P);
#pragma omp parallel for schedule(dynamic, 16)
  for (int i = 0; i < n - 1; ++i) {
    int64_t k = (int64_t)i * n - (int64_t)i * (i + 1) / 2;
    const double xi = v[3 * i], yi = v[3 * i + 1], zi = v[3 * i + 2];
    for (int j = i + 1; j < n; ++j, ++k) {
      tims[3 * k + 0] = v[3 * j + 0] - xi;
      tims[3 * k + 1] = v[3 * j + 1] - yi;
      tims[3 * k + 2] = v[3 * j + 2] - zi;
      if (map) { (*map)[2 * k] = i; (*map)[2 * k + 1] = j; }
    }
  }
}

// ----------------------------------------------------------------------------
// A.7  ScalarTLSEstimator::estimate  (adaptive voting).
// Endpoint order: upstream uses std::sort on the value only (unstable, tie order
// unspecified).  Oracle total order: (value, entry-before-exit, index).
// ----------------------------------------------------------------------------
struct Endpoint { double v; int id; };  // id = +(i+1) entry, -(i+1) exit
inline bool ep_less(const Endpoint& a, const Endpoint& b) {
  if (a.v != b.v) return a.v < b.v;
  const bool ae = a.id > 0, be = b.id > 0;
  if (ae != be) return ae;  // entries first
  return std::abs(a.id) < std::abs(b.id);
}

void tls_estimate(const double* X, const double* ranges, int64_t M, double* estimate, uint8_t* inliers) {
  std::vector<Endpoint> h((size_t)2 * M);
  for (int64_t i = 0; i < M; ++i) {
    h[2 * i] = {X[i] - ranges[i], (int)(i + 1)};
    h[2 * i + 1] = {X[i] + ranges[i], -(int)(i + 1)};
  }
  std::sort(h.begin(), h.end(), ep_less);
  double ranges_sum = 0;
  for (int64_t i = 0; i < M; ++i) ranges_sum += ranges[i];
  double dot_X_w = 0, dot_w = 0, sum_x = 0, sum_xx = 0;
  int64_t card = 0;
  double best_cost = std::numeric_limits<double>::infinity(), best_x = 0;
  bool have = false;
  for (int64_t e = 0; e < 2 * M; ++e) {
    const int64_t idx = std::abs(h[e].id) - 1;
    const int eps = h[e].id > 0 ? 1 : -1;
    const double w = 1.0 / (ranges[idx] * ranges[idx]);
    card += eps;
    dot_w += eps * w;
    dot_X_w += eps * w * X[idx];
    ranges_sum -= eps * ranges[idx];
    sum_x += eps * X[idx];
    sum_xx += eps * X[idx] * X[idx];
    const double xh = dot_X_w / dot_w;
    const double resid = card * xh * xh + sum_xx - 2 * sum_x * xh;
    const double cost = resid + ranges_sum;
    // first minimum wins; NaN (empty consensus set at the last endpoint) never wins
    if (cost < best_cost) { best_cost = cost; best_x = xh; have = true; }
  }
  if (!have) best_x = std::numeric_limits<double>::quiet_NaN();
  if (estimate) *estimate = best_x;
  if (inliers) for (int64_t i = 0; i < M; ++i) inliers[i] = std::fabs(X[i] - best_x) <= ranges[i];
}

// ----------------------------------------------------------------------------
// 3x3 SVD by one-sided (Hestenes) Jacobi.  A = U S V^T, row-major 3x3.
// Upstream uses Eigen::JacobiSVD; any convergent SVD gives the same R = V U^T
// to ~1e-15 when H is not rank-deficient.
// ----------------------------------------------------------------------------
void svd3(const double* A, double* U, double* S, double* V) {
  double a[3][3], v[3][3] = {{1,0,0},{0,1,0},{0,0,1}};
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = A[3 * i + j];
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0;
    for (int p = 0; p < 2; ++p) for (int q = p + 1; q < 3; ++q) {
      double alpha = 0, beta = 0, gamma = 0;
      for (int i = 0; i < 3; ++i) { alpha += a[i][p] * a[i][p]; beta += a[i][q] * a[i][q]; gamma += a[i][p] * a[i][q]; }
      if (gamma == 0) continue;
      off = std::max(off, std::fabs(gamma) / std::sqrt(alpha * beta + 1e-300));
      const double zeta = (beta - alpha) / (2 * gamma);
      const double tt = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1 + zeta * zeta));
      const double c = 1 / std::sqrt(1 + tt * tt), s = c * tt;
      for (int i = 0; i < 3; ++i) {
        const double x = a[i][p], y = a[i][q];
        a[i][p] = c * x - s * y; a[i][q] = s * x + c * y;
        const double vx = v[i][p], vy = v[i][q];
        v[i][p] = c * vx - s * vy; v[i][q] = s * vx + c * vy;
      }
    }
    if (off < 1e-16) break;
  }
  double sv[3]; int ord[3] = {0, 1, 2};
  for (int j = 0; j < 3; ++j) sv[j] = std::sqrt(a[0][j] * a[0][j] + a[1][j] * a[1][j] + a[2][j] * a[2][j]);
  std::sort(ord, ord + 3, [&](int x, int y) { return sv[x] > sv[y]; });
  double u[3][3];
  for (int jj = 0; jj < 3; ++jj) {
    const int j = ord[jj];
    S[jj] = sv[j];
    for (int i = 0; i < 3; ++i) { V[3 * i + jj] = v[i][j]; u[i][jj] = sv[j] > 0 ? a[i][j] / sv[j] : 0; }
  }
  // complete U for (near) rank-deficient input so that it stays orthogonal
  auto norm = [&](int j) { return std::sqrt(u[0][j] * u[0][j] + u[1][j] * u[1][j] + u[2][j] * u[2][j]); };
  const double tiny = 1e-12 * (S[0] > 0 ? 1.0 : 0.0);
  if (norm(0) < 0.5) { u[0][0] = 1; u[1][0] = 0; u[2][0] = 0; }
  if (norm(1) < 0.5 || S[1] <= tiny * S[0]) {
    // any unit vector orthogonal to u0
    int k = std::fabs(u[0][0]) < std::fabs(u[1][0]) ? (std::fabs(u[0][0]) < std::fabs(u[2][0]) ? 0 : 2)
                                                      : (std::fabs(u[1][0]) < std::fabs(u[2][0]) ? 1 : 2);
    double e[3] = {0, 0, 0}; e[k] = 1;
    const double d = e[0] * u[0][0] + e[1] * u[1][0] + e[2] * u[2][0];
    double w[3] = {e[0] - d * u[0][0], e[1] - d * u[1][0], e[2] - d * u[2][0]};
    const double nw = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    for (int i = 0; i < 3; ++i) u[i][1] = w[i] / nw;
  }
  if (norm(2) < 0.5 || S[2] <= tiny * S[0]) {
    u[0][2] = u[1][0] * u[2][1] - u[2][0] * u[1][1];
    u[1][2] = u[2][0] * u[0][1] - u[0][0] * u[2][1];
    u[2][2] = u[0][0] * u[1][1] - u[1][0] * u[0][1];
  }
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) U[3 * i + j] = u[i][j];
}

inline double det3(const double* M) {
  return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// A.6-3a  svdRot: H = sum w x y^T ; R = V diag(1,1,det) U^T
void svd_rot(const double* X, const double* Y, const double* w, int C, double* R) {
  double H[9] = {0};
  for (int i = 0; i < C; ++i)
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) H[3 * r + c] += w[i] * X[3 * i + r] * Y[3 * i + c];
  double U[9], S[3], V[9];
  svd3(H, U, S, V);
  if (det3(U) * det3(V) < 0) { V[2] = -V[2]; V[5] = -V[5]; V[8] = -V[8]; }
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) {
    double s = 0;
    for (int k = 0; k < 3; ++k) s += V[3 * r + k] * U[3 * c + k];
    R[3 * r + c] = s;
  }
}

// ----------------------------------------------------------------------------
// A.6-3  GNC-TLS rotation.
// ----------------------------------------------------------------------------
void gnc_tls(const double* X, const double* Y, int C, double noise_bound, double gnc_factor, int max_iter,
             double cost_thr, double* R, uint8_t* inliers, double* weights_out, int* iters, double* cost_out) {
  double nb2 = noise_bound * noise_bound;
  if (nb2 < 1e-16) nb2 = 1e-2;
  std::vector<double> w((size_t)C, 1.0), r2((size_t)C);
  double mu = 1, prev_cost = std::numeric_limits<double>::infinity(), cost = 0;
  int it = 0;
  for (it = 0; it < max_iter; ++it) {
    svd_rot(X, Y, w.data(), C, R);
    double maxr = 0;
    for (int i = 0; i < C; ++i) {
      const double* x = X + 3 * i; const double* y = Y + 3 * i;
      const double d0 = y[0] - ((R[0] * x[0] + R[1] * x[1]) + R[2] * x[2]);
      const double d1 = y[1] - ((R[3] * x[0] + R[4] * x[1]) + R[5] * x[2]);
      const double d2 = y[2] - ((R[6] * x[0] + R[7] * x[1]) + R[8] * x[2]);
      r2[i] = sqn3(d0, d1, d2);
      maxr = std::max(maxr, r2[i]);
    }
    if (it == 0) {
      mu = 1 / (2 * maxr / nb2 - 1);
      if (mu <= 0) break;  // near-noiseless: weights stay 1
    }
    const double th1 = (mu + 1) / mu * nb2, th2 = mu / (mu + 1) * nb2;
    cost = 0;
    for (int i = 0; i < C; ++i) {
      cost += w[i] * r2[i];
      if (r2[i] >= th1) w[i] = 0;
      else if (r2[i] <= th2) w[i] = 1;
      else w[i] = std::sqrt(nb2 * mu * (mu + 1) / r2[i]) - mu;
    }
    const double cost_diff = std::fabs(cost - prev_cost);
    mu *= gnc_factor;
    prev_cost = cost;
    if (cost_diff < cost_thr) { ++it; break; }
  }
  for (int i = 0; i < C; ++i) inliers[i] = w[i] >= 0.5;
  if (weights_out) std::copy(w.begin(), w.end(), weights_out);
  if (iters) *iters = it;
  if (cost_out) *cost_out = cost;
}

// ----------------------------------------------------------------------------
// A.5  Max clique, PMC-style: k-core -> heuristic -> exact bitset B&B (OpenMP).
// ----------------------------------------------------------------------------
// Batagelj-Zaversnik core decomposition.
void core_numbers(const std::vector<std::vector<int>>& adj, std::vector<int>& core, int& max_core) {
  const int n = (int)adj.size();
  std::vector<int> deg(n), pos(n), vert(n);
  int md = 0;
  for (int v = 0; v < n; ++v) { deg[v] = (int)adj[v].size(); md = std::max(md, deg[v]); }
  std::vector<int> bin(md + 2, 0);
  for (int v = 0; v < n; ++v) bin[deg[v]]++;
  int start = 0;
  for (int d = 0; d <= md; ++d) { int c = bin[d]; bin[d] = start; start += c; }
  for (int v = 0; v < n; ++v) { pos[v] = bin[deg[v]]; vert[pos[v]] = v; bin[deg[v]]++; }
  for (int d = md; d > 0; --d) bin[d] = bin[d - 1];
  bin[0] = 0;
  for (int i = 0; i < n; ++i) {
    const int v = vert[i];
    for (int u : adj[v]) if (deg[u] > deg[v]) {
      const int du = deg[u], pu = pos[u], pw = bin[du], w = vert[pw];
      if (u != w) { pos[u] = pw; vert[pu] = w; pos[w] = pu; vert[pw] = u; }
      bin[du]++; deg[u]--;
    }
  }
  core = deg; max_core = 0;
  for (int v = 0; v < n; ++v) max_core = std::max(max_core, core[v]);
}

// Shared deterministic heuristic (documented deviation from PMC's own heuristic, whose result depends on thread
// timing upstream).  From a seed v, P = N(v); every round computes degP(w) = |N(w) & P| for all members:
//   * members adjacent to every other member (degP == |P|-1) are all taken at once;
//   * otherwise the member with the largest degP (ties: lowest id) is taken and P &= N(u).
// Winner over the seeds: largest clique, ties -> lowest seed id.
void greedy_from_seed(const BitGraph& G, const std::vector<int>& deg, int seed, std::vector<int>& out, int stop_below) {
  (void)deg;
  const int W = G.W;
  std::vector<u64> P(G.row(seed), G.row(seed) + W), U((size_t)W);
  out.clear(); out.push_back(seed);
  int cnt = popc(P.data(), W);
  while (cnt > 0) {
    if ((int)out.size() + cnt < stop_below) { out.clear(); return; }  // cannot even tie
    std::fill(U.begin(), U.end(), 0);
    int best = -1, bd = -1, nU = 0;
    for (int w = 0; w < W; ++w) {
      u64 x = P[w];
      while (x) {
        const int u = w * 64 + __builtin_ctzll(x); x &= x - 1;
        const u64* r = G.row(u);
        int c = 0;
        for (int y = 0; y < W; ++y) c += __builtin_popcountll(r[y] & P[y]);
        if (c > bd) { bd = c; best = u; }
        if (c == cnt - 1) { U[w] |= (u64)1 << (u & 63); ++nU; }
      }
    }
    if (nU > 0) {
      for (int w = 0; w < W; ++w) {
        u64 x = U[w];
        while (x) { out.push_back(w * 64 + __builtin_ctzll(x)); x &= x - 1; }
        P[w] &= ~U[w];
      }
    } else {
      out.push_back(best);
      const u64* r = G.row(best);
      for (int w = 0; w < W; ++w) P[w] &= r[w];
    }
    cnt = popc(P.data(), W);
  }
}

int heuristic_clique(const BitGraph& G, const std::vector<int>& deg, std::vector<int>& best_clique) {
  const int n = G.n;
  int lb = 0; int best_seed = -1;
  std::vector<std::vector<int>> found(n);
  std::atomic<int> lb_a{0};
#pragma omp parallel
  {
    std::vector<int> c;
#pragma omp for schedule(dynamic, 8)
    for (int v = 0; v < n; ++v) {
      const int cur = lb_a.load(std::memory_order_relaxed);
      if (deg[v] + 1 < cur) continue;
      greedy_from_seed(G, deg, v, c, cur);
      if (c.empty()) continue;
      int s = (int)c.size(), old = lb_a.load();
      while (s > old && !lb_a.compare_exchange_weak(old, s)) {}
      found[v] = c;
    }
  }
  for (int v = 0; v < n; ++v) if ((int)found[v].size() > lb) { lb = (int)found[v].size(); best_seed = v; }
  best_clique.clear();
  if (best_seed >= 0) { best_clique = found[best_seed]; std::sort(best_clique.begin(), best_clique.end()); }
  return lb;
}

// Greedy sequential colouring of the set P (BBMC style): returns candidates in
// non-decreasing colour order with their colour numbers.
struct Searcher {
  const BitGraph& G;
  int W;
  std::atomic<int>* best;      // shared incumbent size
  bool decision; int need;     // decision mode: stop at the first clique of size >= need
  bool found = false;
  std::vector<int> cur, best_clique;
  std::vector<std::vector<u64>> Pst;   // per-depth scratch
  std::vector<std::vector<int>> ord, col;
  explicit Searcher(const BitGraph& g, std::atomic<int>* b) : G(g), W(g.W), best(b), decision(false), need(0) {
    // depth never exceeds n; size the per-depth scratch once so references into it stay valid while recursing
    Pst.resize(g.n + 2); ord.resize(g.n + 2); col.resize(g.n + 2);
  }

  void colour(const u64* P, std::vector<int>& order, std::vector<int>& colour_of, std::vector<u64>& Q, std::vector<u64>& S) {
    order.clear(); colour_of.clear();
    Q.assign(P, P + W);
    int k = 0;
    bool any = true;
    while (any) {
      any = false; ++k;
      S = Q;
      for (int w = 0; w < W; ++w) {
        while (S[w]) {
          const int v = w * 64 + __builtin_ctzll(S[w]);
          any = true;
          Q[w] &= ~((u64)1 << (v & 63));
          const u64* r = G.row(v);
          for (int x = w; x < W; ++x) S[x] &= ~r[x];
          S[w] &= ~((u64)1 << (v & 63));
          order.push_back(v); colour_of.push_back(k);
        }
      }
    }
  }

  void expand(std::vector<u64>& P, int depth) {
    std::vector<u64> Q, S;
    std::vector<int>& order = ord[depth]; std::vector<int>& colr = col[depth];
    colour(P.data(), order, colr, Q, S);
    for (int i = (int)order.size() - 1; i >= 0; --i) {
      if (found) return;
      const int bound = (int)cur.size() + colr[i];
      if (decision) { if (bound < need) return; }
      else if (bound <= best->load(std::memory_order_relaxed)) return;
      const int v = order[i];
      cur.push_back(v);
      std::vector<u64>& NP = Pst[depth + 1];
      NP.resize(W);
      const u64* r = G.row(v);
      bool any = false;
      for (int w = 0; w < W; ++w) { NP[w] = P[w] & r[w]; any |= NP[w] != 0; }
      if (decision && (int)cur.size() >= need) { found = true; best_clique = cur; cur.pop_back(); return; }
      if (any) expand(NP, depth + 1);
      else if (!decision) {
        int s = (int)cur.size(), old = best->load();
        while (s > old && !best->compare_exchange_weak(old, s)) {}
        if (s > old) best_clique = cur;
      }
      cur.pop_back();
      P[v >> 6] &= ~((u64)1 << (v & 63));
    }
  }
};

// Does `P` (bitset over G's vertices) contain a clique of size >= need ?
bool exists_clique(const BitGraph& G, const std::vector<u64>& P, int need) {
  if (need <= 0) return true;
  if (popc(P.data(), G.W) < need) return false;
  std::atomic<int> dummy{0};
  Searcher s(G, &dummy);
  s.decision = true; s.need = need;
  std::vector<u64> Pc = P;
  s.expand(Pc, 0);
  return s.found;
}

// Peel P down to its (k)-core inside G: repeatedly drop vertices with < k neighbours in P.
void peel(const BitGraph& G, std::vector<u64>& P, int k) {
  bool changed = true;
  const int W = G.W;
  while (changed) {
    changed = false;
    for (int w = 0; w < W; ++w) {
      u64 x = P[w];
      while (x) {
        const int v = w * 64 + __builtin_ctzll(x); x &= x - 1;
        const u64* r = G.row(v);
        int c = 0;
        for (int y = 0; y < W; ++y) c += __builtin_popcountll(r[y] & P[y]);
        if (c < k) { P[w] &= ~((u64)1 << (v & 63)); changed = true; }
      }
    }
  }
}

bool is_clique(const BitGraph& G, const std::vector<u64>& P) {
  const int W = G.W; const int cnt = popc(P.data(), W);
  for (int w = 0; w < W; ++w) {
    u64 x = P[w];
    while (x) {
      const int v = w * 64 + __builtin_ctzll(x); x &= x - 1;
      const u64* r = G.row(v);
      int c = 0;
      for (int y = 0; y < W; ++y) c += __builtin_popcountll(r[y] & P[y]);
      if (c != cnt - 1) return false;
    }
  }
  return true;
}

// Canonical maximum clique: among all cliques of size omega, the one whose
// sorted vertex list is lexicographically smallest (hard part 3).
void canonical_clique(const BitGraph& G, int omega, std::vector<int>& out) {
  const int W = G.W, n = G.n;
  out.clear();
  std::vector<u64> P((size_t)W, 0);
  for (int v = 0; v < n; ++v) P[v >> 6] |= (u64)1 << (v & 63);
  int need = omega;
  while (need > 0) {
    peel(G, P, need - 1);
    if (popc(P.data(), W) == need && is_clique(G, P)) {
      for (int w = 0; w < W; ++w) { u64 x = P[w]; while (x) { out.push_back(w * 64 + __builtin_ctzll(x)); x &= x - 1; } }
      return;
    }
    bool picked = false;
    for (int w = 0; w < W && !picked; ++w) {
      u64 x = P[w];
      while (x && !picked) {
        const int v = w * 64 + __builtin_ctzll(x); x &= x - 1;
        std::vector<u64> Q((size_t)W, 0);
        const u64* r = G.row(v);
        for (int y = (v >> 6); y < W; ++y) Q[y] = P[y] & r[y];
        Q[v >> 6] &= ~(((u64)2 << (v & 63)) - 1);  // keep ids > v
        if (exists_clique(G, Q, need - 1)) { out.push_back(v); P = Q; --need; picked = true; }
      }
    }
    if (!picked) { out.clear(); return; }  // omega was wrong (never happens when omega is exact)
  }
}

// Exact maximum clique size (and a witness) via parallel B&B over roots.
int exact_max_clique(const BitGraph& G, const std::vector<int>& core, int lb, int nthreads) {
  const int n = G.n, W = G.W;
  // root order: ascending core number (degeneracy-like), ties by id
  std::vector<int> order(n), rank(n);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return core[a] < core[b]; });
  for (int i = 0; i < n; ++i) rank[order[i]] = i;
  std::atomic<int> best{lb};
#pragma omp parallel num_threads(nthreads)
  {
    Searcher s(G, &best);
#pragma omp for schedule(dynamic, 1)
    for (int i = 0; i < n; ++i) {
      const int v = order[i];
      if (core[v] + 1 <= best.load(std::memory_order_relaxed)) continue;  // k-core pruning
      std::vector<u64> P((size_t)W, 0);
      const u64* r = G.row(v);
      int cnt = 0;
      for (int w = 0; w < W; ++w) {
        u64 x = r[w];
        while (x) {
          const int u = w * 64 + __builtin_ctzll(x); x &= x - 1;
          if (rank[u] > i && core[u] + 1 > best.load(std::memory_order_relaxed)) { P[w] |= (u64)1 << (u & 63); ++cnt; }
        }
      }
      if (cnt + 1 <= best.load(std::memory_order_relaxed)) continue;
      s.cur.clear(); s.cur.push_back(v);
      s.expand(P, 0);
    }
  }
  return best.load();
}

void find_max_clique(State& st) {
  const Params& p = st.p;
  const int n = st.n;
  st.clique.clear();
  if (p.inlier_selection_mode == 3) {  // NONE (A.5-7)
    st.clique.resize(n); std::iota(st.clique.begin(), st.clique.end(), 0); return;
  }
  core_numbers(st.adj_list, st.cores, st.max_core);
  std::vector<int> deg(n);
  for (int v = 0; v < n; ++v) deg[v] = (int)st.adj_list[v].size();
  if (p.inlier_selection_mode == 2 && p.kcore_heuristic_threshold != 1 &&
      st.max_core > (int)(p.kcore_heuristic_threshold * (double)n)) {  // A.5-3
    for (int v = 0; v < n; ++v) if (st.cores[v] >= st.max_core) st.clique.push_back(v);
    return;
  }
  std::vector<int> hc;
  st.heuristic_lb = heuristic_clique(st.G, deg, hc);
  if (p.inlier_selection_mode != 0) { st.clique = hc; return; }  // PMC_HEU / KCORE_HEU fall-through
  int nt = p.max_clique_num_threads > 0 ? p.max_clique_num_threads : omp_get_max_threads();
  int omega = st.heuristic_lb;
  if (st.heuristic_lb < st.max_core + 1) omega = exact_max_clique(st.G, st.cores, st.heuristic_lb, nt);  // A.5-4/5
  st.omega_exact = omega;
  canonical_clique(st.G, omega, st.clique);
}

// ----------------------------------------------------------------------------
// A.8  solve()
// ----------------------------------------------------------------------------
void build_graph_from_mask(State& st) {
  const int n = st.n;
  st.adj_list.assign(n, {});
  st.G.init(n);
  st.n_edges = 0;
  if (st.p.materialize_tims) {
    for (int64_t k = 0; k < st.P; ++k) if (st.scale_mask[k]) {  // serial like upstream's addEdge loop
      const int i = st.tims_map[2 * k], j = st.tims_map[2 * k + 1];
      st.adj_list[i].push_back(j); st.adj_list[j].push_back(i);
      st.G.set(i, j); st.G.set(j, i); ++st.n_edges;
    }
  }
}

void solve(State& st, const double* src, const double* dst) {
  const Params& p = st.p;
  const int n = st.n;
  const double t0 = now_s();
  st.P = (int64_t)n * (n - 1) / 2;
  const double beta = 2 * p.noise_bound * std::sqrt(p.cbar2);  // A.4-2
  st.scale = 1.0; st.valid = 0;
  double t1, t2, t3;
  if (p.materialize_tims) {
    compute_tims(src, n, st.src_tims, &st.tims_map);
    compute_tims(dst, n, st.dst_tims, nullptr);
    t1 = now_s();
    st.scale_mask.assign((size_t)st.P, 0);
    if (!p.estimate_scaling) {  // ScaleInliersSelector (A.4-2)
#pragma omp parallel for schedule(static)
      for (int64_t k = 0; k < st.P; ++k) {
        const double* a = &st.src_tims[3 * k]; const double* b = &st.dst_tims[3 * k];
        const double va = std::sqrt(sqn3(a[0], a[1], a[2])), vb = std::sqrt(sqn3(b[0], b[1], b[2]));
        st.scale_mask[k] = std::fabs(va - vb) <= beta;
      }
    } else {  // TLSScaleSolver (A.4-5)
      std::vector<double> raw((size_t)st.P), alphas((size_t)st.P);
#pragma omp parallel for schedule(static)
      for (int64_t k = 0; k < st.P; ++k) {
        const double* a = &st.src_tims[3 * k]; const double* b = &st.dst_tims[3 * k];
        const double va = std::sqrt(sqn3(a[0], a[1], a[2])), vb = std::sqrt(sqn3(b[0], b[1], b[2]));
        raw[k] = vb / va;
        alphas[k] = beta * (1.0 / va);
      }
      tls_estimate(raw.data(), alphas.data(), st.P, &st.scale, st.scale_mask.data());
    }
    t2 = now_s();
    build_graph_from_mask(st);
  } else {
    // streaming variant for large N (same verdicts, no O(N^2) doubles)
    t1 = now_s();
    st.adj_list.assign(n, {}); st.G.init(n);
    st.src_tims.clear(); st.dst_tims.clear(); st.tims_map.clear(); st.scale_mask.clear();
#pragma omp parallel for schedule(dynamic, 8)
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) {
        if (i == j) continue;
        const int lo = std::min(i, j), hi = std::max(i, j);
        const double va = std::sqrt(sqn3(src[3 * hi] - src[3 * lo], src[3 * hi + 1] - src[3 * lo + 1], src[3 * hi + 2] - src[3 * lo + 2]));
        const double vb = std::sqrt(sqn3(dst[3 * hi] - dst[3 * lo], dst[3 * hi + 1] - dst[3 * lo + 1], dst[3 * hi + 2] - dst[3 * lo + 2]));
        if (std::fabs(va - vb) <= beta) { st.G.set(i, j); st.adj_list[i].push_back(j); }
      }
    }
    t2 = now_s();
    int64_t e = 0; for (int i = 0; i < n; ++i) e += (int64_t)st.adj_list[i].size();
    st.n_edges = e / 2;
  }
  t3 = now_s();
  st.timing[0] = t1 - t0; st.timing[1] = t2 - t1; st.timing[2] = t3 - t2;

  find_max_clique(st);
  std::sort(st.clique.begin(), st.clique.end());
  const double t4 = now_s();
  st.timing[3] = t4 - t3;
  const int C = (int)st.clique.size();
  st.rot_mask.clear(); st.trans_mask.clear(); st.rot_src.clear(); st.rot_dst.clear(); st.rot_map.clear();
  if (C <= 1) { st.valid = 0; st.timing[6] = now_s() - t0; return; }  // A.5-6

  // A.6-1 TIMs over the clique
  if (p.rotation_tim_graph == 0) {  // CHAIN
    st.rot_src.resize((size_t)3 * C); st.rot_dst.resize((size_t)3 * C); st.rot_map.resize((size_t)2 * C);
    for (int i = 0; i < C; ++i) {
      const int root = st.clique[i], leaf = st.clique[(i + 1) % C];
      for (int d = 0; d < 3; ++d) {
        st.rot_src[3 * i + d] = src[3 * leaf + d] - src[3 * root + d];
        st.rot_dst[3 * i + d] = dst[3 * leaf + d] - dst[3 * root + d];
      }
      st.rot_map[2 * i] = root; st.rot_map[2 * i + 1] = leaf;
    }
  } else {  // COMPLETE
    const int64_t Pc = (int64_t)C * (C - 1) / 2;
    st.rot_src.resize((size_t)3 * Pc); st.rot_dst.resize((size_t)3 * Pc); st.rot_map.resize((size_t)2 * Pc);
    int64_t k = 0;
    for (int i = 0; i < C - 1; ++i) for (int j = i + 1; j < C; ++j, ++k) {
      const int a = st.clique[i], b = st.clique[j];
      for (int d = 0; d < 3; ++d) {
        st.rot_src[3 * k + d] = src[3 * b + d] - src[3 * a + d];
        st.rot_dst[3 * k + d] = dst[3 * b + d] - dst[3 * a + d];
      }
      st.rot_map[2 * k] = a; st.rot_map[2 * k + 1] = b;
    }
  }
  const int T = (int)(st.rot_src.size() / 3);
  // A.6-2  dst_tims *= 1/scale ; rotation noise bound = nb * 2/scale
  const double inv_s = 1.0 / st.scale;
  for (auto& v : st.rot_dst) v *= inv_s;
  const double rot_nb_base = p.rotation_noise_bound > 0 ? p.rotation_noise_bound : p.noise_bound;
  const double rot_nb = rot_nb_base * (2 / st.scale);
  st.rot_mask.assign((size_t)T, 0); st.gnc_weights.assign((size_t)T, 1.0);
  gnc_tls(st.rot_src.data(), st.rot_dst.data(), T, rot_nb, p.rotation_gnc_factor, p.rotation_max_iterations,
          p.rotation_cost_threshold, st.R, st.rot_mask.data(), st.gnc_weights.data(), &st.gnc_iters, &st.gnc_cost);
  const double t5 = now_s();
  st.timing[4] = t5 - t4;

  // A.8  gather rotation-inlier *points* by max_clique[k], then A.7-4 translation
  std::vector<int> rin;
  for (int k = 0; k < T; ++k) if (st.rot_mask[k]) rin.push_back(k);
  const int M = (int)rin.size();
  st.trans_mask.assign((size_t)M, 1);
  st.t[0] = st.t[1] = st.t[2] = 0;
  if (M >= 2) {
    std::vector<double> raw((size_t)3 * M), rng((size_t)M, p.noise_bound * std::sqrt(p.cbar2));
    for (int i = 0; i < M; ++i) {
      // upstream indexes max_clique_ with the rotation-inlier position (valid for CHAIN: T == C)
      const int pt = st.clique[rin[i] % C];
      const double* x = src + 3 * pt; const double* y = dst + 3 * pt;
      for (int r = 0; r < 3; ++r)
        raw[(size_t)r * M + i] = y[r] - st.scale * ((st.R[3 * r] * x[0] + st.R[3 * r + 1] * x[1]) + st.R[3 * r + 2] * x[2]);
    }
    std::vector<uint8_t> tmp((size_t)M);
    for (int r = 0; r < 3; ++r) {
      tls_estimate(&raw[(size_t)r * M], rng.data(), M, &st.t[r], tmp.data());
      for (int i = 0; i < M; ++i) st.trans_mask[i] = st.trans_mask[i] && tmp[i];
    }
  } else if (M == 1) {
    const int pt = st.clique[rin[0] % C];
    for (int r = 0; r < 3; ++r)
      st.t[r] = dst[3 * pt + r] - st.scale * ((st.R[3 * r] * src[3 * pt] + st.R[3 * r + 1] * src[3 * pt + 1]) + st.R[3 * r + 2] * src[3 * pt + 2]);
  }
  st.valid = 1;
  const double t6 = now_s();
  st.timing[5] = t6 - t5; st.timing[6] = t6 - t0;
}

}  // namespace

// =============================================================================
// C interface (ctypes)
// =============================================================================
extern "C" {

typedef Params orc_params;

void* orc_solve(const double* src, const double* dst, int n, const orc_params* p) {
  State* st = new State();
  st->p = *p; st->n = n;
  solve(*st, src, dst);
  return st;
}
void orc_free(void* h) { delete (State*)h; }

// sizes: [n, P, n_edges, clique, n_rot_tims, n_rot_inliers, n_trans, gnc_iters, heuristic_lb, max_core, omega_exact]
void orc_sizes(void* h, int64_t* out) {
  State* st = (State*)h;
  int64_t nrin = 0; for (auto m : st->rot_mask) nrin += m;
  out[0] = st->n; out[1] = st->P; out[2] = st->n_edges; out[3] = (int64_t)st->clique.size();
  out[4] = (int64_t)st->rot_mask.size(); out[5] = nrin; out[6] = (int64_t)st->trans_mask.size();
  out[7] = st->gnc_iters; out[8] = st->heuristic_lb; out[9] = st->max_core; out[10] = st->omega_exact;
}
void orc_solution(void* h, double* scale, double* R9, double* t3, int* valid, double* gnc_cost) {
  State* st = (State*)h;
  *scale = st->scale; std::memcpy(R9, st->R, sizeof(double) * 9); std::memcpy(t3, st->t, sizeof(double) * 3);
  *valid = st->valid; if (gnc_cost) *gnc_cost = st->gnc_cost;
}
void orc_timing(void* h, double* out7) { std::memcpy(out7, ((State*)h)->timing, sizeof(double) * 7); }
// adjacency as u32 [n, W32] words, bit j of row i = edge(i,j)
void orc_adjacency(void* h, uint32_t* out, int W32) {
  State* st = (State*)h;
  const int n = st->n;
  std::memset(out, 0, sizeof(uint32_t) * (size_t)n * W32);
  for (int i = 0; i < n; ++i) {
    const u64* r = st->G.row(i);
    for (int w = 0; w < st->G.W; ++w) {
      if (2 * w < W32) out[(size_t)i * W32 + 2 * w] = (uint32_t)(r[w] & 0xffffffffu);
      if (2 * w + 1 < W32) out[(size_t)i * W32 + 2 * w + 1] = (uint32_t)(r[w] >> 32);
    }
  }
}
void orc_clique(void* h, int* out) { State* st = (State*)h; std::copy(st->clique.begin(), st->clique.end(), out); }
void orc_rot_mask(void* h, uint8_t* out) { State* st = (State*)h; std::copy(st->rot_mask.begin(), st->rot_mask.end(), out); }
void orc_trans_mask(void* h, uint8_t* out) { State* st = (State*)h; std::copy(st->trans_mask.begin(), st->trans_mask.end(), out); }
void orc_scale_mask(void* h, uint8_t* out) { State* st = (State*)h; std::copy(st->scale_mask.begin(), st->scale_mask.end(), out); }
void orc_gnc_weights(void* h, double* out) { State* st = (State*)h; std::copy(st->gnc_weights.begin(), st->gnc_weights.end(), out); }
void orc_rot_tims(void* h, double* src3xT, double* dst3xT, int* map2xT) {
  State* st = (State*)h;
  std::copy(st->rot_src.begin(), st->rot_src.end(), src3xT);
  std::copy(st->rot_dst.begin(), st->rot_dst.end(), dst3xT);
  std::copy(st->rot_map.begin(), st->rot_map.end(), map2xT);
}
void orc_cores(void* h, int* out) { State* st = (State*)h; std::copy(st->cores.begin(), st->cores.end(), out); }

// ---- component entry points (unit checks) ---------------------------------
void orc_svd3(const double* A, double* U, double* S, double* V) { svd3(A, U, S, V); }
void orc_tls_estimate(const double* X, const double* ranges, int64_t M, double* est, uint8_t* inl) { tls_estimate(X, ranges, M, est, inl); }
void orc_gnc_tls(const double* X, const double* Y, int C, double nb, double factor, int max_iter, double thr,
                 double* R, uint8_t* inl, double* w, int* iters, double* cost) {
  gnc_tls(X, Y, C, nb, factor, max_iter, thr, R, inl, w, iters, cost);
}
// Max clique of an explicit graph given as u32 bit rows; mode as inlier_selection_mode. Returns size.
int orc_max_clique_bits(const uint32_t* adj, int n, int W32, int mode, double kcore_thr, int* out) {
  State st; st.n = n; st.p = Params{}; st.p.inlier_selection_mode = mode; st.p.kcore_heuristic_threshold = kcore_thr;
  st.G.init(n); st.adj_list.assign(n, {});
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j)
    if ((adj[(size_t)i * W32 + (j >> 5)] >> (j & 31)) & 1) { st.G.set(i, j); st.adj_list[i].push_back(j); }
  find_max_clique(st);
  std::sort(st.clique.begin(), st.clique.end());
  std::copy(st.clique.begin(), st.clique.end(), out);
  return (int)st.clique.size();
}
// Edge test only (streaming), u32 rows.  variant 0: (x2+y2)+z2 ; variant 1: x2+(y2+z2) (borderline counter, section 7-2)
void orc_tim_graph(const double* src, const double* dst, int n, double noise_bound, double cbar2, uint32_t* out, int W32, int variant) {
  const double beta = 2 * noise_bound * std::sqrt(cbar2);
  std::memset(out, 0, sizeof(uint32_t) * (size_t)n * W32);
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) {
    if (i == j) continue;
    const int lo = std::min(i, j), hi = std::max(i, j);
    const double ax = src[3 * hi] - src[3 * lo], ay = src[3 * hi + 1] - src[3 * lo + 1], az = src[3 * hi + 2] - src[3 * lo + 2];
    const double bx = dst[3 * hi] - dst[3 * lo], by = dst[3 * hi + 1] - dst[3 * lo + 1], bz = dst[3 * hi + 2] - dst[3 * lo + 2];
    double A, B;
    if (variant == 0) { A = sqn3(ax, ay, az); B = sqn3(bx, by, bz); }
    else { A = ax * ax + (ay * ay + az * az); B = bx * bx + (by * by + bz * bz); }
    if (std::fabs(std::sqrt(A) - std::sqrt(B)) <= beta) out[(size_t)i * W32 + (j >> 5)] |= 1u << (j & 31);
  }
}
int orc_num_threads() { return omp_get_max_threads(); }
// torchrun exports OMP_NUM_THREADS=1 to every rank: the timed CPU arm sets its own thread count (bench.py)
void orc_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

}  // extern "C"
