// CAD depth render on the GPU: replaces what the reference gets from Open3D's Filament
// OffscreenRenderer (no graphics stack here: "unavailable on this box") --
//   PoseEstimator.create_template_from_H     PoseEstimator.py:118-141     render_to_depth_image(z_in_view_space=True) at the pose
//   render_templates                         template_creation.py:188-233 the same from five fixed views
// followed by RGBDImage.create_from_color_and_depth(depth_scale=1.0, depth_trunc=far) + create_from_rgbd_image on the FLOAT depth.
// (paths under /root/reference/src/pose_estimator/EstimHelpers/).
//
// Per pixel the image holds the view-space z of the nearest surface point on the ray ((u-cx)/fx, (v-cy)/fy, 1): one CTA per
// triangle walks the triangle's screen bounding box, every pixel runs the oracle's fp64 Moeller-Trumbore expression (this file
// is compiled with -fmad=false) and the nearest hit wins through an atomicMin on the float bits (positive floats order like
// their bit patterns), so the image equals the oracle's brute force bit for bit.
#include <math_constants.h>

#include "scan.cuh"
#include "solver.cuh"

namespace b2r {
namespace {

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 vcross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ double vdot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
constexpr double kEdgeEps = 1e-9;

struct Cam { double T[12]; double fx, fy, cx, cy, z_near, z_far; int W, H; };

// model -> camera, in the oracle's operation order
__global__ void __launch_bounds__(256) raster_xform_kernel(const double* __restrict__ verts, int nv, Cam c, double* __restrict__ vc) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= nv) return;
  const double x = verts[3 * (size_t)i], y = verts[3 * (size_t)i + 1], z = verts[3 * (size_t)i + 2];
  vc[3 * (size_t)i] = ((c.T[0] * x + c.T[1] * y) + c.T[2] * z) + c.T[3];
  vc[3 * (size_t)i + 1] = ((c.T[4] * x + c.T[5] * y) + c.T[6] * z) + c.T[7];
  vc[3 * (size_t)i + 2] = ((c.T[8] * x + c.T[9] * y) + c.T[10] * z) + c.T[11];
}

__global__ void __launch_bounds__(128) raster_tri_kernel(const double* __restrict__ vc, const int* __restrict__ tris, int nt, Cam c,
                                                         unsigned int* __restrict__ zbuf) {
  const int k = blockIdx.x;
  if (k >= nt) return;
  const int ia = tris[3 * (size_t)k], ib = tris[3 * (size_t)k + 1], ic = tris[3 * (size_t)k + 2];
  const V3 a{vc[3 * (size_t)ia], vc[3 * (size_t)ia + 1], vc[3 * (size_t)ia + 2]};
  const V3 b{vc[3 * (size_t)ib], vc[3 * (size_t)ib + 1], vc[3 * (size_t)ib + 2]};
  const V3 cc{vc[3 * (size_t)ic], vc[3 * (size_t)ic + 1], vc[3 * (size_t)ic + 2]};
  // screen bounding box (one pixel of slack); a vertex at or behind the camera plane -> the whole image
  int u0 = 0, v0 = 0, u1 = c.W - 1, v1 = c.H - 1;
  const double zmin = fmin(a.z, fmin(b.z, cc.z));
  if (zmin > 1e-6) {
    const double ua = c.fx * a.x / a.z + c.cx, ub = c.fx * b.x / b.z + c.cx, uc = c.fx * cc.x / cc.z + c.cx;
    const double va = c.fy * a.y / a.z + c.cy, vb = c.fy * b.y / b.z + c.cy, vcc = c.fy * cc.y / cc.z + c.cy;
    const double lo_u = floor(fmin(ua, fmin(ub, uc))) - 1, hi_u = ceil(fmax(ua, fmax(ub, uc))) + 1;
    const double lo_v = floor(fmin(va, fmin(vb, vcc))) - 1, hi_v = ceil(fmax(va, fmax(vb, vcc))) + 1;
    if (!(hi_u >= 0) || !(hi_v >= 0) || !(lo_u <= c.W - 1) || !(lo_v <= c.H - 1)) return;
    u0 = (int)fmax(lo_u, 0.0); v0 = (int)fmax(lo_v, 0.0);
    u1 = (int)fmin(hi_u, (double)(c.W - 1)); v1 = (int)fmin(hi_v, (double)(c.H - 1));
  } else if (fmax(a.z, fmax(b.z, cc.z)) <= c.z_near) {
    return;  // entirely in front of the near plane
  }
  const V3 e1 = vsub(b, a), e2 = vsub(cc, a);
  const V3 tv{-a.x, -a.y, -a.z};
  const V3 q = vcross(tv, e1);
  const double e2q = vdot(e2, q);
  const int bw = u1 - u0 + 1;
  const long long npx = (long long)bw * (v1 - v0 + 1);
  for (long long p = threadIdx.x; p < npx; p += 128) {
    const int u = u0 + (int)(p % bw), v = v0 + (int)(p / bw);
    const V3 d{((double)u - c.cx) / c.fx, ((double)v - c.cy) / c.fy, 1.0};
    const V3 pv = vcross(d, e2);
    const double det = vdot(e1, pv);
    if (!(fabs(det) > 1e-300)) continue;
    const double inv = 1.0 / det;
    const double bu = vdot(tv, pv) * inv;
    const double bv = vdot(d, q) * inv;
    const double t = e2q * inv;
    if (bu >= -kEdgeEps && bv >= -kEdgeEps && bu + bv <= 1.0 + kEdgeEps && t > c.z_near && t < c.z_far)
      atomicMin(&zbuf[(size_t)v * c.W + u], __float_as_uint((float)t));
  }
}

__global__ void __launch_bounds__(256) raster_clear_kernel(unsigned int* __restrict__ zbuf, int n) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) zbuf[i] = 0x7f800000u;
}

// +inf (0x7f800000, the cleared value) -> 0
__global__ void __launch_bounds__(256) raster_resolve_kernel(unsigned int* __restrict__ zbuf, int n) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n && zbuf[i] >= 0x7f800000u) zbuf[i] = 0u;
}

// float depth image -> ordered cloud (create_from_color_and_depth(depth_scale, depth_trunc) + create_from_rgbd_image)
__global__ void __launch_bounds__(256) backproject32_kernel(const float* __restrict__ depth, int H, int W, double fx, double fy, double cx, double cy,
                                                            float scale, float trunc, int use_trunc, double* __restrict__ pts, int* __restrict__ pix,
                                                            int cap, unsigned long long* ws, int* __restrict__ count) {
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  const int tile = take_ticket(ws, &s_tile);
  const long long i = (long long)tile * 256 + threadIdx.x;
  float d = 0.f;
  if (i < (long long)H * W) {
    d = __fdiv_rn(depth[i], scale);
    if (use_trunc && d >= trunc) d = 0.f;
  }
  const unsigned k = d > 0.f ? 1u : 0u;
  unsigned blk_total;
  const unsigned local = block_scan_256(k, s_warp, &blk_total);
  if (threadIdx.x == 0) {
    s_base = lookback(ws, tile, blk_total);
    if (tile == (int)gridDim.x - 1) *count = (int)(s_base + blk_total);
  }
  __syncthreads();
  if (k) {
    const long long m = (long long)s_base + local;
    if (m < cap) {
      const int u = (int)(i % W), v = (int)(i / W);
      const double z = (double)d;
      pts[3 * m] = __ddiv_rn(__dmul_rn((double)u - cx, z), fx);
      pts[3 * m + 1] = __ddiv_rn(__dmul_rn((double)v - cy, z), fy);
      pts[3 * m + 2] = z;
      if (pix) pix[m] = (int)i;
    }
  }
}

// LiDAR cloud -> image: the selection step of main_seibersdorf.project_and_colorize (main_seibersdorf.py:70-81,149-162):
//   Xc = R p + t;  front = z > z_min;  uv = cv2.projectPoints(float32(Xc), 0, 0, K, D).astype(int32);  inside the image;  mask[v, u]
// One thread per point, ordered compaction of the survivors (ascending point index, like the reference's index arrays).
struct ProjArgs {
  double R[9], t[3], fx, fy, cx, cy, k[8], z_min;
  int nk, W, H;
};
__global__ void __launch_bounds__(256) project_select_kernel(const double* __restrict__ pts, int n, ProjArgs a, const unsigned char* __restrict__ mask,
                                                             int* __restrict__ idx_out, int* __restrict__ uv_out, unsigned long long* ws,
                                                             int* __restrict__ counts /* [0] selected [1] front [2] in image */) {
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  const int tile = take_ticket(ws, &s_tile);
  const int i = tile * 256 + threadIdx.x;
  bool front = false, inimg = false, sel = false;
  int u = 0, v = 0;
  if (i < n) {
    const double x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    const double X = ((a.R[0] * x + a.R[1] * y) + a.R[2] * z) + a.t[0];
    const double Y = ((a.R[3] * x + a.R[4] * y) + a.R[5] * z) + a.t[1];
    const double Z = ((a.R[6] * x + a.R[7] * y) + a.R[8] * z) + a.t[2];
    front = Z > a.z_min;
    if (front) {
      // cv2.projectPoints on float32 input, identity pose: double arithmetic inside, float32 result, then astype(int32)
      const double Xf = (double)(float)X, Yf = (double)(float)Y, Zf = (double)(float)Z;
      const double iz = Zf != 0.0 ? 1.0 / Zf : 1.0;
      const double xn = Xf * iz, yn = Yf * iz;
      double xd = xn, yd = yn;
      if (a.nk > 0) {
        const double r2 = xn * xn + yn * yn, r4 = r2 * r2, r6 = r4 * r2;
        const double a1 = 2 * xn * yn, a2 = r2 + 2 * xn * xn, a3 = r2 + 2 * yn * yn;
        const double cdist = 1 + a.k[0] * r2 + a.k[1] * r4 + a.k[4] * r6;
        const double icdist2 = 1.0 / (1 + a.k[5] * r2 + a.k[6] * r4 + a.k[7] * r6);
        xd = xn * cdist * icdist2 + a.k[2] * a1 + a.k[3] * a2;
        yd = yn * cdist * icdist2 + a.k[2] * a3 + a.k[3] * a1;
      }
      const float uf = (float)(xd * a.fx + a.cx), vf = (float)(yd * a.fy + a.cy);
      // numpy float32 -> int32 truncates towards zero; out-of-range values are rejected by the image test anyway
      u = (fabsf(uf) < 2.0e9f) ? (int)uf : -1;
      v = (fabsf(vf) < 2.0e9f) ? (int)vf : -1;
      inimg = u >= 0 && u < a.W && v >= 0 && v < a.H;
      sel = inimg && (!mask || mask[(size_t)v * a.W + u] != 0);
    }
  }
  const unsigned bf = __ballot_sync(FULL, front), bi = __ballot_sync(FULL, inimg);
  if ((threadIdx.x & 31) == 0) { if (bf) atomicAdd(&counts[1], __popc(bf)); if (bi) atomicAdd(&counts[2], __popc(bi)); }
  unsigned blk_total;
  const unsigned local = block_scan_256(sel ? 1u : 0u, s_warp, &blk_total);
  if (threadIdx.x == 0) {
    s_base = lookback(ws, tile, blk_total);
    if (tile == (int)gridDim.x - 1) counts[0] = (int)(s_base + blk_total);
  }
  __syncthreads();
  if (sel) {
    const size_t m = (size_t)s_base + local;
    idx_out[m] = i;
    if (uv_out) { uv_out[2 * m] = u; uv_out[2 * m + 1] = v; }
  }
}

}  // namespace
}  // namespace b2r

using namespace b2r;

extern "C" {

b2r_status b2r_render_depth(b2r_solver* s, const double* verts_dev, int nv, const int32_t* tris_dev, int nt, const double* T16, double fx,
                            double fy, double cx, double cy, int W, int H, double z_near, double z_far, float* depth_dev) {
  B2R_NVTX("b2r:render_depth");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!verts_dev || !tris_dev || !T16 || !depth_dev || nv < 0 || nt < 0 || W < 1 || H < 1 || !(fx > 0) || !(fy > 0) || !(z_far > z_near))
    return fail(s, B2R_ERR_INVALID_ARG, "render_depth arguments");
  CK(cudaSetDevice(s->device));
  CK(s->d_refine.ensure((size_t)(nv > 0 ? nv : 1) * 24 + 4096));
  double* vc = (double*)s->d_refine.p;
  Cam c;
  for (int k = 0; k < 12; ++k) c.T[k] = T16[k];
  c.fx = fx; c.fy = fy; c.cx = cx; c.cy = cy; c.z_near = z_near; c.z_far = z_far; c.W = W; c.H = H;
  cudaStream_t st = s->stream;
  const int npx = W * H;
  raster_clear_kernel<<<(npx + 255) / 256, 256, 0, st>>>((unsigned int*)depth_dev, npx);   // +inf everywhere
  s->launches += 1;
  if (nv > 0 && nt > 0) {
    raster_xform_kernel<<<(nv + 255) / 256, 256, 0, st>>>(verts_dev, nv, c, vc);
    raster_tri_kernel<<<nt, 128, 0, st>>>(vc, tris_dev, nt, c, (unsigned int*)depth_dev);
    s->launches += 2;
  }
  raster_resolve_kernel<<<(npx + 255) / 256, 256, 0, st>>>((unsigned int*)depth_dev, npx);
  s->launches += 1;
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_backproject_depth32(b2r_solver* s, const float* depth_dev, int H, int W, double fx, double fy, double cx, double cy,
                                   double depth_scale, double depth_trunc, double* points_dev, int32_t* pixel_ids_dev, int capacity,
                                   int* count_out) {
  B2R_NVTX("b2r:backproject_depth32");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!depth_dev || !points_dev || !count_out || H < 1 || W < 1 || capacity < 0 || !(depth_scale > 0))
    return fail(s, B2R_ERR_INVALID_ARG, "backproject_depth32 arguments");
  CK(cudaSetDevice(s->device));
  const int tiles = (int)(((long long)H * W + 255) / 256);
  CK(s->d_refine.ensure((size_t)(tiles + 2) * 8 + 4096));
  unsigned long long* ws = (unsigned long long*)s->d_refine.p;
  int* d_count = (int*)(ws + tiles + 1);
  CK(cudaMemsetAsync(ws, 0, (size_t)(tiles + 2) * 8, s->stream));
  backproject32_kernel<<<tiles, 256, 0, s->stream>>>(depth_dev, H, W, fx, fy, cx, cy, (float)depth_scale, (float)depth_trunc, depth_trunc > 0,
                                                     points_dev, pixel_ids_dev, capacity, ws, d_count);
  s->launches += 1;
  CK(cudaMemcpyAsync(count_out, d_count, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  if (*count_out > capacity) *count_out = capacity;
  return B2R_OK;
}

b2r_status b2r_project_select(b2r_solver* s, const double* pts_dev, int n, const double* R9, const double* t3, const double* K9, const double* D,
                              int nD, int W, int H, const uint8_t* mask_dev, double z_min, int32_t* idx_dev, int32_t* uv_dev, int* n_front_out,
                              int* n_in_image_out, int* n_selected_out) {
  B2R_NVTX("b2r:project_select");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !R9 || !t3 || !K9 || !idx_dev || !n_selected_out || n < 0 || W < 1 || H < 1 || nD < 0 || (nD > 0 && !D))
    return fail(s, B2R_ERR_INVALID_ARG, "project_select arguments");
  CK(cudaSetDevice(s->device));
  ProjArgs a;
  for (int k = 0; k < 9; ++k) a.R[k] = R9[k];
  for (int k = 0; k < 3; ++k) a.t[k] = t3[k];
  a.fx = K9[0]; a.fy = K9[4]; a.cx = K9[2]; a.cy = K9[5]; a.z_min = z_min; a.W = W; a.H = H;
  for (int k = 0; k < 8; ++k) a.k[k] = 0.0;
  a.nk = (nD == 4 || nD == 5 || nD == 8) ? nD : 0;     // main_seibersdorf.py:78: D if D.size in (4, 5, 8) else None
  for (int k = 0; k < a.nk; ++k) a.k[k] = D[k];
  const int tiles = (n + 255) / 256 > 0 ? (n + 255) / 256 : 1;
  CK(s->d_refine.ensure((size_t)(tiles + 4) * 8 + 4096));
  unsigned long long* ws = (unsigned long long*)s->d_refine.p;
  int* counts = (int*)(ws + tiles + 1);
  CK(cudaMemsetAsync(ws, 0, (size_t)(tiles + 4) * 8, s->stream));
  project_select_kernel<<<tiles, 256, 0, s->stream>>>(pts_dev, n, a, mask_dev, idx_dev, uv_dev, ws, counts);
  s->launches += 1;
  int h[3];
  CK(cudaMemcpyAsync(h, counts, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  *n_selected_out = h[0];
  if (n_front_out) *n_front_out = h[1];
  if (n_in_image_out) *n_in_image_out = h[2];
  return B2R_OK;
}

}  // extern "C"
