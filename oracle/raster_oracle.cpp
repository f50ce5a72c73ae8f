// =============================================================================
// oracle/raster_oracle.cpp  --  TEST INFRASTRUCTURE ONLY.  NOT PRODUCT CODE.
//
// CPU restatement of the depth image the reference obtains from Open3D's Filament OffscreenRenderer
// (SURVEY.md section 8f row f-2):
//   PoseEstimator.create_template_from_H     PoseEstimator.py:118-141   (render CAD at the current pose, depth -> cloud)
//   render_templates                         template_creation.py:139-250 (five views of the CAD model -> template clouds)
// (paths under /root/reference/src/pose_estimator/EstimHelpers/).
//
// Filament / OpenGL cannot run here or on the GPU box ("unavailable on this box"): PARITY UNPINNED.  What is restated is the
// geometry both renderers agree on: `render_to_depth_image(z_in_view_space=True)` is, per pixel, the view-space z of the nearest
// surface point on the pixel's viewing ray, and create_from_rgbd_image back-projects pixel (u, v) along ((u-cx)/fx, (v-cy)/fy, 1).
// The ray of pixel (u, v) is taken through exactly that direction, so rendering followed by back-projection returns points ON the
// mesh (a GL rasteriser samples at pixel centres, half a pixel away: documented deviation).
// Brute force on purpose: every pixel against every triangle, Moeller-Trumbore in fp64 with a fixed operation order.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline legs may load this library.
// Build: oracle/Makefile (-O3 -fopenmp -ffp-contract=off).
// =============================================================================
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>
#include <omp.h>

namespace {

struct V3 { double x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }

constexpr double kEdgeEps = 1e-9;   // barycentric slack: shared edges of a closed mesh leave no cracks

// view-space depth at which the ray (dx, dy, 1) from the origin meets triangle (a, b, c); +inf when it does not
inline double ray_tri(double dx, double dy, V3 a, V3 b, V3 c) {
  const V3 d{dx, dy, 1.0};
  const V3 e1 = sub(b, a), e2 = sub(c, a);
  const V3 p = cross(d, e2);
  const double det = dot(e1, p);
  if (!(std::fabs(det) > 1e-300)) return std::numeric_limits<double>::infinity();
  const double inv = 1.0 / det;
  const V3 tv{-a.x, -a.y, -a.z};
  const double u = dot(tv, p) * inv;
  const V3 q = cross(tv, e1);
  const double v = dot(d, q) * inv;
  const double t = dot(e2, q) * inv;
  if (u >= -kEdgeEps && v >= -kEdgeEps && u + v <= 1.0 + kEdgeEps) return t;
  return std::numeric_limits<double>::infinity();
}

}  // namespace

extern "C" {

// verts [nv,3] f64 (model frame), tris [nt,3] i32, T16 = model -> camera (row-major 4x4, OpenCV camera: x right, y down, z forward).
// depth [H,W] f32: view-space z of the nearest hit with near < z < far, 0 where there is none.
void orcx_render_depth(const double* verts, int nv, const int32_t* tris, int nt, const double* T, double fx, double fy, double cx, double cy,
                       int W, int H, double z_near, double z_far, float* depth) {
  std::vector<V3> vc(nv);
  for (int i = 0; i < nv; ++i) {
    const double x = verts[3 * i], y = verts[3 * i + 1], z = verts[3 * i + 2];
    vc[i].x = ((T[0] * x + T[1] * y) + T[2] * z) + T[3];
    vc[i].y = ((T[4] * x + T[5] * y) + T[6] * z) + T[7];
    vc[i].z = ((T[8] * x + T[9] * y) + T[10] * z) + T[11];
  }
#pragma omp parallel for schedule(dynamic, 4)
  for (int v = 0; v < H; ++v)
    for (int u = 0; u < W; ++u) {
      const double dx = ((double)u - cx) / fx, dy = ((double)v - cy) / fy;
      double best = std::numeric_limits<double>::infinity();
      for (int k = 0; k < nt; ++k) {
        const double t = ray_tri(dx, dy, vc[tris[3 * k]], vc[tris[3 * k + 1]], vc[tris[3 * k + 2]]);
        if (t > z_near && t < z_far && t < best) best = t;
      }
      depth[(size_t)v * W + u] = std::isfinite(best) ? (float)best : 0.f;
    }
}

// RGBDImage.create_from_color_and_depth(depth_scale, depth_trunc) + PointCloud.create_from_rgbd_image on a float32 depth image:
// d = depth / scale (float32); d >= trunc -> 0; d > 0: z = double(d), x = (u - cx) z / fx, y = (v - cy) z / fy.  Row-major order.
int orcx_backproject_depth32(const float* depth, int H, int W, double fx, double fy, double cx, double cy, double depth_scale, double depth_trunc,
                             double* pts, int* pix) {
  int m = 0;
  const float sc = (float)depth_scale, tr = (float)depth_trunc;
  for (int v = 0; v < H; ++v)
    for (int u = 0; u < W; ++u) {
      float d = depth[(size_t)v * W + u] / sc;
      if (depth_trunc > 0 && d >= tr) d = 0.f;
      if (d > 0.f) {
        const double z = (double)d;
        pts[3 * m] = (u - cx) * z / fx; pts[3 * m + 1] = (v - cy) * z / fy; pts[3 * m + 2] = z;
        if (pix) pix[m] = v * W + u;
        ++m;
      }
    }
  return m;
}

}  // extern "C"
