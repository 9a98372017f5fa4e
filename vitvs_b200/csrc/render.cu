// Headless planar-scene RGB-D camera: what the Gazebo RealSense plugin hands the servo loop, as a
// CUDA kernel writing device buffers - no graphics API.
//
// replaces: the simulated D435 looking at the textured poster of the reference's Gazebo world - a 0.8 x 0.6 m plane
// (models/viso/meshes/resized.dae:113, texture models/viso/materials/textures/resized.png), pinhole 640 x 480 colour
// camera (_d435.gazebo.xacro:30-37; config.yaml u_max / v_max / f_x) and the depth encoding of
// RealSensePlugin.cpp:252-262 (uint16 millimetres, 0 outside [0.2, 10] m).  Gazebo / OGRE's own rasteriser output is
// "unavailable on this box" and is not emulated pixel for pixel; the oracle is the CPU restatement of the same
// ray-plane maths (oracle/render.py), and this kernel is bit-identical to it: fp64, no fused multiply-adds, the same
// operation order.
//
// World frame: the poster lies in z = 0, centred on the origin, x to the right and y down in the texture; a camera at
// (0, 0, -h) with identity rotation looks along +z at the poster from h metres.  Pose = camera-to-world [R | t].
// One thread per pixel; texture reads through the read-only path (the 5 MB texture is L2 resident).
#include <string>

#include "vvs_internal.h"

namespace vvs {
namespace {

__device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }

__global__ void __launch_bounds__(256)
render_plane_kernel(const uint8_t* __restrict__ tex, int th, int tw, double plane_w, double plane_h,
                    const double* __restrict__ poses, int n, int width, int height, double fx, double fy, double cx,
                    double cy, uint8_t* __restrict__ rgb, uint16_t* __restrict__ depth) {
  const long long idx = blockIdx.x * 256ll + threadIdx.x;
  const long long per = static_cast<long long>(width) * height;
  if (idx >= per * n) return;
  const int cam = static_cast<int>(idx / per);
  const int pix = static_cast<int>(idx - cam * per);
  const int v = pix / width, u = pix - v * width;
  const double* P = poses + cam * 12;     // R row-major [9], t [3]
  const double dx = __ddiv_rn(sub(u, cx), fx), dy = __ddiv_rn(sub(v, cy), fy);
  // world ray direction R (dx, dy, 1)
  const double wx = add(add(mul(P[0], dx), mul(P[1], dy)), P[2]);
  const double wy = add(add(mul(P[3], dx), mul(P[4], dy)), P[5]);
  const double wz = add(add(mul(P[6], dx), mul(P[7], dy)), P[8]);
  uint8_t c[3] = {128, 128, 128};         // background: flat grey
  uint16_t mm = 0;
  if (wz > 1e-12) {
    const double s = __ddiv_rn(-P[11], wz);         // camera-frame depth of the hit (the ray has unit z in camera frame)
    if (s > 0.0) {
      const double px = add(P[9], mul(s, wx)), py = add(P[10], mul(s, wy));
      if (s >= 0.2 && s <= 10.0) mm = static_cast<uint16_t>(floor(add(mul(s, 1000.0), 0.5)));
      // continuous texture coordinates of the hit, texel centres at integers
      const double tx = sub(mul(__ddiv_rn(add(px, mul(0.5, plane_w)), plane_w), tw), 0.5);
      const double ty = sub(mul(__ddiv_rn(add(py, mul(0.5, plane_h)), plane_h), th), 0.5);
      if (tx >= -0.5 && tx <= tw - 0.5 && ty >= -0.5 && ty <= th - 0.5) {
        const double fx0 = floor(tx), fy0 = floor(ty);
        const double ax = sub(tx, fx0), ay = sub(ty, fy0);
        const int x0 = min(max(static_cast<int>(fx0), 0), tw - 1), x1 = min(max(static_cast<int>(fx0) + 1, 0), tw - 1);
        const int y0 = min(max(static_cast<int>(fy0), 0), th - 1), y1 = min(max(static_cast<int>(fy0) + 1, 0), th - 1);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const double c00 = __ldg(tex + (static_cast<size_t>(y0) * tw + x0) * 3 + k);
          const double c01 = __ldg(tex + (static_cast<size_t>(y0) * tw + x1) * 3 + k);
          const double c10 = __ldg(tex + (static_cast<size_t>(y1) * tw + x0) * 3 + k);
          const double c11 = __ldg(tex + (static_cast<size_t>(y1) * tw + x1) * 3 + k);
          const double top = add(c00, mul(ax, sub(c01, c00)));
          const double bot = add(c10, mul(ax, sub(c11, c10)));
          const double val = add(top, mul(ay, sub(bot, top)));
          c[k] = static_cast<uint8_t>(fmin(fmax(floor(add(val, 0.5)), 0.0), 255.0));
        }
      }
    }
  }
  uint8_t* o = rgb + idx * 3;
  o[0] = c[0];
  o[1] = c[1];
  o[2] = c[2];
  depth[idx] = mm;
}

}  // namespace

int launch_render_plane(const uint8_t* tex, int th, int tw, double plane_w, double plane_h, const double* poses, int n,
                        const Camera& cam, uint8_t* rgb, uint16_t* depth, cudaStream_t s, std::string* err) {
  const long long total = static_cast<long long>(n) * cam.width * cam.height;
  render_plane_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, s>>>(
      tex, th, tw, plane_w, plane_h, poses, n, cam.width, cam.height, cam.fx, cam.fy, cam.cx, cam.cy, rgb, depth);
  VVS_CUDA_OK(cudaGetLastError(), err);
  return VVS_OK;
}

}  // namespace vvs
