// seed.cu — RGB-D keyframe back-projection: the point cloud that seeds new Gaussians.
// Replaces, on the keyframe-insertion path that feeds distCUDA2:
//   GaussianModel.create_pcd_from_image / create_pcd_from_image_and_depth
//   (GS/scene/gaussian_model.py:107-156): exposure-corrected colour clamped and quantised to
//   uint8 on the device, moved to the HOST, open3d.geometry.RGBDImage.create_from_color_and_depth
//   (depth_scale 1, depth_trunc 100, no intensity conversion) and
//   open3d.geometry.PointCloud.create_from_rgbd_image(project_valid_depth_only=True), back to
//   the device.  open3d (pinned 0.17.0 in the reference's environment.yml) is a third-party
//   dependency that is not vendored; its published algorithm is restated here and in
//   oracle/seed_oracle.py:
//     depth  >= depth_trunc -> 0;  a pixel is valid iff depth > 0
//     z = depth; x = (u - cx) z / fx; y = (v - cy) z / fy          (double precision)
//     world = inverse(extrinsic) (x, y, z, 1)                       (double precision)
//     colour = uint8 / 255                                          (double precision)
//   and the reference then converts points and colours to float32.
// One thread per pixel, dense outputs + validity byte (the caller compacts and down-samples).
// HBM-bound: 20 B in, 25 B out per pixel.
#include "common.cuh"

namespace b2gs {

struct Cam2World { double m[12]; };   // first three rows of inverse(extrinsic), row-major

__global__ void __launch_bounds__(256)
backproject_kernel(int W, int H, const float* __restrict__ image, const float* __restrict__ exposure_a,
                   const float* __restrict__ exposure_b, const float* __restrict__ depth, double fx,
                   double fy, double cx, double cy, const __grid_constant__ Cam2World c2w,
                   float depth_trunc, float* __restrict__ xyz, float* __restrict__ rgb,
                   uint8_t* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int N = W * H;
  if (i >= N) return;
  const int v = i / W, u = i - v * W;
  float d = depth[i];
  if (d >= depth_trunc) d = 0.f;
  const bool ok = d > 0.f;           // NaN -> invalid
  valid[i] = ok ? 1 : 0;
  // colour: clamp(exp(a) * img + b, 0, 1) in float32 (torch op sequence), * 255, truncated to uint8
  const float ea = exposure_a ? expf(exposure_a[0]) : 1.f;
  const float eb = exposure_b ? exposure_b[0] : 0.f;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float x = __fadd_rn(__fmul_rn(ea, image[(size_t)c * N + i]), eb);
    x = fminf(fmaxf(x, 0.f), 1.f);
    const unsigned q = (unsigned)__fmul_rn(x, 255.f);
    rgb[3 * (size_t)i + c] = (float)((double)q / 255.0);
  }
  if (!ok) {
    xyz[3 * (size_t)i] = xyz[3 * (size_t)i + 1] = xyz[3 * (size_t)i + 2] = 0.f;
    return;
  }
  const double z = (double)d;
  const double x = __ddiv_rn(__dmul_rn((double)u - cx, z), fx);
  const double y = __ddiv_rn(__dmul_rn((double)v - cy, z), fy);
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    // left-to-right, separately rounded (what the oracle does; Eigen may differ in the last bit of
    // the double, which the float32 conversion hides)
    double acc = __dmul_rn(c2w.m[4 * r], x);
    acc = __dadd_rn(acc, __dmul_rn(c2w.m[4 * r + 1], y));
    acc = __dadd_rn(acc, __dmul_rn(c2w.m[4 * r + 2], z));
    acc = __dadd_rn(acc, c2w.m[4 * r + 3]);
    xyz[3 * (size_t)i + r] = (float)acc;
  }
}

}  // namespace b2gs

extern "C" int b2gs_backproject_rgbd(int32_t W, int32_t H, const float* image, const float* exposure_a,
                                     const float* exposure_b, const float* depth, double fx, double fy,
                                     double cx, double cy, const double* cam2world, float depth_trunc,
                                     float* xyz, float* rgb, uint8_t* valid, void* stream) {
  using namespace b2gs;
  if (W <= 0 || H <= 0) return B2GS_OK;
  if (!image || !depth || !cam2world || !xyz || !rgb || !valid) {
    set_error("b2gs_backproject_rgbd: NULL pointer");
    return B2GS_ERR_INVALID;
  }
  if (!(fx != 0.0) || !(fy != 0.0)) { set_error("b2gs_backproject_rgbd: zero focal length"); return B2GS_ERR_INVALID; }
  Cam2World m;
  for (int k = 0; k < 12; ++k) m.m[k] = cam2world[k];
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long N = (long long)W * H;
  B2GS_PROF("backproject", st);
  backproject_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(W, H, image, exposure_a, exposure_b, depth,
                                                                 fx, fy, cx, cy, m, depth_trunc, xyz, rgb,
                                                                 valid);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}
