// preprocess_fwd.cu — per-Gaussian forward: cull, project, EWA cov2D, conic, radius,
// tile rect, SH -> RGB.  Replaces preprocessCUDA<3> (CR/forward.cu:157-258) and
// checkFrustum (CR/rasterizer_impl.cu:54-66).
//
// HBM-bound: one thread per Gaussian, outputs packed as 16 B records
// (conic+opacity, rgb+depth) so that the render kernels fetch a splat with two
// 128-bit loads.  Writes the depth sort key (float bits) and the identity
// payload for the depth-ordering sort in the same pass.
#include "common.cuh"
#include "gsmath.cuh"
#include "expand.cuh"

namespace b2gs {

// coefficient k, channel c: DC from `dc`, the rest from `rest` (packed layout: rest = dc + 3)
#define B2GS_SHK(k, c) ((k) == 0 ? dc[(c)] : rest[3 * ((k) - 1) + (c)])
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float* __restrict__ dc,
                                            const float* __restrict__ rest, float3 pos,
                                            float3 campos, uint8_t& clamp_bits) {
  // CR/forward.cu:22-73
  float3 dir = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
  float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  dir.x = dir.x / len;
  dir.y = dir.y / len;
  dir.z = dir.z / len;
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float r = B2GS_SH_C0 * B2GS_SHK(0, c);
    if (deg > 0) {
      float x = dir.x, y = dir.y, z = dir.z;
      r = r - B2GS_SH_C1 * y * B2GS_SHK(1, c) + B2GS_SH_C1 * z * B2GS_SHK(2, c) - B2GS_SH_C1 * x * B2GS_SHK(3, c);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z;
        float xy = x * y, yz = y * z, xz = x * z;
        r = r + B2GS_SH_C2_0 * xy * B2GS_SHK(4, c) + B2GS_SH_C2_1 * yz * B2GS_SHK(5, c) +
            B2GS_SH_C2_2 * (2.0f * zz - xx - yy) * B2GS_SHK(6, c) + B2GS_SH_C2_3 * xz * B2GS_SHK(7, c) +
            B2GS_SH_C2_4 * (xx - yy) * B2GS_SHK(8, c);
        if (deg > 2) {
          r = r + B2GS_SH_C3_0 * y * (3.0f * xx - yy) * B2GS_SHK(9, c) +
              B2GS_SH_C3_1 * xy * z * B2GS_SHK(10, c) +
              B2GS_SH_C3_2 * y * (4.0f * zz - xx - yy) * B2GS_SHK(11, c) +
              B2GS_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * B2GS_SHK(12, c) +
              B2GS_SH_C3_4 * x * (4.0f * zz - xx - yy) * B2GS_SHK(13, c) +
              B2GS_SH_C3_5 * z * (xx - yy) * B2GS_SHK(14, c) +
              B2GS_SH_C3_6 * x * (xx - 3.0f * yy) * B2GS_SHK(15, c);
        }
      }
    }
    r += 0.5f;
    if (r < 0.f) clamp_bits |= (uint8_t)(1u << c);
    res[c] = fmaxf(r, 0.0f);
  }
  return make_float3(res[0], res[1], res[2]);
}

// RAW is a template parameter so that the RAW=false instantiation is, instruction for
// instruction, the kernel whose roundings are bit-identical to the reference's.
template <bool RAW>
__global__ void __launch_bounds__(256, 5)
preprocess_fwd_kernel(int P, int D, int M, const float* __restrict__ means3D,
                      const float* __restrict__ scales, float scale_modifier,
                      const float* __restrict__ rotations, const float* __restrict__ opacities,
                      const float* __restrict__ shs, const float* __restrict__ cov3D_precomp,
                      const float* __restrict__ colors_precomp, const float* __restrict__ viewmatrix,
                      const float* __restrict__ projmatrix, const float* __restrict__ campos_p, int W,
                      int H, float tan_fovx, float tan_fovy, float focal_x, float focal_y, int gx,
                      int gy, int prefiltered,
                      const float* __restrict__ shs_rest, int32_t* __restrict__ radii,
                      int32_t* __restrict__ n_touched, uint32_t* __restrict__ tile_count, GeomWS g,
                      float* __restrict__ acc_zero) {
  __shared__ float s_view[16], s_proj[16];
  __shared__ WarpRects s_rects[8];
  pdl_wait();
  if (threadIdx.x < 16) s_view[threadIdx.x] = viewmatrix[threadIdx.x];
  else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = projmatrix[threadIdx.x - 16];
  __syncthreads();

  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = idx < P;
  float3 p_orig = {0.f, 0.f, 0.f};
  // every per-Gaussian input is requested up front, visible or not: the kernel is bound by the
  // latency of dependent loads (means -> cull test -> rotation / scale -> ... -> SH), not by bytes;
  // 30 % of the rows are culled later and their extra 44 B are the price of one round trip less
  float4 q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  float3 s_in = {0.f, 0.f, 0.f}, dc_in = {0.f, 0.f, 0.f};
  float op_in = 0.f;
  const bool sr_path = cov3D_precomp == nullptr;
  const bool sh_path = colors_precomp == nullptr;
  if (valid) {
    p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    if (sr_path) {
      q_in = reinterpret_cast<const float4*>(rotations)[idx];
      s_in = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
    }
    op_in = opacities[idx];
    if (sh_path) {
      const float* dc = RAW ? shs + (size_t)idx * 3 : shs + (size_t)idx * M * 3;
      dc_in = {dc[0], dc[1], dc[2]};
    }
    n_touched[idx] = 0;  // render_fwd accumulates into it (CR/forward.cu:369-371)
  }

  bool visible = valid;
  int my_radius_i = 0;
  uint32_t touched = 0;
  float2 point_image = {0.f, 0.f};
  float4 con_o = {0.f, 0.f, 0.f, 0.f};
  float3 p_view = xform_point_4x3(p_orig, s_view);
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;

  if (valid && p_view.z <= 0.2f) {  // near cull only (CR/auxiliary.h:153)
    if (prefiltered) {
      printf("Point is filtered although prefiltered is set. This shouldn't happen!");
      __trap();
    }
    visible = false;
  }

  if (visible) {
    float4 p_hom = xform_point_4x4(p_orig, s_proj);
    float p_w = 1.0f / (p_hom.w + 0.0000001f);
    float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    float cov6[6];
    if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int i = 0; i < 6; ++i) cov6[i] = cov3D_precomp[6 * idx + i];
    } else {
      float4 q = q_in;
      float sx = s_in.x, sy = s_in.y, sz = s_in.z;
      if (RAW) {  // fused activations: exp / normalize
        float qn;
        q = act_normalize(q, qn);
        sx = expf(sx); sy = expf(sy); sz = expf(sz);
      }
      cov3d_from_scale_rot(sx, sy, sz, scale_modifier, q, cov6);
    }

    Ewa e;
    ewa_setup(e, p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov6, s_view);
    Mat3 c2 = ewa_cov2d(e);
    c2.m[0][0] += 0.3f;  // low-pass: at least ~one pixel wide (CR/forward.cu:111-113)
    c2.m[1][1] += 0.3f;
    float3 cov = {c2.m[0][0], c2.m[0][1], c2.m[1][1]};

    float det = (cov.x * cov.z - cov.y * cov.y);
    if (det == 0.0f) {
      visible = false;
    } else {
      float det_inv = 1.f / det;
      float3 conic = {cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv};
      float mid = 0.5f * (cov.x + cov.z);
      float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
      float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
      float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
      point_image = {ndc_to_pix(p_proj.x, W), ndc_to_pix(p_proj.y, H)};
      my_radius_i = (int)my_radius;
      tile_rect(point_image, my_radius_i, gx, gy, x0, y0, x1, y1);
      touched = (uint32_t)((x1 - x0) * (y1 - y0));
      if (touched == 0) visible = false;
      con_o = {conic.x, conic.y, conic.z, RAW ? act_sigmoid(op_in) : op_in};
    }
  }

  if (valid && !visible) {
    radii[idx] = 0;
    g.tiles_touched[idx] = 0;
    touched = 0;
  }
  if (visible) {
    float3 rgb;
    uint8_t clamp_bits = 0;
    if (colors_precomp == nullptr) {
      const float3 campos = {campos_p[0], campos_p[1], campos_p[2]};
      const float dc[3] = {dc_in.x, dc_in.y, dc_in.z};
      const float* rest = RAW ? shs_rest + (size_t)idx * (M - 1) * 3 : shs + (size_t)idx * M * 3 + 3;
      rgb = sh_to_rgb(D, dc, rest, p_orig, campos, clamp_bits);
    } else {
      rgb = {colors_precomp[3 * idx], colors_precomp[3 * idx + 1], colors_precomp[3 * idx + 2]};
    }
    radii[idx] = my_radius_i;
    g.tiles_touched[idx] = touched;
    g.xy[idx] = point_image;
    g.conic_o[idx] = con_o;
    g.rgbd[idx] = make_float4(rgb.x, rgb.y, rgb.z, p_view.z);
    g.rect[idx] = make_ushort4((unsigned short)x0, (unsigned short)y0, (unsigned short)x1,
                               (unsigned short)y1);
    g.clamped[idx] = clamp_bits;
    if (acc_zero) {   // the render backward accumulates into this row; culled rows are never read
      float4* row = reinterpret_cast<float4*>(acc_zero + (size_t)idx * ACC_STRIDE);
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      row[0] = z; row[1] = z; row[2] = z;
    }
  }

  // instances per tile: the T counters become the tile ranges (binning.cu) — replaces the
  // P-wide InclusiveSum + identifyTileRanges (CR/rasterizer_impl.cu:280,116-138)
  const int lane = threadIdx.x & 31;
  warp_expand_rects<4>(
      s_rects[threadIdx.x >> 5], lane, touched, (uint32_t)x0 | ((uint32_t)y0 << 16),
      (uint32_t)(x1 - x0), (uint32_t)gx,
      [&](uint32_t tile) { atomicAdd(&tile_count[tile * CURSOR_STRIDE], 1u); return 0u; },
      [](uint32_t, uint32_t, int) {});
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D,
                                    const float* __restrict__ viewmatrix,
                                    uint8_t* __restrict__ present) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
  float3 v = xform_point_4x3(p, viewmatrix);
  present[idx] = v.z > 0.2f ? 1 : 0;
}

int launch_preprocess_fwd(const B2gsScene& s, const GeomWS& g, const ImageWS& im, int32_t* radii,
                          int32_t* n_touched, float* acc_zero, cudaStream_t st) {
  // focal lengths exactly as the reference derives them (CR/rasterizer_impl.cu:225-226)
  const float focal_y = s.H / (2.0f * s.tan_fovy);
  const float focal_x = s.W / (2.0f * s.tan_fovx);
  const int gx = (s.W + TILE - 1) / TILE, gy = (s.H + TILE - 1) / TILE;
  const int blocks = (s.P + 255) / 256;
  B2GS_PROF("preprocess_fwd", st);
  auto kern = s.raw_params ? preprocess_fwd_kernel<true> : preprocess_fwd_kernel<false>;
  B2GS_LAUNCH(kern, blocks, 256, st,
      s.P, s.D, s.M, s.means3D, s.scales, s.scale_modifier, s.rotations, s.opacities, s.shs,
      s.cov3D_precomp, s.colors_precomp, s.viewmatrix, s.projmatrix, s.campos, s.W, s.H,
      s.tan_fovx, s.tan_fovy, focal_x, focal_y, gx, gy, s.prefiltered, s.shs_rest, radii,
      n_touched, im.cursor, g, acc_zero);
  B2GS_LAUNCH_CHECK(st, s.debug);
  return B2GS_OK;
}

}  // namespace b2gs

extern "C" int b2gs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                                 const float* projmatrix, uint8_t* present, void* stream) {
  (void)projmatrix;  // the reference computes p_hom but only tests view-space z
  if (P < 0) {
    b2gs::set_error("b2gs_mark_visible: negative P");
    return B2GS_ERR_INVALID;
  }
  if (P == 0) return B2GS_OK;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  b2gs::mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, viewmatrix, present);
  B2GS_LAUNCH_CHECK(st, 0);
  return B2GS_OK;
}
