// preprocess_bwd.cu — fused per-Gaussian backward.
// Replaces computeCov2DCUDA (CR/backward.cu:150-346), preprocessCUDA backward
// (:418-539) with SH backward (:21-145) and cov3D backward (:350-413), and the
// torch.sum over dL_dtau[P,6] (DGR/diff_gaussian_rasterization/__init__.py:152).
//
// One pass over the Gaussians: reads the packed 48 B accumulator row written by
// render_bwd, recomputes the EWA intermediates (cov3D is recomputed from
// scale/rotation instead of being stored and re-read), writes every per-Gaussian
// gradient row exactly once (zeros for culled Gaussians, so the caller needs no
// memset of the outputs), and reduces the pose gradient dL_dtau in-kernel:
// warp shuffle -> shared -> one partial row per block -> a fixed-order final sum
// (deterministic given the accumulator).
//
// Provenance note: the closed-form derivative chains in here RESTATE the reference's formulas —
// the conic / cov2D -> cov3D -> T -> J -> t chain follows CR/backward.cu:204-271 term by term
// (same intermediate names, so the two can be read side by side) and the SH gradient block
// follows CR/backward.cu:59-128.  Everything around them is new: the fusion of three reference
// kernels, the packed accumulator row, the cross-product form of the pose Jacobian, the fused
// activations, the in-kernel tau reduction, the accumulate / statistics / regulariser sinks.
#include "common.cuh"
#include "gsmath.cuh"

namespace b2gs {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}

// d|v|^-1 v / dv applied to dv (CR/auxiliary.h:107-118)
__device__ __forceinline__ V3 dnormvdv3(V3 v, V3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  V3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

// SH backward for one Gaussian: writes dL_dsh rows [0, M), returns dL/d(mean) through
// the view direction (CR/backward.cu:21-145).
// dc / rest: coefficient 0 and coefficients 1..M-1 (packed layout: rest = dc + 3), same split
// for the outputs; `accumulate` adds instead of writing.
#define B2GS_SHK(k, c) ((k) == 0 ? dc[(c)] : rest[3 * ((k) - 1) + (c)])
__device__ __forceinline__ V3 sh_backward(int deg, int M, const float* __restrict__ dc,
                                          const float* __restrict__ rest, V3 pos, V3 campos,
                                          uint8_t clamp_bits, V3 dL_dRGB, float* __restrict__ d_dc,
                                          float* __restrict__ d_rest, bool accumulate,
                                          float* __restrict__ stage_row = nullptr) {
  V3 dir_orig = {pos.x - campos.x, pos.y - campos.y, pos.z - campos.z};
  float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  dL_dRGB.x *= (clamp_bits & 1) ? 0.f : 1.f;
  dL_dRGB.y *= (clamp_bits & 2) ? 0.f : 1.f;
  dL_dRGB.z *= (clamp_bits & 4) ? 0.f : 1.f;
  const float g[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};

  float ddir[3] = {0.f, 0.f, 0.f};  // dL/d(dir)
  // the constant (DC) basis function on its own: degree 0 — the mapping configuration — then never
  // touches the run-time indexed basis table below (a 64 B local-memory array)
  if (d_dc != nullptr) {
    if (accumulate) {
      red_add3(d_dc, B2GS_SH_C0 * g[0], B2GS_SH_C0 * g[1], B2GS_SH_C0 * g[2]);
    } else {
      d_dc[0] = B2GS_SH_C0 * g[0]; d_dc[1] = B2GS_SH_C0 * g[1]; d_dc[2] = B2GS_SH_C0 * g[2];
    }
  }
  if (deg == 0) {
    for (int k = 0; stage_row != nullptr && k < 3 * (M - 1); ++k) stage_row[k] = 0.f;
    for (int k = 1; d_dc != nullptr && !accumulate && k < M; ++k) {
      float* o = d_rest + 3 * (k - 1);
      o[0] = 0.f; o[1] = 0.f; o[2] = 0.f;
    }
    return dnormvdv3(dir_orig, V3{0.f, 0.f, 0.f});
  }
  float basis[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) basis[i] = 0.f;
  basis[0] = B2GS_SH_C0;
  {
    basis[1] = -B2GS_SH_C1 * y;
    basis[2] = B2GS_SH_C1 * z;
    basis[3] = -B2GS_SH_C1 * x;
    float xx = 0, yy = 0, zz = 0, xy = 0, yz = 0, xz = 0;
    if (deg > 1) {
      xx = x * x; yy = y * y; zz = z * z; xy = x * y; yz = y * z; xz = x * z;
      basis[4] = B2GS_SH_C2_0 * xy;
      basis[5] = B2GS_SH_C2_1 * yz;
      basis[6] = B2GS_SH_C2_2 * (2.f * zz - xx - yy);
      basis[7] = B2GS_SH_C2_3 * xz;
      basis[8] = B2GS_SH_C2_4 * (xx - yy);
      if (deg > 2) {
        basis[9] = B2GS_SH_C3_0 * y * (3.f * xx - yy);
        basis[10] = B2GS_SH_C3_1 * xy * z;
        basis[11] = B2GS_SH_C3_2 * y * (4.f * zz - xx - yy);
        basis[12] = B2GS_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
        basis[13] = B2GS_SH_C3_4 * x * (4.f * zz - xx - yy);
        basis[14] = B2GS_SH_C3_5 * z * (xx - yy);
        basis[15] = B2GS_SH_C3_6 * x * (xx - 3.f * yy);
      }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float dx = -B2GS_SH_C1 * B2GS_SHK(3, c);
      float dy = -B2GS_SH_C1 * B2GS_SHK(1, c);
      float dz = B2GS_SH_C1 * B2GS_SHK(2, c);
      if (deg > 1) {
        dx += B2GS_SH_C2_0 * y * B2GS_SHK(4, c) + B2GS_SH_C2_2 * 2.f * -x * B2GS_SHK(6, c) + B2GS_SH_C2_3 * z * B2GS_SHK(7, c) +
              B2GS_SH_C2_4 * 2.f * x * B2GS_SHK(8, c);
        dy += B2GS_SH_C2_0 * x * B2GS_SHK(4, c) + B2GS_SH_C2_1 * z * B2GS_SHK(5, c) + B2GS_SH_C2_2 * 2.f * -y * B2GS_SHK(6, c) +
              B2GS_SH_C2_4 * 2.f * -y * B2GS_SHK(8, c);
        dz += B2GS_SH_C2_1 * y * B2GS_SHK(5, c) + B2GS_SH_C2_2 * 2.f * 2.f * z * B2GS_SHK(6, c) + B2GS_SH_C2_3 * x * B2GS_SHK(7, c);
        if (deg > 2) {
          dx += (B2GS_SH_C3_0 * B2GS_SHK(9, c) * 3.f * 2.f * xy + B2GS_SH_C3_1 * B2GS_SHK(10, c) * yz +
                 B2GS_SH_C3_2 * B2GS_SHK(11, c) * -2.f * xy + B2GS_SH_C3_3 * B2GS_SHK(12, c) * -3.f * 2.f * xz +
                 B2GS_SH_C3_4 * B2GS_SHK(13, c) * (-3.f * xx + 4.f * zz - yy) +
                 B2GS_SH_C3_5 * B2GS_SHK(14, c) * 2.f * xz + B2GS_SH_C3_6 * B2GS_SHK(15, c) * 3.f * (xx - yy));
          dy += (B2GS_SH_C3_0 * B2GS_SHK(9, c) * 3.f * (xx - yy) + B2GS_SH_C3_1 * B2GS_SHK(10, c) * xz +
                 B2GS_SH_C3_2 * B2GS_SHK(11, c) * (-3.f * yy + 4.f * zz - xx) +
                 B2GS_SH_C3_3 * B2GS_SHK(12, c) * -3.f * 2.f * yz + B2GS_SH_C3_4 * B2GS_SHK(13, c) * -2.f * xy +
                 B2GS_SH_C3_5 * B2GS_SHK(14, c) * -2.f * yz + B2GS_SH_C3_6 * B2GS_SHK(15, c) * -3.f * 2.f * xy);
          dz += (B2GS_SH_C3_1 * B2GS_SHK(10, c) * xy + B2GS_SH_C3_2 * B2GS_SHK(11, c) * 4.f * 2.f * yz +
                 B2GS_SH_C3_3 * B2GS_SHK(12, c) * 3.f * (2.f * zz - xx - yy) +
                 B2GS_SH_C3_4 * B2GS_SHK(13, c) * 4.f * 2.f * xz + B2GS_SH_C3_5 * B2GS_SHK(14, c) * (xx - yy));
        }
      }
      ddir[0] += dx * g[c];
      ddir[1] += dy * g[c];
      ddir[2] += dz * g[c];
    }
  }
  const int active = (deg + 1) * (deg + 1);
  for (int k = 1; d_dc != nullptr && k < M; ++k) {
    const float bk = (k < active && k < 16) ? basis[k] : 0.f;
    float* o = d_rest + 3 * (k - 1);
    if (stage_row != nullptr) {      // the warp adds its 32 rows to the arena together (kernel tail)
      stage_row[3 * (k - 1)] = bk * g[0]; stage_row[3 * (k - 1) + 1] = bk * g[1]; stage_row[3 * (k - 1) + 2] = bk * g[2];
    } else if (accumulate) {
      if (bk != 0.f) red_add3(o, bk * g[0], bk * g[1], bk * g[2]);
    } else {
      o[0] = bk * g[0]; o[1] = bk * g[1]; o[2] = bk * g[2];
    }
  }
  return dnormvdv3(dir_orig, V3{ddir[0], ddir[1], ddir[2]});
}

template <bool RAW>
__global__ void __launch_bounds__(128, 6)
preprocess_bwd_kernel(int P, int D, int M, const float* __restrict__ means3D,
                      const int32_t* __restrict__ radii, const float* __restrict__ shs,
                      const uint8_t* __restrict__ clamped, const float* __restrict__ scales,
                      const float* __restrict__ rotations, float scale_modifier,
                      const float* __restrict__ cov3D_precomp, const float* __restrict__ viewmatrix,
                      const float* __restrict__ projmatrix, const float* __restrict__ projmatrix_raw,
                      const float* __restrict__ campos_p, float h_x, float h_y, float tan_fovx,
                      float tan_fovy, const float* __restrict__ acc,
                      float* __restrict__ dL_dmeans3D, float* __restrict__ dL_dmeans2D,
                      float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors,
                      float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                      float* __restrict__ dL_drots, float* __restrict__ dL_dcov3D,
                      float* __restrict__ tau_partials,
                      const float* __restrict__ shs_rest, const float* __restrict__ opacities,
                      int accumulate, float* __restrict__ dL_dsh_rest,
                      float* __restrict__ st_grad_accum, float* __restrict__ st_denom,
                      float* __restrict__ st_max_radii, float iso_w, int iso_stride, int iso_offset,
                      const int32_t* __restrict__ gate0, const int32_t* __restrict__ gate1,
                      const int32_t* __restrict__ gate2, const float4* __restrict__ conic_o, float half_w,
                      float half_h, int stage_floats) {
  __shared__ float s_view[16], s_proj[16];
  __shared__ float s_tau[4][7];
  // accumulate mode, SH degree > 0: the dL_dsh_rest rows of a warp's 32 Gaussians are one contiguous,
  // 16-byte aligned span of the arena, so they are staged here and added with red.v4 by the whole
  // warp — 11 reduction lane-operations per Gaussian at degree 3 instead of 45 scalar ones (the
  // kernel was bound by their number at cfg5: 0.98 -> 0.78 ms with v2 triplets alone)
  // DYNAMIC shared memory, 128 * 3 (M - 1) floats, handed over by the launch only when this path
  // runs: as a static 23 KB array it took the L1 carve-out of six resident blocks at degree 0 too
  // and cost the benched SH0 configuration 2 us per view.
  extern __shared__ __align__(16) float s_rest[];
  pdl_wait();
  // statistics gates (B2gsBackwardIO.stats_gate): this view overflowed / the previous step was
  // discarded / the host has already counted this view for the step being re-run
  if ((gate0 && *gate0) || (gate1 && *gate1) || (gate2 && *gate2)) {
    st_grad_accum = nullptr; st_denom = nullptr; st_max_radii = nullptr;
  }
  if (threadIdx.x < 16) s_view[threadIdx.x] = viewmatrix[threadIdx.x];
  else if (threadIdx.x < 32) s_proj[threadIdx.x - 16] = projmatrix[threadIdx.x - 16];
  __syncthreads();
  const float* view = s_view;
  const float* proj = s_proj;

  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float tau[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  // all per-Gaussian inputs are requested before the first dependent use (radius -> accumulator row
  // -> mean -> rotation / scale -> opacity ... would otherwise be a chain of DRAM round trips; the
  // kernel is latency-bound).  Rows of culled Gaussians are fetched in vain: ~30 % of 100 B.
  int radius_in = 0;
  float4 co_in = {0.f, 0.f, 0.f, 0.f};
  float4 a0 = {0.f, 0.f, 0.f, 0.f}, a1 = a0, a2 = a0, q_in = make_float4(1.f, 0.f, 0.f, 0.f);
  float3 mean = {0.f, 0.f, 0.f}, s_in = {0.f, 0.f, 0.f};
  float op_raw = 0.f;
  uint8_t clamp_in = 0;
  if (idx < P) {
    radius_in = radii[idx];
    co_in = conic_o[idx];
    const float4* arow = reinterpret_cast<const float4*>(acc + (size_t)idx * ACC_STRIDE);
    a0 = arow[0]; a1 = arow[1]; a2 = arow[2];
    mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    if (cov3D_precomp == nullptr) {
      q_in = reinterpret_cast<const float4*>(rotations)[idx];
      s_in = {scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]};
    }
    if (RAW) op_raw = opacities[idx];
    if (shs != nullptr) clamp_in = clamped[idx];
  }
  const bool live = radius_in > 0;
  const int R3 = 3 * (M - 1);
  const bool coop = stage_floats >= 128 * R3 && RAW && accumulate && dL_dsh != nullptr && dL_dsh_rest != nullptr &&
                    shs != nullptr && M > 1 && D > 0 && blockDim.x == 128 &&
                    (reinterpret_cast<uintptr_t>(dL_dsh_rest) & 15u) == 0;
  float* const stage_row = coop ? s_rest + threadIdx.x * R3 : nullptr;
  if (coop && !live) {
    for (int k = 0; k < R3; ++k) stage_row[k] = 0.f;
  }

  if (idx < P && !live && !accumulate) {
    // culled: every gradient row is zero (reference: torch::zeros, DGR/rasterize_points.cu:160-170)
    if (dL_dmeans3D) { dL_dmeans3D[3 * idx] = 0.f; dL_dmeans3D[3 * idx + 1] = 0.f; dL_dmeans3D[3 * idx + 2] = 0.f; }
    if (dL_dmeans2D) { dL_dmeans2D[3 * idx] = 0.f; dL_dmeans2D[3 * idx + 1] = 0.f; dL_dmeans2D[3 * idx + 2] = 0.f; }
    if (dL_dopacity) dL_dopacity[idx] = 0.f;
    if (dL_dcolors) { dL_dcolors[3 * idx] = 0.f; dL_dcolors[3 * idx + 1] = 0.f; dL_dcolors[3 * idx + 2] = 0.f; }
    if (dL_dsh && !RAW) for (int k = 0; k < 3 * M; ++k) dL_dsh[(size_t)idx * M * 3 + k] = 0.f;
    if (dL_dsh && RAW) {
      for (int k = 0; k < 3; ++k) dL_dsh[(size_t)idx * 3 + k] = 0.f;
      for (int k = 0; k < 3 * (M - 1); ++k) dL_dsh_rest[(size_t)idx * (M - 1) * 3 + k] = 0.f;
    }
    if (dL_dscales) { dL_dscales[3 * idx] = 0.f; dL_dscales[3 * idx + 1] = 0.f; dL_dscales[3 * idx + 2] = 0.f; }
    if (dL_drots) reinterpret_cast<float4*>(dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (dL_dcov3D) for (int k = 0; k < 6; ++k) dL_dcov3D[6 * idx + k] = 0.f;
  }

  if (live) {
    // render_bwd leaves the moments of q = G * dL_dalpha about the splat centre (render_bwd.cu,
    // accumulator row layout); the factors CR/backward.cu:735-757 applies per pixel — conic,
    // opacity, the viewport scale — are applied to the sums here
    const float dmean2D_x = half_w * co_in.w * -fmaf(co_in.x, a0.x, co_in.y * a0.y);
    const float dmean2D_y = half_h * co_in.w * -fmaf(co_in.z, a0.y, co_in.y * a0.x);
    const float nho = -0.5f * co_in.w;
    const float3 dL_dconic = {nho * a0.z, nho * a0.w, nho * a1.x};
    const float dL_dop = a1.y;
    const V3 dL_dcol = {a1.z, a1.w, a2.x};
    const float dL_ddepth = a2.y;

    float cov6[6];
    float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
    float sx = 0.f, sy = 0.f, sz = 0.f, q_norm = 1.f;
    if (cov3D_precomp != nullptr) {
#pragma unroll
      for (int i = 0; i < 6; ++i) cov6[i] = cov3D_precomp[6 * idx + i];
    } else {
      q = q_in;
      sx = s_in.x; sy = s_in.y; sz = s_in.z;
      if (RAW) {  // same activations as the forward (gsmath.cuh)
        q = act_normalize(q, q_norm);
        sx = expf(sx); sy = expf(sy); sz = expf(sz);
      }
      cov3d_from_scale_rot(sx, sy, sz, scale_modifier, q, cov6);
    }

    // ---------------- cov2D / conic backward (CR/backward.cu:150-346) ----------------
    Ewa e;
    ewa_setup(e, mean, h_x, h_y, tan_fovx, tan_fovy, cov6, view);
    const float3 t = e.t;
    const float x_grad_mul = (e.txtz < -e.limx || e.txtz > e.limx) ? 0.f : 1.f;
    const float y_grad_mul = (e.tytz < -e.limy || e.tytz > e.limy) ? 0.f : 1.f;
    const Mat3& J = e.J;
    const Mat3& Wm = e.W;
    const Mat3& T = e.T;
    const Mat3& Vrk = e.Vrk;
    Mat3 cov2D = ewa_cov2d(e);
    const float a = cov2D.m[0][0] += 0.3f;
    const float b = cov2D.m[0][1];
    const float c = cov2D.m[1][1] += 0.3f;

    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
      dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
      dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);

      dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
      dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
      dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
      dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][1] * dL_dc;
      dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][2] * dL_dc;
      dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db + 2 * T.m[1][1] * T.m[1][2] * dL_dc;
    }

    // dL/dT (upper 2x3)
    const float dL_dT00 = 2 * (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_da +
                          (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_db;
    const float dL_dT01 = 2 * (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_da +
                          (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_db;
    const float dL_dT02 = 2 * (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_da +
                          (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_db;
    const float dL_dT10 = 2 * (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_dc +
                          (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_db;
    const float dL_dT11 = 2 * (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_dc +
                          (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_db;
    const float dL_dT12 = 2 * (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_dc +
                          (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_db;

    const float dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
    const float dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
    const float dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
    const float dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;

    const float tz = 1.f / t.z;
    const float tz2 = tz * tz;
    const float tz3 = tz2 * tz;

    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 +
                         (2 * h_y * t.y) * tz3 * dL_dJ12;

    // pose: rho gets dL/dt, theta gets (-[t]x)^T dL/dt = t x dL/dt  (t is the CLAMPED mean, quirk q4)
    {
      const V3 gt = {dL_dtx, dL_dty, dL_dtz};
      const V3 th = cross3(V3{t.x, t.y, t.z}, gt);
      tau[0] += gt.x; tau[1] += gt.y; tau[2] += gt.z;
      tau[3] += th.x; tau[4] += th.y; tau[5] += th.z;
    }

    // dL/dmean through the covariance path (assign)
    V3 dmean;
    dmean.x = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
    dmean.y = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
    dmean.z = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;

    // pose through W (rotation columns), CR/backward.cu:299-343
    {
      const float dL_dW00 = J.m[0][0] * dL_dT00;
      const float dL_dW01 = J.m[0][0] * dL_dT01;
      const float dL_dW02 = J.m[0][0] * dL_dT02;
      const float dL_dW10 = J.m[1][1] * dL_dT10;
      const float dL_dW11 = J.m[1][1] * dL_dT11;
      const float dL_dW12 = J.m[1][1] * dL_dT12;
      const float dL_dW20 = J.m[0][2] * dL_dT00 + J.m[1][2] * dL_dT10;
      const float dL_dW21 = J.m[0][2] * dL_dT01 + J.m[1][2] * dL_dT11;
      const float dL_dW22 = J.m[0][2] * dL_dT02 + J.m[1][2] * dL_dT12;
      const V3 c1 = {view[0], view[1], view[2]};
      const V3 c2 = {view[4], view[5], view[6]};
      const V3 c3 = {view[8], view[9], view[10]};
      const V3 w1 = {dL_dW00, dL_dW10, dL_dW20};
      const V3 w2 = {dL_dW01, dL_dW11, dL_dW21};
      const V3 w3 = {dL_dW02, dL_dW12, dL_dW22};
      const V3 k1 = cross3(c1, w1), k2 = cross3(c2, w2), k3 = cross3(c3, w3);
      tau[3] += k1.x + k2.x + k3.x;
      tau[4] += k1.y + k2.y + k3.y;
      tau[5] += k1.z + k2.z + k3.z;
    }

    // ---------------- projection backward (CR/backward.cu:446-528) ----------------
    const float4 m_hom = xform_point_4x4(mean, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    {
      const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
      const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
      dmean.x += (proj[0] * m_w - proj[3] * mul1) * dmean2D_x + (proj[1] * m_w - proj[3] * mul2) * dmean2D_y;
      dmean.y += (proj[4] * m_w - proj[7] * mul1) * dmean2D_x + (proj[5] * m_w - proj[7] * mul2) * dmean2D_y;
      dmean.z += (proj[8] * m_w - proj[11] * mul1) * dmean2D_x + (proj[9] * m_w - proj[11] * mul2) * dmean2D_y;
    }
    // camera-frame point p_C = R m + t
    V3 p_C;
    p_C.x = view[0] * mean.x + view[4] * mean.y + view[8] * mean.z + view[12];
    p_C.y = view[1] * mean.x + view[5] * mean.y + view[9] * mean.z + view[13];
    p_C.z = view[2] * mean.x + view[6] * mean.y + view[10] * mean.z + view[14];
    {
      const float alpha = 1.0f * m_w;
      const float beta = -m_hom.x * m_w * m_w;
      const float gamma = -m_hom.y * m_w * m_w;
      const float pa = projmatrix_raw[0];
      const float pb = projmatrix_raw[5];
      const float pe = projmatrix_raw[11];
      const V3 d1 = {alpha * pa, 0.f, beta * pe};
      const V3 d2 = {0.f, alpha * pb, gamma * pe};
      const V3 d1t = cross3(p_C, d1);
      const V3 d2t = cross3(p_C, d2);
      tau[0] += dmean2D_x * d1.x + dmean2D_y * d2.x;
      tau[1] += dmean2D_x * d1.y + dmean2D_y * d2.y;
      tau[2] += dmean2D_x * d1.z + dmean2D_y * d2.z;
      tau[3] += dmean2D_x * d1t.x + dmean2D_y * d2t.x;
      tau[4] += dmean2D_x * d1t.y + dmean2D_y * d2t.y;
      tau[5] += dmean2D_x * d1t.z + dmean2D_y * d2t.z;
    }
    // depth = p_C.z
    {
      const float dL_dpCz = dL_ddepth;
      dmean.x += dL_dpCz * view[2];
      dmean.y += dL_dpCz * view[6];
      dmean.z += dL_dpCz * view[10];
      tau[2] += dL_dpCz;
      tau[3] += dL_dpCz * p_C.y;
      tau[4] += dL_dpCz * -p_C.x;
    }

    // ---------------- colour ----------------
    if (shs != nullptr) {
      const V3 campos = {campos_p[0], campos_p[1], campos_p[2]};
      const float* dc = RAW ? shs + (size_t)idx * 3 : shs + (size_t)idx * M * 3;
      const float* rest = RAW ? shs_rest + (size_t)idx * (M - 1) * 3 : dc + 3;
      float* d_dc = nullptr;
      float* d_rest = nullptr;
      if (dL_dsh) {
        d_dc = RAW ? dL_dsh + (size_t)idx * 3 : dL_dsh + (size_t)idx * M * 3;
        d_rest = RAW ? dL_dsh_rest + (size_t)idx * (M - 1) * 3 : d_dc + 3;
      }
      const V3 dm = sh_backward(D, M, dc, rest, V3{mean.x, mean.y, mean.z}, campos, clamp_in,
                                dL_dcol, d_dc, d_rest, accumulate != 0, stage_row);
      dmean.x += dm.x; dmean.y += dm.y; dmean.z += dm.z;
      tau[0] += -dm.x; tau[1] += -dm.y; tau[2] += -dm.z;
      if (dL_dcolors) { dL_dcolors[3 * idx] = dL_dcol.x; dL_dcolors[3 * idx + 1] = dL_dcol.y; dL_dcolors[3 * idx + 2] = dL_dcol.z; }
    } else if (dL_dcolors) {
      dL_dcolors[3 * idx] = dL_dcol.x; dL_dcolors[3 * idx + 1] = dL_dcol.y; dL_dcolors[3 * idx + 2] = dL_dcol.z;
    }

    // ---------------- cov3D -> scale, rotation (CR/backward.cu:350-413) ----------------
    if (scales != nullptr && (dL_dscales || dL_drots)) {
      const float r = q.x, x = q.y, y = q.z, z = q.w;
      const Mat3 R = quat_to_Rcols(r, x, y, z);
      Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
      const float s0 = scale_modifier * sx, s1 = scale_modifier * sy, s2 = scale_modifier * sz;
      S.m[0][0] = s0; S.m[1][1] = s1; S.m[2][2] = s2;
      const Mat3 Mm = mat3_mul(S, R);
      const Mat3 dL_dSigma = mat3_cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2],
                                       0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                                       0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
      Mat3 M2;
#pragma unroll
      for (int cc = 0; cc < 3; ++cc)
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) M2.m[cc][rr] = 2.0f * Mm.m[cc][rr];
      const Mat3 dL_dM = mat3_mul(M2, dL_dSigma);
      const Mat3 Rt = mat3_transpose(R);
      Mat3 dL_dMt = mat3_transpose(dL_dM);
      if (dL_dscales) {
        float ds0 = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
        float ds1 = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
        float ds2 = Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] + Rt.m[2][2] * dL_dMt.m[2][2];
        if (RAW) { ds0 *= sx; ds1 *= sy; ds2 *= sz; }  // d exp(x) = exp(x) dx
        if (accumulate) {
          // red.global.add: fire-and-forget, no load latency at the tail of the kernel (rows are
          // private to this thread, so there is no contention)
          red_add3(&dL_dscales[3 * (size_t)idx], ds0, ds1, ds2);
        } else {
          dL_dscales[3 * idx + 0] = ds0; dL_dscales[3 * idx + 1] = ds1; dL_dscales[3 * idx + 2] = ds2;
        }
      }
#pragma unroll
      for (int rr = 0; rr < 3; ++rr) {
        dL_dMt.m[0][rr] *= s0;
        dL_dMt.m[1][rr] *= s1;
        dL_dMt.m[2][rr] *= s2;
      }
      if (dL_drots) {
        float4 dq;
        dq.x = 2 * z * (dL_dMt.m[0][1] - dL_dMt.m[1][0]) + 2 * y * (dL_dMt.m[2][0] - dL_dMt.m[0][2]) + 2 * x * (dL_dMt.m[1][2] - dL_dMt.m[2][1]);
        dq.y = 2 * y * (dL_dMt.m[1][0] + dL_dMt.m[0][1]) + 2 * z * (dL_dMt.m[2][0] + dL_dMt.m[0][2]) + 2 * r * (dL_dMt.m[1][2] - dL_dMt.m[2][1]) - 4 * x * (dL_dMt.m[2][2] + dL_dMt.m[1][1]);
        dq.z = 2 * x * (dL_dMt.m[1][0] + dL_dMt.m[0][1]) + 2 * r * (dL_dMt.m[2][0] - dL_dMt.m[0][2]) + 2 * z * (dL_dMt.m[1][2] + dL_dMt.m[2][1]) - 4 * y * (dL_dMt.m[2][2] + dL_dMt.m[0][0]);
        dq.w = 2 * r * (dL_dMt.m[0][1] - dL_dMt.m[1][0]) + 2 * x * (dL_dMt.m[2][0] + dL_dMt.m[0][2]) + 2 * y * (dL_dMt.m[1][2] + dL_dMt.m[2][1]) - 4 * z * (dL_dMt.m[1][1] + dL_dMt.m[0][0]);
        // dq is the gradient wrt the quaternion the kernel used (quirk q3: the rasterizer never
        // back-propagates its own normalisation); with RAW the F.normalize backward
        // that autograd would run (GS/scene/gaussian_model.py:80-82) is applied here
        if (RAW) {
          const float dot = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;
          const float inv = 1.0f / q_norm;
          dq.x = (dq.x - q.x * dot) * inv; dq.y = (dq.y - q.y * dot) * inv;
          dq.z = (dq.z - q.z * dot) * inv; dq.w = (dq.w - q.w * dot) * inv;
        }
        if (accumulate) {
          float* out = dL_drots + 4 * (size_t)idx;
          if ((reinterpret_cast<uintptr_t>(out) & 15u) == 0) {      // uniform: rows are 16 B
            red_add_v4(out, dq.x, dq.y, dq.z, dq.w);
          } else {
            atomicAdd(out, dq.x); atomicAdd(out + 1, dq.y); atomicAdd(out + 2, dq.z); atomicAdd(out + 3, dq.w);
          }
        } else {
          reinterpret_cast<float4*>(dL_drots)[idx] = dq;
        }
      }
    } else if (!accumulate) {
      if (dL_dscales) { dL_dscales[3 * idx] = 0.f; dL_dscales[3 * idx + 1] = 0.f; dL_dscales[3 * idx + 2] = 0.f; }
      if (dL_drots) reinterpret_cast<float4*>(dL_drots)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
    }

    if (dL_dcov3D) {
#pragma unroll
      for (int k = 0; k < 6; ++k) dL_dcov3D[6 * idx + k] = dcov[k];
    }
    float d_op = dL_dop;
    if (RAW) {  // d sigmoid(x) = o (1 - o) dx
      const float o = act_sigmoid(op_raw);
      d_op *= o * (1.f - o);
    }
    // (pose-only callers — tracking — pass NULL for the per-Gaussian outputs they do not want)
    if (accumulate) {
      if (dL_dmeans3D) {
        red_add3(&dL_dmeans3D[3 * (size_t)idx], dmean.x, dmean.y, dmean.z);
      }
      if (dL_dopacity) atomicAdd(&dL_dopacity[idx], d_op);
    } else {
      if (dL_dmeans3D) {
        dL_dmeans3D[3 * idx] = dmean.x;
        dL_dmeans3D[3 * idx + 1] = dmean.y;
        dL_dmeans3D[3 * idx + 2] = dmean.z;
      }
      if (dL_dopacity) dL_dopacity[idx] = d_op;
    }
    if (dL_dmeans2D) {
      dL_dmeans2D[3 * idx] = dmean2D_x;
      dL_dmeans2D[3 * idx + 1] = dmean2D_y;
      dL_dmeans2D[3 * idx + 2] = 0.f;
    }
    // densification statistics of this view (GS/scene/gaussian_model.py:691-695,
    // utils/slam_backend.py:277-284), visibility_filter = radii > 0
    if (st_grad_accum) atomicAdd(&st_grad_accum[idx], sqrtf(dmean2D_x * dmean2D_x + dmean2D_y * dmean2D_y));
    if (st_denom) atomicAdd(&st_denom[idx], 1.f);
    // radii >= 0: the int order of the bit patterns is the float order; atomic so that views
    // running on concurrent streams can share the statistics arrays
    if (st_max_radii) atomicMax(reinterpret_cast<int*>(st_max_radii) + idx, __float_as_int((float)radius_in));
  }

  if (coop) {
    __syncwarp();
    const int wl = threadIdx.x & 31, ww = threadIdx.x >> 5;
    const size_t span0 = (size_t)(blockIdx.x * 128 + ww * 32) * (size_t)R3;   // the warp's first float in dL_dsh_rest
    const size_t total = (size_t)P * (size_t)R3;
    const float4* src = reinterpret_cast<const float4*>(s_rest + ww * 32 * R3);
    for (int q = wl; q < 8 * R3; q += 32) {
      const float4 v = src[q];
      if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) {
        const size_t f = span0 + 4 * (size_t)q;
        if (f + 4 <= total) {
          red_add_v4(dL_dsh_rest + f, v.x, v.y, v.z, v.w);
        } else {                        // the last chunk of the array: stay inside it
          const float e[4] = {v.x, v.y, v.z, v.w};
          for (int k = 0; k < 4; ++k)
            if (f + k < total && e[k] != 0.f) atomicAdd(dL_dsh_rest + f + k, e[k]);
        }
      }
    }
  }

  // ---- isotropic-scale regulariser (utils/slam_backend.py:229-231), every Gaussian, visible or
  // not: value partial + gradient wrt the raw log-scales added to the gradient arena ----
  float reg = 0.f;
  if (RAW && iso_w != 0.f && idx < P && (iso_stride <= 1 || idx % iso_stride == iso_offset)) {
    const float s0 = expf(s_in.x), s1 = expf(s_in.y), s2 = expf(s_in.z);
    const float m = (s0 + s1 + s2) / 3.0f;
    const float d0 = s0 - m, d1 = s1 - m, d2 = s2 - m;
    reg = fabsf(d0) + fabsf(d1) + fabsf(d2);
    const float g0 = (float)(d0 > 0.f) - (float)(d0 < 0.f), g1 = (float)(d1 > 0.f) - (float)(d1 < 0.f),
                g2 = (float)(d2 > 0.f) - (float)(d2 < 0.f);
    const float gm = (g0 + g1 + g2) / 3.0f;          // through the mean
    // d exp(x) = exp(x) dx
    red_add3(&dL_dscales[3 * (size_t)idx], iso_w * (g0 - gm) * s0, iso_w * (g1 - gm) * s1, iso_w * (g2 - gm) * s2);
  }

  // ---- dL_dtau (+ the regulariser partial): warp -> block -> one partial row per block ----
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    float v = k < 6 ? tau[k] : reg;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_tau[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 7) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) v += s_tau[w][threadIdx.x];
    tau_partials[(size_t)blockIdx.x * 8 + threadIdx.x] = v;
  }
}

// fixed-order final sum of the per-block partials -> dL_dtau[6]
__global__ void __launch_bounds__(256) tau_finalize_kernel(int blocks, const float* __restrict__ partials,
                                                           float* __restrict__ dL_dtau) {
  __shared__ float s[8][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int b = threadIdx.x; b < blocks; b += 256) {
#pragma unroll
    for (int k = 0; k < 6; ++k) v[k] += partials[(size_t)b * 8 + k];
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) s[warp][k] = x;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float x = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) x += s[w][threadIdx.x];
    dL_dtau[threadIdx.x] = x;
  }
}

int launch_preprocess_bwd(const B2gsScene& s, const GeomWS& g, const B2gsBackwardIO& io,
                          const float* acc, cudaStream_t st) {
  const float focal_y = s.H / (2.0f * s.tan_fovy);
  const float focal_x = s.W / (2.0f * s.tan_fovx);
  const int blocks = (s.P + 127) / 128;
  float* partials = const_cast<float*>(acc) + (size_t)s.P * ACC_STRIDE;  // scratch tail
  B2GS_PROF("preprocess_bwd", st);
  auto kern = s.raw_params ? preprocess_bwd_kernel<true> : preprocess_bwd_kernel<false>;
  // staging buffer of the dL_dsh_rest rows (kernel: `coop`): only where that path can run
  const int stage_floats = (s.raw_params && io.accumulate && io.dL_dsh && io.dL_dsh_rest && s.shs && s.M > 1 &&
                            s.M <= 16 && s.D > 0) ? 128 * 3 * (s.M - 1) : 0;
  B2GS_LAUNCH_SMEM(kern, blocks, 128, sizeof(float) * (size_t)stage_floats, st,
      s.P, s.D, s.M, s.means3D, io.radii, s.shs, g.clamped, s.scales, s.rotations, s.scale_modifier,
      s.cov3D_precomp, s.viewmatrix, s.projmatrix, s.projmatrix_raw, s.campos, focal_x, focal_y,
      s.tan_fovx, s.tan_fovy, acc, io.dL_dmeans3D, io.dL_dmeans2D, io.dL_dopacity, io.dL_dcolors,
      s.shs ? io.dL_dsh : nullptr, io.dL_dscales, io.dL_drotations, io.dL_dcov3D, partials,
      s.shs_rest, s.opacities, io.accumulate, io.dL_dsh_rest, io.densify_grad_accum,
      io.densify_denom, io.max_radii2D,
      io.iso_reg_weight != 0.f ? io.iso_reg_weight / (3.0f * (float)s.P) : 0.f, io.iso_reg_stride,
      io.iso_reg_offset, io.stats_gate[0], io.stats_gate[1], io.stats_gate[2], g.conic_o,
      0.5f * (float)s.W, 0.5f * (float)s.H, stage_floats);
  B2GS_LAUNCH_CHECK(st, s.debug);
  if (io.dL_dtau) {   // defer_tau callers hand the partial rows to b2gs_window_finalize instead
    tau_finalize_kernel<<<1, 256, 0, st>>>(blocks, partials, io.dL_dtau);
    B2GS_LAUNCH_CHECK(st, s.debug);
  }
  return B2GS_OK;
}

}  // namespace b2gs
