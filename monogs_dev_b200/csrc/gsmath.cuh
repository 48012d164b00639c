// gsmath.cuh — small fp32 helpers shared by the per-Gaussian kernels.
//
// The expression SHAPES (operand order, explicit zero terms) follow the
// reference's so that nvcc's default FMA contraction produces the same roundings:
// keys, radii and tile rects must be bit-identical to the reference's
// (SURVEY.md §7.3 item 1).  Column-major 3x3 with m[col][row], product term
// order k = 0,1,2 (the order the reference's GLM dependency uses).
#pragma once
#include <cuda_runtime.h>

namespace b2gs {

struct Mat3 {
  float m[3][3];  // m[col][row]
};

__device__ __forceinline__ Mat3 mat3_cols(float c00, float c01, float c02, float c10, float c11,
                                          float c12, float c20, float c21, float c22) {
  Mat3 r;
  r.m[0][0] = c00; r.m[0][1] = c01; r.m[0][2] = c02;
  r.m[1][0] = c10; r.m[1][1] = c11; r.m[1][2] = c12;
  r.m[2][0] = c20; r.m[2][1] = c21; r.m[2][2] = c22;
  return r;
}

__device__ __forceinline__ Mat3 mat3_mul(const Mat3& a, const Mat3& b) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int w = 0; w < 3; ++w) {
      r.m[c][w] = a.m[0][w] * b.m[c][0] + a.m[1][w] * b.m[c][1] + a.m[2][w] * b.m[c][2];
    }
  }
  return r;
}

__device__ __forceinline__ Mat3 mat3_transpose(const Mat3& a) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int w = 0; w < 3; ++w) r.m[c][w] = a.m[w][c];
  return r;
}

// view-space point: rows of the column-major 4x4 (CR/auxiliary.h:58-66)
__device__ __forceinline__ float3 xform_point_4x3(const float3 p, const float* __restrict__ mt) {
  float3 t;
  t.x = mt[0] * p.x + mt[4] * p.y + mt[8] * p.z + mt[12];
  t.y = mt[1] * p.x + mt[5] * p.y + mt[9] * p.z + mt[13];
  t.z = mt[2] * p.x + mt[6] * p.y + mt[10] * p.z + mt[14];
  return t;
}

__device__ __forceinline__ float4 xform_point_4x4(const float3 p, const float* __restrict__ mt) {
  float4 t;
  t.x = mt[0] * p.x + mt[4] * p.y + mt[8] * p.z + mt[12];
  t.y = mt[1] * p.x + mt[5] * p.y + mt[9] * p.z + mt[13];
  t.z = mt[2] * p.x + mt[6] * p.y + mt[10] * p.z + mt[14];
  t.w = mt[3] * p.x + mt[7] * p.y + mt[11] * p.z + mt[15];
  return t;
}

// NDC -> pixel, evaluated in double like the reference (CR/auxiliary.h:41-44).
__device__ __forceinline__ float ndc_to_pix(float v, int S) {
  return ((v + 1.0) * S - 1.0) * 0.5;
}

// tile rectangle of a splat (CR/auxiliary.h:46-56)
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, int gx, int gy, int& x0,
                                          int& y0, int& x1, int& y1) {
  x0 = min(gx, max(0, (int)((p.x - max_radius) / TILE)));
  y0 = min(gy, max(0, (int)((p.y - max_radius) / TILE)));
  x1 = min(gx, max(0, (int)((p.x + max_radius + TILE - 1) / TILE)));
  y1 = min(gy, max(0, (int)((p.y + max_radius + TILE - 1) / TILE)));
}

// Rotation (columns as the reference builds them from the raw, un-normalised
// quaternion, CR/forward.cu:129-141) and M = S * R.
__device__ __forceinline__ Mat3 quat_to_Rcols(float r, float x, float y, float z) {
  return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// Sigma = (S R)^T (S R), upper triangle (CR/forward.cu:120-154)
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod,
                                                     float4 q, float* cov6) {
  Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
  S.m[0][0] = mod * sx;
  S.m[1][1] = mod * sy;
  S.m[2][2] = mod * sz;
  Mat3 R = quat_to_Rcols(q.x, q.y, q.z, q.w);
  Mat3 M = mat3_mul(S, R);
  Mat3 Sigma = mat3_mul(mat3_transpose(M), M);
  cov6[0] = Sigma.m[0][0];
  cov6[1] = Sigma.m[0][1];
  cov6[2] = Sigma.m[0][2];
  cov6[3] = Sigma.m[1][1];
  cov6[4] = Sigma.m[1][2];
  cov6[5] = Sigma.m[2][2];
}

// EWA projection pieces shared by forward and backward (CR/forward.cu:76-115,
// CR/backward.cu:173-206): clamped view-space mean, J, W, T = W*J.
struct Ewa {
  float3 t;        // clamped
  float txtz, tytz;
  float limx, limy;
  Mat3 J, W, T, Vrk;
};

__device__ __forceinline__ void ewa_setup(Ewa& e, const float3 mean, float fx, float fy,
                                          float tan_fovx, float tan_fovy, const float* cov6,
                                          const float* __restrict__ view) {
  float3 t = xform_point_4x3(mean, view);
  e.limx = 1.3f * tan_fovx;
  e.limy = 1.3f * tan_fovy;
  e.txtz = t.x / t.z;
  e.tytz = t.y / t.z;
  t.x = min(e.limx, max(-e.limx, e.txtz)) * t.z;
  t.y = min(e.limy, max(-e.limy, e.tytz)) * t.z;
  e.t = t;
  e.J = mat3_cols(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z),
                  0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z),
                  0.f, 0.f, 0.f);
  e.W = mat3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  e.T = mat3_mul(e.W, e.J);
  e.Vrk = mat3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
}

// cov2D = T^T Vrk^T T (before the 0.3 low-pass)
__device__ __forceinline__ Mat3 ewa_cov2d(const Ewa& e) {
  return mat3_mul(mat3_mul(mat3_transpose(e.T), mat3_transpose(e.Vrk)), e.T);
}

// ---- fused parameter activations (GS/scene/gaussian_model.py:76-101) ---------------------
// scaling_activation = exp, opacity_activation = sigmoid, rotation_activation = F.normalize
__device__ __forceinline__ float act_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float4 act_normalize(float4 q, float& norm) {
  norm = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
  return make_float4(q.x / norm, q.y / norm, q.z / norm, q.w / norm);
}

// SH basis constants (CR/auxiliary.h:22-39)
#define B2GS_SH_C0 0.28209479177387814f
#define B2GS_SH_C1 0.4886025119029199f
#define B2GS_SH_C2_0 1.0925484305920792f
#define B2GS_SH_C2_1 -1.0925484305920792f
#define B2GS_SH_C2_2 0.31539156525252005f
#define B2GS_SH_C2_3 -1.0925484305920792f
#define B2GS_SH_C2_4 0.5462742152960396f
#define B2GS_SH_C3_0 -0.5900435899266435f
#define B2GS_SH_C3_1 2.890611442640554f
#define B2GS_SH_C3_2 -0.4570457994644658f
#define B2GS_SH_C3_3 0.3731763325901154f
#define B2GS_SH_C3_4 -0.4570457994644658f
#define B2GS_SH_C3_5 1.445305721320277f
#define B2GS_SH_C3_6 -0.5900435899266435f

}  // namespace b2gs
