// pose_math.cuh — update_pose (utils/pose_utils.py:76-93) as a single-thread device function,
// shared by b2gs_pose_update (loss.cu) and b2gs_pose_step (window.cu).
#pragma once
#include "common.cuh"

namespace b2gs {

// state: [0:9] R row-major, [9:12] T, [16:32] viewmatrix, [32:48] full projection,
// [48:51] camera centre.  One thread: the arithmetic is ~200 flops.
__device__ inline void pose_update_device(float* __restrict__ state, const float* __restrict__ proj,
                                          float* __restrict__ rot_delta, float* __restrict__ trans_delta,
                                          float* __restrict__ tau_norm) {
  const float th[3] = {rot_delta[0], rot_delta[1], rot_delta[2]};
  const float rho[3] = {trans_delta[0], trans_delta[1], trans_delta[2]};
  // skew(theta), its square (utils/pose_utils.py:12-22)
  const float Wm[3][3] = {{0.f, -th[2], th[1]}, {th[2], 0.f, -th[0]}, {-th[1], th[0], 0.f}};
  float W2[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      float s = 0.f;
      for (int k = 0; k < 3; ++k) s += Wm[i][k] * Wm[k][j];
      W2[i][j] = s;
    }
  const float angle = sqrtf(th[0] * th[0] + th[1] * th[1] + th[2] * th[2]);
  float ra, rb, va, vb;  // R = I + ra W + rb W2 ; V = I + va W + vb W2  (:25-58)
  if (angle < 1e-5f) {
    ra = 1.f; rb = 0.5f; va = 0.5f; vb = 1.0f / 6.0f;
  } else {
    const float s = sinf(angle), c = cosf(angle);
    ra = s / angle;
    rb = (1.f - c) / (angle * angle);
    va = (1.0f - c) / (angle * angle);
    vb = (angle - s) / (angle * angle * angle);
  }
  float dR[3][3], V[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const float I = i == j ? 1.f : 0.f;
      dR[i][j] = I + ra * Wm[i][j] + rb * W2[i][j];
      V[i][j] = I + va * Wm[i][j] + vb * W2[i][j];
    }
  float t[3];
  for (int i = 0; i < 3; ++i) t[i] = V[i][0] * rho[0] + V[i][1] * rho[1] + V[i][2] * rho[2];
  // new_w2c = Exp(tau) @ T_w2c  (:76-93)
  float R[3][3], T[3], nR[3][3], nT[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) R[i][j] = state[3 * i + j];
    T[i] = state[9 + i];
  }
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) nR[i][j] = dR[i][0] * R[0][j] + dR[i][1] * R[1][j] + dR[i][2] * R[2][j];
    nT[i] = dR[i][0] * T[0] + dR[i][1] * T[1] + dR[i][2] * T[2] + t[i];
  }
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) state[3 * i + j] = nR[i][j];
    state[9 + i] = nT[i];
  }
  // viewmatrix = w2c^T stored row-major: view[4*c + r] = w2c[r][c] (utils/camera_utils.py:94-96)
  float view[16];
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) view[4 * c + r] = nR[r][c];
    view[12 + r] = nT[r];
  }
  view[3] = view[7] = view[11] = 0.f;
  view[15] = 1.f;
  for (int i = 0; i < 16; ++i) state[16 + i] = view[i];
  // full_proj = view @ proj as stored (utils/camera_utils.py:98-104)
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float s = 0.f;
      for (int k = 0; k < 4; ++k) s += view[4 * i + k] * proj[4 * k + j];
      state[32 + 4 * i + j] = s;
    }
  // camera centre = inverse(view)[3,:3] = -R^T T (utils/camera_utils.py:106-107)
  for (int j = 0; j < 3; ++j)
    state[48 + j] = -(nR[0][j] * nT[0] + nR[1][j] * nT[1] + nR[2][j] * nT[2]);
  if (tau_norm) {
    float n2 = 0.f;
    for (int i = 0; i < 3; ++i) n2 += th[i] * th[i] + rho[i] * rho[i];
    tau_norm[0] = sqrtf(n2);  // converged = tau.norm() < threshold, evaluated by the caller
  }
  for (int i = 0; i < 3; ++i) {
    rot_delta[i] = 0.f;
    trans_delta[i] = 0.f;
  }
}


}  // namespace b2gs
