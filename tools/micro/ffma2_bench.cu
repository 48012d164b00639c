// Micro-benchmark: issue rate of FFMA vs FFMA2 (packed fp32x2) on sm_100a — dev tool.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, float a, float b, int iters) {
  float2 x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = make_float2(threadIdx.x * 0.001f + i, i * 0.5f);
  const float2 aa = make_float2(a, a), bb = make_float2(b, b);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { x[i].x = fmaf(x[i].x, a, b); x[i].y = fmaf(x[i].y, a, b); }
      else if (MODE == 1) x[i] = __ffma2_rn(x[i], aa, bb);
      else if (MODE == 2) { x[i] = __fmul2_rn(x[i], aa); x[i] = __fadd2_rn(x[i], bb); }
      else { x[i].x = x[i].x * a; x[i].y = x[i].y * a; x[i].x = x[i].x + b; x[i].y = x[i].y + b; }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i].x + x[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* d) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 4096, blocks = 148 * 8, threads = 256;
  k<MODE><<<blocks, threads>>>(d, 0.999f, 0.001f, 16);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(d, 0.999f, 0.001f, iters);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double fl = (double)blocks * threads * iters * 16 * 2;  // 16 scalar fma-equivalents per iter
  printf("%-22s %.3f ms  %.1f TFLOP/s (fma-equivalent)\n", name, ms, fl / ms * 1e-9);
}
int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  run<0>("FFMA scalar", d); run<1>("FFMA2 packed", d); run<2>("FMUL2+FADD2", d); run<3>("FMUL+FADD scalar", d);
  return 0;
}
