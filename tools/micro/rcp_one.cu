// rcp.approx.ftz(1.0f) and the Newton-refined reciprocal must be exactly 1 — dev check.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* o, float x) {
  float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  o[0] = r; o[1] = fmaf(r, fmaf(-x, r, 1.0f), r);
  float2 a = make_float2(3.f, 5.f), b = make_float2(x, 2.f);
  float2 c = __fmul2_rn(a, b); o[2] = c.x; o[3] = c.y;
}
int main() { float* d; cudaMalloc(&d, 16); k<<<1,1>>>(d, 1.0f); float h[4]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("rcp(1)=%.9g newton=%.9g exact=%d  mul2=(%g,%g)\n", h[0], h[1], h[0]==1.0f && h[1]==1.0f, h[2], h[3]); return 0; }
