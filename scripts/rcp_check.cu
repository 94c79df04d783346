// rcp.approx.ftz.f32(1.0f) must be exactly 1.0f: render_backward divides T by (1 - alpha) with alpha = 0 for the
// lanes of a visit that do not contribute (nvcc -arch=sm_100a scripts/rcp_check.cu -o /tmp/rcp_check && /tmp/rcp_check)
#include <cstdio>
__global__ void k(float* o) {
  float x = 1.0f, y, z;
  asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(z) : "f"(0.0f));
  o[0] = y; o[1] = z;
  float t = 0.3f;
  for (int i = 0; i < 1000; i++) { float r; asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f - 0.0f * t)); t = t * r; }
  o[2] = t;
}
int main() {
  float* d; float h[3];
  cudaMalloc(&d, 12); k<<<1, 1>>>(d); cudaMemcpy(h, d, 12, cudaMemcpyDeviceToHost);
  printf("rcp(1)=%.9g bits %08x  ex2(0)=%.9g  T after 1000 no-op steps %.9g (0.3f = %.9g)\n", h[0], *(unsigned*)&h[0], h[1], h[2], 0.3f);
  return (h[0] == 1.0f && h[2] == 0.3f) ? 0 : 1;
}
