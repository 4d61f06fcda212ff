// dependent-issue latency of FFMA vs FFMA2 (one chain per thread, one warp per scheduler)
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 8192
template <int MODE> __global__ void k(float* out, float a, float b, long long* cyc) {
  float x = threadIdx.x * 0.001f;
  unsigned long long xx = ((unsigned long long)__float_as_uint(x) << 32) | __float_as_uint(1.0f);
  unsigned long long ab = ((unsigned long long)__float_as_uint(a) << 32) | __float_as_uint(b);
  long long t0 = clock64();
#pragma unroll 16
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(a), "f"(b));
    if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(xx) : "l"(ab));
    if (MODE == 2) asm volatile("add.f32 %0, %0, %1;" : "+f"(x) : "f"(a));
    if (MODE == 3) asm volatile("add.f32x2 %0, %0, %1;" : "+l"(xx) : "l"(ab));
    if (MODE == 4) { asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(xx) : "l"(ab)); float lo = __uint_as_float((unsigned)xx); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(lo) : "f"(a), "f"(b)); xx = (xx & 0xffffffff00000000ull) | __float_as_uint(lo); }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x + __uint_as_float((unsigned)(xx >> 32)) + __uint_as_float((unsigned)xx);
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int MODE> void run(const char* n, int per) {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 128 * 4); cudaMalloc(&cyc, 148 * 8);
  k<MODE><<<148, 128>>>(out, 1.0001f, 0.5f, cyc); cudaDeviceSynchronize();
  k<MODE><<<148, 128>>>(out, 1.0001f, 0.5f, cyc); cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost);
  double c = 0; for (int i = 0; i < 148; ++i) c += h[i]; c /= 148;
  printf("%-22s %.2f cycles per dependent step (%d instr)\n", n, c / ITERS, per);
}
int main() { run<0>("FFMA chain", 1); run<1>("FFMA2 chain", 1); run<2>("FADD chain", 1); run<3>("FADD2 chain", 1); run<4>("FFMA2->FFMA(lo) chain", 2); return 0; }
