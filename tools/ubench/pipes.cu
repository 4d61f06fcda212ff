// Micro-benchmark: issue rates of the instruction classes the sweep kernel is made of (sm_100a).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu ; prints warp-instr/clk/SM.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define CHAINS 8

template <int MODE> __global__ void __launch_bounds__(256) k(float* out, float a, float b, long long* cyc) {
  float x[CHAINS];
  unsigned long long xx[CHAINS / 2 + 4];
  __shared__ float sm[256 * 4 + 64];
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) x[i] = threadIdx.x * 0.001f + i;
#pragma unroll
  for (int i = 0; i < CHAINS / 2 + 4; ++i) xx[i] = ((unsigned long long)__float_as_uint(threadIdx.x * 0.001f + i) << 32) | __float_as_uint(1.0f + i);
  unsigned long long ab = ((unsigned long long)__float_as_uint(a) << 32) | __float_as_uint(b);
  sm[threadIdx.x] = a; sm[threadIdx.x + 256] = b; sm[threadIdx.x + 512] = a; sm[threadIdx.x + 768] = b;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < CHAINS / 2 + 4; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(xx[i]) : "l"(ab));
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(a));
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < CHAINS / 2 + 4; ++i) asm volatile("add.f32x2 %0, %0, %1;" : "+l"(xx[i]) : "l"(ab));
    } else if (MODE == 4) {
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("mul.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(a));
    } else if (MODE == 5) {
#pragma unroll
      for (int i = 0; i < CHAINS / 2 + 4; ++i) asm volatile("mul.f32x2 %0, %0, %1;" : "+l"(xx[i]) : "l"(ab));
    } else if (MODE == 6) {   // fma + int alu mix 1:1
#pragma unroll
      for (int i = 0; i < CHAINS / 2; ++i) {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
        unsigned u = __float_as_uint(x[i + 4]);
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(__float_as_uint(a)), "r"(__float_as_uint(b)));
        x[i + 4] = __uint_as_float(u);
      }
    } else if (MODE == 7) {   // shuffles
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) x[i] = __shfl_up_sync(0xffffffffu, x[i], 1);
    } else if (MODE == 8) {   // MUFU rcp
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
    } else if (MODE == 9) {   // LDS.32
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(&sm[(threadIdx.x + i * 32 + (__float_as_uint(x[i]) & 0)) & 1023])) : "memory"); x[i] = v; }
    } else if (MODE == 10) {  // LDS.128
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { float4 v; asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"((unsigned)__cvta_generic_to_shared(&sm[((threadIdx.x * 4) + i * 128) & 1023])) : "memory"); x[i] = v.x + v.w; }
    } else if (MODE == 11) {  // FMNMX (alu pipe)
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { unsigned u = __float_as_uint(x[i]); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(__float_as_uint(a) + it), "r"(__float_as_uint(b))); x[i] = __uint_as_float(u); }
    } else if (MODE == 16) {  // IADD3
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { int u = __float_as_int(x[i]); asm volatile("add.s32 %0, %0, %1;" : "+r"(u) : "r"(it)); x[i] = __int_as_float(u); }
    } else if (MODE == 17) {  // FSEL via setp+selp (2 alu ops)
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("{.reg .pred p; setp.gt.f32 p, %0, %1; selp.f32 %0, %2, %0, p;}" : "+f"(x[i]) : "f"(a), "f"(b));
    } else if (MODE == 18) {  // FMNMX alternating min/max so nothing is idempotent
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(x[(i + 1) % CHAINS])); }
    } else if (MODE == 19) {  // MOV chain? use prmt (alu)
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { unsigned u = __float_as_uint(x[i]); asm volatile("prmt.b32 %0, %0, %1, 0x1230;" : "+r"(u) : "r"(it)); x[i] = __uint_as_float(u); }
    } else if (MODE == 20) {  // FFMA + LOP3 2:1
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i + 4]) : "f"(a), "f"(b));
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) { unsigned u = (unsigned)xx[i]; asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(it), "r"(__float_as_uint(b))); xx[i] = (xx[i] & 0xffffffff00000000ull) | u; }
    } else if (MODE == 12) {  // fma + fmnmx mix 1:1
#pragma unroll
      for (int i = 0; i < CHAINS / 2; ++i) {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
        asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i + 4]) : "f"(a));
      }
    } else if (MODE == 13) {  // ffma2 + ffma 1:1
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(xx[i]) : "l"(ab));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(a), "f"(b));
      }
    } else if (MODE == 14) {  // STS.32
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) asm volatile("st.shared.f32 [%0], %1;" :: "r"((unsigned)__cvta_generic_to_shared(&sm[(threadIdx.x + i * 32) & 1023])), "f"(x[i]) : "memory");
    } else if (MODE == 15) {  // imad
#pragma unroll
      for (int i = 0; i < CHAINS; ++i) { int u = __float_as_int(x[i]); asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(u) : "r"(__float_as_int(a)), "r"(__float_as_int(b))); x[i] = __int_as_float(u); }
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += x[i];
#pragma unroll
  for (int i = 0; i < CHAINS / 2 + 4; ++i) s += __uint_as_float((unsigned)(xx[i] >> 32)) + __uint_as_float((unsigned)xx[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE> void run(const char* name, int per_iter, int blocks_per_sm) {
  int nb = 148 * blocks_per_sm;
  float* out; long long* cyc;
  cudaMalloc(&out, nb * 256 * sizeof(float)); cudaMalloc(&cyc, nb * sizeof(long long));
  k<MODE><<<nb, 256>>>(out, 1.0001f, 0.5f, cyc);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<nb, 256>>>(out, 1.0001f, 0.5f, cyc);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[148 * 8]; cudaMemcpy(h, cyc, nb * sizeof(long long), cudaMemcpyDeviceToHost);
  double c = 0; for (int i = 0; i < nb; ++i) c += h[i]; c /= nb;
  // warp-instr per clk per SM = blocks_per_sm * 8 warps * ITERS * per_iter / cycles
  printf("%-28s bps=%d  %.3f warp-instr/clk/SM  (%.1f us, %.0f cyc)\n", name, blocks_per_sm,
         (double)blocks_per_sm * 8 * ITERS * per_iter / c, ms * 1e3, c);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int bps : {1, 2}) {
    run<0>("FFMA", 8, bps); run<1>("FFMA2", 8, bps); run<2>("FADD", 8, bps); run<3>("FADD2", 8, bps);
    run<4>("FMUL", 8, bps); run<5>("FMUL2", 8, bps); run<6>("FFMA+LOP3 1:1", 8, bps); run<7>("SHFL", 8, bps);
    run<8>("MUFU.RCP", 8, bps); run<9>("LDS.32(+FADD)", 8, bps); run<10>("LDS.128(+FADD)", 8, bps);
    run<11>("LOP3", 8, bps); run<16>("IADD3", 8, bps); run<17>("FSETP+SELP (2)", 16, bps); run<18>("FMNMX dep", 8, bps); run<19>("PRMT", 8, bps); run<20>("FFMA+LOP3 2:1 (12)", 12, bps); run<12>("FFMA+FMNMX 1:1", 8, bps); run<13>("FFMA2+FFMA 1:1", 8, bps);
    run<14>("STS.32", 8, bps); run<15>("IMAD", 8, bps);
  }
  return 0;
}
