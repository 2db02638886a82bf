// Micro-benchmark 4: 8 FFMA2 interleaved with 8 ops of another kind, 1/2/3 warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 bc(float t) { u64 c; asm("mov.b64 %0,{%1,%1};" : "=l"(c) : "f"(t)); return c; }
template <int KIND> __device__ __forceinline__ float other(float a, float t) {
  float r;
  if (KIND == 0) asm volatile("max.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(t));
  if (KIND == 1) asm volatile("set.eq.f32.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(t));          // FSET.BF
  if (KIND == 2) asm volatile("{.reg .pred p; setp.gt.f32 p,%1,%2; selp.f32 %0,%1,%2,p;}" : "=f"(r) : "f"(a), "f"(t));  // FSETP+FSEL
  if (KIND == 3) asm volatile("add.sat.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(t));             // FADD.SAT
  if (KIND == 4) { unsigned x; asm volatile("lop3.b32 %0,%1,%2,%3,0x96;" : "=r"(x) : "r"(__float_as_uint(a)), "r"(__float_as_uint(t)), "r"(0x12345u)); r = __uint_as_float(x); }
  if (KIND == 5) asm volatile("mul.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(t));                 // FMUL scalar
  if (KIND == 6) asm volatile("ex2.approx.ftz.f32 %0,%1;" : "=f"(r) : "f"(a));                 // MUFU
  return r;
}
template <int KIND, int NP, int NO>
__global__ void __launch_bounds__(512) k(float* out, int iters, float seed) {
  const float t = seed + threadIdx.x * 1e-7f;
  u64 p[NP > 0 ? NP : 1]; float a[NO > 0 ? NO : 1]; const u64 c = bc(t);
  for (int i = 0; i < NP; ++i) p[i] = bc(t + i);
  for (int i = 0; i < NO; ++i) a[i] = t + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < 8; ++i) { if (i < NP) p[i] = ffma2(p[i], c, c); if (i < NO) a[i] = other<KIND>(a[i], t); }
    }
  }
  float acc = 0;
  for (int i = 0; i < NP; ++i) { float lo, hi; asm("mov.b64 {%0,%1},%2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); acc += lo + hi; }
  for (int i = 0; i < NO; ++i) acc += a[i];
  if (acc == 12345.f) out[0] = acc;
}
template <int KIND, int NP, int NO>
void run(const char* name) {
  float* d; cudaMalloc(&d, 4);
  const int iters = 4096;
  printf("%-26s", name);
  for (int w = 1; w <= 4; ++w) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<KIND, NP, NO><<<148, w * 128>>>(d, 4, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<KIND, NP, NO><<<148, w * 128>>>(d, iters, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double cyc = ms * 1e-3 * 1.965e9 / (iters * 4.0) / w;
    printf("  w=%d: %6.2f cyc/step", w, cyc);
  }
  printf("   (step = %d FFMA2 + %d other)\n", NP, NO);
  cudaFree(d);
}
int main() {
  run<0, 8, 0>("FFMA2 only");
  run<0, 0, 8>("FMNMX only");   run<0, 8, 8>("FFMA2 + FMNMX");
  run<1, 0, 8>("FSET.BF only"); run<1, 8, 8>("FFMA2 + FSET.BF");
  run<2, 0, 8>("FSETP+FSEL only"); run<2, 8, 8>("FFMA2 + FSETP+FSEL");
  run<3, 0, 8>("FADD.SAT only"); run<3, 8, 8>("FFMA2 + FADD.SAT");
  run<4, 0, 8>("LOP3 only"); run<4, 8, 8>("FFMA2 + LOP3");
  run<5, 0, 8>("FMUL only"); run<5, 8, 8>("FFMA2 + FMUL");
  run<6, 0, 2>("2 MUFU only"); run<6, 8, 2>("FFMA2 + 2 MUFU");
  return 0;
}
