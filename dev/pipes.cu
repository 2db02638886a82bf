// Micro-benchmark: throughput of packed (f32x2) vs scalar fp32 ops and mixes on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0,%1,%2,%3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float fsel(float a, float b, float c) { float r; asm volatile("{.reg .pred p; setp.gt.f32 p,%1,%2; selp.f32 %0,%1,%3,p;}" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float fmnmx(float a, float b) { float r; asm volatile("max.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float mufu(float a) { float r; asm volatile("ex2.approx.ftz.f32 %0,%1;" : "=f"(r) : "f"(a)); return r; }

// MODE: NP packed + NS scalar + NA alu(max) + NM mufu ops per inner step, each on its own independent chain
template <int NP, int NS, int NA, int NM>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
  u64 p[NP > 0 ? NP : 1]; float s[NS > 0 ? NS : 1]; float a[NA > 0 ? NA : 1]; float m[NM > 0 ? NM : 1];
  const float t = seed + threadIdx.x * 1e-7f;
  u64 c; asm("mov.b64 %0,{%1,%1};" : "=l"(c) : "f"(t));
  for (int i = 0; i < NP; ++i) p[i] = c;
  for (int i = 0; i < NS; ++i) s[i] = t;
  for (int i = 0; i < NA; ++i) a[i] = t;
  for (int i = 0; i < NM; ++i) m[i] = t;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < NP; ++i) p[i] = ffma2(p[i], c, c);
#pragma unroll
      for (int i = 0; i < NS; ++i) s[i] = ffma(s[i], t, t);
#pragma unroll
      for (int i = 0; i < NA; ++i) a[i] = fmnmx(a[i], t);
#pragma unroll
      for (int i = 0; i < NM; ++i) m[i] = mufu(m[i]);
    }
  }
  float acc = 0;
  for (int i = 0; i < NP; ++i) { float lo, hi; asm("mov.b64 {%0,%1},%2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); acc += lo + hi; }
  for (int i = 0; i < NS; ++i) acc += s[i];
  for (int i = 0; i < NA; ++i) acc += a[i];
  for (int i = 0; i < NM; ++i) acc += m[i];
  if (acc == 12345.f) out[0] = acc;
}
template <int NP, int NS, int NA, int NM>
void run(const char* name, int warps_per_smsp) {
  float* d; cudaMalloc(&d, 4);
  const int iters = 4096;
  dim3 grid(148 * (warps_per_smsp * 4 * 32 / 256 > 0 ? 1 : 1)), block(warps_per_smsp * 128);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<NP, NS, NA, NM><<<grid, block>>>(d, 16, 1.0f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<NP, NS, NA, NM><<<grid, block>>>(d, iters, 1.0f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  // cycles per SMSP per inner step at 1.965 GHz; per step each warp issues NP+NS+NA+NM instrs, warps_per_smsp warps
  double cyc = ms * 1e-3 * 1.965e9 / (iters * 4.0);
  printf("%-28s warps/SMSP %d: %.2f cycles per step per SMSP (%d instr/step/warp) -> %.2f cyc per warp-instr-group\n", name, warps_per_smsp,
         cyc, NP + NS + NA + NM, cyc / warps_per_smsp);
  cudaFree(d);
}
int main() {
  for (int w = 1; w <= 4; w *= 2) {
    run<8, 0, 0, 0>("8 FFMA2", w);
    run<0, 8, 0, 0>("8 FFMA", w);
    run<0, 16, 0, 0>("16 FFMA", w);
    run<4, 8, 0, 0>("4 FFMA2 + 8 FFMA", w);
    run<4, 4, 0, 0>("4 FFMA2 + 4 FFMA", w);
    run<6, 4, 0, 0>("6 FFMA2 + 4 FFMA", w);
    run<0, 0, 8, 0>("8 FMNMX", w);
    run<0, 8, 8, 0>("8 FFMA + 8 FMNMX", w);
    run<8, 0, 8, 0>("8 FFMA2 + 8 FMNMX", w);
    run<4, 4, 4, 0>("4 FFMA2+4 FFMA+4 FMNMX", w);
    run<0, 0, 0, 4>("4 MUFU", w);
    run<4, 0, 0, 2>("4 FFMA2 + 2 MUFU", w);
    run<4, 4, 4, 1>("4 FFMA2+4 FFMA+4 FMNMX+1 MUFU", w);
  }
  return 0;
}
