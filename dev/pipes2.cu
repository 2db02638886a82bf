// Micro-benchmark 2: packed ops with distinct register operands (register-file port pressure).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0,%1,%2,%3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float fmul(float a, float b) { float r; asm volatile("mul.rn.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 bc(float t) { u64 c; asm("mov.b64 %0,{%1,%1};" : "=l"(c) : "f"(t)); return c; }

// MODE 0: p = fma2(p, q_i, r_i)  three distinct reg pairs     1: p = mul2(p, q_i)   2: p = fma2(p, c, c)
// MODE 3: scalar s = fma(s, q_i, r_i) distinct   4: scalar s = fmul(s, q_i)   5: p = fma2(p, imm, q_i)
template <int MODE, int N>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
  u64 p[N], q[N], r[N]; float s[N], sq[N], sr[N];
  const float t = seed + threadIdx.x * 1e-7f;
  for (int i = 0; i < N; ++i) { p[i] = bc(t + i); q[i] = bc(1.0f + 1e-7f * i + t * 1e-9f); r[i] = bc(t * 1e-9f + i * 1e-8f); s[i] = t; sq[i] = 1.0f + 1e-7f * i + t * 1e-9f; sr[i] = t * 1e-9f + i * 1e-8f; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
#pragma unroll
      for (int i = 0; i < N; ++i) {
        if (MODE == 0) p[i] = ffma2(p[i], q[i], r[i]);
        if (MODE == 1) p[i] = fmul2(p[i], q[i]);
        if (MODE == 2) p[i] = ffma2(p[i], q[0], q[0]);
        if (MODE == 3) s[i] = ffma(s[i], sq[i], sr[i]);
        if (MODE == 4) s[i] = fmul(s[i], sq[i]);
        if (MODE == 5) { u64 x; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(x) : "l"(p[i]), "l"(0x3f8000013f800001ull), "l"(q[i])); p[i] = x; }
        if (MODE == 6) s[i] = ffma(s[i], 1.0000001f, sr[i]);
        if (MODE == 7) { p[i] = fmul2(p[i], q[i]); s[i] = fmul(s[i], sq[i]); }
      }
    }
  }
  float acc = 0;
  for (int i = 0; i < N; ++i) { float lo, hi; asm("mov.b64 {%0,%1},%2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); acc += lo + hi + s[i]; }
  if (acc == 12345.f) out[0] = acc;
}
template <int MODE, int N>
void run(const char* name, int w) {
  float* d; cudaMalloc(&d, 4);
  const int iters = 4096;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE, N><<<148, w * 128>>>(d, 16, 1.0f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE, N><<<148, w * 128>>>(d, iters, 1.0f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double cyc = ms * 1e-3 * 1.965e9 / (iters * 4.0) / w / N;
  printf("%-44s warps/SMSP %d: %.2f cycles per op per SMSP  (%s)\n", name, w, cyc, cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
}
int main() {
  for (int w = 1; w <= 2; ++w) {
    run<0, 8>("FFMA2 3 distinct reg pairs", w);
    run<1, 8>("FMUL2 2 distinct reg pairs", w);
    run<2, 8>("FFMA2 p,c,c", w);
    run<5, 8>("FFMA2 p,imm,q", w);
    run<3, 8>("FFMA 3 distinct regs", w);
    run<4, 8>("FMUL 2 distinct regs", w);
    run<6, 8>("FFMA s,imm,r", w);
    run<7, 8>("FMUL2 + FMUL pair (per 2 ops)", w);
  }
  return 0;
}
