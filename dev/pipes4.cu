// Micro-benchmark 5: does the register-operand count of packed / ALU ops limit their overlap?
// step = 8 FFMA2 (operand forms below) [+ 8 other ops], w warps per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0,%1,%2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 bc(float t) { u64 c; asm("mov.b64 %0,{%1,%1};" : "=l"(c) : "f"(t)); return c; }
__device__ __forceinline__ float fset(float a, float b) { float r; asm volatile("set.eq.f32.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float fsel(float a, float b) { float r; asm volatile("{.reg .pred p; setp.gt.f32 p,%1,%2; selp.f32 %0,%1,%2,p;}" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float fclamp(float a, float b) { float r; asm volatile("max.f32 %0,%1,%2; min.f32 %0,%0,0f3F800000;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float fsat(float a, float b) { float r; asm volatile("add.sat.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
// PF: 0 none, 1 ffma2(p,c,c), 2 ffma2(p,q_i,r_i), 3 fmul2(p,q_i), 4 ffma2(p,q_i,c)
// OF: 0 none, 1 fset(a,t), 2 fset(a,b_i), 3 fsetp+fsel(a,b_i), 4 add.sat(a,b_i)
template <int PF, int OF>
__global__ void __launch_bounds__(512) k(float* out, int iters, float seed) {
  const float t = seed + threadIdx.x * 1e-7f;
  u64 p[8], q[8], r[8]; float a[8], b[8]; const u64 c = bc(t);
  for (int i = 0; i < 8; ++i) { p[i] = bc(t + i); q[i] = bc(t * 0.5f + i); r[i] = bc(t * 0.25f + i); a[i] = t + i; b[i] = t * 3 + i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (PF == 1) p[i] = ffma2(p[i], c, c);
        if (PF == 2) p[i] = ffma2(p[i], q[i], r[i]);
        if (PF == 3) p[i] = fmul2(p[i], q[i]);
        if (PF == 4) p[i] = ffma2(p[i], q[i], c);
        if (OF == 1) a[i] = fset(a[i], t);
        if (OF == 2) a[i] = fset(a[i], b[i]);
        if (OF == 3) a[i] = fsel(a[i], b[i]);
        if (OF == 4) a[i] = fsat(a[i], b[i]);
        if (OF == 5) a[i] = fclamp(a[i], b[i]);
      }
    }
  }
  float acc = 0;
  for (int i = 0; i < 8; ++i) { float lo, hi; asm("mov.b64 {%0,%1},%2;" : "=f"(lo), "=f"(hi) : "l"(p[i] ^ q[i] ^ r[i])); acc += lo + hi + a[i] + b[i]; }
  if (acc == 12345.f) out[0] = acc;
}
template <int PF, int OF>
void run(const char* name) {
  float* d; cudaMalloc(&d, 4);
  const int iters = 4096;
  printf("%-34s", name);
  for (int w = 1; w <= 4; ++w) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<PF, OF><<<148, w * 128>>>(d, 4, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<PF, OF><<<148, w * 128>>>(d, iters, 1.0f);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double cyc = ms * 1e-3 * 1.965e9 / (iters * 4.0) / w;
    printf("  w=%d: %6.2f", w, cyc);
  }
  printf("   cycles per step of 8\n");
  cudaFree(d);
}
int main() {
  run<1, 0>("ffma2(p,c,c)");
  run<4, 0>("ffma2(p,q_i,c)");
  run<2, 0>("ffma2(p,q_i,r_i)");
  run<3, 0>("fmul2(p,q_i)");
  run<0, 1>("fset(a,t)");
  run<0, 2>("fset(a,b_i)");
  run<0, 3>("fsetp+fsel(a,b_i)");
  run<0, 4>("add.sat(a,b_i)");
  run<1, 1>("ffma2(p,c,c) + fset(a,t)");
  run<2, 1>("ffma2(p,q_i,r_i) + fset(a,t)");
  run<2, 2>("ffma2(p,q_i,r_i) + fset(a,b_i)");
  run<3, 2>("fmul2(p,q_i) + fset(a,b_i)");
  run<2, 3>("ffma2(p,q_i,r_i) + fsetp+fsel");
  run<2, 4>("ffma2(p,q_i,r_i) + add.sat");
  run<3, 4>("fmul2(p,q_i) + add.sat");
  run<0, 5>("max+min(a,b_i)");
  run<2, 5>("ffma2(p,q_i,r_i) + max+min");
  run<3, 5>("fmul2(p,q_i) + max+min");
  return 0;
}
