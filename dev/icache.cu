// Micro-benchmark 3: does a long straight-line loop body (instruction footprint) lower IPC?
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fmnmx(float a, float b) { float r; asm volatile("max.f32 %0,%1,%2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 bc(float t) { u64 c; asm("mov.b64 %0,{%1,%1};" : "=l"(c) : "f"(t)); return c; }
// body: REP x (4 FFMA2 + 4 FMNMX) on 8 independent chains => 8*REP instructions per loop iteration
template <int REP>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
  const float t = seed + threadIdx.x * 1e-7f;
  u64 p[4]; float a[4]; const u64 c = bc(t);
  for (int i = 0; i < 4; ++i) { p[i] = bc(t + i); a[i] = t + i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < REP; ++r) {
#pragma unroll
      for (int i = 0; i < 4; ++i) { p[i] = ffma2(p[i], c, c); a[i] = fmnmx(a[i], t); }
    }
  }
  float acc = 0;
  for (int i = 0; i < 4; ++i) { float lo, hi; asm("mov.b64 {%0,%1},%2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); acc += lo + hi + a[i]; }
  if (acc == 12345.f) out[0] = acc;
}
template <int REP>
void run(int w) {
  float* d; cudaMalloc(&d, 4);
  const int total = 1 << 22;            // instructions per warp
  const int iters = total / (8 * REP);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<REP><<<148, w * 128>>>(d, 4, 1.0f);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<REP><<<148, w * 128>>>(d, iters, 1.0f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double cyc = ms * 1e-3 * 1.965e9;
  printf("body %5d instrs (%6.1f KB)  warps/SMSP %d: IPC per SMSP %.3f\n", 8 * REP, 8 * REP * 16 / 1024.0, w, (double)iters * 8 * REP * w / cyc);
  cudaFree(d);
}
int main() {
  for (int w = 1; w <= 2; ++w) {
    run<4>(w); run<16>(w); run<32>(w); run<64>(w); run<96>(w); run<128>(w); run<192>(w); run<256>(w); run<512>(w); run<1024>(w);
  }
  return 0;
}
