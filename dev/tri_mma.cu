// DEV PROTOTYPE (not part of the product library): the TriSpace backward's coefficient-gradient contraction
//     D[k][c] = sum_px T[k][px] * G[c][px]      (k: 126 monomials, c: 3 channels, px: pixels)
// on the 5th-generation tensor cores instead of FFMA2 over a shared-memory transpose.
//
//   * A = monomial tile, M = 128 rows (126 monomials + 2 zero rows), K = pixels; B = upstream gradient, N = 16 rows
//     (N = 16 is the smallest N tcgen05.mma takes at M = 128); D = 128 x 16 fp32 in tensor memory, accumulated
//     over all of a block's pixels and read back once (tcgen05.ld) at the end.
//   * fp32 accuracy from bf16 operands: every value is split v = hi + lo with hi = bf16(v), lo = bf16(v - hi)
//     (16 mantissa bits together).  The gradient's hi and lo parts sit in rows c and 3 + c of ONE B tile, so
//     A_hi B and A_lo B give hi*hi | hi*lo and lo*hi | lo*lo in columns 0-2 | 3-5: two tcgen05.mma.kind::f16 per
//     16 pixels, and the monomial tile (the big operand) is fetched from shared memory twice.
//     -DUSE_TF32: one pass of kind::tf32 on the unsplit fp32 values (10 mantissa bits: measured 5.7e-4, fails 1e-4).
//   * canonical K-major, no-swizzle operand layout (core matrix = 8 rows x 16 B); the stride between 16-byte
//     pixel chunks (LBO) is an ODD number of 16-byte units, which makes the pixel-pair stores of a warp conflict-free.
//   * two threads per pixel pair: the monomial tree's root branches are independent -- warps 0-1 walk the branch
//     of x0 (56 monomials), warps 2-3 the branches of x1..x4 (69).
//
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -I curl_b200/csrc -o dev/variants/tri_mma dev/tri_mma.cu
// Run  :  dev/variants/tri_mma [n_pixels]     prints accuracy against a double-precision host reference and timings.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <vector>

#include "curl_poly.cuh"

using namespace curl;

constexpr int NV = 5, DEG = 4;
constexpr int NC = poly_num_coeffs(NV, DEG);      // 126
constexpr int BT = 128;                            // threads per block
constexpr int CH = 128;                            // pixels per chunk = K of one batch of MMAs (64 pixel pairs)
constexpr int M_ROWS = 128, N_ROWS = 16;
#ifdef USE_TF32
constexpr int PX16 = 4;                            // pixels per 16 bytes
#else
constexpr int PX16 = 8;
#endif
constexpr int KCH = CH / PX16;                     // 16-byte pixel chunks per batch
constexpr int LBO_A = M_ROWS + 1;                  // 16-byte units between consecutive pixel chunks of A (odd: conflict-free)
constexpr int LBO_B = N_ROWS + 1;
constexpr int SBO = 8;                             // 16-byte units between 8-row groups: rows are 16 bytes apart
constexpr int A_BYTES = KCH * LBO_A * 16;
constexpr int B_BYTES = KCH * LBO_B * 16;
constexpr int SMEM_BYTES = 2 * A_BYTES + B_BYTES;
constexpr int TMEM_COLS_P = 32;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

// 64-bit shared-memory matrix descriptor (layout of cute::UMMA::SmemDescriptor): start address, leading / stride
// byte offsets in 16-byte units, version 1 (Blackwell), no swizzle
__device__ __forceinline__ unsigned long long make_desc(unsigned addr, unsigned lbo, unsigned sbo) {
  return (unsigned long long)((addr >> 4) & 0x3fffu) | ((unsigned long long)(lbo & 0x3fffu) << 16) |
         ((unsigned long long)(sbo & 0x3fffu) << 32) | (1ull << 46);
}
// 32-bit instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (bit 4), A / B format at bits 7 / 10
// (1 = BF16, 2 = TF32), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
#ifdef USE_TF32
constexpr unsigned IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((unsigned)(N_ROWS >> 3) << 17) | ((unsigned)(M_ROWS >> 4) << 24);
#define MMA_KIND "tf32"
#else
constexpr unsigned IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(N_ROWS >> 3) << 17) | ((unsigned)(M_ROWS >> 4) << 24);
#define MMA_KIND "f16"
#endif

__device__ __forceinline__ void mma_ss(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::" MMA_KIND " [%0], %1, %2, %3, p;\n}" ::"r"(tmem_d), "l"(da), "l"(db), "r"(IDESC), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(
          smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// v = hi + lo with hi = bf16(v) (round to nearest), lo = bf16(v - hi); both pixels of the pair in one 32-bit word
__device__ __forceinline__ void split_store(F2 v, unsigned a_hi, unsigned a_lo) {
#ifdef USE_TF32
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a_hi), "f"(v.lo()), "f"(v.hi()) : "memory");
  (void)a_lo;
#else
  unsigned hw, lw;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(hw) : "f"(v.hi()), "f"(v.lo()));      // {upper: second pixel, lower: first pixel}
  const F2 h(__uint_as_float(hw << 16), __uint_as_float(hw & 0xffff0000u));
  const F2 l = vsub(v, h);
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(lw) : "f"(l.hi()), "f"(l.lo()));
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(a_hi), "r"(hw) : "memory");
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(a_lo), "r"(lw) : "memory");
#endif
}
struct SplitStore {
  unsigned hi, lo;     // shared-space byte address of row 0 in this thread's pixel-pair column
  template <int RANK>
  __device__ __forceinline__ void operator()(F2 m) { split_store(m, hi + 16 * RANK, lo + 16 * RANK); }
};

// vars: [n][5] pixel variables, g: [n][3] upstream gradient, D: [126][3] (+= by float atomics: prototype)
__global__ void __launch_bounds__(BT, 3) contract_mma(const float* __restrict__ vars, const float* __restrict__ g, float* D, int n) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ unsigned s_tmem;
  unsigned char* A_hi = smem;
  unsigned char* A_lo = smem + A_BYTES;
  unsigned char* B_t = smem + 2 * A_BYTES;
  const int t = threadIdx.x;
  for (int i = t; i < SMEM_BYTES / 16; i += BT) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);   // zero rows stay zero
  if (t == 0) { mbar_init(&s_bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (t < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(TMEM_COLS_P) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned tmem = s_tmem;
  const unsigned long long dA_hi = make_desc(smem_u32(A_hi), LBO_A, SBO), dA_lo = make_desc(smem_u32(A_lo), LBO_A, SBO);
  const unsigned long long dB = make_desc(smem_u32(B_t), LBO_B, SBO);

  const int n_chunks = (n + CH - 1) / CH;
  unsigned parity = 0, first = 1;
  for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    if (!first) { mbar_wait(&s_bar, parity); parity ^= 1u; }       // the previous batch of MMAs has read the tiles
    {
      const int pr = t & (CH / 2 - 1), role = t >> 6;
      const int px = chunk * CH + 2 * pr;
      F2 v[5];
#pragma unroll
      for (int q = 0; q < 5; ++q) v[q] = F2(px < n ? vars[(size_t)px * 5 + q] : 0.f, px + 1 < n ? vars[(size_t)(px + 1) * 5 + q] : 0.f);
      // column of this pixel pair inside its 16-byte pixel chunk
#ifdef USE_TF32
      const unsigned colA = (unsigned)(pr >> 1) * (LBO_A * 16) + (unsigned)(pr & 1) * 8;
      const unsigned colB = (unsigned)(pr >> 1) * (LBO_B * 16) + (unsigned)(pr & 1) * 8;
#else
      const unsigned colA = (unsigned)(pr >> 2) * (LBO_A * 16) + (unsigned)(pr & 3) * 4;
      const unsigned colB = (unsigned)(pr >> 2) * (LBO_B * 16) + (unsigned)(pr & 3) * 4;
#endif
      SplitStore st{smem_u32(A_hi) + colA, smem_u32(A_lo) + colA};
      if (role == 0) {
        F2 gg[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) gg[c] = F2(px < n ? g[(size_t)px * 3 + c] : 0.f, px + 1 < n ? g[(size_t)(px + 1) * 3 + c] : 0.f);
        st.template operator()<0>(F2(1.0f));
        st.template operator()<mono_rank(NV, 1, 0, 0, 0, 0)>(v[0]);
        MonoWalk<F2, NV, DEG - 1, 0, 1, 0, 0, 0, 0>::template from<0>(v, v[0], st);
#pragma unroll
        for (int c = 0; c < 3; ++c) split_store(gg[c], smem_u32(B_t) + colB + 16 * c, smem_u32(B_t) + colB + 16 * (3 + c));
      } else {
        MonoWalk<F2, NV, DEG, 0, 0, 0, 0, 0, 0>::template from<1>(v, F2(1.0f), st);
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy stores -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (t == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int k = 0; k < KCH / 2; ++k) {                             // two 16-byte pixel chunks per instruction (K = 16 bf16 / 8 tf32)
        const unsigned long long oa = (unsigned long long)(2 * k * LBO_A), ob = (unsigned long long)(2 * k * LBO_B);   // start-address field, 16-byte units
        mma_ss(tmem, dA_hi + oa, dB + ob, (first && k == 0) ? 0u : 1u);
#ifndef USE_TF32
        mma_ss(tmem, dA_lo + oa, dB + ob, 1u);
#endif
      }
      mma_commit(&s_bar);
    }
    first = 0;
  }
  if (!first) mbar_wait(&s_bar, parity);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (!first) {
    float d[16];
    const unsigned taddr = tmem + (((unsigned)t >> 5) << 21);         // lane quarter of this warp
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;"
        : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3]), "=f"(d[4]), "=f"(d[5]), "=f"(d[6]), "=f"(d[7]), "=f"(d[8]), "=f"(d[9]),
          "=f"(d[10]), "=f"(d[11]), "=f"(d[12]), "=f"(d[13]), "=f"(d[14]), "=f"(d[15])
        : "r"(taddr) : "memory");
    if (t < NC) { atomicAdd(&D[t * 3 + 0], d[0] + d[3]); atomicAdd(&D[t * 3 + 1], d[1] + d[4]); atomicAdd(&D[t * 3 + 2], d[2] + d[5]); }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (t < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS_P) : "memory");
}

// ---- host driver ----------------------------------------------------------------------------------------
// monomials in the reference's order: by degree, lexicographically descending inside a degree
static void ref_monomials(const double* x, std::vector<double>& out) {
  out.clear();
  for (int d = 0; d <= DEG; ++d) {
    int e[5];
    for (e[0] = d; e[0] >= 0; --e[0]) for (e[1] = d - e[0]; e[1] >= 0; --e[1]) for (e[2] = d - e[0] - e[1]; e[2] >= 0; --e[2])
      for (e[3] = d - e[0] - e[1] - e[2]; e[3] >= 0; --e[3]) {
        e[4] = d - e[0] - e[1] - e[2] - e[3];
        double m = 1.0;
        for (int q = 0; q < 5; ++q) for (int r = 0; r < e[q]; ++r) m *= x[q];
        out.push_back(m);
      }
  }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

int main(int argc, char** argv) {
  const int n_big = argc > 1 ? atoi(argv[1]) : 32 * 512 * 512;
  const int n_small = 70001;       // not a multiple of the chunk: exercises the tail
  std::vector<float> hv((size_t)n_big * 5), hg((size_t)n_big * 3);
  unsigned long long s = 88172645463325252ull;
  auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (float)((s >> 11) * (1.0 / 9007199254740992.0)); };
  for (auto& x : hv) x = rnd();
  for (auto& x : hg) x = 2.f * rnd() - 1.f;
  float *dv, *dg, *dD;
  CK(cudaMalloc(&dv, hv.size() * 4)); CK(cudaMalloc(&dg, hg.size() * 4)); CK(cudaMalloc(&dD, NC * 3 * 4));
  CK(cudaMemcpy(dv, hv.data(), hv.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dg, hg.data(), hg.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaFuncSetAttribute(contract_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  CK(cudaFuncSetAttribute(contract_mma, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  printf("dynamic smem per block %d B (%s operands)\n", SMEM_BYTES, MMA_KIND);

  // accuracy on the small case against a double-precision host reference
  CK(cudaMemset(dD, 0, NC * 3 * 4));
  contract_mma<<<sms * 3, BT, SMEM_BYTES>>>(dv, dg, dD, n_small);
  CK(cudaDeviceSynchronize());
  std::vector<float> hD(NC * 3);
  CK(cudaMemcpy(hD.data(), dD, NC * 3 * 4, cudaMemcpyDeviceToHost));
  std::vector<double> ref(NC * 3, 0.0), mono;
  for (int i = 0; i < n_small; ++i) {
    double x[5]; for (int q = 0; q < 5; ++q) x[q] = hv[(size_t)i * 5 + q];
    ref_monomials(x, mono);
    for (int k = 0; k < NC; ++k) for (int c = 0; c < 3; ++c) ref[k * 3 + c] += mono[k] * (double)hg[(size_t)i * 3 + c];
  }
  double mx = 0, err = 0; int worst = 0;
  for (int k = 0; k < NC * 3; ++k) { mx = fmax(mx, fabs(ref[k])); if (fabs(hD[k] - ref[k]) > err) { err = fabs(hD[k] - ref[k]); worst = k; } }
  printf("accuracy (n = %d): max |D - ref| = %.3e, max |ref| = %.3e, relative to max-norm %.3e (worst entry %d: %.6f vs %.6f)\n",
         n_small, err, mx, err / mx, worst, hD[worst], ref[worst]);

  // timing on the big case
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int grid_mult : {3, 6, 12}) {
    for (int it = 0; it < 2; ++it) contract_mma<<<sms * grid_mult, BT, SMEM_BYTES>>>(dv, dg, dD, n_big);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int it = 0; it < 5; ++it) contract_mma<<<sms * grid_mult, BT, SMEM_BYTES>>>(dv, dg, dD, n_big);
    cudaEventRecord(e1); CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    printf("contract_mma  n = %d, grid %d x SMs: %.3f ms per space = %.2f Gpx/s (x3 spaces: %.3f ms)\n", n_big, grid_mult, ms, n_big / ms / 1e6, 3 * ms);
  }
  return 0;
}
