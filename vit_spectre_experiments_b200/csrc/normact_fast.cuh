// Device helpers of the fast bf16 row-wise tails (normact_fast_fwd.cu / normact_fast_bwd.cu): packed fp32x2 arithmetic
// (FFMA2 / FMUL2 / FADD2: two lanes per issue slot -- the generic kernels are issue-bound, profiles/r01_ncu_normact_v2.txt),
// mixed-precision FHADD.BF16 / FHFMA.BF16 (a bf16 operand straight out of a packed register half, fp32 result: no
// unpack instruction), sub-warp / multi-warp row sums.
#pragma once

#include "normact_common.cuh"

namespace sb200 {
namespace na {

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(float lo, float hi) {
  u64 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 splat(float x) { return pk(x, x); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
  u64 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
  u64 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float hsum(u64 v) {
  float lo, hi;
  upk(v, lo, hi);
  return lo + hi;
}
// fp32 = bf16 + fp32 / bf16 * bf16 + fp32 (FHADD.BF16 / FHFMA.BF16)
__device__ __forceinline__ float fhadd(unsigned short h, float c) {
  float r;
  asm("add.rn.f32.bf16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(c));
  return r;
}
__device__ __forceinline__ float fhfma(unsigned short a, unsigned short b, float c) {
  float r;
  asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(r) : "h"(a), "h"(b), "f"(c));
  return r;
}
__device__ __forceinline__ unsigned short lo16(unsigned w) { return static_cast<unsigned short>(w & 0xffffu); }
__device__ __forceinline__ unsigned short hi16(unsigned w) { return static_cast<unsigned short>(w >> 16); }
// packed bf16 pair -> packed fp32 pair (two ALU ops), and back (one F2FP)
__device__ __forceinline__ u64 bf2(unsigned w) { return pk(bf_lo(w), bf_hi(w)); }
__device__ __forceinline__ unsigned bfpack2(u64 v) {
  float lo, hi;
  upk(v, lo, hi);
  return bf_pack(lo, hi);
}
// (bf16 pair + c) as a packed fp32 pair
__device__ __forceinline__ u64 bf2_add(unsigned w, float c) { return pk(fhadd(lo16(w), c), fhadd(hi16(w), c)); }
__device__ __forceinline__ u64 bf2_add(unsigned w, u64 c) {
  float lo, hi;
  upk(c, lo, hi);
  return pk(fhadd(lo16(w), lo), fhadd(hi16(w), hi));
}
__device__ __forceinline__ u64 tanh2(u64 v) {
  float lo, hi;
  upk(v, lo, hi);
  return pk(tanh_approx(lo), tanh_approx(hi));
}
__device__ __forceinline__ u64 ex2_2(u64 v) {
  float lo, hi;
  upk(v, lo, hi);
  return pk(ex2_approx(lo), ex2_approx(hi));
}
__device__ __forceinline__ u64 ldg2(const float* p) {   // two consecutive floats as a packed pair
  const float2 v = __ldg(reinterpret_cast<const float2*>(p));
  return pk(v.x, v.y);
}

// Sum of (a, b) over the TPR threads of a row; every thread of the row gets the totals.
// TPR <= 32: rows are aligned lane groups, shuffles only.  TPR > 32: warp totals meet in shared memory behind a named
// barrier of the row's own warps (ids 1 .. RPC); `slot` alternates between calls, the CTA-wide barrier that ends every
// iteration orders the reuse.
template <int TPR, int RW>
__device__ __forceinline__ void row_sum2(float& a, float& b, float (*red)[2][RW], int slot, int rg) {
  constexpr int kTop = TPR >= 32 ? 16 : TPR / 2;
#pragma unroll
  for (int o = kTop; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  if constexpr (TPR > 32) {
    constexpr int WPR = TPR / 32;
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
      red[slot][0][w] = a;
      red[slot][1][w] = b;
    }
    ptx::named_bar_sync(1 + rg, TPR);   // only the warps of this row meet here; rows of the CTA run on independently
    float sa = red[slot][0][rg * WPR], sb = red[slot][1][rg * WPR];
#pragma unroll
    for (int k = 1; k < WPR; ++k) {
      sa += red[slot][0][rg * WPR + k];
      sb += red[slot][1][rg * WPR + k];
    }
    a = sa;
    b = sb;
  }
}

// Shape of a fast launch: N = 8 * VPT * TPR exactly (no idle lanes, no column predicates)
struct FastShape {
  int vpt, tpr;
};
inline bool fast_shape(int N, bool bwd, FastShape* fs) {
  // forward: 3 vectors per thread where N allows (fewest threads per row); the backward keeps its gamma / beta /
  // column-sum accumulators in registers, 2 vectors per thread bound that
  struct Row { int N, vpt, tpr; };
  static const Row fwd[] = {{384, 3, 16}, {768, 3, 32}, {1536, 3, 64}, {3072, 3, 128},
                            {256, 2, 16}, {512, 2, 32}, {1024, 2, 64}, {2048, 2, 128}};
  static const Row bw[] = {{384, 3, 16}, {768, 3, 32}, {1536, 2, 96}, {3072, 2, 192},
                           {256, 2, 16}, {512, 2, 32}, {1024, 2, 64}, {2048, 2, 128}};
  for (const Row& r : (bwd ? bw : fwd))
    if (r.N == N) {
      fs->vpt = r.vpt;
      fs->tpr = r.tpr;
      return true;
    }
  return false;
}

}  // namespace na
}  // namespace sb200
