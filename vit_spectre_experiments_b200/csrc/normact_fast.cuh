// Device helpers of the fast bf16 row-wise tails (normact_fast_fwd.cu / normact_fast_bwd.cu): packed fp32x2 arithmetic
// (FFMA2 / FMUL2 / FADD2: two lanes per issue slot -- the generic kernels are issue-bound, profiles/r01_ncu_normact_v2.txt),
// mixed-precision FHADD.BF16 / FHFMA.BF16 (a bf16 operand straight out of a packed register half, fp32 result: no
// unpack instruction), sub-warp / multi-warp row sums.
#pragma once

#include <cstdlib>

#include "normact_common.cuh"

namespace sb200 {
namespace na {

typedef unsigned long long u64;

// Ring hand-back.  0 (default): the ring is consumed in groups of two row groups behind one CTA-wide barrier.
// 1 (-DSB200_NA_ASYNC_RING=1): every warp returns a stage through a per-stage "empty" mbarrier as soon as it has read
// it and thread 0 refills opportunistically (try_wait; it only blocks when its own next stage has not been issued yet):
// no CTA-wide barrier inside the row loop, the rows of a CTA drift apart by up to the ring depth.  Measured SLOWER on
// every shape (profiles/r02_kernel_bench_na_v10_ring_ab.txt: N = 1536 backward 1.117 vs 1.011 ms, forward 0.690 vs
// 0.662 ms; LayerNorm forward 0.162 vs 0.154 ms; C4 step 75.6 vs 73.6 ms on the same box) -- as the barrier-free variant
// of the round-1 kernels was: warps in lock step share the stage in L1 / shared memory and poll less.
#ifndef SB200_NA_ASYNC_RING
#define SB200_NA_ASYNC_RING 0
#endif
constexpr bool kAsyncRing = SB200_NA_ASYNC_RING != 0;

// thread 0's refill state for the asynchronous ring
struct RingProducer {
  int next = 0;   // next iteration whose loads have not been issued
  // issue everything that can be issued without blocking, up to S iterations ahead of `it`; the loads of `it` itself
  // are waited for (their stage was last read S iterations ago by every warp: all of them get there without thread 0)
  template <class Issue>
  __device__ __forceinline__ void pump(int it, int n_it, int S, uint64_t* empty, Issue&& issue) {
    while (next < n_it && next < it + S) {
      const int s = next % S, use = next / S;
      if (use > 0) {
        const uint32_t parity = static_cast<uint32_t>(use - 1) & 1u;
        if (next == it) ptx::mbar_wait(&empty[s], parity);
        else if (!ptx::mbar_try_wait(&empty[s], parity)) break;
      }
      issue(next, s);
      ++next;
    }
  }
};

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

// Skip modes that only the fast kernels know (the generic kernels treat these ratios as table-driven "general" pooling):
// AdaptiveAvgPool1d between widths in ratio 3:2, the 512 <-> 768 of the reference's CIFAR configs
// (configs/spectre_vit_cifar100.py:17-19).  Up (N = 3/2 Ns):  out[3k] = s[2k], out[3k+1] = (s[2k] + s[2k+1]) / 2,
// out[3k+2] = s[2k+1];  down (N = 2/3 Ns):  out[2k] = (s[3k] + s[3k+1]) / 2, out[2k+1] = (s[3k+1] + s[3k+2]) / 2.
// The pattern has period 3 resp. 2 in the output, so a thread owns VPT CONSECUTIVE vectors (24 resp. 16 columns facing
// 16 resp. 24 skip columns) and every index below is a compile-time constant.
enum : int { kFastUp32 = 100, kFastDown32 = 101 };
constexpr bool contig_mode(int mode) { return mode == kFastUp32 || mode == kFastDown32; }
constexpr int skip_width(int mode, int N) {
  return mode == kSkipIdentity ? N : mode == kSkipUp4 ? N / 4 : mode == kFastUp32 ? N / 3 * 2 : mode == kFastDown32 ? N / 2 * 3 : 0;
}
// vector owned by thread t as its i-th: strided by the threads of the row, or consecutive for the 3:2 modes
template <int MODE, int VPT, int TPR>
__device__ __forceinline__ int vec_of(int t, int i) { return contig_mode(MODE) ? t * VPT + i : t + i * TPR; }
// pooled skip value facing local output column jl (0 .. 8 VPT) of a thread, from its local skip columns sk[]
template <int MODE>
__device__ __forceinline__ float skip32_val(const float* sk, int jl) {
  if constexpr (MODE == kFastUp32) {
    const int k = jl / 3, m = jl % 3;
    return m == 0 ? sk[2 * k] : m == 2 ? sk[2 * k + 1] : 0.5f * (sk[2 * k] + sk[2 * k + 1]);
  } else {
    const int k = jl / 2, m = jl % 2;
    return 0.5f * (sk[3 * k + m] + sk[3 * k + m + 1]);
  }
}
// transpose: scatter the gradient d of local output column jl into the local skip-gradient accumulators ds[]
template <int MODE>
__device__ __forceinline__ void skip32_scatter(float* ds, int jl, float d) {
  if constexpr (MODE == kFastUp32) {
    const int k = jl / 3, m = jl % 3;
    if (m == 0) ds[2 * k] += d;
    else if (m == 2) ds[2 * k + 1] += d;
    else { ds[2 * k] = fmaf(0.5f, d, ds[2 * k]); ds[2 * k + 1] = fmaf(0.5f, d, ds[2 * k + 1]); }
  } else {
    const int k = jl / 2, m = jl % 2;
    ds[3 * k + m] = fmaf(0.5f, d, ds[3 * k + m]);
    ds[3 * k + m + 1] = fmaf(0.5f, d, ds[3 * k + m + 1]);
  }
}
__device__ __forceinline__ void unpack8(uint4 q, float* v) {
  v[0] = bf_lo(q.x); v[1] = bf_hi(q.x); v[2] = bf_lo(q.y); v[3] = bf_hi(q.y);
  v[4] = bf_lo(q.z); v[5] = bf_hi(q.z); v[6] = bf_lo(q.w); v[7] = bf_hi(q.w);
}
// fast mode of a launch: the generic skip mode, or one of the 3:2 modes when the widths say so; -1: not covered
inline int fast_mode(int skip_mode, int N, int Ns) {
  if (skip_mode == kSkipNone || skip_mode == kSkipIdentity || skip_mode == kSkipUp4) return skip_mode;
  if (skip_mode == kSkipGeneral && 2 * N == 3 * Ns && N % 24 == 0) return kFastUp32;
  if (skip_mode == kSkipGeneral && 3 * N == 2 * Ns && N % 16 == 0) return kFastDown32;
  return -1;
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
  if (bwd && (N == 1536 || N == 3072)) {
    // the wide backward: 3 vectors per thread (384 fat threads, 168 registers) beat 2 (480 threads, 128 registers) --
    // 1.009 vs 1.056 ms at N = 1536, 0.523 vs 0.572 ms at N = 3072 (profiles/r02_kernel_bench_na_v9_ab.txt): the per-row
    // overhead is spread over 24 columns instead of 16.  SB200_NA_BWD_VPT3=0 selects the 2-vector shape.
    static const bool v2 = [] { const char* e = getenv("SB200_NA_BWD_VPT3"); return e && e[0] == '0'; }();
    if (!v2) {
      fs->vpt = 3;
      fs->tpr = N / 24;
      return true;
    }
  }
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
