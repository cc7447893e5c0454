// Shared pieces of the row-wise tail kernels (normact.cu: generic widths / fp32; normact_fast.cu: the bf16
// training stream at exact-fit widths, packed fp32x2 arithmetic): kernel arguments, the shared-memory input ring,
// GELU, the counter-based dropout mask.
#pragma once

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace sb200 {
namespace na {

// skip variants; every kernel instantiation carries exactly one (lean inner loops, small code)
enum : int { kSkipNone = 0, kSkipIdentity = 1, kSkipUp2 = 2, kSkipUp4 = 3, kSkipUp8 = 4, kSkipDown2 = 5, kSkipDown4 = 6,
             kSkipDown8 = 7, kSkipGeneral = 8, kSkipModes = 9 };
constexpr int up_k(int mode) { return mode == kSkipUp2 ? 2 : mode == kSkipUp4 ? 4 : mode == kSkipUp8 ? 8 : 0; }
constexpr int down_k(int mode) { return mode == kSkipDown2 ? 2 : mode == kSkipDown4 ? 4 : mode == kSkipDown8 ? 8 : 0; }

constexpr int kMaxStages = 6;
constexpr int kMaxSmem = 200 * 1024;   // opt-in dynamic shared memory of every kernel here

struct Args {
  const void* in0; const void* in1; const float* gamma; const float* beta; const void* skip;
  void* out; float* mean; float* rstd;
  void* out_pool;             // optional [rows, N / pool_k]: average of every pool_k consecutive columns of `out`
  const void* dout; const void* dout2;   // upstream gradient (+ optional second addend)
  const void* d_out_pool;     // optional [rows, N / pool_k]: gradient that arrived through out_pool
  void* d_in0; void* d_skip; float* part;   // part: [grid][nacc][N] partials: dgamma, dbeta (, column sums of d_in0)
  int nacc;
  long long rows; int N; int Ns; int act; int skip_mode; int pool_k;
  int tpr; int rpc;                                                 // threads per row, rows per CTA
  int stages; int stage_bytes;                                      // shared-memory ring of input row groups
  float eps; float drop_p; unsigned seed_lo; unsigned seed_hi;
  const unsigned long long* seed_dev;   // optional device word added to the seed (CUDA-graph replays: the host seed is frozen)
  int reverse;   // sweep the row groups from the last to the first (the caller knows where the producer left L2 hot)
  // counter-based dropout (drop_mask8): thresholds derived from drop_p on the host
  unsigned drop_thresh;       // p * 2^32: an element is dropped iff its 32-bit draw is below this
  unsigned drop_q32;          // P(a vector of 8 has at least one drop) * 2^32  (prefilter)
  unsigned drop_cum[7];       // P(first dropped index <= k | at least one drop) * 2^32, k = 0..6
  int drop_direct;            // large p: test every element (the prefilter would almost always fail)
  float keep_scale;           // 1 / (1 - p)
};

// host: fill the dropout fields of Args from p
inline void set_dropout(Args& a, float p) {
  a.drop_p = p;
  a.drop_thresh = a.drop_q32 = 0u;
  a.drop_direct = 0;
  a.keep_scale = 1.0f;
  for (unsigned& c : a.drop_cum) c = 0xffffffffu;
  if (!(p > 0.f)) return;
  const double pd = p < 0.999999f ? p : 0.999999;
  auto to_u32 = [](double x) { x *= 4294967296.0; return x >= 4294967295.0 ? 0xffffffffu : static_cast<unsigned>(x); };
  a.drop_thresh = to_u32(pd);
  a.keep_scale = static_cast<float>(1.0 / (1.0 - pd));
  double keep8 = 1.0;
  for (int k = 0; k < 8; ++k) keep8 *= 1.0 - pd;
  const double q = 1.0 - keep8;
  a.drop_q32 = to_u32(q);
  a.drop_direct = q > 0.25 ? 1 : 0;
  double keepk = 1.0;
  for (int k = 0; k < 7; ++k) {
    keepk *= 1.0 - pd;                       // (1 - p)^(k + 1)
    a.drop_cum[k] = to_u32((1.0 - keepk) / q);
  }
}

// ------------------------------------------------------------------ bf16 <-> fp32
__device__ __forceinline__ float bf_lo(unsigned u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(unsigned u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ unsigned bf_pack(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const unsigned*>(&h);
}

// 1-D bulk async copy global -> shared (16-byte aligned, size a multiple of 16), completion on an mbarrier
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(ptx::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(ptx::smem_u32(bar)) : "memory");
}

__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float tanh_approx(float x) {
  float r;
  asm("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// ------------------------------------------------------------------ GELU (erf form, nn.GELU() of layers.py:89)
// fp32 streams: erf(|u|) = 1 - (a1 t + ... + a5 t^5) exp(-u^2),  t = 1 / (1 + p |u|)   (Abramowitz & Stegun 7.1.26,
// |err| < 2e-7).  Returns Phi(x) = 0.5 (1 + erf(x / sqrt 2)) and, through *gauss, exp(-x^2 / 2).
__device__ __forceinline__ float norm_cdf(float x, float* gauss) {
  const float t = rcp_approx(fmaf(0.3275911f * 0.70710678118654752f, fabsf(x), 1.0f));
  const float e = ex2_approx(x * x * -0.72134752044448170f);     // exp(-x^2 / 2)
  float poly = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);   // coefficients pre-halved
  poly = fmaf(poly, t, 0.5f * 1.421413741f);
  poly = fmaf(poly, t, 0.5f * -0.284496736f);
  poly = fmaf(poly, t, 0.5f * 0.254829592f);
  const float half_tail = poly * (t * e);             // 0.5 * erfc(|x| / sqrt 2)
  *gauss = e;
  return x >= 0.f ? 1.0f - half_tail : half_tail;
}
// bf16 streams: Phi(x) = 0.5 (1 + tanh(x (c0 + c1 x^2 + c2 x^4))) -- the odd function atanh(erf(x / sqrt 2)) fitted by a
// minimax polynomial (|x Phi(x) - gelu(x)| < 2.6e-5 over the reals, tools/fit_gelu.py) on the hardware tanh
// (MUFU.TANH, relative error 2^-11): 2^-12 of |gelu|, 16x below the bf16 rounding of the result.
constexpr float kGeluC0 = 0.7975052754f, kGeluC1 = 0.03700801997f, kGeluC2 = -0.0003519023915f;
constexpr float kInvSqrt2Pi = 0.39894228040143268f, kHalfLog2e = 0.72134752044448170f;
__device__ __forceinline__ float norm_cdf_fast(float x, float* gauss) {
  const float x2 = x * x;
  const float t = tanh_approx(x * fmaf(fmaf(kGeluC2, x2, kGeluC1), x2, kGeluC0));
  *gauss = ex2_approx(x2 * -kHalfLog2e);
  return fmaf(0.5f, t, 0.5f);
}
template <bool FAST>
__device__ __forceinline__ float gelu(float x) {
  if constexpr (FAST) {
    const float x2 = x * x, hx = 0.5f * x;
    return fmaf(hx, tanh_approx(x * fmaf(fmaf(kGeluC2, x2, kGeluC1), x2, kGeluC0)), hx);
  } else {
    float g;
    return x * norm_cdf(x, &g);
  }
}
template <bool FAST>
__device__ __forceinline__ float gelu_grad(float x) {
  float g;
  const float cdf = FAST ? norm_cdf_fast(x, &g) : norm_cdf(x, &g);
  return fmaf(x * kInvSqrt2Pi, g, cdf);
}

// ------------------------------------------------------------------ counter-based dropout
// The mask of a vector of 8 consecutive columns is a pure function of (seed, row, vector index); forward and backward
// both call drop_mask8().  Elements are i.i.d. Bernoulli(p), sampled vector-wise so that the common case costs one
// hash and one compare per 8 elements: a first draw decides "no drop in this vector" (probability (1-p)^8); otherwise
// the first dropped index j follows its conditional law (geometric, tabulated in drop_cum) and the elements after j are
// dropped independently with probability p.  Large p (prefilter rarely passes) tests every element directly.
__device__ __forceinline__ unsigned mix32(unsigned x) {
  x ^= x >> 16; x *= 0x7feb352du;
  x ^= x >> 15; x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}
__device__ __forceinline__ unsigned drop_elems(unsigned r, unsigned thresh, int first) {
  constexpr unsigned kMul[8] = {0x9E3779B1u, 0x85EBCA77u, 0xC2B2AE3Du, 0x27D4EB2Fu, 0x165667B1u, 0xD3A2646Du, 0xFD7046C5u, 0xB55A4F09u};
  unsigned m = 0u;
#pragma unroll
  for (int e = 0; e < 8; ++e) m |= (e >= first && r * kMul[e] < thresh) ? (1u << e) : 0u;
  return m;
}
__device__ __forceinline__ unsigned drop_mask8_slow(const Args& a, unsigned vh) {
  const unsigned r = mix32(vh ^ 0x68bc21ebu);
  int j = 0;
#pragma unroll
  for (int k = 0; k < 7; ++k) j += r >= a.drop_cum[k] ? 1 : 0;
  return (1u << j) | drop_elems(mix32(vh + 0x2c1b3c6du), a.drop_thresh, j + 1);
}
// bit e set <=> element e of vector `vec` of the row with seed `rseed` is DROPPED
__device__ __forceinline__ unsigned drop_mask8(const Args& a, unsigned rseed, unsigned vec) {
  unsigned x = vec * 0x9E3779B1u + rseed;
  x ^= x >> 15;
  x *= 0x85EBCA77u;
  x ^= x >> 13;
  if (a.drop_direct) return drop_elems(mix32(x), a.drop_thresh, 0);
  if (x >= a.drop_q32) return 0u;
  return drop_mask8_slow(a, x);
}
// call seed (+ the optional device-resident word: a captured graph bumps it between replays), read once per thread
struct Seed {
  unsigned lo, hi;
};
__device__ __forceinline__ Seed load_seed(const Args& a) {
  unsigned long long s = static_cast<unsigned long long>(a.seed_hi) << 32 | a.seed_lo;
  if (a.seed_dev) s += __ldg(a.seed_dev) * 0x9E3779B97F4A7C15ull;
  return {static_cast<unsigned>(s), static_cast<unsigned>(s >> 32)};
}
// seed of a row (one 32-bit mix of the row index under the 64-bit call seed; drop_mask8 mixes the vector index in)
__device__ __forceinline__ unsigned row_seed(const Seed& sd, long long row) {
  unsigned x = static_cast<unsigned>(row) * 0x9E3779B1u + sd.lo;
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  return x + (sd.hi ^ static_cast<unsigned>(row >> 32)) * 0x846ca68bu;
}

// ------------------------------------------------------------------ fast path (normact_fast.cu)
// Returns true if it launched (or failed: *rc != 0); false if the shape is not covered and the generic kernels must run.
bool fast_fwd_launch(const Args& a, cudaStream_t st, int* rc);
bool fast_bwd_launch(const Args& a, size_t workspace_bytes, cudaStream_t st, int* rc, int* grid_out);
int fast_bwd_grid_bound();   // upper bound of the backward grid (sizes the partial-sum workspace)

}  // namespace na
}  // namespace sb200
