// Fused row-wise tails around the mixer and the SpectreLinear GEMMs:
//
//   out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + pool(skip) )
//
// which is, with the right operands,
//   x1 = LN1(mix(x)) + x                      spectre_vit/models/spectre/spectre.py:66       (in0 = mix, skip = x)
//   a  = GELU(LN(Linear(x))) + avg_pool(x)    spectre_vit/models/spectre/layers.py:95-101    (in0 = Linear(x), skip = x)
//   x2 = LN2(x1 + ff)                         spectre.py:67                                  (in0 = x1, in1 = ff)
// and its backward (d_in0 = d_in1, d_skip, dgamma, dbeta) in ONE pass over the rows each way, instead of
// the 6-10 elementwise / reduction launches of the stock composition (LayerNorm backward's gamma/beta
// reduction and the adaptive-pool kernels were 50 % of the ViT-S/16 step: profiles/r01_step_profile_c4.txt).
//
// Every activation tensor of a call has ONE element type T (float32, or bfloat16 for the bf16 training
// stream); statistics, gamma/beta and all arithmetic are fp32.
//
// Mapping: a row is owned by TPR threads (a whole number of warps); a thread owns VPT (1-3) vectors of 8
// consecutive columns (16-byte bf16 / 2x16-byte fp32 accesses, coalesced).  A CTA of up to 256 threads
// carries RPC = 256 / TPR rows at a time (narrow rows: one warp per row, no barrier inside a row) and strides
// over the row groups (ascending, or descending with `reverse`).  LayerNorm statistics take one pass: sum and
// sum of squares about a pivot from the row itself.  The backward keeps the per-column gamma / beta /
// column-sum accumulators of every row group in shared memory (a thread only touches its own columns) and
// folds them in a fixed order into ONE partial per CTA, reduced by reduce_gamma_beta_kernel: deterministic.
// Loads are decoupled from the math: the rows of a group are contiguous in every operand, so ONE thread
// issues one bulk async copy (cp.async.bulk, the 1-D TMA path) per operand per group into a 2-4 stage
// shared-memory ring, completion on an mbarrier -- tens of KB in flight per SM at any time without
// holding registers for them.  Results go straight from registers to HBM with 16-byte stores.
// The kernels are instruction-bound next to the GELU (erf) and the dropout mask, hence: erf by the
// Abramowitz-Stegun 7.1.26 rational (|err| < 2e-7, one exp shared with the Gaussian of the derivative),
// one odd multiply of a per-vector hash per element for the mask, pooling ratios resolved at compile time
// (general widths: window tables built once per CTA in shared memory, transpose fused into the backward).
#include <cstdarg>
#include <type_traits>
#include <utility>

#include "normact_common.cuh"

namespace sb200 {
namespace na {

__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 q = *reinterpret_cast<const uint4*>(p);
  v[0] = bf_lo(q.x); v[1] = bf_hi(q.x); v[2] = bf_lo(q.y); v[3] = bf_hi(q.y);
  v[4] = bf_lo(q.z); v[5] = bf_hi(q.z); v[6] = bf_lo(q.w); v[7] = bf_hi(q.w);
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) =
      make_uint4(bf_pack(v[0], v[1]), bf_pack(v[2], v[3]), bf_pack(v[4], v[5]), bf_pack(v[6], v[7]));
}
// n in {1, 2, 4} consecutive elements (n * sizeof(T) aligned)
template <int n>
__device__ __forceinline__ void ldn(const float* p, float (&v)[n]) {
  if constexpr (n == 4) { const float4 a = *reinterpret_cast<const float4*>(p); v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
  else if constexpr (n == 2) { const float2 a = *reinterpret_cast<const float2*>(p); v[0] = a.x; v[1] = a.y; }
  else v[0] = *p;
}
template <int n>
__device__ __forceinline__ void ldn(const __nv_bfloat16* p, float (&v)[n]) {
  if constexpr (n == 4) { const uint2 a = *reinterpret_cast<const uint2*>(p); v[0] = bf_lo(a.x); v[1] = bf_hi(a.x); v[2] = bf_lo(a.y); v[3] = bf_hi(a.y); }
  else if constexpr (n == 2) { const unsigned a = *reinterpret_cast<const unsigned*>(p); v[0] = bf_lo(a); v[1] = bf_hi(a); }
  else v[0] = __bfloat162float(*p);
}
template <int n>
__device__ __forceinline__ void stn(float* p, const float (&v)[n]) {
  if constexpr (n == 4) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  else if constexpr (n == 2) *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
  else *p = v[0];
}
template <int n>
__device__ __forceinline__ void stn(__nv_bfloat16* p, const float (&v)[n]) {
  if constexpr (n == 4) *reinterpret_cast<uint2*>(p) = make_uint2(bf_pack(v[0], v[1]), bf_pack(v[2], v[3]));
  else if constexpr (n == 2) *reinterpret_cast<unsigned*>(p) = bf_pack(v[0], v[1]);
  else *p = __float2bfloat16_rn(v[0]);
}
__device__ __forceinline__ float ld1(const float* p) { return *p; }   // also reads the shared-memory ring: no __ldg
__device__ __forceinline__ float ld1(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void st1(float* p, float v) { *p = v; }
__device__ __forceinline__ void st1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// sum over the threads of one row of NV values.  One warp per row: shuffles only.  Several warps per row:
// one CTA-wide barrier (every row group of the CTA executes the same trip count); `slot` alternates so
// consecutive calls need one barrier each.
template <int NV>
__device__ __forceinline__ void row_sum(float (&v)[NV], float (*red)[NV][8], int slot, int wpr, int rg, int tpr = 32) {
  // narrow rows (tpr = 8 or 16 threads: several rows per warp) only fold the lanes of their own aligned group
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    if (o < tpr) {
#pragma unroll
      for (int i = 0; i < NV; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
    }
  if (wpr > 1) {
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
#pragma unroll
      for (int i = 0; i < NV; ++i) red[slot][i][w] = v[i];
    __syncthreads();
    // wpr <= 8 partials per row: lane (l & 7) fetches one, three butterfly steps leave the sum in every lane
    // (branch-free; a runtime-length loop here compiled to ~60 instructions of remainder handling)
    const int k = lane & 7;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float s = k < wpr ? red[slot][i][rg * wpr + k] : 0.f;
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      v[i] = s;
    }
  }
}

// adaptive average pool Ns -> N (torch semantics), general widths.  Output column j averages the skip columns
// [floor(j Ns / N), ceil((j+1) Ns / N)); skip column c feeds exactly the outputs [floor(c N / Ns), ceil((c+1) N / Ns)).
// Both maps are tabulated once per CTA in shared memory (first | count << 16), so the row loop has no divisions.
__device__ __forceinline__ void pool_windows(unsigned* tab, int n_out, int n_in) {
  for (int j = threadIdx.x; j < n_out; j += blockDim.x) {
    const int s = static_cast<int>((static_cast<long long>(j) * n_in) / n_out);
    const int e = static_cast<int>((static_cast<long long>(j + 1) * n_in + n_out - 1) / n_out);
    tab[j] = static_cast<unsigned>(s) | static_cast<unsigned>(e - s) << 16;
  }
}
// o[0..8) += pool(skip)[j0 .. j0+8)
// `short_win`: every window has one or two elements (N >= Ns): branch-free, (first + last) / 2
template <class T>
__device__ __forceinline__ void skip_general_add(const T* srow, const unsigned* win, int j0, bool short_win, float (&o)[8]) {
  const uint4 wa = *reinterpret_cast<const uint4*>(win + j0), wb = *reinterpret_cast<const uint4*>(win + j0 + 4);
  const unsigned w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int s = static_cast<int>(w[e] & 0xffffu), n = static_cast<int>(w[e] >> 16);
    if (short_win) {
      o[e] = fmaf(0.5f, ld1(srow + s) + ld1(srow + s + n - 1), o[e]);
    } else {
      float acc = ld1(srow + s);
      for (int k = 1; k < n; ++k) acc += ld1(srow + s + k);
      o[e] = fmaf(acc, rcp_approx(static_cast<float>(n)), o[e]);
    }
  }
}
// d_skip[c0 .. c0+8) of one row from the row's window-normalised gradients dn[j] = d[j] / len(j) (shared memory)
// `short_cov`: every skip column feeds one or two outputs (N <= Ns): branch-free
template <class T>
__device__ __forceinline__ void skip_general_bwd(T* drow, const float* dn, const unsigned* cov, int c0, bool short_cov) {
  const uint4 wa = *reinterpret_cast<const uint4*>(cov + c0), wb = *reinterpret_cast<const uint4*>(cov + c0 + 4);
  const unsigned w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
  float g[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int s = static_cast<int>(w[e] & 0xffffu), n = static_cast<int>(w[e] >> 16);
    float acc = dn[s];
    if (short_cov) {
      const float b = dn[s + n - 1];
      acc = n == 2 ? acc + b : acc;
    } else {
      for (int k = 1; k < n; ++k) acc += dn[s + k];
    }
    g[e] = acc;
  }
  st8(drow + c0, g);
}

// the same, one skip column per lane: for N > Ns (a skip column feeds ceil(N / Ns) + 1 outputs) the vector form above
// runs 8 runtime-length loops per lane on Ns / 8 lanes of the row -- 43 % of the executed instructions of the C2 backward
// (N = 256, Ns = 48: 6 busy lanes of 32, profiles/r02_ncu_src_c2_v13.txt); here every lane of the row takes columns
template <class T>
__device__ __forceinline__ void skip_general_bwd_cols(T* drow, const float* dn, const unsigned* cov, int t, int tpr, int Ns) {
  for (int c = t; c < Ns; c += tpr) {
    const unsigned w = cov[c];
    const int s = static_cast<int>(w & 0xffffu), n = static_cast<int>(w >> 16);
    float acc = dn[s];
    for (int k = 1; k < n; ++k) acc += dn[s + k];
    st1(drow + c, acc);
  }
}

// o[0..8) += pool(skip)[j0 .. j0+8) for integer ratios resolved at compile time
template <class T, int K>
__device__ __forceinline__ void skip_up_add(const T* srow, int j0, float (&o)[8]) {
  float s[8 / K];
  ldn<8 / K>(srow + j0 / K, s);
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] += s[e / K];
}
template <class T, int K>
__device__ __forceinline__ void skip_down_add(const T* srow, int j0, float (&o)[8]) {
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll
  for (int v = 0; v < K; ++v) {
    float s[8];
    ld8(srow + static_cast<long long>(j0) * K + v * 8, s);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[(v * 8 + i) / K] += s[i];
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = fmaf(acc[e], 1.0f / K, o[e]);
}
// transposes: d_skip from the (masked) upstream gradient d[0..8) of columns j0 .. j0+8
template <class T, int K>
__device__ __forceinline__ void skip_up_bwd(T* drow, int j0, const float (&d)[8]) {
  float s[8 / K];
#pragma unroll
  for (int i = 0; i < 8 / K; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < K; ++r) acc += d[i * K + r];
    s[i] = acc;
  }
  stn<8 / K>(drow + j0 / K, s);
}
template <class T, int K>
__device__ __forceinline__ void skip_down_bwd(T* drow, int j0, const float (&d)[8]) {
#pragma unroll
  for (int v = 0; v < K; ++v) {
    float s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = d[(v * 8 + i) / K] * (1.0f / K);
    st8(drow + static_cast<long long>(j0) * K + v * 8, s);
  }
}

// second output: pooled[j0 / K + i] = mean(o[i K .. (i + 1) K)), K in {2, 4, 8}; and its transpose in the backward
template <class T, int K>
__device__ __forceinline__ void pool_store_k(T* prow, int j0, const float (&o)[8]) {
  float s[8 / K];
#pragma unroll
  for (int i = 0; i < 8 / K; ++i) {
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < K; ++r) acc += o[i * K + r];
    s[i] = acc * (1.0f / K);
  }
  stn<8 / K>(prow + j0 / K, s);
}
template <class T>
__device__ __forceinline__ void pool_store(T* prow, int j0, int k, const float (&o)[8]) {
  if (k == 4) pool_store_k<T, 4>(prow, j0, o);
  else if (k == 2) pool_store_k<T, 2>(prow, j0, o);
  else pool_store_k<T, 8>(prow, j0, o);
}
template <class T, int K>
__device__ __forceinline__ void pool_grad_add_k(const T* prow, int j0, float (&d)[8]) {
  float s[8 / K];
  ldn<8 / K>(prow + j0 / K, s);
#pragma unroll
  for (int e = 0; e < 8; ++e) d[e] = fmaf(s[e / K], 1.0f / K, d[e]);
}
template <class T>
__device__ __forceinline__ void pool_grad_add(const T* prow, int j0, int k, float (&d)[8]) {
  if (k == 4) pool_grad_add_k<T, 4>(prow, j0, d);
  else if (k == 2) pool_grad_add_k<T, 2>(prow, j0, d);
  else pool_grad_add_k<T, 8>(prow, j0, d);
}

// ------------------------------------------------------------------ forward
template <class T, int VPT, int MODE, int ACT>
__global__ void __launch_bounds__(256, VPT == 3 ? 2 : 3) norm_act_fwd_kernel(const Args a) {
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t full[kMaxStages];
  __shared__ float red[2][2][8];
  const int tpr = a.tpr, wpr = tpr >> 5;
  const int rg = threadIdx.x / tpr, t = threadIdx.x - rg * tpr;
  const int nvec = a.N >> 3, S = a.stages;
  const float inv_n = 1.0f / static_cast<float>(a.N);
  const bool drop = a.drop_p > 0.f;
  const Seed sd = drop ? load_seed(a) : Seed{0u, 0u};
  const float keep_scale = a.keep_scale;
  constexpr bool kFastGelu = std::is_same<T, __nv_bfloat16>::value;   // bf16 stream: tanh-form Phi (normact_common.cuh)
  const long long groups = (a.rows + a.rpc - 1) / a.rpc;
  const int n_it = static_cast<int>((groups - blockIdx.x + gridDim.x - 1) / gridDim.x);
  auto group_of = [&](int it) {   // row group of iteration it: ascending, or descending from the last group
    const long long g = blockIdx.x + static_cast<long long>(it) * gridDim.x;
    return a.reverse ? groups - 1 - g : g;
  };
  const int off_in1 = a.rpc * a.N * static_cast<int>(sizeof(T));
  const int off_skip = off_in1 * (a.in1 ? 2 : 1);
  unsigned* win = reinterpret_cast<unsigned*>(ring + S * a.stage_bytes);   // general widths: pooling windows [N]
  if constexpr (MODE == kSkipGeneral) pool_windows(win, a.N, a.Ns);

  auto issue = [&](int it, int stg) {   // one thread: the group's rows are contiguous in every operand
    const long long row0 = group_of(it) * a.rpc;
    const long long left = a.rows - row0;
    const uint32_t nr = static_cast<uint32_t>(left < a.rpc ? left : a.rpc);
    uint8_t* st = ring + stg * a.stage_bytes;
    uint64_t* bar = &full[stg];
    const uint32_t bmain = nr * a.N * sizeof(T), bskip = MODE != kSkipNone ? nr * a.Ns * static_cast<uint32_t>(sizeof(T)) : 0u;
    ptx::mbar_arrive_expect_tx(bar, bmain * (a.in1 ? 2u : 1u) + bskip);
    bulk_load(st, static_cast<const T*>(a.in0) + row0 * a.N, bmain, bar);
    if (a.in1) bulk_load(st + off_in1, static_cast<const T*>(a.in1) + row0 * a.N, bmain, bar);
    if (MODE != kSkipNone) bulk_load(st + off_skip, static_cast<const T*>(a.skip) + row0 * a.Ns, bskip, bar);
  };
  if (threadIdx.x == 0) {
    for (int k = 0; k < S; ++k) ptx::mbar_init(&full[k], 1);
    ptx::mbar_fence_init();
    for (int it = 0; it < S - 1 && it < n_it; ++it) issue(it, it);
  }
  __syncthreads();

  int stg = 0;
  uint32_t phase = 0;
  for (int it = 0; it < n_it; ++it) {
    if (threadIdx.x == 0 && it + S - 1 < n_it) issue(it + S - 1, stg == 0 ? S - 1 : stg - 1);   // the stage freed by it-1
    ptx::mbar_wait(&full[stg], phase);
    const long long row = group_of(it) * a.rpc + rg;
    const bool live = row < a.rows;
    const uint8_t* st = ring + stg * a.stage_bytes;
    const T* r0 = reinterpret_cast<const T*>(st) + rg * a.N;
    float z[VPT][8];
    // one pass: sum and sum of squares about a pivot taken from the row itself (its first element), which keeps
    // the E[d^2] - E[d]^2 form free of cancellation; one cross-thread reduction instead of two
    float pivot = 0.f;
    if (live) {
      pivot = ld1(r0);
      if (a.in1) pivot += ld1(reinterpret_cast<const T*>(st + off_in1) + rg * a.N);
    }
    float s[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * tpr;
      if (live && v < nvec) {
        ld8(r0 + v * 8, z[i]);
        if (a.in1) {
          float w[8];
          ld8(reinterpret_cast<const T*>(st + off_in1) + rg * a.N + v * 8, w);
#pragma unroll
          for (int e = 0; e < 8; ++e) z[i][e] += w[e];
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float d = z[i][e] - pivot;
          s[0] += d;
          s[1] = fmaf(d, d, s[1]);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) z[i][e] = 0.f;
      }
    }
    row_sum<2>(s, red, it & 1, wpr, rg, tpr);
    const float dmean = s[0] * inv_n;
    const float mean = pivot + dmean;
    const float var = fmaxf(fmaf(-dmean, dmean, s[1] * inv_n), 0.f);
    if (live) {
      const float rstd = rsqrtf(var + a.eps);
      if (t == 0) {
        a.mean[row] = mean;
        a.rstd[row] = rstd;
      }
      const float shift = -mean * rstd;
      const T* srow = reinterpret_cast<const T*>(st + off_skip) + rg * a.Ns;
      const unsigned rseed = drop ? row_seed(sd, row) : 0u;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = t + i * tpr;
        if (v >= nvec) continue;
        const int j0 = v * 8;
        float g[8], b[8], o[8];
        ld8(a.gamma + j0, g);
        ld8(a.beta + j0, b);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float u = fmaf(fmaf(z[i][e], rstd, shift), g[e], b[e]);
          o[e] = ACT ? gelu<kFastGelu>(u) : u;
        }
        if constexpr (MODE == kSkipIdentity) {
          float sk[8];
          ld8(srow + j0, sk);
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] += sk[e];
        } else if constexpr (up_k(MODE) > 0) {
          skip_up_add<T, up_k(MODE)>(srow, j0, o);
        } else if constexpr (down_k(MODE) > 0) {
          skip_down_add<T, down_k(MODE)>(srow, j0, o);
        } else if constexpr (MODE == kSkipGeneral) {
          skip_general_add<T>(srow, win, j0, a.N >= a.Ns, o);
        }
        if (drop) {
          const unsigned dm = drop_mask8(a, rseed, static_cast<unsigned>(v));
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = (dm >> e & 1u) ? 0.f : o[e] * keep_scale;
        }
        st8(static_cast<T*>(a.out) + row * a.N + j0, o);
        if (a.out_pool) pool_store<T>(static_cast<T*>(a.out_pool) + row * (a.N / a.pool_k), j0, a.pool_k, o);
      }
    }
    __syncthreads();   // every thread is done with this stage: the producer may refill it
    if (++stg == S) { stg = 0; phase ^= 1u; }
  }
}

// ------------------------------------------------------------------ backward
// d = dout * dropout_mask ; d_skip = pool^T(d) ; g = d * act'(u) ; LayerNorm backward of g.
template <class T, int VPT, int MODE, int ACT>
__global__ void __launch_bounds__(256, 3) norm_act_bwd_kernel(const Args a) {
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t full[kMaxStages];
  __shared__ float red[2][2][8];
  const int tpr = a.tpr, wpr = tpr >> 5;
  const int rg = threadIdx.x / tpr, t = threadIdx.x - rg * tpr;
  const int nvec = a.N >> 3, S = a.stages;
  // per-row-group gamma/beta (/ column-sum) accumulators [rpc][nacc][N] (a thread only ever touches its own columns
  // of its own row group: no conflicts, and 32 registers less than keeping them in the register file)
  float* accs = reinterpret_cast<float*>(ring + S * a.stage_bytes);
  const int nacc = a.nacc;                                            // 2, or 3 with the column sums of d_in0
  float* sgam = accs + a.rpc * nacc * a.N;                            // [N] gamma, [N] beta
  float* myacc = accs + rg * nacc * a.N;
  // general skip widths: 1 / len(j) per output column [N], covering outputs per skip column [Ns], and the
  // row group's window-normalised gradients [rpc][N]
  float* sinv = sgam + 2 * a.N;
  unsigned* cov = reinterpret_cast<unsigned*>(sinv + a.N);
  float* dn = reinterpret_cast<float*>(cov + a.Ns) + rg * a.N;
  const float inv_n = 1.0f / static_cast<float>(a.N);
  const bool drop = a.drop_p > 0.f;
  const Seed sd = drop ? load_seed(a) : Seed{0u, 0u};
  const float keep_scale = a.keep_scale;
  constexpr bool kFastGelu = std::is_same<T, __nv_bfloat16>::value;   // bf16 stream: tanh-form Phi (normact_common.cuh)
  const long long groups = (a.rows + a.rpc - 1) / a.rpc;
  const int n_it = static_cast<int>((groups - blockIdx.x + gridDim.x - 1) / gridDim.x);
  auto group_of = [&](int it) {
    const long long g = blockIdx.x + static_cast<long long>(it) * gridDim.x;
    return a.reverse ? groups - 1 - g : g;
  };
  const int off_in0 = a.rpc * a.N * static_cast<int>(sizeof(T)), off_in1 = 2 * off_in0;
  const int off_d2 = off_in0 * (a.in1 ? 3 : 2);

  auto issue = [&](int it, int stg) {   // dout, in0 (, in1) of one row group: contiguous, one bulk copy each
    const long long row0 = group_of(it) * a.rpc;
    const long long left = a.rows - row0;
    const uint32_t nr = static_cast<uint32_t>(left < a.rpc ? left : a.rpc);
    uint8_t* st = ring + stg * a.stage_bytes;
    uint64_t* bar = &full[stg];
    const uint32_t bmain = nr * a.N * sizeof(T);
    ptx::mbar_arrive_expect_tx(bar, bmain * (2u + (a.in1 ? 1u : 0u) + (a.dout2 ? 1u : 0u)));
    bulk_load(st, static_cast<const T*>(a.dout) + row0 * a.N, bmain, bar);
    bulk_load(st + off_in0, static_cast<const T*>(a.in0) + row0 * a.N, bmain, bar);
    if (a.in1) bulk_load(st + off_in1, static_cast<const T*>(a.in1) + row0 * a.N, bmain, bar);
    if (a.dout2) bulk_load(st + off_d2, static_cast<const T*>(a.dout2) + row0 * a.N, bmain, bar);
  };
  if (threadIdx.x == 0) {
    for (int k = 0; k < S; ++k) ptx::mbar_init(&full[k], 1);
    ptx::mbar_fence_init();
    for (int it = 0; it < S - 1 && it < n_it; ++it) issue(it, it);
  }
  for (int i = threadIdx.x; i < a.N; i += blockDim.x) {
    sgam[i] = a.gamma[i];
    sgam[a.N + i] = a.beta[i];
  }
  for (int i = threadIdx.x; i < a.rpc * nacc * a.N; i += blockDim.x) accs[i] = 0.f;
  if constexpr (MODE == kSkipGeneral) {
    pool_windows(cov, a.Ns, a.N);
    for (int j = threadIdx.x; j < a.N; j += blockDim.x) {
      const int s0 = static_cast<int>((static_cast<long long>(j) * a.Ns) / a.N);
      const int e0 = static_cast<int>((static_cast<long long>(j + 1) * a.Ns + a.N - 1) / a.N);
      sinv[j] = 1.0f / static_cast<float>(e0 - s0);
    }
  }
  __syncthreads();

  // row statistics of the first row (prefetched one iteration ahead afterwards)
  float mean_n = 0.f, rstd_n = 0.f;
  {
    const long long row = group_of(0) * a.rpc + rg;
    if (n_it > 0 && row < a.rows) { mean_n = a.mean[row]; rstd_n = a.rstd[row]; }
  }
  int stg = 0;
  uint32_t phase = 0;
  for (int it = 0; it < n_it; ++it) {
    if (threadIdx.x == 0 && it + S - 1 < n_it) issue(it + S - 1, stg == 0 ? S - 1 : stg - 1);
    const long long row = group_of(it) * a.rpc + rg;
    const bool live = row < a.rows;
    const float mean = mean_n, rstd = rstd_n;
    {
      const long long rown = group_of(it + 1) * a.rpc + rg;
      if (it + 1 < n_it && rown < a.rows) { mean_n = a.mean[rown]; rstd_n = a.rstd[rown]; }
    }
    ptx::mbar_wait(&full[stg], phase);
    const uint8_t* st = ring + stg * a.stage_bytes;
    const float shift = -mean * rstd;
    const unsigned rseed = drop ? row_seed(sd, row) : 0u;
    float xh[VPT][8], gy[VPT][8];
    float s[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * tpr;
      if (!live || v >= nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) xh[i][e] = gy[i][e] = 0.f;
        continue;
      }
      const int j0 = v * 8;
      float d[8], z[8];
      ld8(reinterpret_cast<const T*>(st) + rg * a.N + j0, d);
      ld8(reinterpret_cast<const T*>(st + off_in0) + rg * a.N + j0, z);
      if (a.in1) {
        float w[8];
        ld8(reinterpret_cast<const T*>(st + off_in1) + rg * a.N + j0, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) z[e] += w[e];
      }
      if (a.dout2) {
        float w[8];
        ld8(reinterpret_cast<const T*>(st + off_d2) + rg * a.N + j0, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) d[e] += w[e];
      }
      if (a.d_out_pool) pool_grad_add<T>(static_cast<const T*>(a.d_out_pool) + row * (a.N / a.pool_k), j0, a.pool_k, d);
      if (drop) {
        const unsigned dm = drop_mask8(a, rseed, static_cast<unsigned>(v));
#pragma unroll
        for (int e = 0; e < 8; ++e) d[e] = (dm >> e & 1u) ? 0.f : d[e] * keep_scale;
      }
      // ---- skip path: transpose of the pooling
      if constexpr (MODE == kSkipIdentity) {
        st8(static_cast<T*>(a.d_skip) + row * a.Ns + j0, d);
      } else if constexpr (up_k(MODE) > 0) {
        skip_up_bwd<T, up_k(MODE)>(static_cast<T*>(a.d_skip) + row * a.Ns, j0, d);
      } else if constexpr (down_k(MODE) > 0) {
        skip_down_bwd<T, down_k(MODE)>(static_cast<T*>(a.d_skip) + row * a.Ns, j0, d);
      } else if constexpr (MODE == kSkipGeneral) {
        float w[8], q[8];
        ld8(sinv + j0, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) q[e] = d[e] * w[e];
        st8(dn + j0, q);     // gathered per skip column after the row sums' barrier
      }
      // ---- main path
      float gam[8], bet[8], dg[8], db[8];
      ld8(sgam + j0, gam);
      if (ACT) ld8(sgam + a.N + j0, bet);
      ld8(myacc + j0, dg);
      ld8(myacc + a.N + j0, db);
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x = fmaf(z[e], rstd, shift);
        float g = d[e];
        if (ACT) g *= gelu_grad<kFastGelu>(fmaf(x, gam[e], bet[e]));
        db[e] += g;
        dg[e] = fmaf(g, x, dg[e]);
        const float y = g * gam[e];
        xh[i][e] = x;
        gy[i][e] = y;
        if (e & 1) { s1 = fmaf(y, x, s1); s[0] += y; } else { s[1] = fmaf(y, x, s[1]); s0 += y; }
      }
      s[0] += s0;
      s[1] += s1;
      st8(myacc + j0, dg);
      st8(myacc + a.N + j0, db);
    }
    row_sum<2>(s, red, it & 1, wpr, rg, tpr);
    if constexpr (MODE == kSkipGeneral) {
      if (wpr <= 1) __syncwarp();   // at most one warp per row: dn is written and read inside the same warp
      if (live) {
        if (a.N <= a.Ns) {
          for (int u = t; u < (a.Ns >> 3); u += tpr) skip_general_bwd<T>(static_cast<T*>(a.d_skip) + row * a.Ns, dn, cov, u * 8, true);
        } else {
          skip_general_bwd_cols<T>(static_cast<T*>(a.d_skip) + row * a.Ns, dn, cov, t, tpr, a.Ns);
        }
      }
    }
    if (live) {
      const float m1 = s[0] * inv_n, m2 = s[1] * inv_n;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        const int v = t + i * tpr;
        if (v >= nvec) continue;
        float dz[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) dz[e] = rstd * (gy[i][e] - m1 - xh[i][e] * m2);
        st8(static_cast<T*>(a.d_in0) + row * a.N + v * 8, dz);
        if (nacc == 3) {   // column sums of d_in0: the bias gradient of the Linear that produced in0
          float cs[8];
          ld8(myacc + 2 * a.N + v * 8, cs);
#pragma unroll
          for (int e = 0; e < 8; ++e) cs[e] += dz[e];
          st8(myacc + 2 * a.N + v * 8, cs);
        }
      }
    }
    __syncthreads();   // stage consumed
    if (++stg == S) { stg = 0; phase ^= 1u; }
  }
  // fold the row groups of the CTA in a fixed order: one partial [2][N] per CTA  (the loop's last barrier
  // already ordered every accumulator update)
  float* pg = a.part + static_cast<size_t>(blockIdx.x) * nacc * a.N;
  for (int c = threadIdx.x; c < nacc * a.N; c += blockDim.x) {
    float v = accs[c];
    for (int r = 1; r < a.rpc; ++r) v += accs[r * nacc * a.N + c];
    pg[c] = v;
  }
}

// out_k[c] = sum_s part[s][k][c] for k < nacc (dgamma, dbeta, column sums).  CTA = 8 columns x 32 slab lanes (the
// partials are few MB but the reduction is latency-bound: many short threads); lane sums are combined in a fixed
// order (deterministic).
__global__ void __launch_bounds__(256) reduce_gamma_beta_kernel(const float* __restrict__ part, float* __restrict__ dgamma,
                                                                float* __restrict__ dbeta, float* __restrict__ dsum, int N,
                                                                int nacc, int slabs) {
  __shared__ float sm[32][9];
  pdl_launch_dependents();
  pdl_wait();
  const int c8 = threadIdx.x & 7, ln = threadIdx.x >> 3;
  const int col = blockIdx.x * 8 + c8;
  const int ncol = nacc * N;
  float tot = 0.f;
  if (col < ncol) {
    // up to 8 independent loads in flight per thread (a 32-byte sector per slab row and 8 columns)
    float acc[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc[u] = 0.f;
    for (int k = ln; k < slabs; k += 256) {
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (k + 32 * u < slabs) acc[u] += part[static_cast<size_t>(k + 32 * u) * ncol + col];
    }
    tot = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
  }
  sm[ln][c8] = tot;
  __syncthreads();
  if (ln == 0 && col < ncol) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) s += sm[k][c8];
    if (col < N) dgamma[col] = s;
    else if (col < 2 * N) dbeta[col - N] = s;
    else dsum[col - 2 * N] = s;
  }
}

// ------------------------------------------------------------------ host side
// SB200_NA_FAST=0 keeps the bf16 stream on the generic kernels (A/B measurements); default: normact_fast.cu
static bool fast_enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_NA_FAST");
    return !(e && e[0] == '0');
  }();
  return on;
}

static int pick_skip(int N, int Ns, bool has_skip) {
  if (!has_skip) return kSkipNone;
  if (Ns == N) return kSkipIdentity;
  const int up = N % Ns == 0 ? N / Ns : 0, down = Ns % N == 0 ? Ns / N : 0;
  if (up == 2) return kSkipUp2;
  if (up == 4) return kSkipUp4;
  if (up == 8) return kSkipUp8;
  if (down == 2) return kSkipDown2;
  if (down == 4) return kSkipDown4;
  if (down == 8) return kSkipDown8;
  return kSkipGeneral;
}

// threads per row (whole warps), rows per CTA, vectors per thread, shared-memory ring
struct Shape {
  int vpt, tpr, rpc, stages, stage_bytes;
  size_t smem;   // dynamic shared memory of the launch
};
static Shape launch_shape(int N, int Ns, bool has_in1, bool has_skip, int elem, bool bwd, bool has_dout2 = false,
                          int nacc = 2, bool general = false) {
  Shape sh{};
  const int nvec = N / 8;
  if (nvec <= 32) {   // one vector per thread; rows narrower than a warp share it (N = 48: 8 threads per row, 4 rows per warp)
    sh.vpt = 1;
    sh.tpr = nvec <= 8 ? 8 : nvec <= 16 ? 16 : 32;
  }
  else {
    // 2 or 3 vectors per thread: fewest idle lanes first, then the larger count (fewer warps per row: cheaper
    // row sums, more work per thread between barriers)
    const int t2 = ((nvec + 1) / 2 + 31) / 32 * 32, t3 = ((nvec + 2) / 3 + 31) / 32 * 32;
    // (forward only: the backward's shared-memory accumulators scale with rows per CTA and would leave 8 warps per SM)
    if (!bwd && t3 * 3 <= t2 * 2) { sh.vpt = 3; sh.tpr = t3; }
    else { sh.vpt = 2; sh.tpr = t2; }
  }
  // bytes of one row in the ring: forward in0 (+ in1) + skip; backward dout + in0 (+ in1)
  const int row_bytes = bwd ? N * elem * (2 + (has_in1 ? 1 : 0) + (has_dout2 ? 1 : 0))
                            : N * elem * (has_in1 ? 2 : 1) + (has_skip ? Ns * elem : 0);
  sh.rpc = 256 / sh.tpr;
  while (sh.rpc > 1 && sh.rpc * row_bytes > 48 * 1024) sh.rpc /= 2;
  if (sh.rpc < 1) sh.rpc = 1;
  sh.stage_bytes = (sh.rpc * row_bytes + 127) / 128 * 128;
  // ring depth: whatever keeps three CTAs resident (~74 KB each) next to the backward's accumulators
  size_t fixed = bwd ? static_cast<size_t>(nacc * sh.rpc + 2) * N * sizeof(float) : 0;
  if (general) fixed += bwd ? (static_cast<size_t>(sh.rpc + 1) * N + Ns) * sizeof(float) : N * sizeof(float);
  const long long room = (sh.vpt == 3 ? 108 : 74) * 1024 - static_cast<long long>(fixed);   // vpt 3: two CTAs per SM
  sh.stages = room > 0 ? static_cast<int>(room / sh.stage_bytes) : 0;
  if (sh.stages < 2) sh.stages = 2;
  if (sh.stages > kMaxStages) sh.stages = kMaxStages;
  sh.smem = static_cast<size_t>(sh.stages) * sh.stage_bytes + fixed;
  return sh;
}

static int validate(const char* fn, const spectre_norm_act_t* d, bool bwd) {
  if (!d) return fail(SPECTRE_E_NULL, "%s: null descriptor", fn);
  if (!d->in0 || !d->gamma || !d->beta || !d->mean || !d->rstd) return fail(SPECTRE_E_NULL, "%s: null pointer", fn);
  if (d->rows < 1 || d->N < 8 || (d->N % 8) || d->N > 4096)
    return fail(SPECTRE_E_SHAPE, "%s: need rows >= 1 and N a multiple of 8 in [8, 4096] (rows=%lld N=%d)", fn, d->rows, d->N);
  const bool has_skip = bwd ? d->d_skip != nullptr : d->skip != nullptr;
  if (has_skip && (d->Ns < 8 || (d->Ns % 8) || d->Ns > 16384))
    return fail(SPECTRE_E_SHAPE, "%s: skip width must be a multiple of 8 in [8, 16384] (Ns=%d)", fn, d->Ns);
  if (d->act_dtype != SPECTRE_DTYPE_F32 && d->act_dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, d->act_dtype);
  if (d->drop_p < 0.f || d->drop_p >= 1.f) return fail(SPECTRE_E_SHAPE, "%s: drop_p=%g outside [0,1)", fn, d->drop_p);
  const void* ptrs[] = {d->in0, d->in1, d->gamma, d->beta, d->skip, d->out, d->dout, d->dout2,
                        d->d_in0, d->d_skip, d->dgamma, d->dbeta, d->dcolsum, d->workspace};
  for (const void* q : ptrs)
    if (!aligned16(q)) return fail(SPECTRE_E_ALIGN, "%s: pointers must be 16-byte aligned", fn);
  if (!bwd && !d->out) return fail(SPECTRE_E_NULL, "%s: null out", fn);
  if ((bwd ? d->d_out_pool != nullptr : d->out_pool != nullptr)) {
    if ((d->pool_k != 2 && d->pool_k != 4 && d->pool_k != 8) || d->N % (8 * d->pool_k))
      return fail(SPECTRE_E_SHAPE, "%s: pooled output needs pool_k in {2, 4, 8} and N a multiple of 8 * pool_k (pool_k=%d N=%d)",
                  fn, d->pool_k, d->N);
    if (!aligned16(d->out_pool) || !aligned16(d->d_out_pool))
      return fail(SPECTRE_E_ALIGN, "%s: pointers must be 16-byte aligned", fn);
  }
  if (bwd && (!d->dout || !d->dgamma || !d->dbeta || !d->d_in0)) return fail(SPECTRE_E_NULL, "%s: null gradient pointer", fn);
  return SPECTRE_OK;
}

static Args to_args(const spectre_norm_act_t& d, bool bwd) {
  Args a{};
  a.in0 = d.in0; a.in1 = d.in1; a.gamma = d.gamma; a.beta = d.beta; a.skip = d.skip;
  a.out = d.out; a.mean = d.mean; a.rstd = d.rstd;
  a.dout = d.dout; a.dout2 = bwd ? d.dout2 : nullptr; a.d_in0 = d.d_in0; a.d_skip = d.d_skip;
  a.nacc = (bwd && d.dcolsum) ? 3 : 2;
  a.part = static_cast<float*>(d.workspace);
  const bool has_skip = bwd ? d.d_skip != nullptr : d.skip != nullptr;
  a.out_pool = bwd ? nullptr : d.out_pool;
  a.d_out_pool = bwd ? d.d_out_pool : nullptr;
  a.pool_k = (a.out_pool || a.d_out_pool) ? d.pool_k : 0;
  a.rows = d.rows; a.N = d.N; a.Ns = has_skip ? d.Ns : d.N; a.act = d.act; a.eps = d.eps;
  set_dropout(a, d.drop_p);
  a.seed_lo = static_cast<unsigned>(d.seed); a.seed_hi = static_cast<unsigned>(d.seed >> 32);
  a.seed_dev = d.seed_dev;
  a.skip_mode = pick_skip(a.N, a.Ns, has_skip);
  a.reverse = d.reverse ? 1 : 0;
  return a;
}

static int grid_for(long long rows, const Shape& sh) {
  const long long groups = (rows + sh.rpc - 1) / sh.rpc;
  const int block = sh.tpr * sh.rpc;
  // resident threads per SM by registers: 768 (80 registers each); 512 for the 3-vector variants (<= 128 each)
  int per_sm = (sh.vpt == 3 ? 512 : 768) / block;
  if (per_sm < 1) per_sm = 1;
  if (per_sm > 6) per_sm = 6;
  const int by_smem = static_cast<int>((220 * 1024) / (sh.smem + 1024));
  if (per_sm > by_smem) per_sm = by_smem < 1 ? 1 : by_smem;
  const long long cap = static_cast<long long>(device_sm_count()) * per_sm;
  return static_cast<int>(groups < cap ? groups : cap);
}

// kernel table: [bf16][vpt - 1][act][mode]
typedef void (*KernelFn)(const Args);
template <class T, int VPT, int ACT, int... M>
static void fill_row(KernelFn* fwd, KernelFn* bwd, std::integer_sequence<int, M...>) {
  ((fwd[M] = norm_act_fwd_kernel<T, VPT, M, ACT>), ...);
  if constexpr (VPT <= 2) ((bwd[M] = norm_act_bwd_kernel<T, VPT, M, ACT>), ...);   // 3 vectors per thread: forward only
}
struct KernelTable {
  KernelFn fwd[2][3][2][kSkipModes], bwd[2][3][2][kSkipModes];
  KernelTable() : fwd{}, bwd{} {
    using Seq = std::make_integer_sequence<int, kSkipModes>;
    fill_row<float, 1, 0>(fwd[0][0][0], bwd[0][0][0], Seq{});
    fill_row<float, 1, 1>(fwd[0][0][1], bwd[0][0][1], Seq{});
    fill_row<float, 2, 0>(fwd[0][1][0], bwd[0][1][0], Seq{});
    fill_row<float, 2, 1>(fwd[0][1][1], bwd[0][1][1], Seq{});
    fill_row<float, 3, 0>(fwd[0][2][0], bwd[0][2][0], Seq{});
    fill_row<float, 3, 1>(fwd[0][2][1], bwd[0][2][1], Seq{});
    fill_row<__nv_bfloat16, 1, 0>(fwd[1][0][0], bwd[1][0][0], Seq{});
    fill_row<__nv_bfloat16, 1, 1>(fwd[1][0][1], bwd[1][0][1], Seq{});
    fill_row<__nv_bfloat16, 2, 0>(fwd[1][1][0], bwd[1][1][0], Seq{});
    fill_row<__nv_bfloat16, 2, 1>(fwd[1][1][1], bwd[1][1][1], Seq{});
    fill_row<__nv_bfloat16, 3, 0>(fwd[1][2][0], bwd[1][2][0], Seq{});
    fill_row<__nv_bfloat16, 3, 1>(fwd[1][2][1], bwd[1][2][1], Seq{});
  }
};
static const KernelTable& table() {
  static const KernelTable t;
  return t;
}

}  // namespace na
}  // namespace sb200

using namespace sb200;
using namespace sb200::na;

extern "C" {

int spectre_norm_act_fwd(const spectre_norm_act_t* d, void* stream) {
  if (int rc = validate("spectre_norm_act_fwd", d, false)) return rc;
  Args a = to_args(*d, false);
  const bool bf = d->act_dtype == SPECTRE_DTYPE_BF16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (bf && fast_enabled()) {
    int rc = SPECTRE_OK;
    if (fast_fwd_launch(a, st, &rc)) {
      if (rc != SPECTRE_OK) return rc;
      last_launch_count() = 1;
      last_error().clear();
      return SPECTRE_OK;
    }
  }
  const Shape sh = launch_shape(a.N, a.Ns, a.in1 != nullptr, a.skip_mode != kSkipNone, bf ? 2 : 4, false, false, 2,
                                a.skip_mode == kSkipGeneral);
  a.tpr = sh.tpr; a.rpc = sh.rpc; a.stages = sh.stages; a.stage_bytes = sh.stage_bytes;
  const int grid = grid_for(a.rows, sh), block = sh.tpr * sh.rpc;
  const KernelFn kern = table().fwd[bf ? 1 : 0][sh.vpt - 1][a.act ? 1 : 0][a.skip_mode];
  if (sh.smem > kMaxSmem || ensure_dynamic_smem(reinterpret_cast<const void*>(kern), kMaxSmem) != cudaSuccess)
    return fail(SPECTRE_E_SHAPE, "spectre_norm_act_fwd: N=%d Ns=%d needs %zu B of shared memory (limit %d)", a.N, a.Ns, sh.smem, kMaxSmem);
  kern<<<grid, block, sh.smem, st>>>(a);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "norm_act_fwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

int spectre_norm_act_supported(int N, int Ns, int act_dtype) {
  if (N < 8 || (N % 8) || N > 4096 || Ns < 8 || (Ns % 8) || Ns > 16384) return 0;
  if (act_dtype != SPECTRE_DTYPE_F32 && act_dtype != SPECTRE_DTYPE_BF16) return 0;
  const int elem = act_dtype == SPECTRE_DTYPE_BF16 ? 2 : 4;
  const bool general = pick_skip(N, Ns, true) == kSkipGeneral;
  // the widest launches of each direction: forward with in1 and a skip, backward with in1, dout2 and column sums
  const Shape f = launch_shape(N, Ns, true, true, elem, false, false, 2, general);
  const Shape b = launch_shape(N, Ns, true, false, elem, true, true, 3, general);
  return f.smem <= static_cast<size_t>(kMaxSmem) && b.smem <= static_cast<size_t>(kMaxSmem) ? 1 : 0;
}

// upper bound over element types / optional operands (the query has no such arguments)
size_t spectre_norm_act_bwd_workspace_bytes(long long rows, int N) {
  if (rows < 1 || N < 8) return 0;
  size_t best = 0;
  for (int elem : {2, 4})
    for (int extra = 0; extra < 3; ++extra)
      for (int nacc : {2, 3}) {
        const Shape sh = launch_shape(N, N, extra >= 1, false, elem, true, extra >= 2, nacc);
        const size_t b = static_cast<size_t>(grid_for(rows, sh)) * 3 * N * sizeof(float);
        if (b > best) best = b;
      }
  const size_t fast = static_cast<size_t>(fast_bwd_grid_bound()) * 3 * N * sizeof(float);
  return fast > best ? fast : best;
}

int spectre_norm_act_bwd(const spectre_norm_act_t* d, void* stream) {
  if (int rc = validate("spectre_norm_act_bwd", d, true)) return rc;
  Args a = to_args(*d, true);
  const bool bf = d->act_dtype == SPECTRE_DTYPE_BF16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (bf && fast_enabled()) {
    int rc = SPECTRE_OK, fgrid = 0;
    if (fast_bwd_launch(a, d->workspace ? d->workspace_bytes : 0, st, &rc, &fgrid)) {
      if (rc != SPECTRE_OK) return rc;
      launch_pdl(reduce_gamma_beta_kernel, dim3((a.nacc * a.N + 7) / 8), dim3(256), 0, st, static_cast<const float*>(a.part),
                 d->dgamma, d->dbeta, d->dcolsum, a.N, a.nacc, fgrid);
      last_launch_count() = 2;
      const cudaError_t err = cudaGetLastError();
      if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "norm_act_bwd launch: %s", cudaGetErrorString(err));
      last_error().clear();
      return SPECTRE_OK;
    }
  }
  const Shape sh = launch_shape(a.N, a.Ns, a.in1 != nullptr, false, bf ? 2 : 4, true, a.dout2 != nullptr, a.nacc,
                                a.skip_mode == kSkipGeneral);
  a.tpr = sh.tpr; a.rpc = sh.rpc; a.stages = sh.stages; a.stage_bytes = sh.stage_bytes;
  const int grid = grid_for(a.rows, sh), block = sh.tpr * sh.rpc;
  if (!d->workspace || d->workspace_bytes < static_cast<size_t>(grid) * a.nacc * a.N * sizeof(float))
    return fail(SPECTRE_E_WORKSPACE, "spectre_norm_act_bwd: workspace %zu B < required %zu B", d->workspace_bytes,
                static_cast<size_t>(grid) * a.nacc * a.N * sizeof(float));
  const KernelFn kern = table().bwd[bf ? 1 : 0][sh.vpt - 1][a.act ? 1 : 0][a.skip_mode];
  if (sh.smem > kMaxSmem || ensure_dynamic_smem(reinterpret_cast<const void*>(kern), kMaxSmem) != cudaSuccess)
    return fail(SPECTRE_E_SHAPE, "spectre_norm_act_bwd: N=%d Ns=%d needs %zu B of shared memory (limit %d)", a.N, a.Ns, sh.smem, kMaxSmem);
  kern<<<grid, block, sh.smem, st>>>(a);
  int launches = 1;
  reduce_gamma_beta_kernel<<<(a.nacc * a.N + 7) / 8, 256, 0, st>>>(a.part, d->dgamma, d->dbeta, d->dcolsum, a.N, a.nacc, grid);
  ++launches;
  last_launch_count() = launches;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "norm_act_bwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

}  // extern "C"
