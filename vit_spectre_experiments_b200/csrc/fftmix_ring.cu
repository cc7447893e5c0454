// FFT token mixer for fp32 activations, second generation: TMA-fed, in place in shared memory (sm_100a).
//
// Same operator as fftmix.cu -- y = irfft2(rfft2(x) * filter), norm="ortho", and its backward
// (spectre_vit/models/spectre/spectre.py:50,66,136) -- for the fp32 stream
// of the CIFAR / MNIST configs and for the fp32 kernels that the reference's fp16 autocast falls back to
// (repl/train.py:205,219: torch.fft up-casts to fp32 there).  What changed against fftmix.cu, after ncu showed it
// stalled on its own global loads and on three CTA-wide barriers per image group at 15-44 % occupancy
// (profiles/r02_ncu_fft32_v11.txt):
//
//   * the activation tile [G images][H*W tokens][tile channels] arrives by ONE 3-D TMA box per tensor and
//     iteration (cp.async.bulk.tensor + mbarrier), issued an iteration ahead by a single thread: no warp ever
//     waits on a global load;
//   * the 2-D transform runs IN PLACE in that tile.  A real row of W tokens becomes its half-complex
//     spectrum stored in the same W slots (slot 0 = Re X0, slot k = Re Xk, slot W-k = Im Xk, slot W/2 = Re
//     X_{W/2}), every thread touching only its own channel lane -- so there is no separate spectrum buffer
//     (shared memory per CTA drops 2.2x, more CTAs per SM) and no write/read race inside a pass;
//   * the kw = 0 and kw = W/2 columns hold REAL data before the column pass and only their real part
//     survives the c2r row pass (torch.fft.irfft2 = c2c over H, then c2r over W, which ignores Im of the
//     DC / Nyquist bins).  real(ifft_H(A * F)) for a Hermitian A equals irfft_H(A * Fh) with the
//     Hermitian-symmetrised filter column Fh[k] = (F[k] + conj(F[(H-k) % H])) / 2, so those columns go
//     through length-H REAL transforms, and one thread owns both of them: W/2 equally heavy column items
//     per (image, lane) instead of W/2 + 1;
//   * the backward keeps X-hat in registers and accumulates the filter gradient of the real columns on
//     their Hermitian half only (the other half is its conjugate);
//   * a lane is two adjacent channels (float2 arithmetic, 8x8 / 7x7 / 4x4 grids) or ONE channel (14x14: the
//     backward holds three length-14 complex vectors per thread, 84 registers instead of 168);
//   * every index a thread needs (its row items, its column item) is computed once, before the image loop.
//
// HBM traffic is unchanged: x read once, y written once (+ the filter tile per CTA, an L2 hit).
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"
#include "fft_codelets.cuh"
#include "sm100_ptx.cuh"

namespace sb200 {
namespace ring {

using namespace ptx;

// ------------------------------------------------------------------ lane types
// VEC = 2: V = float2 (two channels), a complex filter / accumulator entry is a float4 (re0, re1, im0, im1)
// VEC = 1: V = float,                 ...                                    a float2 (re, im)
template <int VEC> struct Lane;
template <> struct Lane<2> { using V = float2; using F = float4; };
template <> struct Lane<1> { using V = float;  using F = float2; };

__device__ __forceinline__ Cx<float2> unpackf(const float4& v) { return {make_float2(v.x, v.y), make_float2(v.z, v.w)}; }
__device__ __forceinline__ Cx<float> unpackf(const float2& v) { return {v.x, v.y}; }
__device__ __forceinline__ float4 packf(const Cx<float2>& z) { return make_float4(z.re.x, z.re.y, z.im.x, z.im.y); }
__device__ __forceinline__ float2 packf(const Cx<float>& z) { return make_float2(z.re, z.im); }
__device__ __forceinline__ void addf(float4& a, const float4& b) { a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
__device__ __forceinline__ void addf(float2& a, const float2& b) { a.x += b.x; a.y += b.y; }
__device__ __forceinline__ void stv(float* p, float2 v) { __stcs(reinterpret_cast<float2*>(p), v); }
__device__ __forceinline__ void stv(float* p, float v) { __stcs(p, v); }
// filter memory order is (re, im) per channel: two channels = (re0, im0, re1, im1)
__device__ __forceinline__ float4 ldfilt(const float* p, float4*) {
  const float4 g = __ldg(reinterpret_cast<const float4*>(p));
  return make_float4(g.x, g.z, g.y, g.w);
}
__device__ __forceinline__ float2 ldfilt(const float* p, float2*) { return __ldg(reinterpret_cast<const float2*>(p)); }
__device__ __forceinline__ void stslab(float* p, const float4& t, float w) {
  *reinterpret_cast<float4*>(p) = make_float4(t.x * w, t.z * w, t.y * w, t.w * w);
}
__device__ __forceinline__ void stslab(float* p, const float2& t, float w) {
  *reinterpret_cast<float2*>(p) = make_float2(t.x * w, t.y * w);
}

template <int H_, int W_, int P_, int G_, int VEC_>
struct Cfg {
  static constexpr int H = H_, W = W_, P = P_, G = G_, VEC = VEC_;
  using V = typename Lane<VEC>::V;
  using F = typename Lane<VEC>::F;
  static constexpr int kTileC = P * VEC;      // channels per CTA tile
  static constexpr int HW = H * W;
  static constexpr int Wh = W / 2 + 1;
  static constexpr int Hh = H / 2 + 1;
  static constexpr int NCX = (W - 1) / 2;     // complex columns kw = 1 .. NCX
  static constexpr bool kNyq = (W % 2 == 0);  // a second real column at kw = W/2
  static constexpr int NRE = kNyq ? 2 : 1;
  static constexpr int NS = NCX + 1;          // column items per (image, lane): item 0 = the real columns
  static constexpr int kColItems = NS * G * P;
  static constexpr int kRowItems = G * H * P;
  static constexpr int NT = (kColItems + 31) / 32 * 32;
  static constexpr int kRowRounds = (kRowItems + NT - 1) / NT;
  static constexpr int kStageBytes = G * HW * P * static_cast<int>(sizeof(V));  // per tensor and stage
  static constexpr int kFiltCx = H * NCX * P;     // F entries, [kh][kw-1][p]
  static constexpr int kFiltRe = NRE * Hh * P;    // F entries, [col][k][p]
  static constexpr int kFiltBytes = (kFiltCx + kFiltRe) * static_cast<int>(sizeof(F));
  static constexpr int kAccCx = H;                // accumulators of a complex item
  static constexpr int kAccRe = NRE * Hh;         // ... of the real item
  static constexpr int kAcc = kAccCx > kAccRe ? kAccCx : kAccRe;
  static constexpr int kBarBytes = 128;
  // register budgets: 64 per thread forward, 128 backward (170 for the three length-16 complex vectors of 16x16)
  static constexpr int kMinBlocksFwd = 1024 / NT;
  static constexpr int kMinBlocksBwd = (H >= 16 ? 384 : 512) / NT;
  static_assert(kStageBytes % 128 == 0, "TMA destinations must stay 128-byte aligned");
  static_assert(kFiltBytes % 16 == 0, "barriers follow the filter tile");
  static constexpr size_t smem(int stages, int tensors) {
    return static_cast<size_t>(stages) * tensors * kStageBytes + kFiltBytes + kBarBytes;
  }
};

// ------------------------------------------------------------------ filter tile
// complex columns: fcx[(kh*NCX + kw-1)*P + p] = scale * filt[kh, kw, lane channels]
// real columns:    fre[(j*Hh + k)*P + p]     = scale * (F[k, kwj] + conj(F[(H-k)%H, kwj])) / 2,  kwj in {0, W/2}
template <class C>
__device__ __forceinline__ void load_filter(typename C::F* fcx, typename C::F* fre, const float* __restrict__ filt,
                                            int c0, int E, float scale) {
  // All global loads of a batch are issued before their first use: one dependent load -> scale -> store chain per
  // entry costs a serial L2 round trip each (11 per thread on the 14x14 grid).
  using F = typename C::F;
  using V = typename C::V;
  using O = VOps<V>;
  constexpr int H = C::H, P = C::P, Wh = C::Wh;
  constexpr int kCx = (C::kFiltCx + C::NT - 1) / C::NT, kRe = (C::kFiltRe + C::NT - 1) / C::NT;
  {
    F raw[kCx];
#pragma unroll
    for (int k = 0; k < kCx; ++k) {
      const int i = threadIdx.x + k * C::NT, p = i % P, e = i / P, kw = e % C::NCX + 1, kh = e / C::NCX;
      const int c = c0 + C::VEC * p;
      raw[k] = packf(Cx<V>{O::zero(), O::zero()});
      if (i < C::kFiltCx && c < E)
        raw[k] = ldfilt(filt + (static_cast<size_t>(kh * Wh + kw) * E + c) * 2, static_cast<F*>(nullptr));
    }
#pragma unroll
    for (int k = 0; k < kCx; ++k) {
      const int i = threadIdx.x + k * C::NT;
      Cx<V> z = unpackf(raw[k]);
      z.re = O::muls(z.re, scale);
      z.im = O::muls(z.im, scale);
      if (i < C::kFiltCx) fcx[i] = packf(z);
    }
  }
  {
    F ra[kRe], rb[kRe];
#pragma unroll
    for (int k = 0; k < kRe; ++k) {
      const int i = threadIdx.x + k * C::NT, p = i % P, e = i / P, kk = e % C::Hh, j = e / C::Hh;
      const int c = c0 + C::VEC * p, kw = j == 0 ? 0 : C::W / 2, k2 = (H - kk) % H;
      ra[k] = rb[k] = packf(Cx<V>{O::zero(), O::zero()});
      if (i < C::kFiltRe && c < E) {
        ra[k] = ldfilt(filt + (static_cast<size_t>(kk * Wh + kw) * E + c) * 2, static_cast<F*>(nullptr));
        rb[k] = ldfilt(filt + (static_cast<size_t>(k2 * Wh + kw) * E + c) * 2, static_cast<F*>(nullptr));
      }
    }
#pragma unroll
    for (int k = 0; k < kRe; ++k) {
      const int i = threadIdx.x + k * C::NT;
      const Cx<V> a = unpackf(ra[k]), b = unpackf(rb[k]);
      Cx<V> z;
      z.re = O::muls(O::add(a.re, b.re), 0.5f * scale);
      z.im = O::muls(O::sub(a.im, b.im), 0.5f * scale);
      if (i < C::kFiltRe) fre[i] = packf(z);
    }
  }
}

// ------------------------------------------------------------------ row passes (in place)
// The row items of a thread (image slot g, grid row h, lane p) do not change over the image loop: their
// shared-memory and global offsets are computed once.
template <class C>
struct RowPlan {
  int soff[C::kRowRounds];  // V elements from the start of the stage
  int goff[C::kRowRounds];  // floats from the start of image b0 in global memory
  int g[C::kRowRounds];     // image slot of the item, -1: no item in this round
  __device__ __forceinline__ void init(int c0, int E, int cls) {
    const int tokens = C::HW + cls;
#pragma unroll
    for (int r = 0; r < C::kRowRounds; ++r) {
      const unsigned ri = threadIdx.x + r * C::NT;
      const unsigned p = ri % C::P, gh = ri / C::P, gg = gh / C::H, h = gh % C::H;
      const bool live = ri < static_cast<unsigned>(C::kRowItems);
      soff[r] = static_cast<int>(gh * C::W * C::P + p);
      goff[r] = static_cast<int>((gg * tokens + cls + h * C::W) * E + c0 + C::VEC * p);
      g[r] = live ? static_cast<int>(gg) : -1;
    }
  }
};

template <class C>
__device__ __forceinline__ void rows_forward(typename C::V* st, const RowPlan<C>& plan) {
  using V = typename C::V;
  constexpr int W = C::W, P = C::P;
#pragma unroll
  for (int r = 0; r < C::kRowRounds; ++r) {
    if (plan.g[r] < 0) continue;
    V* base = st + plan.soff[r];
    V xin[W];
#pragma unroll
    for (int w = 0; w < W; ++w) xin[w] = base[w * P];
    Cx<V> X[C::Wh];
    RFFT<W, V>::run(xin, X);
    base[0] = X[0].re;
#pragma unroll
    for (int k = 1; k <= C::NCX; ++k) {
      base[k * P] = X[k].re;
      base[(W - k) * P] = X[k].im;
    }
    if constexpr (C::kNyq) base[(W / 2) * P] = X[W / 2].re;
  }
}

template <class C>
__device__ __forceinline__ void rows_inverse(const typename C::V* st, const RowPlan<C>& plan, float* __restrict__ y_b0,
                                             int images_left, int E) {
  using V = typename C::V;
  using O = VOps<V>;
  constexpr int W = C::W, P = C::P;
#pragma unroll
  for (int r = 0; r < C::kRowRounds; ++r) {
    if (plan.g[r] < 0) continue;
    const V* base = st + plan.soff[r];
    Cx<V> X[C::Wh];
    X[0] = {base[0], O::zero()};
#pragma unroll
    for (int k = 1; k <= C::NCX; ++k) X[k] = {base[k * P], base[(W - k) * P]};
    if constexpr (C::kNyq) X[W / 2] = {base[(W / 2) * P], O::zero()};
    V out[W];
    IRFFT<W, V>::run(X, out);
    if (plan.g[r] < images_left) {
      float* dst = y_b0 + plan.goff[r];
#pragma unroll
      for (int w = 0; w < W; ++w) stv(dst + static_cast<size_t>(w) * E, out[w]);
    }
  }
}

// The CLS token passes through the mixer unchanged (spectre.py:66 mixes the patch grid only).  Its G * kTileC floats
// are loaded at the top of an iteration and stored at its end: a store right behind the load would park the warp for a
// DRAM round trip in every iteration.
template <class C>
struct ClsCopy {
  static constexpr int kPer = (C::G * C::kTileC + C::NT - 1) / C::NT;
  float v[kPer];
  __device__ __forceinline__ void load(const float* __restrict__ src, int b0, int B, int c0, int E) {
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = threadIdx.x + k * C::NT, b = b0 + i / C::kTileC, c = c0 + i % C::kTileC;
      if (i < C::G * C::kTileC && b < B && c < E) v[k] = __ldcs(src + static_cast<size_t>(b) * (C::HW + 1) * E + c);
    }
  }
  __device__ __forceinline__ void store(float* __restrict__ dst, int b0, int B, int c0, int E) const {
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      const int i = threadIdx.x + k * C::NT, b = b0 + i / C::kTileC, c = c0 + i % C::kTileC;
      if (i < C::G * C::kTileC && b < B && c < E) __stcs(dst + static_cast<size_t>(b) * (C::HW + 1) * E + c, v[k]);
    }
  }
};

// ------------------------------------------------------------------ pipeline bookkeeping
// One thread feeds the ring: iteration `it` of this CTA is image group blockIdx.y + it*gridDim.y, its tiles
// land in stage it % S and complete full[it % S] (phase (it / S) & 1).
template <class C, int TENSORS>
struct Feeder {
  unsigned char* stages;
  uint64_t* full;
  int S, c0, n_it;
  __device__ __forceinline__ typename C::V* stage(int it, int tensor) const {
    return reinterpret_cast<typename C::V*>(stages + static_cast<size_t>((it % S) * TENSORS + tensor) * C::kStageBytes);
  }
  __device__ __forceinline__ int image0(int it) const { return (blockIdx.y + it * gridDim.y) * C::G; }
  __device__ __forceinline__ void issue(int it, const CUtensorMap* m0, const CUtensorMap* m1) const {
    if (it >= n_it) return;
    uint64_t* bar = full + it % S;
    mbar_arrive_expect_tx(bar, TENSORS * C::kStageBytes);
    tma_load_3d(stage(it, 0), m0, bar, c0, 0, image0(it));
    if constexpr (TENSORS == 2) tma_load_3d(stage(it, 1), m1, bar, c0, 0, image0(it));
  }
  __device__ __forceinline__ void wait(int it) const { mbar_wait(full + it % S, (it / S) & 1); }
};

// ------------------------------------------------------------------ forward
template <class C>
__global__ void __launch_bounds__(C::NT, C::kMinBlocksFwd) fftmix_ring_fwd_kernel(const __grid_constant__ CUtensorMap tmx,
                                                                const float* __restrict__ x,
                                                                const float* __restrict__ filt, float* __restrict__ y,
                                                                int B, int E, int cls, int S) {
  using V = typename C::V;
  using F = typename C::F;
  constexpr int H = C::H, W = C::W, P = C::P, G = C::G;
  extern __shared__ __align__(128) unsigned char smem[];
  F* fcx = reinterpret_cast<F*>(smem + static_cast<size_t>(S) * C::kStageBytes);
  F* fre = fcx + C::kFiltCx;
  uint64_t* full = reinterpret_cast<uint64_t*>(fre + C::kFiltRe);
  const int c0 = blockIdx.x * C::kTileC;
  const int groups = (B + G - 1) / G;
  const int n_it = (groups - static_cast<int>(blockIdx.y) + static_cast<int>(gridDim.y) - 1) / static_cast<int>(gridDim.y);
  const Feeder<C, 1> feed{smem, full, S, c0, n_it};

  if (threadIdx.x == 0) {
    prefetch_tensormap(&tmx);
    for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
    mbar_fence_init();
    for (int s = 0; s < S; ++s) feed.issue(s, &tmx, nullptr);
  }
  load_filter<C>(fcx, fre, filt, c0, E, 1.0f / static_cast<float>(H * W));
  RowPlan<C> plan;
  plan.init(c0, E, cls);
  const unsigned q = threadIdx.x, p = q % P, g = (q / P) % G, s = q / (P * G);
  const bool col_live = q < static_cast<unsigned>(C::kColItems);
  const int ioff = static_cast<int>(g * C::HW * P + p);
  const size_t img_floats = static_cast<size_t>(C::HW + cls) * E;
  __syncthreads();

#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    V* st = feed.stage(it, 0);
    const int b0 = feed.image0(it);
    ClsCopy<C> cc;
    if (cls) cc.load(x, b0, B, c0, E);
    feed.wait(it);
    rows_forward<C>(st, plan);
    __syncthreads();
    // every thread is past the previous iteration's last read of its stage: refill it
    if (threadIdx.x == 0 && it > 0) feed.issue(it - 1 + S, &tmx, nullptr);

    if (col_live) {
      V* img = st + ioff;
      if (s == 0) {
#pragma unroll
        for (int j = 0; j < C::NRE; ++j) {
          V* col = img + (j == 0 ? 0 : W / 2) * P;
          const F* f = fre + j * C::Hh * P + p;
          V a[H];
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = col[h * W * P];
          Cx<V> A[C::Hh];
          RFFT<H, V>::run(a, A);
#pragma unroll
          for (int k = 0; k < C::Hh; ++k) A[k] = cmul(A[k], unpackf(f[k * P]));
          IRFFT<H, V>::run(A, a);
#pragma unroll
          for (int h = 0; h < H; ++h) col[h * W * P] = a[h];
        }
      } else {
        V* re = img + s * P;
        V* im = img + (W - s) * P;
        const F* f = fcx + (s - 1) * P + p;
        Cx<V> z[H];
#pragma unroll
        for (int h = 0; h < H; ++h) z[h] = {re[h * W * P], im[h * W * P]};
        CFFT<H, -1, V>::run(z);
#pragma unroll
        for (int kh = 0; kh < H; ++kh) z[kh] = cmul(z[kh], unpackf(f[kh * C::NCX * P]));
        CFFT<H, 1, V>::run(z);
#pragma unroll
        for (int h = 0; h < H; ++h) {
          re[h * W * P] = z[h].re;
          im[h * W * P] = z[h].im;
        }
      }
    }
    fence_proxy_async_smem();  // this stage is overwritten by the async proxy (TMA) later on
    __syncthreads();
    rows_inverse<C>(st, plan, y + static_cast<size_t>(b0) * img_floats, B - b0, E);
    if (cls) cc.store(y, b0, B, c0, E);
  }
}

// ------------------------------------------------------------------ backward
// part: [gridDim.y][H][Wh][E][2] fp32 partial filter gradients (one slab per blockIdx.y)
template <class C>
__global__ void __launch_bounds__(C::NT, C::kMinBlocksBwd) fftmix_ring_bwd_kernel(const __grid_constant__ CUtensorMap tmdy,
                                                                const __grid_constant__ CUtensorMap tmx,
                                                                const float* __restrict__ dy,
                                                                const float* __restrict__ filt, float* __restrict__ dx,
                                                                float* __restrict__ part, int B, int E, int cls, int S) {
  using V = typename C::V;
  using F = typename C::F;
  using O = VOps<V>;
  constexpr int H = C::H, W = C::W, P = C::P, G = C::G;
  extern __shared__ __align__(128) unsigned char smem[];
  F* fcx = reinterpret_cast<F*>(smem + static_cast<size_t>(S) * 2 * C::kStageBytes);
  F* fre = fcx + C::kFiltCx;
  uint64_t* full = reinterpret_cast<uint64_t*>(fre + C::kFiltRe);
  const int c0 = blockIdx.x * C::kTileC;
  const int groups = (B + G - 1) / G;
  const int n_it = (groups - static_cast<int>(blockIdx.y) + static_cast<int>(gridDim.y) - 1) / static_cast<int>(gridDim.y);
  const Feeder<C, 2> feed{smem, full, S, c0, n_it};

  if (threadIdx.x == 0) {
    prefetch_tensormap(&tmdy);
    prefetch_tensormap(&tmx);
    for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
    mbar_fence_init();
    for (int s = 0; s < S; ++s) feed.issue(s, &tmdy, &tmx);
  }
  load_filter<C>(fcx, fre, filt, c0, E, 1.0f / static_cast<float>(H * W));
  RowPlan<C> plan;
  plan.init(c0, E, cls);
  const unsigned q = threadIdx.x, p = q % P, g = (q / P) % G, s = q / (P * G);
  const bool col_live = q < static_cast<unsigned>(C::kColItems);
  const int ioff = static_cast<int>(g * C::HW * P + p);
  const size_t img_floats = static_cast<size_t>(C::HW + cls) * E;

  Cx<V> acc[C::kAcc];
#pragma unroll
  for (int i = 0; i < C::kAcc; ++i) acc[i] = {O::zero(), O::zero()};
  __syncthreads();

#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    V* sy = feed.stage(it, 0);
    V* sx = feed.stage(it, 1);
    const int b0 = feed.image0(it);
    if (cls) {   // (backward: the iteration is long and registers are tight -- copied on the spot, measured faster)
      ClsCopy<C> cc;
      cc.load(dy, b0, B, c0, E);
      cc.store(dx, b0, B, c0, E);
    }
    feed.wait(it);
    rows_forward<C>(sy, plan);
    rows_forward<C>(sx, plan);
    __syncthreads();
    if (threadIdx.x == 0 && it > 0) feed.issue(it - 1 + S, &tmdy, &tmx);

    if (col_live) {
      if (s == 0) {
#pragma unroll
        for (int j = 0; j < C::NRE; ++j) {
          const int coff = ioff + (j == 0 ? 0 : W / 2) * P;
          const F* f = fre + j * C::Hh * P + p;
          V a[H];
          Cx<V> AX[C::Hh], AY[C::Hh];
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = sx[coff + h * W * P];
          RFFT<H, V>::run(a, AX);
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = sy[coff + h * W * P];
          RFFT<H, V>::run(a, AY);
#pragma unroll
          for (int k = 0; k < C::Hh; ++k) {
            cmac_conj(acc[j * C::Hh + k], AX[k], AY[k]);       // acc += conj(X) * DY
            AY[k] = cmul_conj(AY[k], unpackf(f[k * P]));       // DY * conj(Fh)
          }
          IRFFT<H, V>::run(AY, a);
#pragma unroll
          for (int h = 0; h < H; ++h) sy[coff + h * W * P] = a[h];
        }
      } else {
        const int roff = ioff + s * P, moff = ioff + (W - s) * P;
        const F* f = fcx + (s - 1) * P + p;
        Cx<V> zx[H], zy[H];
#pragma unroll
        for (int h = 0; h < H; ++h) zx[h] = {sx[roff + h * W * P], sx[moff + h * W * P]};
        CFFT<H, -1, V>::run(zx);
#pragma unroll
        for (int h = 0; h < H; ++h) zy[h] = {sy[roff + h * W * P], sy[moff + h * W * P]};
        CFFT<H, -1, V>::run(zy);
#pragma unroll
        for (int kh = 0; kh < H; ++kh) {
          cmac_conj(acc[kh], zx[kh], zy[kh]);
          zy[kh] = cmul_conj(zy[kh], unpackf(f[kh * C::NCX * P]));
        }
        CFFT<H, 1, V>::run(zy);
#pragma unroll
        for (int h = 0; h < H; ++h) {
          sy[roff + h * W * P] = zy[h].re;
          sy[moff + h * W * P] = zy[h].im;
        }
      }
    }
    fence_proxy_async_smem();
    __syncthreads();
    rows_inverse<C>(sy, plan, dx + static_cast<size_t>(b0) * img_floats, B - b0, E);
  }

  // ---- partial filter gradient of this CTA: park the accumulators in the (now idle) ring as
  // [g][kh][kw][p], sum the G image slots, apply the half-spectrum weights and the ortho 1/(H*W).
  __syncthreads();
  F* scr = reinterpret_cast<F*>(smem);
  constexpr int Wh = C::Wh;
  static_assert(static_cast<size_t>(G) * H * Wh * P * sizeof(F) <= 4u * C::kStageBytes, "accumulators fit the ring");
  if (col_live) {
    F* mine = scr + g * H * Wh * P + p;
    if (s == 0) {
#pragma unroll
      for (int j = 0; j < C::NRE; ++j) {
        const int kw = j == 0 ? 0 : W / 2;
#pragma unroll
        for (int k = 0; k < C::Hh; ++k) {
          const Cx<V> a = acc[j * C::Hh + k];
          mine[(k * Wh + kw) * P] = packf(a);
          if (k > 0 && H - k != k) mine[((H - k) * Wh + kw) * P] = packf(Cx<V>{a.re, O::neg(a.im)});
        }
      }
    } else {
#pragma unroll
      for (int kh = 0; kh < H; ++kh) mine[(kh * Wh + s) * P] = packf(acc[kh]);
    }
  }
  __syncthreads();
  float* slab = part + static_cast<size_t>(blockIdx.y) * H * Wh * E * 2;
  for (int i = threadIdx.x; i < H * Wh * P; i += C::NT) {
    const int pp = i % P, e = i / P, kw = e % Wh, c = c0 + C::VEC * pp;
    if (c >= E) continue;
    F t = scr[i];
#pragma unroll
    for (int gg = 1; gg < G; ++gg) addf(t, scr[gg * H * Wh * P + i]);
    const bool edge = (kw == 0) || (W % 2 == 0 && kw == W / 2);
    stslab(slab + (static_cast<size_t>(e) * E + c) * 2, t, (edge ? 1.0f : 2.0f) / static_cast<float>(H * W));
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess) f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// [B][HW tokens][E] fp32 view of the patch grid (CLS skipped); boxes of `chans` channels x HW tokens x G images,
// dense in shared memory; channels past E and images past B read as zero.
static int make_map(CUtensorMap* map, const float* base, int B, int HW, int E, int cls, int chans, int G) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
  bind_primary_context();
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(E), static_cast<cuuint64_t>(HW), static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(E) * 4, static_cast<cuuint64_t>(HW + cls) * E * 4};
  const cuuint32_t box[3] = {static_cast<cuuint32_t>(chans), static_cast<cuuint32_t>(HW), static_cast<cuuint32_t>(G)};
  const cuuint32_t estr[3] = {1, 1, 1};
  void* gptr = const_cast<float*>(base + static_cast<size_t>(cls) * E);
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, gptr, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled(fp32 ring) failed (%d)", static_cast<int>(r));
  return SPECTRE_OK;
}

static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && e[0]) ? atoi(e) : dflt;
}

// Stages of the ring.  Two: the tile of iteration it+1 is in flight while iteration it computes, and the shared memory
// a third stage would take buys more as another resident CTA (measured, profiles/r02_kernel_bench_fft32_v11.txt:
// 8x8 E=512 B=1024 forward 0.0635 -> 0.0573 ms, backward 0.0952 -> 0.0809 ms with 2 instead of 3 stages).
template <class C>
static int pick_stages(int tensors) {
  const int forced = env_int("SB200_FFT_RING_STAGES", 0);
  if (forced >= 2 && forced <= 4 && C::smem(forced, tensors) <= 200 * 1024) return forced;
  return 2;
}

template <class K>
static int pick_slabs(K kernel, int threads, size_t smem, int tiles, int groups) {
  const int per_sm = cached_blocks_per_sm(reinterpret_cast<const void*>(kernel), threads, smem);
  int ny = (device_sm_count() * per_sm) / tiles;
  if (ny < 1) ny = 1;
  if (ny > groups) ny = groups;
  return ny;
}

template <class C>
static int fwd_launch(const float* x, const float* filt, float* y, int B, int E, int cls, cudaStream_t st) {
  auto kern = fftmix_ring_fwd_kernel<C>;
  const int S = pick_stages<C>(1);
  const size_t smem = C::smem(S, 1);
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem)) != cudaSuccess)
    return fail(SPECTRE_E_CUDA, "cudaFuncSetAttribute(ring fwd): %s", cudaGetErrorString(cudaGetLastError()));
  CUtensorMap tmx;
  if (int rc = make_map(&tmx, x, B, C::HW, E, cls, C::kTileC, C::G)) return rc;
  const int tiles = E / C::kTileC;
  const int groups = (B + C::G - 1) / C::G;
  const int ny = pick_slabs(kern, C::NT, smem, tiles, groups);
  kern<<<dim3(tiles, ny), C::NT, smem, st>>>(tmx, x, filt, y, B, E, cls, S);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_ring_fwd launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

template <class C>
static int bwd_slabs(int B, int E) {
  auto kern = fftmix_ring_bwd_kernel<C>;
  const int S = pick_stages<C>(2);
  const size_t smem = C::smem(S, 2);
  ensure_dynamic_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem));
  const int groups = (B + C::G - 1) / C::G;
  return pick_slabs(kern, C::NT, smem, E / C::kTileC, groups);
}

template <class C>
static int bwd_launch(const float* dy, const float* x, const float* filt, float* dx, float* dfilt, void* ws,
                      size_t ws_bytes, int B, int E, int cls, cudaStream_t st) {
  auto kern = fftmix_ring_bwd_kernel<C>;
  const int S = pick_stages<C>(2);
  const size_t smem = C::smem(S, 2);
  const int tiles = E / C::kTileC;
  const int ny = bwd_slabs<C>(B, E);
  const size_t slab_floats = static_cast<size_t>(C::H) * C::Wh * E * 2;
  if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
    return fail(SPECTRE_E_WORKSPACE, "fftmix_bwd: workspace %zu B < required %zu B", ws_bytes,
                slab_floats * sizeof(float) * ny);
  CUtensorMap tmdy, tmx;
  if (int rc = make_map(&tmdy, dy, B, C::HW, E, cls, C::kTileC, C::G)) return rc;
  if (int rc = make_map(&tmx, x, B, C::HW, E, cls, C::kTileC, C::G)) return rc;
  float* part = (ny > 1) ? static_cast<float*>(ws) : dfilt;
  kern<<<dim3(tiles, ny), C::NT, smem, st>>>(tmdy, tmx, dy, filt, dx, part, B, E, cls, S);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_ring_bwd launch: %s", cudaGetErrorString(err));
  last_launch_count() = 1;
  if (ny > 1) {
    launch_reduce_partials(part, dfilt, slab_floats, ny, st);
    err = cudaGetLastError();
    if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_ring_bwd reduce launch: %s", cudaGetErrorString(err));
    last_launch_count() = 2;
  }
  return SPECTRE_OK;
}

// ------------------------------------------------------------------ shape table
// (H, W, lanes per tile P, images per iteration G, channels per lane VEC); tile = P * VEC channels, E a multiple of it.
#define SB200_RING_SHAPES(F)                                                                             \
  F(8, 8, 16, 2, 2) F(8, 8, 16, 4, 2) F(8, 8, 8, 4, 2) F(7, 7, 16, 2, 2) F(7, 7, 8, 2, 2) F(7, 7, 8, 4, 2) \
  F(7, 7, 8, 8, 2) F(4, 4, 16, 4, 2) F(14, 14, 16, 1, 1) F(14, 14, 16, 2, 1) F(16, 16, 16, 1, 1)

struct Pick {
  int P, G, VEC;
};

// Two-channel lanes on 32-channel tiles (128-byte token rows) where E allows, 16-channel tiles for the narrow
// CIFAR / MNIST embeddings (E = 48: three exact tiles); one-channel lanes for the 14x14 and 16x16 grids.
// Images per iteration: the forward likes wider CTAs (8 warps), the backward -- 128 registers per thread -- more of
// them; small batches take the smaller group so that there are CTAs for every SM.
static Pick pick(int H, int W, int E, int B, bool bwd) {
  Pick k = {0, 0, 2};
  if (H >= 14) {
    k.VEC = 1;
    k.P = E % 16 == 0 ? 16 : 0;
    k.G = (H == 14 && static_cast<long long>(B) * E / 16 >= 8LL * device_sm_count()) ? 2 : 1;
  } else {
    k.P = E % 32 == 0 ? 16 : (E % 16 == 0 ? 8 : 0);
    if (k.P == 0) return k;
    const bool small = static_cast<long long>(B) * E / (2 * k.P) < 8LL * device_sm_count();
    if (H == 8 && W == 8) k.G = k.P == 16 ? ((bwd || small) ? 2 : 4) : 4;
    else if (H == 7 && W == 7) k.G = k.P == 16 ? 2 : (small ? 2 : 4);
    else k.G = 4;
  }
  const int forced = env_int(bwd ? "SB200_FFT_RING_G_BWD" : "SB200_FFT_RING_G_FWD", env_int("SB200_FFT_RING_G", 0));
  if (forced > 0) k.G = forced;
  return k;
}

static bool enabled() {
  static const bool on = env_int("SB200_FFT_RING", 1) != 0;
  return on;
}

}  // namespace ring

int fftmix_ring_supported(int H, int W, int E, const void* a, const void* b) {
  if (!ring::enabled() || E % 4 != 0) return 0;
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) return 0;
  // B = 1 / B = huge span the group sizes pick() can return for this (H, W, E)
  for (int bwd = 0; bwd < 2; ++bwd)
    for (int B : {1, 1 << 30}) {
      const ring::Pick k = ring::pick(H, W, E, B, bwd != 0);
      bool ok = false;
#define F(h, w, p, g, v) if (H == h && W == w && k.P == p && k.G == g && k.VEC == v) ok = true;
      SB200_RING_SHAPES(F)
#undef F
      if (!ok) return 0;
    }
  return 1;
}

int fftmix_ring_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls,
                             cudaStream_t st) {
  const ring::Pick k = ring::pick(H, W, E, B, false);
#define F(h, w, p, g, v)                                                                                        \
  if (H == h && W == w && k.P == p && k.G == g && k.VEC == v)                                                   \
    return ring::fwd_launch<ring::Cfg<h, w, p, g, v>>(static_cast<const float*>(x), filt, static_cast<float*>(y), B, \
                                                      E, cls, st);
  SB200_RING_SHAPES(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_ring: no kernel for grid %dx%d E=%d", H, W, E);
}

int fftmix_ring_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                             size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st) {
  const ring::Pick k = ring::pick(H, W, E, B, true);
#define F(h, w, p, g, v)                                                                                           \
  if (H == h && W == w && k.P == p && k.G == g && k.VEC == v)                                                      \
    return ring::bwd_launch<ring::Cfg<h, w, p, g, v>>(static_cast<const float*>(dy), static_cast<const float*>(x), \
                                                      filt, static_cast<float*>(dx), dfilt, ws, ws_bytes, B, E, cls, st);
  SB200_RING_SHAPES(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_ring: no kernel for grid %dx%d E=%d", H, W, E);
}

size_t fftmix_ring_bwd_workspace(int B, int H, int W, int E) {
  if (!fftmix_ring_supported(H, W, E, nullptr, nullptr)) return 0;
  const ring::Pick k = ring::pick(H, W, E, B, true);
  int ny = 0;
#define F(h, w, p, g, v) \
  if (H == h && W == w && k.P == p && k.G == g && k.VEC == v) ny = ring::bwd_slabs<ring::Cfg<h, w, p, g, v>>(B, E);
  SB200_RING_SHAPES(F)
#undef F
  return ny <= 1 ? 0 : static_cast<size_t>(ny) * H * (W / 2 + 1) * E * 2 * sizeof(float);
}

}  // namespace sb200
