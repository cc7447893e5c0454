// FFT token mixer for fp32 activations, second generation: TMA-fed, in place in shared memory (sm_100a).
//
// Same operator as fftmix.cu -- y = irfft2(rfft2(x) * filter), norm="ortho", and its backward
// (spectre_vit/models/spectre/spectre.py:50,66,136) -- for the fp32 stream
// of the CIFAR / MNIST configs (and the fp32 kernels that fp16 autocast falls back to).  What changed against
// fftmix.cu, after ncu showed it stalled on its own global loads and on three CTA-wide barriers per image
// group at 15-44 % occupancy (profiles/r02_ncu_fft32_v11.txt):
//
//   * the activation tile [G images][H*W tokens][2P channels] arrives by ONE 3-D TMA box per tensor and
//     iteration (cp.async.bulk.tensor + mbarrier), issued one to two iterations ahead by a single thread:
//     no warp ever waits on a global load;
//   * the 2-D transform runs IN PLACE in that tile.  A real row of W tokens becomes its half-complex
//     spectrum stored in the same W slots (slot 0 = Re X0, slot k = Re Xk, slot W-k = Im Xk, slot W/2 = Re
//     X_{W/2}), every thread touching only its own channel pair -- so there is no separate spectrum buffer
//     (shared memory per CTA drops 2.2x, more CTAs per SM) and no write/read race inside a pass;
//   * the kw = 0 and kw = W/2 columns hold REAL data before the column pass and only their real part
//     survives the c2r row pass (torch.fft.irfft2 = c2c over H, then c2r over W, which ignores Im of the
//     DC / Nyquist bins).  real(ifft_H(A * F)) for a Hermitian A equals irfft_H(A * Fh) with the
//     Hermitian-symmetrised filter column Fh[k] = (F[k] + conj(F[(H-k) % H])) / 2, so those columns go
//     through length-H REAL transforms, and one thread owns both of them: W/2 equally heavy column items
//     per (image, pair) instead of W/2 + 1;
//   * the backward keeps X-hat in registers and accumulates the filter gradient of the real columns on
//     their Hermitian half only (the other half is its conjugate).
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
using V = float2;
using C2 = Cx<float2>;

__device__ __forceinline__ C2 unpack4(const float4& v) { return {make_float2(v.x, v.y), make_float2(v.z, v.w)}; }
__device__ __forceinline__ float4 pack4(const C2& z) { return make_float4(z.re.x, z.re.y, z.im.x, z.im.y); }

template <int H_, int W_, int P_, int G_>
struct Cfg {
  static constexpr int H = H_, W = W_, P = P_, G = G_;
  static constexpr int HW = H * W;
  static constexpr int Wh = W / 2 + 1;
  static constexpr int Hh = H / 2 + 1;
  static constexpr int NCX = (W - 1) / 2;     // complex columns kw = 1 .. NCX
  static constexpr bool kNyq = (W % 2 == 0);  // a second real column at kw = W/2
  static constexpr int NRE = kNyq ? 2 : 1;
  static constexpr int NS = NCX + 1;          // column items per (image, pair): item 0 = the real columns
  static constexpr int kColItems = NS * G * P;
  static constexpr int kRowItems = G * H * P;
  static constexpr int NT = (kColItems + 31) / 32 * 32;
  static constexpr int kStageElems = G * HW * P;  // float2 per tensor and stage
  static constexpr int kStageBytes = kStageElems * 8;
  static constexpr int kFiltCx = H * NCX * P;     // float4 entries, [kh][kw-1][p]
  static constexpr int kFiltRe = NRE * Hh * P;    // float4 entries, [col][k][p]
  static constexpr int kFiltBytes = (kFiltCx + kFiltRe) * 16;
  static constexpr int kAccCx = H;                // accumulators of a complex item
  static constexpr int kAccRe = NRE * Hh;         // ... of the real item
  static constexpr int kAcc = kAccCx > kAccRe ? kAccCx : kAccRe;
  static constexpr int kBarBytes = 128;
  static_assert(kStageBytes % 128 == 0, "TMA destinations must stay 128-byte aligned");
  static constexpr size_t smem(int stages, int tensors) {
    return static_cast<size_t>(stages) * tensors * kStageBytes + kFiltBytes + kBarBytes;
  }
};

// ------------------------------------------------------------------ filter tile
// complex columns: fcx[(kh*NCX + kw-1)*P + p] = scale * filt[kh, kw, c .. c+1] as (re0, re1, im0, im1)
// real columns:    fre[(j*Hh + k)*P + p]     = scale * (F[k, kwj] + conj(F[(H-k)%H, kwj])) / 2,  kwj in {0, W/2}
template <class C>
__device__ __forceinline__ void load_filter(float4* fcx, float4* fre, const float* __restrict__ filt, int c0, int E,
                                            float scale) {
  constexpr int H = C::H, P = C::P, Wh = C::Wh;
  for (int i = threadIdx.x; i < C::kFiltCx; i += C::NT) {
    const int p = i % P, e = i / P, kw = e % C::NCX + 1, kh = e / C::NCX, c = c0 + 2 * p;
    float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < E) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(filt + (static_cast<size_t>(kh * Wh + kw) * E + c) * 2));
      f = make_float4(g.x * scale, g.z * scale, g.y * scale, g.w * scale);
    }
    fcx[i] = f;
  }
  for (int i = threadIdx.x; i < C::kFiltRe; i += C::NT) {
    const int p = i % P, e = i / P, k = e % C::Hh, j = e / C::Hh, c = c0 + 2 * p;
    const int kw = j == 0 ? 0 : C::W / 2, k2 = (H - k) % H;
    float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < E) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(filt + (static_cast<size_t>(k * Wh + kw) * E + c) * 2));
      const float4 b = __ldg(reinterpret_cast<const float4*>(filt + (static_cast<size_t>(k2 * Wh + kw) * E + c) * 2));
      const float h = 0.5f * scale;
      f = make_float4((a.x + b.x) * h, (a.z + b.z) * h, (a.y - b.y) * h, (a.w - b.w) * h);
    }
    fre[i] = f;
  }
}

// ------------------------------------------------------------------ row passes (in place)
template <class C>
__device__ __forceinline__ void rows_forward(float2* st) {
  constexpr int W = C::W, P = C::P;
#pragma unroll 1
  for (int r = threadIdx.x; r < C::kRowItems; r += C::NT) {
    const int p = r % P, gh = r / P;
    float2* base = st + gh * W * P + p;
    V xin[W];
#pragma unroll
    for (int w = 0; w < W; ++w) xin[w] = base[w * P];
    C2 X[C::Wh];
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
__device__ __forceinline__ void rows_inverse(const float2* st, float* __restrict__ y, int b0, int B, int c0, int E,
                                             int cls) {
  constexpr int W = C::W, P = C::P, H = C::H;
  const int tokens = C::HW + cls;
#pragma unroll 1
  for (int r = threadIdx.x; r < C::kRowItems; r += C::NT) {
    const int p = r % P, gh = r / P, g = gh / H, h = gh % H;
    const float2* base = st + gh * W * P + p;
    C2 X[C::Wh];
    X[0] = {base[0], make_float2(0.f, 0.f)};
#pragma unroll
    for (int k = 1; k <= C::NCX; ++k) X[k] = {base[k * P], base[(W - k) * P]};
    if constexpr (C::kNyq) X[W / 2] = {base[(W / 2) * P], make_float2(0.f, 0.f)};
    V out[W];
    IRFFT<W, V>::run(X, out);
    const int b = b0 + g, c = c0 + 2 * p;
    if (b < B && c < E) {
      float* dst = y + (static_cast<size_t>(b) * tokens + cls + h * W) * E + c;
#pragma unroll
      for (int w = 0; w < W; ++w) PairIO<float>::st(dst + static_cast<size_t>(w) * E, out[w]);
    }
  }
}

// The CLS token passes through the mixer unchanged (spectre.py:66 mixes the patch grid only).
template <class C>
__device__ __forceinline__ void copy_cls(const float* __restrict__ src, float* __restrict__ dst, int b0, int B, int c0,
                                         int E) {
  for (int i = threadIdx.x; i < C::G * C::P; i += C::NT) {
    const int p = i % C::P, b = b0 + i / C::P, c = c0 + 2 * p;
    if (b < B && c < E) {
      const size_t off = static_cast<size_t>(b) * (C::HW + 1) * E + c;
      PairIO<float>::st(dst + off, PairIO<float>::ld(src + off));
    }
  }
}

// ------------------------------------------------------------------ pipeline bookkeeping
// One thread feeds the ring: iteration `it` of this CTA is image group blockIdx.y + it*gridDim.y, its tiles
// land in stage it % S and complete full[it % S] (phase (it / S) & 1).
template <class C, int TENSORS>
struct Feeder {
  unsigned char* stages;
  uint64_t* full;
  int S, c0, n_it;
  __device__ __forceinline__ float2* stage(int it, int tensor) const {
    return reinterpret_cast<float2*>(stages + static_cast<size_t>((it % S) * TENSORS + tensor) * C::kStageBytes);
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
__global__ void __launch_bounds__(C::NT) fftmix_ring_fwd_kernel(const __grid_constant__ CUtensorMap tmx,
                                                                const float* __restrict__ x,
                                                                const float* __restrict__ filt, float* __restrict__ y,
                                                                int B, int E, int cls, int S) {
  constexpr int H = C::H, W = C::W, P = C::P, G = C::G;
  extern __shared__ __align__(128) unsigned char smem[];
  float4* fcx = reinterpret_cast<float4*>(smem + static_cast<size_t>(S) * C::kStageBytes);
  float4* fre = fcx + C::kFiltCx;
  uint64_t* full = reinterpret_cast<uint64_t*>(fre + C::kFiltRe);
  const int c0 = blockIdx.x * 2 * P;
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
  __syncthreads();

#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    float2* st = feed.stage(it, 0);
    const int b0 = feed.image0(it);
    if (cls) copy_cls<C>(x, y, b0, B, c0, E);
    feed.wait(it);
    rows_forward<C>(st);
    __syncthreads();
    // every thread is past the previous iteration's last read of its stage: refill it
    if (threadIdx.x == 0 && it > 0) feed.issue(it - 1 + S, &tmx, nullptr);

    if (threadIdx.x < C::kColItems) {
      const int q = threadIdx.x, p = q % P, g = (q / P) % G, s = q / (P * G);
      float2* img = st + g * C::HW * P + p;
      if (s == 0) {
#pragma unroll
        for (int j = 0; j < C::NRE; ++j) {
          float2* col = img + (j == 0 ? 0 : W / 2) * P;
          const float4* f = fre + j * C::Hh * P + p;
          V a[H];
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = col[h * W * P];
          C2 A[C::Hh];
          RFFT<H, V>::run(a, A);
#pragma unroll
          for (int k = 0; k < C::Hh; ++k) A[k] = cmul(A[k], unpack4(f[k * P]));
          IRFFT<H, V>::run(A, a);
#pragma unroll
          for (int h = 0; h < H; ++h) col[h * W * P] = a[h];
        }
      } else {
        float2* re = img + s * P;
        float2* im = img + (W - s) * P;
        const float4* f = fcx + (s - 1) * P + p;
        C2 z[H];
#pragma unroll
        for (int h = 0; h < H; ++h) z[h] = {re[h * W * P], im[h * W * P]};
        CFFT<H, -1, V>::run(z);
#pragma unroll
        for (int kh = 0; kh < H; ++kh) z[kh] = cmul(z[kh], unpack4(f[kh * C::NCX * P]));
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
    rows_inverse<C>(st, y, b0, B, c0, E, cls);
  }
}

// ------------------------------------------------------------------ backward
// part: [gridDim.y][H][Wh][E][2] fp32 partial filter gradients (one slab per blockIdx.y)
template <class C>
__global__ void __launch_bounds__(C::NT) fftmix_ring_bwd_kernel(const __grid_constant__ CUtensorMap tmdy,
                                                                const __grid_constant__ CUtensorMap tmx,
                                                                const float* __restrict__ dy,
                                                                const float* __restrict__ filt, float* __restrict__ dx,
                                                                float* __restrict__ part, int B, int E, int cls, int S) {
  constexpr int H = C::H, W = C::W, P = C::P, G = C::G;
  extern __shared__ __align__(128) unsigned char smem[];
  float4* fcx = reinterpret_cast<float4*>(smem + static_cast<size_t>(S) * 2 * C::kStageBytes);
  float4* fre = fcx + C::kFiltCx;
  uint64_t* full = reinterpret_cast<uint64_t*>(fre + C::kFiltRe);
  const int c0 = blockIdx.x * 2 * P;
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

  const int q = threadIdx.x, p = q % P, g = (q / P) % G, s = q / (P * G);
  C2 acc[C::kAcc];
#pragma unroll
  for (int i = 0; i < C::kAcc; ++i) acc[i] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
  __syncthreads();

#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    float2* sy = feed.stage(it, 0);
    float2* sx = feed.stage(it, 1);
    const int b0 = feed.image0(it);
    if (cls) copy_cls<C>(dy, dx, b0, B, c0, E);
    feed.wait(it);
    rows_forward<C>(sy);
    rows_forward<C>(sx);
    __syncthreads();
    if (threadIdx.x == 0 && it > 0) feed.issue(it - 1 + S, &tmdy, &tmx);

    if (q < C::kColItems) {
      const int ioff = g * C::HW * P + p;
      if (s == 0) {
#pragma unroll
        for (int j = 0; j < C::NRE; ++j) {
          const int coff = ioff + (j == 0 ? 0 : W / 2) * P;
          const float4* f = fre + j * C::Hh * P + p;
          V a[H];
          C2 AX[C::Hh], AY[C::Hh];
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = sx[coff + h * W * P];
          RFFT<H, V>::run(a, AX);
#pragma unroll
          for (int h = 0; h < H; ++h) a[h] = sy[coff + h * W * P];
          RFFT<H, V>::run(a, AY);
#pragma unroll
          for (int k = 0; k < C::Hh; ++k) {
            cmac_conj(acc[j * C::Hh + k], AX[k], AY[k]);       // acc += conj(X) * DY
            AY[k] = cmul_conj(AY[k], unpack4(f[k * P]));       // DY * conj(Fh)
          }
          IRFFT<H, V>::run(AY, a);
#pragma unroll
          for (int h = 0; h < H; ++h) sy[coff + h * W * P] = a[h];
        }
      } else {
        const int roff = ioff + s * P, moff = ioff + (W - s) * P;
        const float4* f = fcx + (s - 1) * P + p;
        C2 zx[H], zy[H];
#pragma unroll
        for (int h = 0; h < H; ++h) zx[h] = {sx[roff + h * W * P], sx[moff + h * W * P]};
        CFFT<H, -1, V>::run(zx);
#pragma unroll
        for (int h = 0; h < H; ++h) zy[h] = {sy[roff + h * W * P], sy[moff + h * W * P]};
        CFFT<H, -1, V>::run(zy);
#pragma unroll
        for (int kh = 0; kh < H; ++kh) {
          cmac_conj(acc[kh], zx[kh], zy[kh]);
          zy[kh] = cmul_conj(zy[kh], unpack4(f[kh * C::NCX * P]));
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
    rows_inverse<C>(sy, dx, b0, B, c0, E, cls);
  }

  // ---- partial filter gradient of this CTA: park the accumulators in the (now idle) ring as
  // [g][kh][kw][p], sum the G image slots, apply the half-spectrum weights and the ortho 1/(H*W).
  __syncthreads();
  float4* scr = reinterpret_cast<float4*>(smem);
  constexpr int Wh = C::Wh;
  if (q < C::kColItems) {
    float4* mine = scr + g * H * Wh * P + p;
    if (s == 0) {
#pragma unroll
      for (int j = 0; j < C::NRE; ++j) {
        const int kw = j == 0 ? 0 : W / 2;
#pragma unroll
        for (int k = 0; k < C::Hh; ++k) {
          const C2 a = acc[j * C::Hh + k];
          mine[(k * Wh + kw) * P] = pack4(a);
          if (k > 0 && H - k != k) mine[((H - k) * Wh + kw) * P] = make_float4(a.re.x, a.re.y, -a.im.x, -a.im.y);
        }
      }
    } else {
#pragma unroll
      for (int kh = 0; kh < H; ++kh) mine[(kh * Wh + s) * P] = pack4(acc[kh]);
    }
  }
  __syncthreads();
  float* slab = part + static_cast<size_t>(blockIdx.y) * H * Wh * E * 2;
  for (int i = threadIdx.x; i < H * Wh * P; i += C::NT) {
    const int pp = i % P, e = i / P, kw = e % Wh, c = c0 + 2 * pp;
    if (c >= E) continue;
    float4 t = scr[i];
#pragma unroll
    for (int gg = 1; gg < G; ++gg) {
      const float4 u = scr[gg * H * Wh * P + i];
      t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
    }
    const bool edge = (kw == 0) || (W % 2 == 0 && kw == W / 2);
    const float wgt = (edge ? 1.0f : 2.0f) / static_cast<float>(H * W);
    *reinterpret_cast<float4*>(slab + (static_cast<size_t>(e) * E + c) * 2) =
        make_float4(t.x * wgt, t.z * wgt, t.y * wgt, t.w * wgt);
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

// [B][HW tokens][E] fp32 view of the patch grid (CLS skipped); boxes of 2P channels x HW tokens x G images,
// dense in shared memory; channels past E and images past B read as zero.
static int make_map(CUtensorMap* map, const float* base, int B, int HW, int E, int cls, int chans, int G) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
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
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem);
  if (per_sm < 1) per_sm = 1;
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
  if (int rc = make_map(&tmx, x, B, C::HW, E, cls, 2 * C::P, C::G)) return rc;
  const int tiles = (E + 2 * C::P - 1) / (2 * C::P);
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
  const int tiles = (E + 2 * C::P - 1) / (2 * C::P);
  const int groups = (B + C::G - 1) / C::G;
  return pick_slabs(kern, C::NT, smem, tiles, groups);
}

template <class C>
static int bwd_launch(const float* dy, const float* x, const float* filt, float* dx, float* dfilt, void* ws,
                      size_t ws_bytes, int B, int E, int cls, cudaStream_t st) {
  auto kern = fftmix_ring_bwd_kernel<C>;
  const int S = pick_stages<C>(2);
  const size_t smem = C::smem(S, 2);
  const int tiles = (E + 2 * C::P - 1) / (2 * C::P);
  const int ny = bwd_slabs<C>(B, E);
  const size_t slab_floats = static_cast<size_t>(C::H) * C::Wh * E * 2;
  if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
    return fail(SPECTRE_E_WORKSPACE, "fftmix_bwd: workspace %zu B < required %zu B", ws_bytes,
                slab_floats * sizeof(float) * ny);
  CUtensorMap tmdy, tmx;
  if (int rc = make_map(&tmdy, dy, B, C::HW, E, cls, 2 * C::P, C::G)) return rc;
  if (int rc = make_map(&tmx, x, B, C::HW, E, cls, 2 * C::P, C::G)) return rc;
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
// (H, W, pairs per tile, images per iteration).  P = 16 (128-byte token rows) where E is a multiple of 32,
// P = 8 for the narrow CIFAR/MNIST embeddings that are multiples of 16 only (E = 48: three exact tiles).
#define SB200_RING_SHAPES(F) F(8, 8, 16, 2) F(8, 8, 16, 4) F(8, 8, 8, 4) F(7, 7, 16, 2) F(7, 7, 8, 2) F(7, 7, 8, 4) F(7, 7, 8, 8) F(7, 7, 24, 2) F(7, 7, 24, 4) F(4, 4, 16, 4)

static int pick_pairs(int E) {
  const int forced = env_int("SB200_FFT_RING_P", 0);
  if (forced > 0) return E % (2 * forced) == 0 ? forced : 0;
  return E % 32 == 0 ? 16 : (E % 16 == 0 ? 8 : 0);
}

// images per iteration: the forward likes wider CTAs (8 warps), the backward -- 128 registers per thread -- more of them
// (small batches: the smaller group, so that there are CTAs for every SM)
static int pick_images(int H, int W, int P, bool bwd, int B, int E) {
  const int forced = env_int(bwd ? "SB200_FFT_RING_G_BWD" : "SB200_FFT_RING_G_FWD", env_int("SB200_FFT_RING_G", 0));
  if (forced > 0) return forced;
  const bool small = static_cast<long long>(B) * E / (2 * P) < 8LL * device_sm_count();  // < 2 CTAs of 4 images per SM
  if (H == 8 && W == 8) return P == 16 ? ((bwd || small) ? 2 : 4) : 4;
  if (H == 7 && W == 7) return P == 16 ? 2 : (small ? 2 : 4);   // P = 8 and P = 24 (E = 48 as ONE tile of 192-byte rows)
  return 4;
}

static bool enabled() {
  static const bool on = env_int("SB200_FFT_RING", 1) != 0;
  return on;
}

}  // namespace ring

int fftmix_ring_supported(int H, int W, int E, const void* a, const void* b) {
  if (!ring::enabled() || E % 4 != 0) return 0;
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) return 0;
  const int P = ring::pick_pairs(E);
  if (!P) return 0;
  // every group size pick_images() can return for a compiled (H, W, P) is compiled (a forced one is checked here)
  const int forced = ring::env_int("SB200_FFT_RING_G", 0);
#define F(h, w, p, g) if (H == h && W == w && P == p && (forced <= 0 || forced == g)) return 1;
  SB200_RING_SHAPES(F)
#undef F
  return 0;
}

int fftmix_ring_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls,
                             cudaStream_t st) {
  const int P = ring::pick_pairs(E), G = ring::pick_images(H, W, P, false, B, E);
#define F(h, w, p, g)                                                                                       \
  if (H == h && W == w && P == p && G == g)                                                                 \
    return ring::fwd_launch<ring::Cfg<h, w, p, g>>(static_cast<const float*>(x), filt, static_cast<float*>(y), B, E, \
                                                   cls, st);
  SB200_RING_SHAPES(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_ring: no kernel for grid %dx%d E=%d", H, W, E);
}

int fftmix_ring_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                             size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st) {
  const int P = ring::pick_pairs(E), G = ring::pick_images(H, W, P, true, B, E);
#define F(h, w, p, g)                                                                                              \
  if (H == h && W == w && P == p && G == g)                                                                        \
    return ring::bwd_launch<ring::Cfg<h, w, p, g>>(static_cast<const float*>(dy), static_cast<const float*>(x), filt, \
                                                   static_cast<float*>(dx), dfilt, ws, ws_bytes, B, E, cls, st);
  SB200_RING_SHAPES(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_ring: no kernel for grid %dx%d E=%d", H, W, E);
}

size_t fftmix_ring_bwd_workspace(int B, int H, int W, int E) {
  if (!ring::enabled() || E % 4 != 0) return 0;
  const int P = ring::pick_pairs(E);
  if (!P) return 0;
  const int G = ring::pick_images(H, W, P, true, B, E);
  int ny = 0;
#define F(h, w, p, g) if (H == h && W == w && P == p && G == g) ny = ring::bwd_slabs<ring::Cfg<h, w, p, g>>(B, E);
  SB200_RING_SHAPES(F)
#undef F
  return ny <= 1 ? 0 : static_cast<size_t>(ny) * H * (W / 2 + 1) * E * 2 * sizeof(float);
}

}  // namespace sb200
