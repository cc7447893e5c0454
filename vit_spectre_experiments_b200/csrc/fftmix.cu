// FFT token mixer, forward and backward, as one memory-bound pass each (sm_100a).
//
// Replaces the rfft2 -> (* filter) -> irfft2 composition that occupies the
// reference's mix_layer slot (spectre_vit/models/spectre/spectre.py:50,66;
// conventions of spectre.py:136 and repl/orthogonal_permut.py:23-28) and its
// autograd backward.
//
// Work decomposition
//   CTA tile   = 32 channels (16 channel PAIRS; a thread owns a pair, so a bf16 pair
//                is one 32-bit access and an fp32 pair one 64-bit access) x G images.
//   blockIdx.x = channel tile, blockIdx.y strides over image groups (persistent in y):
//                the filter tile is loaded into shared memory once per CTA and the
//                backward's filter-gradient lives in registers across the image loop.
//   row pass   : thread (image, h, pair) loads W tokens straight from HBM (coalesced
//                across pairs), real FFT in registers, writes W/2+1 bins to smem.
//   col pass   : thread (image, kw, pair) reads H bins from smem, complex FFT, multiplies
//                by the filter, inverse complex FFT, writes back in place.
//   row pass^-1: thread (image, h, pair) reads W/2+1 bins, C2R, stores W tokens to HBM.
// HBM traffic is exactly: x read once, y written once (+ the filter once per CTA, L2-hit).
//
// Shared-memory layout: spectrum entry = float4 (re.c0, re.c1, im.c0, im.c1) for one pair,
// index ((g*H + h)*Wh + kw)*16 + pair, so the 16 pairs of a tile are 256 contiguous bytes
// and every quarter-warp LDS.128/STS.128 covers all 32 banks exactly once.
#include <cstdarg>
#include <cstdlib>

#include "common.cuh"
#include "fft_codelets.cuh"

namespace sb200 {

constexpr int kPairs = 16;  // channel pairs per CTA tile
constexpr int kTileC = 2 * kPairs;

using V = float2;
using C2 = Cx<float2>;

__device__ __forceinline__ float4 pack(const C2& z) { return make_float4(z.re.x, z.re.y, z.im.x, z.im.y); }
__device__ __forceinline__ C2 unpack(const float4& v) { return {make_float2(v.x, v.y), make_float2(v.z, v.w)}; }

// Per-grid launch shape: G images per CTA iteration, NT threads.  Chosen so the column
// pass (G*Wh*16 items) fills whole warps and the row pass (G*H*16 items) takes <= 2 rounds.
template <int H, int W>
struct Shape {
  static constexpr int Wh = W / 2 + 1;
  static constexpr int G_fwd = (H * W <= 64) ? 2 : 1;
  static constexpr int G_bwd = (H * W <= 64) ? 2 : 1;
  static constexpr int round32(int n) { return (n + 31) / 32 * 32; }
  static constexpr int NT_fwd = round32(G_fwd * Wh * kPairs);
  static constexpr int NT_bwd = round32(G_bwd * Wh * kPairs);
};

// ---------------------------------------------------------------- filter tile -> smem
// fs[(kh*Wh + kw)*16 + p] = scale * filt[kh, kw, c0+2p .. c0+2p+1] as (re0, re1, im0, im1)
template <int H, int W, int NT>
__device__ __forceinline__ void load_filter_tile(float4* fs, const float* __restrict__ filt, int c0, int E, float scale) {
  constexpr int Wh = W / 2 + 1;
  for (int i = threadIdx.x; i < H * Wh * kPairs; i += NT) {
    const int p = i % kPairs, e = i / kPairs, c = c0 + 2 * p;
    float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < E) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(filt + (static_cast<size_t>(e) * E + c) * 2));
      f = make_float4(g.x * scale, g.z * scale, g.y * scale, g.w * scale);
    }
    fs[i] = f;
  }
}

// ---------------------------------------------------------------- row passes
template <int H, int W, int G, int NT, class T>
__device__ __forceinline__ void rows_forward(float4* spec, const T* __restrict__ x, int b0, int B, int c0, int E, int cls) {
  constexpr int Wh = W / 2 + 1;
  const int tokens = H * W + cls;
  for (int r = threadIdx.x; r < G * H * kPairs; r += NT) {
    const int p = r % kPairs, gh = r / kPairs, g = gh / H, h = gh % H;
    const int c = c0 + 2 * p, b = b0 + g;
    if (b >= B || c >= E) continue;
    const T* src = x + (static_cast<size_t>(b) * tokens + cls + h * W) * E + c;
    V xin[W];
#pragma unroll
    for (int w = 0; w < W; ++w) xin[w] = PairIO<T>::ld(src + static_cast<size_t>(w) * E);
    C2 X[Wh];
    RFFT<W, V>::run(xin, X);
    float4* dst = spec + (gh * Wh) * kPairs + p;
#pragma unroll
    for (int kw = 0; kw < Wh; ++kw) dst[kw * kPairs] = pack(X[kw]);
  }
}

template <int H, int W, int G, int NT, class T>
__device__ __forceinline__ void rows_inverse(const float4* spec, T* __restrict__ y, int b0, int B, int c0, int E, int cls) {
  constexpr int Wh = W / 2 + 1;
  const int tokens = H * W + cls;
  for (int r = threadIdx.x; r < G * H * kPairs; r += NT) {
    const int p = r % kPairs, gh = r / kPairs, g = gh / H, h = gh % H;
    const int c = c0 + 2 * p, b = b0 + g;
    if (b >= B || c >= E) continue;
    const float4* srcs = spec + (gh * Wh) * kPairs + p;
    C2 X[Wh];
#pragma unroll
    for (int kw = 0; kw < Wh; ++kw) X[kw] = unpack(srcs[kw * kPairs]);
    V out[W];
    IRFFT<W, V>::run(X, out);
    T* dst = y + (static_cast<size_t>(b) * tokens + cls + h * W) * E + c;
#pragma unroll
    for (int w = 0; w < W; ++w) PairIO<T>::st(dst + static_cast<size_t>(w) * E, out[w]);
  }
}

template <int G, int NT, class T>
__device__ __forceinline__ void copy_cls(const T* __restrict__ src, T* __restrict__ dst, int b0, int B, int c0, int E, int tokens) {
  for (int i = threadIdx.x; i < G * kPairs; i += NT) {
    const int p = i % kPairs, b = b0 + i / kPairs, c = c0 + 2 * p;
    if (b < B && c < E) {
      const size_t off = static_cast<size_t>(b) * tokens * E + c;
      PairIO<T>::st(dst + off, PairIO<T>::ld(src + off));
    }
  }
}

// ---------------------------------------------------------------- forward kernel
template <int H, int W, int G, int NT, class T>
__global__ void __launch_bounds__(NT) fftmix_fwd_kernel(const T* __restrict__ x, const float* __restrict__ filt,
                                                        T* __restrict__ y, int B, int E, int cls) {
  constexpr int Wh = W / 2 + 1;
  extern __shared__ float4 smem[];
  float4* fs = smem;                        // [H*Wh][16]
  float4* spec = smem + H * Wh * kPairs;    // [G][H][Wh][16]
  const int c0 = blockIdx.x * kTileC;
  const int tokens = H * W + cls;

  load_filter_tile<H, W, NT>(fs, filt, c0, E, 1.0f / static_cast<float>(H * W));

  for (int b0 = blockIdx.y * G; b0 < B; b0 += gridDim.y * G) {
    rows_forward<H, W, G, NT, T>(spec, x, b0, B, c0, E, cls);
    if (cls) copy_cls<G, NT, T>(x, y, b0, B, c0, E, tokens);
    __syncthreads();

    for (int q = threadIdx.x; q < G * Wh * kPairs; q += NT) {
      const int p = q % kPairs, gk = q / kPairs, kw = gk % Wh, g = gk / Wh;
      if (b0 + g >= B || c0 + 2 * p >= E) continue;
      float4* col = spec + (g * H * Wh + kw) * kPairs + p;
      const float4* fcol = fs + kw * kPairs + p;
      C2 z[H];
#pragma unroll
      for (int h = 0; h < H; ++h) z[h] = unpack(col[h * Wh * kPairs]);
      CFFT<H, -1, V>::run(z);
#pragma unroll
      for (int kh = 0; kh < H; ++kh) z[kh] = cmul(z[kh], unpack(fcol[kh * Wh * kPairs]));
      CFFT<H, 1, V>::run(z);
#pragma unroll
      for (int h = 0; h < H; ++h) col[h * Wh * kPairs] = pack(z[h]);
    }
    __syncthreads();

    rows_inverse<H, W, G, NT, T>(spec, y, b0, B, c0, E, cls);
    __syncthreads();
  }
}

// ---------------------------------------------------------------- backward kernel
// part: [gridDim.y][H][Wh][E][2] fp32 partial filter gradients (one slab per blockIdx.y)
template <int H, int W, int G, int NT, class T>
__global__ void __launch_bounds__(NT) fftmix_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                        const float* __restrict__ filt, T* __restrict__ dx,
                                                        float* __restrict__ part, int B, int E, int cls) {
  constexpr int Wh = W / 2 + 1;
  constexpr int kSpec = G * H * Wh * kPairs;
  constexpr int kColItems = G * Wh * kPairs;
  constexpr int kItems = (kColItems + NT - 1) / NT;  // column items owned by one thread
  extern __shared__ float4 smem[];
  float4* fs = smem;                  // [H*Wh][16]
  float4* specx = fs + H * Wh * kPairs;
  float4* specy = specx + kSpec;
  const int c0 = blockIdx.x * kTileC;
  const int tokens = H * W + cls;

  load_filter_tile<H, W, NT>(fs, filt, c0, E, 1.0f / static_cast<float>(H * W));

  C2 acc[kItems][H];
#pragma unroll
  for (int it = 0; it < kItems; ++it)
#pragma unroll
    for (int kh = 0; kh < H; ++kh) acc[it][kh] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};

  for (int b0 = blockIdx.y * G; b0 < B; b0 += gridDim.y * G) {
    rows_forward<H, W, G, NT, T>(specy, dy, b0, B, c0, E, cls);
    rows_forward<H, W, G, NT, T>(specx, x, b0, B, c0, E, cls);
    if (cls) copy_cls<G, NT, T>(dy, dx, b0, B, c0, E, tokens);
    __syncthreads();

#pragma unroll
    for (int it = 0; it < kItems; ++it) {
      const int q = threadIdx.x + it * NT;
      if (q >= kColItems) continue;
      const int p = q % kPairs, gk = q / kPairs, kw = gk % Wh, g = gk / Wh;
      if (b0 + g >= B || c0 + 2 * p >= E) continue;
      const int off = (g * H * Wh + kw) * kPairs + p;
      float4* colx = specx + off;
      float4* coly = specy + off;
      const float4* fcol = fs + kw * kPairs + p;
      C2 z[H];
      // X column: transform in registers, park it back in shared memory (this thread is
      // the only reader), so that only one column + the accumulators are live at a time.
#pragma unroll
      for (int h = 0; h < H; ++h) z[h] = unpack(colx[h * Wh * kPairs]);
      CFFT<H, -1, V>::run(z);
#pragma unroll
      for (int kh = 0; kh < H; ++kh) colx[kh * Wh * kPairs] = pack(z[kh]);
      // dY column
#pragma unroll
      for (int h = 0; h < H; ++h) z[h] = unpack(coly[h * Wh * kPairs]);
      CFFT<H, -1, V>::run(z);
#pragma unroll
      for (int kh = 0; kh < H; ++kh) {
        cmac_conj(acc[it][kh], unpack(colx[kh * Wh * kPairs]), z[kh]);   // acc += conj(X) * DY
        z[kh] = cmul_conj(z[kh], unpack(fcol[kh * Wh * kPairs]));        // DY * conj(filt)
      }
      CFFT<H, 1, V>::run(z);
#pragma unroll
      for (int h = 0; h < H; ++h) coly[h * Wh * kPairs] = pack(z[h]);
    }
    __syncthreads();

    rows_inverse<H, W, G, NT, T>(specy, dx, b0, B, c0, E, cls);
    __syncthreads();
  }

  // ---- partial filter gradient of this CTA: sum the G image slots through smem, apply the
  // half-spectrum weights c[kw] and the ortho 1/(H*W), write one slab.
#pragma unroll
  for (int it = 0; it < kItems; ++it) {
    const int q = threadIdx.x + it * NT;
    if (q >= kColItems) continue;
    const int p = q % kPairs, gk = q / kPairs, kw = gk % Wh, g = gk / Wh;
    float4* col = specx + (g * H * Wh + kw) * kPairs + p;
#pragma unroll
    for (int kh = 0; kh < H; ++kh) col[kh * Wh * kPairs] = pack(acc[it][kh]);
  }
  __syncthreads();
  float* slab = part + static_cast<size_t>(blockIdx.y) * H * Wh * E * 2;
  for (int i = threadIdx.x; i < H * Wh * kPairs; i += NT) {
    const int p = i % kPairs, e = i / kPairs, kw = e % Wh, c = c0 + 2 * p;
    if (c >= E) continue;
    float4 s = specx[i];
#pragma unroll
    for (int g = 1; g < G; ++g) {
      const float4 t = specx[g * H * Wh * kPairs + i];
      s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
    }
    const bool edge = (kw == 0) || (W % 2 == 0 && kw == W / 2);
    const float wgt = (edge ? 1.0f : 2.0f) / static_cast<float>(H * W);
    // (re0, re1, im0, im1) -> memory order (re0, im0, re1, im1)
    *reinterpret_cast<float4*>(slab + (static_cast<size_t>(e) * E + c) * 2) =
        make_float4(s.x * wgt, s.z * wgt, s.y * wgt, s.w * wgt);
  }
}

// ---------------------------------------------------------------- host side
template <int H, int W>
struct Smem {
  static constexpr int Wh = W / 2 + 1;
  static constexpr size_t fwd = sizeof(float4) * (1 + Shape<H, W>::G_fwd) * H * Wh * kPairs;
  static constexpr size_t bwd = sizeof(float4) * (1 + 2 * Shape<H, W>::G_bwd) * H * Wh * kPairs;
};

// Number of blockIdx.y slabs: fill every SM with as many co-resident CTAs as fit, never
// more slabs than image groups.
template <class K>
static int pick_slabs(K kernel, int threads, size_t smem, int tiles, int groups) {
  const int per_sm = cached_blocks_per_sm(reinterpret_cast<const void*>(kernel), threads, smem);
  int ny = (device_sm_count() * per_sm) / tiles;
  if (ny < 1) ny = 1;
  if (ny > groups) ny = groups;
  return ny;
}

template <int H, int W, class T>
static int fwd_launch(const void* x, const float* filt, void* y, int B, int E, int cls, cudaStream_t st) {
  using S = Shape<H, W>;
  auto kern = fftmix_fwd_kernel<H, W, S::G_fwd, S::NT_fwd, T>;
  constexpr size_t smem = Smem<H, W>::fwd;
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem)) != cudaSuccess)
    return fail(SPECTRE_E_CUDA, "cudaFuncSetAttribute(fwd): %s", cudaGetErrorString(cudaGetLastError()));
  const int tiles = (E + kTileC - 1) / kTileC;
  const int groups = (B + S::G_fwd - 1) / S::G_fwd;
  const int ny = pick_slabs(kern, S::NT_fwd, smem, tiles, groups);
  kern<<<dim3(tiles, ny), S::NT_fwd, smem, st>>>(static_cast<const T*>(x), filt, static_cast<T*>(y), B, E, cls);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_fwd launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

template <int H, int W, class T>
static int bwd_slabs(int B, int E) {
  using S = Shape<H, W>;
  auto kern = fftmix_bwd_kernel<H, W, S::G_bwd, S::NT_bwd, T>;
  ensure_dynamic_smem(reinterpret_cast<const void*>(kern), static_cast<int>(Smem<H, W>::bwd));
  const int tiles = (E + kTileC - 1) / kTileC;
  const int groups = (B + S::G_bwd - 1) / S::G_bwd;
  return pick_slabs(kern, S::NT_bwd, Smem<H, W>::bwd, tiles, groups);
}

template <int H, int W, class T>
static int bwd_launch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                      size_t ws_bytes, int B, int E, int cls, cudaStream_t st) {
  using S = Shape<H, W>;
  constexpr int Wh = W / 2 + 1;
  auto kern = fftmix_bwd_kernel<H, W, S::G_bwd, S::NT_bwd, T>;
  const int tiles = (E + kTileC - 1) / kTileC;
  const int ny = bwd_slabs<H, W, T>(B, E);
  const size_t slab_floats = static_cast<size_t>(H) * Wh * E * 2;
  if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
    return fail(SPECTRE_E_WORKSPACE, "fftmix_bwd: workspace %zu B < required %zu B", ws_bytes,
                slab_floats * sizeof(float) * ny);
  float* part = (ny > 1) ? static_cast<float*>(ws) : dfilt;
  kern<<<dim3(tiles, ny), S::NT_bwd, Smem<H, W>::bwd, st>>>(static_cast<const T*>(dy), static_cast<const T*>(x), filt,
                                                            static_cast<T*>(dx), part, B, E, cls);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_bwd launch: %s", cudaGetErrorString(err));
  last_launch_count() = 1;
  if (ny > 1) {
    launch_reduce_partials(part, dfilt, slab_floats, ny, st);
    err = cudaGetLastError();
    if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_bwd reduce launch: %s", cudaGetErrorString(err));
    last_launch_count() = 2;
  }
  return SPECTRE_OK;
}

// (H, W) grids with compiled kernels
#define SB200_FFT_GRIDS(F) F(4, 4) F(7, 7) F(8, 8) F(14, 14) F(16, 16)

int fftmix_supported(int H, int W) {
#define F(h, w) if (H == h && W == w) return 1;
  SB200_FFT_GRIDS(F)
#undef F
  return 0;
}

// Tensor-core (tcgen05) variants for bf16 activations live in fftmix_tc.cu.
int fftmix_tc_supported(int H, int W, int E);
int fftmix_tc_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls,
                           cudaStream_t st);
int fftmix_tc_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                           size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st);
size_t fftmix_tc_bwd_workspace(int B, int H, int W, int E);

// TMA-fed in-place variants for fp32 activations live in fftmix_ring.cu (SB200_FFT_RING=0 keeps the kernels above).
int fftmix_ring_supported(int H, int W, int E, const void* a, const void* b);
int fftmix_ring_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls,
                             cudaStream_t st);
int fftmix_ring_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                             size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st);
size_t fftmix_ring_bwd_workspace(int B, int H, int W, int E);

// SB200_FFT_TC=0 keeps bf16 on the SIMT butterflies (A/B measurements); default is the tensor-core path.
static bool tc_enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_FFT_TC");
    return !(e && e[0] == '0');
  }();
  return on;
}

int fftmix_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls, int dtype,
                        cudaStream_t st) {
  if (dtype == SPECTRE_DTYPE_BF16 && tc_enabled() && fftmix_tc_supported(H, W, E))
    return fftmix_tc_fwd_dispatch(x, filt, y, B, H, W, E, cls, st);
  if (dtype == SPECTRE_DTYPE_F32 && fftmix_ring_supported(H, W, E, x, y))
    return fftmix_ring_fwd_dispatch(x, filt, y, B, H, W, E, cls, st);
#define F(h, w)                                                                                     \
  if (H == h && W == w)                                                                             \
    return dtype == SPECTRE_DTYPE_F32 ? fwd_launch<h, w, float>(x, filt, y, B, E, cls, st)          \
                                      : fwd_launch<h, w, __nv_bfloat16>(x, filt, y, B, E, cls, st);
  SB200_FFT_GRIDS(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix: no kernel compiled for grid %dx%d", H, W);
}

int fftmix_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                        size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st) {
  if (dtype == SPECTRE_DTYPE_BF16 && tc_enabled() && fftmix_tc_supported(H, W, E))
    return fftmix_tc_bwd_dispatch(dy, x, filt, dx, dfilt, ws, ws_bytes, B, H, W, E, cls, st);
  if (dtype == SPECTRE_DTYPE_F32 && fftmix_ring_supported(H, W, E, dy, x))
    return fftmix_ring_bwd_dispatch(dy, x, filt, dx, dfilt, ws, ws_bytes, B, H, W, E, cls, st);
#define F(h, w)                                                                                                      \
  if (H == h && W == w)                                                                                              \
    return dtype == SPECTRE_DTYPE_F32 ? bwd_launch<h, w, float>(dy, x, filt, dx, dfilt, ws, ws_bytes, B, E, cls, st) \
                                      : bwd_launch<h, w, __nv_bfloat16>(dy, x, filt, dx, dfilt, ws, ws_bytes, B, E, cls, st);
  SB200_FFT_GRIDS(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix: no kernel compiled for grid %dx%d", H, W);
}

// Upper bound over both activation dtypes so the caller can size once per shape.
size_t fftmix_bwd_workspace(int B, int H, int W, int E) {
  int ny = 0;
#define F(h, w)                                                                                 \
  if (H == h && W == w) {                                                                       \
    const int a = bwd_slabs<h, w, float>(B, E), b = bwd_slabs<h, w, __nv_bfloat16>(B, E);       \
    ny = a > b ? a : b;                                                                         \
  }
  SB200_FFT_GRIDS(F)
#undef F
  size_t need = ny <= 1 ? 0 : static_cast<size_t>(ny) * H * (W / 2 + 1) * E * 2 * sizeof(float);
  if (tc_enabled() && fftmix_tc_supported(H, W, E)) {
    const size_t t = fftmix_tc_bwd_workspace(B, H, W, E);
    if (t > need) need = t;
  }
  const size_t r = fftmix_ring_bwd_workspace(B, H, W, E);
  if (r > need) need = r;
  return need;
}

}  // namespace sb200
