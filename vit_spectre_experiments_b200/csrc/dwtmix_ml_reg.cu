// Multi-level Haar DWT token mixer on 4x4 / 8x8 grids with the WHOLE grid of a channel in one thread's registers
// (sm_100a).  Same operator and coefficient layout as dwtmix_ml.cu -- J-level orthonormal 2-D Haar over the token grid
// (DWTForward(J=3, wave="haar", mode="zero"), spectre_vit/repl/dwt_experiments.py:56), a learned real weight per
// coefficient, inverse transform; SURVEY.md section 8f row N4 -- for the CIFAR-shaped grids where H*W <= 64.
//
// dwtmix_ml.cu keeps the tile in shared memory, one element per thread and access, a CTA barrier per level and
// direction, 2-byte global loads for bf16: 13-31 % of the HBM peak (profiles/r01_kernel_bench_dwtml_v10.jsonl).  Here:
//   * the tile [G images][H*W tokens][128 bytes of channels] arrives by ONE TMA box per tensor and iteration into a
//     3-stage ring (cp.async.bulk.tensor + mbarrier); results leave straight from registers (bf16) or by a TMA store
//     from the same slots (fp32) -- whichever measured faster, SB200_DWTML_STG=0/1 forces one;
//   * a thread owns ONE channel of ONE image: it pulls its H*W values into registers (conflict-free: a warp reads 32
//     adjacent channels of a token), runs all levels forward, multiplies by the weights, runs all levels back and
//     stores -- no thread ever reads another thread's element, so the only CTA barrier of an iteration is the one
//     that frees the stage for the next TMA load;
//   * the weight tile is staged with its global loads batched ahead of their use (a dependent load -> scale -> store
//     chain per element was 38 % of all stall samples of the first version: 32 serial L2 round trips per CTA);
//   * butterflies are un-normalised (add / subtract only): a coefficient finalised at level l went through l halvings
//     each way, so the weight tile in shared memory is pre-scaled by 4^-l (and the weight gradient once at the end);
//   * the backward holds U x, U dy and the weight-gradient accumulators of its channel in registers (3 x 64).
// HBM traffic: activations once in, once out; weights once per CTA.
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace sb200 {
namespace mlreg {

using namespace ptx;

template <class T> struct Io;
template <> struct Io<float> {
  static constexpr CUtensorMapDataType kType = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  static __device__ __forceinline__ float ld(const float* p) { return *p; }
  static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
};
template <> struct Io<__nv_bfloat16> {
  static constexpr CUtensorMapDataType kType = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) {
    return __uint_as_float(static_cast<unsigned>(*reinterpret_cast<const unsigned short*>(p)) << 16);
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
};

template <int N_, class T_, int G_, int S_ = 3, int D_ = 0>
struct Cfg {
  static constexpr bool kDirectStore = D_ != 0;   // results leave by per-thread global stores instead of a TMA store
  using T = T_;
  static constexpr int N = N_, G = G_, HW = N * N;
  static constexpr int TC = 128 / static_cast<int>(sizeof(T));  // channels per tile: 128-byte token rows
  static constexpr int NT = G * TC;                             // one thread per (image slot, channel)
  static constexpr int kStageBytes = G * HW * TC * static_cast<int>(sizeof(T));  // per tensor
  static constexpr int kWgtBytes = HW * TC * 4;
  static constexpr int S = S_;
  static_assert(kStageBytes % 128 == 0, "TMA slots stay 128-byte aligned");
  static constexpr size_t smem(int tensors) { return static_cast<size_t>(S) * tensors * kStageBytes + kWgtBytes + 128; }
};

// number of levels that touch grid point (r, q): it is a block corner of level l iff r, q are multiples of 2^(l-1)
// (branch-free; grids here have at most 3 levels)
__host__ __device__ inline int levels_of(int r, int q, int levels) {
  const int rq = r | q;
  return 1 + ((levels >= 2 && (rq & 1) == 0) ? 1 : 0) + ((levels >= 3 && (rq & 3) == 0) ? 1 : 0);
}

// one un-normalised level on the register grid; layout of the results as in dwtmix_ml.cu:
//   forward: (ll, lowW/highH, highW/lowH, hh) -> (a, c, b, d);   inverse reads (a, c, b, d) and writes (a, b, c, d)
template <int N, int L, bool INV>
__device__ __forceinline__ void level(float (&v)[N * N]) {
  constexpr int step = 1 << L, half = step >> 1;
#pragma unroll
  for (int r0 = 0; r0 < N; r0 += step)
#pragma unroll
    for (int q0 = 0; q0 < N; q0 += step) {
      const int pa = r0 * N + q0, pb = pa + half, pc = pa + half * N, pd = pc + half;
      const float a = v[pa], b = v[INV ? pc : pb], c = v[INV ? pb : pc], d = v[pd];
      const float p = a + b, q = a - b, r = c + d, s = c - d;
      v[pa] = p + r;
      v[INV ? pb : pc] = p - r;
      v[INV ? pc : pb] = q + s;
      v[pd] = q - s;
    }
}

template <int N>
__device__ __forceinline__ void forward_levels(float (&v)[N * N], int levels) {
  level<N, 1, false>(v);
  if constexpr (N >= 4) { if (levels >= 2) level<N, 2, false>(v); }
  if constexpr (N >= 8) { if (levels >= 3) level<N, 3, false>(v); }
}
template <int N>
__device__ __forceinline__ void inverse_levels(float (&v)[N * N], int levels) {
  if constexpr (N >= 8) { if (levels >= 3) level<N, 3, true>(v); }
  if constexpr (N >= 4) { if (levels >= 2) level<N, 2, true>(v); }
  level<N, 1, true>(v);
}

// part: [gridDim.y][H*W][E] fp32 partial weight gradients (one slab per blockIdx.y), backward only
template <class C, bool BWD>
__global__ void __launch_bounds__(C::NT, BWD ? 2 : 4)
dwtml_reg_kernel(const __grid_constant__ CUtensorMap tm_in, const __grid_constant__ CUtensorMap tm_x,
                 const __grid_constant__ CUtensorMap tm_out, const typename C::T* __restrict__ in,
                 const float* __restrict__ wgt, typename C::T* __restrict__ out, float* __restrict__ part, int B, int E,
                 int cls, int levels, int subband) {
  using T = typename C::T;
  constexpr int N = C::N, HW = C::HW, TC = C::TC, G = C::G, S = C::S;
  constexpr int TENS = BWD ? 2 : 1;
  extern __shared__ __align__(128) unsigned char smem[];
  float* ws = reinterpret_cast<float*>(smem + static_cast<size_t>(S) * TENS * C::kStageBytes);  // [HW][TC], pre-scaled
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(S) * TENS * C::kStageBytes + C::kWgtBytes);
  const int c0 = blockIdx.x * TC;
  // weight / weight-gradient row of grid point pos: in place on the lattice [H, W, E], or (single-level mixer of
  // dwtmix.cu, levels == 1) sub-band major [4, H/2, W/2, E] with (ll, lowW/highH, highW/lowH, hh) = block points (a, c, b, d)
  auto wrow = [&](int pos) {
    if (!subband) return pos;
    const int r = pos / N, q = pos % N;
    return (((r & 1) + 2 * (q & 1)) * (N / 2) + (r >> 1)) * (N / 2) + (q >> 1);
  };
  const int groups = (B + G - 1) / G;
  const int n_it = (groups - static_cast<int>(blockIdx.y) + static_cast<int>(gridDim.y) - 1) / static_cast<int>(gridDim.y);
  auto stage = [&](int it, int tensor) {
    return reinterpret_cast<T*>(smem + static_cast<size_t>((it % S) * TENS + tensor) * C::kStageBytes);
  };
  auto image0 = [&](int it) { return (static_cast<int>(blockIdx.y) + it * static_cast<int>(gridDim.y)) * G; };
  auto issue = [&](int it) {
    if (it >= n_it) return;
    uint64_t* bar = full + it % S;
    mbar_arrive_expect_tx(bar, TENS * C::kStageBytes);
    tma_load_3d(stage(it, 0), &tm_in, bar, c0, 0, image0(it));
    if constexpr (BWD) tma_load_3d(stage(it, 1), &tm_x, bar, c0, 0, image0(it));
  };

  if (threadIdx.x == 0) {
    prefetch_tensormap(&tm_in);
    prefetch_tensormap(&tm_out);
    if constexpr (BWD) prefetch_tensormap(&tm_x);
    for (int s = 0; s < S; ++s) mbar_init(full + s, 1);
    mbar_fence_init();
    for (int s = 0; s < S; ++s) issue(s);
  }
  {   // weight tile -> shared memory, pre-scaled.  Loads in batches of 8 per thread BEFORE their first use: one at a
      // time (a dependent load -> multiply -> store chain per element) cost 32 serial L2 round trips per CTA, 38 % of
      // all stall samples of the kernel (profiles/r02_ncu_dwtml_v11.txt)
    constexpr int kPer = HW * TC / C::NT, kBatch = kPer % 8 == 0 ? 8 : 4;
    static_assert(HW * TC % C::NT == 0 && kPer % kBatch == 0, "weight tile splits evenly over the threads");
#pragma unroll 1
    for (int k0 = 0; k0 < kPer; k0 += kBatch) {
      float wv[kBatch];
#pragma unroll
      for (int k = 0; k < kBatch; ++k) {
        const int i = threadIdx.x + (k0 + k) * C::NT, pos = i / TC, cc = c0 + i % TC;
        wv[k] = cc < E ? __ldg(wgt + static_cast<size_t>(wrow(pos)) * E + cc) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < kBatch; ++k) {
        const int i = threadIdx.x + (k0 + k) * C::NT, pos = i / TC;
        const int l = levels_of(pos / N, pos % N, levels);
        ws[i] = wv[k] * (1.0f / static_cast<float>(1 << (2 * l)));
      }
    }
  }
  const int g = threadIdx.x / TC, c = threadIdx.x % TC;
  const int toff = g * HW * TC + c;   // element offset of this thread's (image slot, channel) in a stage
  const size_t img_elems = static_cast<size_t>(HW + cls) * E;
  float acc[BWD ? HW : 1];
  if constexpr (BWD) {
#pragma unroll
    for (int i = 0; i < HW; ++i) acc[i] = 0.f;
  }
  __syncthreads();

#pragma unroll 1
  for (int it = 0; it < n_it; ++it) {
    T* sa = stage(it, 0);
    const int b0 = image0(it);
    // CLS passes through: the load is issued here, its store waits until the end of the iteration (a store right
    // behind the load would park the warp for a DRAM round trip in every iteration)
    const bool cls_live = cls && b0 + g < B && c0 + c < E;
    const size_t cls_off = static_cast<size_t>(b0 + g) * img_elems + c0 + c;
    T cls_val;
    if (cls_live) cls_val = in[cls_off];
    mbar_wait(full + it % S, (it / S) & 1);

    float v[HW];
    if constexpr (BWD) {
      const T* sx = stage(it, 1);
      float u[HW];
#pragma unroll
      for (int i = 0; i < HW; ++i) u[i] = Io<T>::ld(sx + toff + i * TC);
      forward_levels<N>(u, levels);
#pragma unroll
      for (int i = 0; i < HW; ++i) v[i] = Io<T>::ld(sa + toff + i * TC);
      forward_levels<N>(v, levels);
#pragma unroll
      for (int i = 0; i < HW; ++i) acc[i] = fmaf(u[i], v[i], acc[i]);
    } else {
#pragma unroll
      for (int i = 0; i < HW; ++i) v[i] = Io<T>::ld(sa + toff + i * TC);
      forward_levels<N>(v, levels);
    }
#pragma unroll
    for (int i = 0; i < HW; ++i) v[i] *= ws[i * TC + c];
    inverse_levels<N>(v, levels);
    if constexpr (C::kDirectStore) {
      if (b0 + g < B) {
        T* dst = out + static_cast<size_t>(b0 + g) * img_elems + static_cast<size_t>(cls) * E + c0 + c;
#pragma unroll
        for (int i = 0; i < HW; ++i) Io<T>::st(dst + static_cast<size_t>(i) * E, v[i]);
      }
      if (cls_live) out[cls_off] = cls_val;
      __syncthreads();                       // every thread has read its slots: the stage is free
      if (threadIdx.x == 0) issue(it + S);
    } else {
#pragma unroll
      for (int i = 0; i < HW; ++i) Io<T>::st(sa + toff + i * TC, v[i]);
      fence_proxy_async_smem();   // generic-proxy writes above, async-proxy (TMA store) reads below
      if (cls_live) out[cls_off] = cls_val;
      __syncthreads();
      if (threadIdx.x == 0) {
        tma_store_3d(&tm_out, sa, c0, 0, b0);
        tma_store_commit();
        if (it >= 1) {
          tma_store_wait_read<1>();   // every store but the one just issued has read its slots: stage (it-1) % S is free
          issue(it - 1 + S);
        }
      }
    }
  }
  if (threadIdx.x == 0) tma_store_wait_all<0>();

  if constexpr (BWD) {
    // sum the G image slots through the idle ring, undo the 4^-l of the un-normalised butterflies, write one slab
    __syncthreads();
    float* scr = reinterpret_cast<float*>(smem);   // [G][HW][TC]
    static_assert(static_cast<size_t>(G) * HW * TC * 4 <= static_cast<size_t>(S) * 2 * C::kStageBytes, "scratch fits the ring");
#pragma unroll
    for (int i = 0; i < HW; ++i) scr[(g * HW + i) * TC + c] = acc[i];
    __syncthreads();
    float* slab = part + static_cast<size_t>(blockIdx.y) * HW * E;
    for (int i = threadIdx.x; i < HW * TC; i += C::NT) {
      const int pos = i / TC, cc = c0 + i % TC;
      if (cc >= E) continue;
      float t = scr[i];
#pragma unroll
      for (int gg = 1; gg < G; ++gg) t += scr[gg * HW * TC + i];
      const int l = levels_of(pos / N, pos % N, levels);
      slab[static_cast<size_t>(wrow(pos)) * E + cc] = t * (1.0f / static_cast<float>(1 << (2 * l)));
    }
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

// [B][HW tokens][E] view of the patch grid (CLS skipped); boxes of TC channels x HW tokens x G images, dense in shared
// memory; images past B read as zero and are dropped on store.
template <class C>
static int make_map(CUtensorMap* map, const void* base, int B, int E, int cls) {
  using T = typename C::T;
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
  bind_primary_context();
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(E), static_cast<cuuint64_t>(C::HW), static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(E) * sizeof(T),
                                 static_cast<cuuint64_t>(C::HW + cls) * E * sizeof(T)};
  const cuuint32_t box[3] = {static_cast<cuuint32_t>(C::TC), static_cast<cuuint32_t>(C::HW), static_cast<cuuint32_t>(C::G)};
  const cuuint32_t estr[3] = {1, 1, 1};
  void* gptr = const_cast<T*>(static_cast<const T*>(base) + static_cast<size_t>(cls) * E);
  const CUresult r = enc(map, Io<T>::kType, 3, gptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled(dwtml) failed (%d)", static_cast<int>(r));
  return SPECTRE_OK;
}

template <class C, bool BWD>
static int slabs(int B, int E, int max_slabs) {
  auto kern = dwtml_reg_kernel<C, BWD>;
  const size_t smem = C::smem(BWD ? 2 : 1);
  ensure_dynamic_smem(reinterpret_cast<const void*>(kern), static_cast<int>(smem));
  const int per_sm = cached_blocks_per_sm(reinterpret_cast<const void*>(kern), C::NT, smem);
  const int tiles = E / C::TC, groups = (B + C::G - 1) / C::G;
  int ny = device_sm_count() * per_sm / tiles;
  if (ny < 1) ny = 1;
  if (ny > groups) ny = groups;
  if (ny > max_slabs) ny = max_slabs;   // the workspace the caller sized with spectre_dwtmix_ml_bwd_workspace_bytes
  return ny;
}

template <class C>
static int launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                  size_t ws_bytes, int B, int E, int cls, int levels, int max_slabs, int subband, cudaStream_t st) {
  using T = typename C::T;
  // runtime-API calls first: they bind the primary context to this thread (an autograd worker may not have one yet),
  // which the driver-API tensor-map encoder needs (CUDA_ERROR_INVALID_CONTEXT otherwise)
  const int ny = backward ? slabs<C, true>(B, E, max_slabs) : slabs<C, false>(B, E, 1 << 30);
  CUtensorMap tin, tx, tout;
  if (int rc = make_map<C>(&tin, in, B, E, cls)) return rc;
  if (int rc = make_map<C>(&tout, out, B, E, cls)) return rc;
  if (int rc = make_map<C>(&tx, backward ? xsave : in, B, E, cls)) return rc;
  const int tiles = E / C::TC;
  last_launch_count() = 1;
  if (!backward) {
    dwtml_reg_kernel<C, false><<<dim3(tiles, ny), C::NT, C::smem(1), st>>>(
        tin, tx, tout, static_cast<const T*>(in), w, static_cast<T*>(out), nullptr, B, E, cls, levels, subband);
  } else {
    const size_t slab_floats = static_cast<size_t>(C::HW) * E;
    if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
      return fail(SPECTRE_E_WORKSPACE, "dwtmix_ml_bwd: workspace %zu B < required %zu B", ws_bytes,
                  slab_floats * sizeof(float) * ny);
    float* part = ny > 1 ? static_cast<float*>(ws) : dw;
    dwtml_reg_kernel<C, true><<<dim3(tiles, ny), C::NT, C::smem(2), st>>>(
        tin, tx, tout, static_cast<const T*>(in), w, static_cast<T*>(out), part, B, E, cls, levels, subband);
    if (ny > 1) {
      launch_reduce_partials(part, dw, slab_floats, ny, st);
      last_launch_count() = 2;
    }
  }
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "dwtmix_ml (register kernel) launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

static bool enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_DWTML_REG");
    return !(e && e[0] == '0');
  }();
  return on;
}

}  // namespace mlreg

// 4x4 / 8x8 grids, E a multiple of the 128-byte tile (64 bf16 / 32 fp32 channels)
int dwtml_reg_supported(int H, int W, int E, int dtype) {
  if (!mlreg::enabled() || H != W || (H != 4 && H != 8)) return 0;
  return E % (dtype == SPECTRE_DTYPE_F32 ? 32 : 64) == 0;
}

// image slabs (partial weight-gradient slabs) the backward would like for this shape
int dwtml_reg_bwd_slabs(int B, int H, int E, int dtype) {
  using namespace mlreg;
  if (dtype == SPECTRE_DTYPE_F32)
    return H == 8 ? slabs<Cfg<8, float, 4, 3, 0>, true>(B, E, 1 << 30) : slabs<Cfg<4, float, 4, 3, 0>, true>(B, E, 1 << 30);
  return H == 8 ? slabs<Cfg<8, __nv_bfloat16, 2, 3, 1>, true>(B, E, 1 << 30)
                : slabs<Cfg<4, __nv_bfloat16, 2, 3, 1>, true>(B, E, 1 << 30);
}

int dwtml_reg_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                     size_t ws_bytes, int B, int H, int E, int cls, int levels, int dtype, int max_slabs, int subband,
                     cudaStream_t st_) {
  using namespace mlreg;
  // results leave by per-thread global stores (bf16: forward 0.0328 vs 0.0338 ms) or by a TMA store from the stage
  // (fp32: 0.0524 / 0.0745 vs 0.0532 / 0.0767 ms); SB200_DWTML_STG=0/1 forces one; profiles/r02_kernel_bench_dwtml_v11.txt
  static const int forced = [] { const char* e = getenv("SB200_DWTML_STG"); return e ? atoi(e) : -1; }();
  const int direct = forced >= 0 ? forced : (dtype == SPECTRE_DTYPE_BF16 ? 1 : 0);
#define GO(n, T, g)                                                                                                       \
  return direct ? launch<Cfg<n, T, g, 3, 1>>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, E, cls, levels, max_slabs, subband, st_) \
                : launch<Cfg<n, T, g, 3, 0>>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, E, cls, levels, max_slabs, subband, st_)
  if (dtype == SPECTRE_DTYPE_F32) {
    if (H == 8) GO(8, float, 4);
    GO(4, float, 4);
  }
  if (H == 8) GO(8, __nv_bfloat16, 2);
  GO(4, __nv_bfloat16, 2);
#undef GO
}

}  // namespace sb200
