// Multi-level Haar DWT token mixer, forward and backward (sm_100a).
//
// J-level orthonormal 2-D Haar over the token grid per channel (the reference's only DWT call site is
// DWTForward(J=3, wave="haar", mode="zero"), spectre_vit/repl/dwt_experiments.py:56), multiply every coefficient by a
// learned real weight, inverse transform; CLS passes through.
//
// Coefficients live IN PLACE on the token lattice: level l (l = 1..J) works on the 2x2 blocks of the points whose
// coordinates are multiples of 2^(l-1); with a b / c d the block's points (b = next along W, c = next along H) it
// leaves LL at a, "low W / high H" at c, "high W / low H" at b, HH at d, and the next level recurses on the LL points.
// The weight (and its gradient) is float32 [H, W, E] in that layout.  The transform of a level is the same symmetric
// orthogonal 4x4 butterfly as the single-level kernel, so the inverse is the same code run from level J down to 1,
// and the backward is dx = T^-1(w * T dy), dw = sum_b (T x_b) * (T dy_b).
//
// One CTA owns a 32-channel tile of one image at a time: the H*W x 32 grid tile sits in shared memory as fp32 (8 KB
// for 8x8, 32 KB for 16x16; the backward holds x and dy: twice that), levels are separated by CTA barriers, weight
// gradients accumulate in registers (a thread always meets the same (position, channel) elements) and leave as one
// partial slab per image slab, reduced in a fixed order by launch_reduce_partials (deterministic).
#include <type_traits>

#include "common.cuh"

namespace sb200 {

// register-resident variant for 4x4 / 8x8 grids (dwtmix_ml_reg.cu); SB200_DWTML_REG=0 keeps the kernels below
int dwtml_reg_supported(int H, int W, int E, int dtype);
int dwtml_reg_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                     size_t ws_bytes, int B, int H, int E, int cls, int levels, int dtype, int max_slabs, int subband,
                     cudaStream_t st);

namespace {

constexpr int kMlTile = 32;      // channels per CTA
constexpr int kMlThreads = 256;
constexpr int kMlMaxPerThread = 32;   // H*W*32 / 256 with H*W <= 256 (kernels are instantiated for 4, 8, 16, 32)

template <class T>
__device__ __forceinline__ float ml_ld(const T* p);
template <>
__device__ __forceinline__ float ml_ld<float>(const float* p) { return __ldcs(p); }
template <>
__device__ __forceinline__ float ml_ld<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __uint_as_float(static_cast<unsigned>(__ldcs(reinterpret_cast<const unsigned short*>(p))) << 16);
}
__device__ __forceinline__ void ml_st(float* p, float v) { __stcs(p, v); }
__device__ __forceinline__ void ml_st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// one level of the in-place transform on buffer(s) in shared memory; `inverse` reads (ll, s1, s2, hh) from (a, c, b, d)
template <bool kTwo>
__device__ __forceinline__ void ml_level(float* A, float* Bf, int H, int W, int level, bool inverse) {
  const int step = 1 << level, half = step >> 1;
  const int bw = W >> level, items = (H >> level) * bw * kMlTile;
  for (int it = threadIdx.x; it < items; it += kMlThreads) {
    const int c = it % kMlTile, blk = it / kMlTile;
    const int br = blk / bw;
    const int r0 = br << level, q0 = (blk - br * bw) << level;
    const int pa = (r0 * W + q0) * kMlTile + c, pb = pa + half * kMlTile, pc = pa + half * W * kMlTile,
              pd = pc + half * kMlTile;
    const int i1 = inverse ? pc : pb, i2 = inverse ? pb : pc;   // second / third operand of the butterfly
#pragma unroll
    for (int t = 0; t < (kTwo ? 2 : 1); ++t) {
      float* M = t ? Bf : A;
      const float a = M[pa], b = M[i1], cc = M[i2], d = M[pd];
      const float p = a + b, q = a - b, r = cc + d, s = cc - d;
      const float o0 = 0.5f * (p + r), o1 = 0.5f * (p - r), o2 = 0.5f * (q + s), o3 = 0.5f * (q - s);
      // forward: (ll, s1, s2, hh) -> (a, c, b, d);  inverse: (a, b, c, d) -> (a, b, c, d)
      M[pa] = o0;
      M[inverse ? pb : pc] = o1;
      M[inverse ? pc : pb] = o2;
      M[pd] = o3;
    }
  }
  __syncthreads();
}

template <class T, bool kBackward, int KMAX>   // KMAX >= elements per thread = H*W*32 / 256
__global__ void __launch_bounds__(kMlThreads, KMAX <= 8 ? 4 : 2) dwtml_kernel(const T* __restrict__ in, const T* __restrict__ xsave,
                                                           const float* __restrict__ wgt, T* __restrict__ out,
                                                           float* __restrict__ part, int B, int H, int W, int E, int cls,
                                                           int levels) {
  extern __shared__ float sm[];
  const int HW = H * W, n = HW * kMlTile, tokens = HW + cls;
  float* A = sm;                       // the tile being mixed (x forward, dy backward)
  float* X = sm + n;                   // backward: the saved x
  const int c0 = blockIdx.x * kMlTile;
  float acc[KMAX];
  if (kBackward) {
#pragma unroll
    for (int k = 0; k < KMAX; ++k) acc[k] = 0.f;
  }

  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    const size_t base = static_cast<size_t>(b) * tokens * E;
    {   // all of the thread's loads are issued before the first shared-memory store (one DRAM round trip per image,
        // not one per element)
      float v[KMAX], vx[kBackward ? KMAX : 1];
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const int idx = threadIdx.x + k * kMlThreads;
        const int c = c0 + idx % kMlTile, pos = idx / kMlTile;
        const bool live = idx < n && c < E;
        const size_t off = base + static_cast<size_t>(cls + pos) * E + c;
        v[k] = live ? ml_ld<T>(in + off) : 0.f;
        if (kBackward) vx[k] = live ? ml_ld<T>(xsave + off) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const int idx = threadIdx.x + k * kMlThreads;
        if (idx < n) {
          A[idx] = v[k];
          if (kBackward) X[idx] = vx[k];
        }
      }
    }
    if (cls && threadIdx.x < kMlTile && c0 + threadIdx.x < E)
      ml_st(out + base + c0 + threadIdx.x, ml_ld<T>(in + base + c0 + threadIdx.x));
    __syncthreads();
    for (int l = 1; l <= levels; ++l) ml_level<kBackward>(A, X, H, W, l, false);
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const int idx = threadIdx.x + k * kMlThreads;
      if (idx < n) {
        const int c = c0 + idx % kMlTile, pos = idx / kMlTile;
        const float wv = c < E ? __ldg(wgt + static_cast<size_t>(pos) * E + c) : 0.f;
        const float v = A[idx];
        if (kBackward) acc[k] = fmaf(X[idx], v, acc[k]);
        A[idx] = v * wv;
      }
    }
    __syncthreads();
    for (int l = levels; l >= 1; --l) ml_level<false>(A, X, H, W, l, true);
    for (int idx = threadIdx.x; idx < n; idx += kMlThreads) {
      const int c = c0 + idx % kMlTile, pos = idx / kMlTile;
      if (c < E) ml_st(out + base + static_cast<size_t>(cls + pos) * E + c, A[idx]);
    }
    __syncthreads();   // the tile is reloaded for the next image
  }

  if (kBackward) {
    float* slab = part + static_cast<size_t>(blockIdx.y) * HW * E;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) {
      const int idx = threadIdx.x + k * kMlThreads;
      if (idx < n) {
        const int c = c0 + idx % kMlTile, pos = idx / kMlTile;
        if (c < E) slab[static_cast<size_t>(pos) * E + c] = acc[k];
      }
    }
  }
}

int ml_slabs(int B, int E) {
  const int ctas_x = (E + kMlTile - 1) / kMlTile;
  int ny = (device_sm_count() * 8 + ctas_x - 1) / ctas_x;
  if (ny < 1) ny = 1;
  if (ny > B) ny = B;
  return ny;
}

template <class T>
int ml_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
              size_t ws_bytes, int B, int H, int W, int E, int cls, int levels, cudaStream_t st) {
  const int ny = ml_slabs(B, E);
  const dim3 grid((E + kMlTile - 1) / kMlTile, ny);
  const size_t smem = static_cast<size_t>(backward ? 2 : 1) * H * W * kMlTile * sizeof(float);
  last_launch_count() = 1;
  const int per = (H * W * kMlTile + kMlThreads - 1) / kMlThreads;
  float* part = nullptr;
  if (backward) {
    const size_t slab_floats = static_cast<size_t>(H) * W * E;
    if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
      return fail(SPECTRE_E_WORKSPACE, "dwtmix_ml_bwd: workspace %zu B < required %zu B", ws_bytes,
                  slab_floats * sizeof(float) * ny);
    part = ny > 1 ? static_cast<float*>(ws) : dw;
  }
  auto go = [&](auto kmax) {
    constexpr int K = decltype(kmax)::value;
    if (backward) {
      ensure_dynamic_smem(reinterpret_cast<const void*>(dwtml_kernel<T, true, K>), 64 * 1024);
      dwtml_kernel<T, true, K><<<grid, kMlThreads, smem, st>>>(static_cast<const T*>(in), static_cast<const T*>(xsave), w,
                                                               static_cast<T*>(out), part, B, H, W, E, cls, levels);
    } else {
      ensure_dynamic_smem(reinterpret_cast<const void*>(dwtml_kernel<T, false, K>), 64 * 1024);
      dwtml_kernel<T, false, K><<<grid, kMlThreads, smem, st>>>(static_cast<const T*>(in), nullptr, w, static_cast<T*>(out),
                                                                nullptr, B, H, W, E, cls, levels);
    }
  };
  if (per <= 4) go(std::integral_constant<int, 4>{});
  else if (per <= 8) go(std::integral_constant<int, 8>{});
  else if (per <= 16) go(std::integral_constant<int, 16>{});
  else go(std::integral_constant<int, kMlMaxPerThread>{});
  if (backward && ny > 1) {
    launch_reduce_partials(part, dw, static_cast<size_t>(H) * W * E, ny, st);
    last_launch_count() = 2;
  }
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "dwtmix_ml launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

int ml_check(const char* fn, int B, int H, int W, int E, int has_cls, int levels, int dtype) {
  if (B < 1 || H < 2 || W < 2 || E < 4 || (E & 3) || (has_cls != 0 && has_cls != 1))
    return fail(SPECTRE_E_SHAPE, "%s: bad shape B=%d H=%d W=%d E=%d has_cls=%d (need B>=1, E a multiple of 4)", fn, B, H, W,
                E, has_cls);
  if (levels < 1 || levels > 4 || (H % (1 << levels)) || (W % (1 << levels)))
    return fail(SPECTRE_E_SHAPE, "%s: levels=%d needs 1 <= levels <= 4 and H, W multiples of 2^levels (got %dx%d)", fn,
                levels, H, W);
  if (H * W > 256) return fail(SPECTRE_E_SHAPE, "%s: grid %dx%d exceeds 256 tokens", fn, H, W);
  if (dtype != SPECTRE_DTYPE_F32 && dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, dtype);
  return SPECTRE_OK;
}

}  // namespace
}  // namespace sb200

using namespace sb200;

extern "C" {

int spectre_dwtmix_ml_fwd(const void* x, const float* w, void* y, int B, int H, int W, int E, int has_cls, int levels,
                          int act_dtype, void* stream) {
  if (!x || !w || !y) return fail(SPECTRE_E_NULL, "spectre_dwtmix_ml_fwd: null pointer");
  if (int rc = ml_check("spectre_dwtmix_ml_fwd", B, H, W, E, has_cls, levels, act_dtype)) return rc;
  if (!aligned16(x) || !aligned16(w) || !aligned16(y))
    return fail(SPECTRE_E_ALIGN, "spectre_dwtmix_ml_fwd: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dwtml_reg_supported(H, W, E, act_dtype)) {
    const int rc = dwtml_reg_launch(false, x, nullptr, w, y, nullptr, nullptr, 0, B, H, E, has_cls, levels, act_dtype,
                                    ml_slabs(B, E), 0, st);
    if (rc == SPECTRE_OK) last_error().clear();
    return rc;
  }
  const int rc = act_dtype == SPECTRE_DTYPE_F32
                     ? ml_launch<float>(false, x, nullptr, w, y, nullptr, nullptr, 0, B, H, W, E, has_cls, levels, st)
                     : ml_launch<__nv_bfloat16>(false, x, nullptr, w, y, nullptr, nullptr, 0, B, H, W, E, has_cls, levels, st);
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

int spectre_dwtmix_ml_bwd(const void* dy, const void* x, const float* w, void* dx, float* dw, void* workspace,
                          size_t workspace_bytes, int B, int H, int W, int E, int has_cls, int levels, int act_dtype,
                          void* stream) {
  if (!dy || !x || !w || !dx || !dw) return fail(SPECTRE_E_NULL, "spectre_dwtmix_ml_bwd: null pointer");
  if (int rc = ml_check("spectre_dwtmix_ml_bwd", B, H, W, E, has_cls, levels, act_dtype)) return rc;
  if (!aligned16(dy) || !aligned16(x) || !aligned16(w) || !aligned16(dx) || !aligned16(dw) || !aligned16(workspace))
    return fail(SPECTRE_E_ALIGN, "spectre_dwtmix_ml_bwd: pointers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t wsb = workspace ? workspace_bytes : 0;
  if (dwtml_reg_supported(H, W, E, act_dtype)) {
    const int rc = dwtml_reg_launch(true, dy, x, w, dx, dw, workspace, wsb, B, H, E, has_cls, levels, act_dtype,
                                    ml_slabs(B, E), 0, st);
    if (rc == SPECTRE_OK) last_error().clear();
    return rc;
  }
  const int rc = act_dtype == SPECTRE_DTYPE_F32
                     ? ml_launch<float>(true, dy, x, w, dx, dw, workspace, wsb, B, H, W, E, has_cls, levels, st)
                     : ml_launch<__nv_bfloat16>(true, dy, x, w, dx, dw, workspace, wsb, B, H, W, E, has_cls, levels, st);
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

size_t spectre_dwtmix_ml_bwd_workspace_bytes(int B, int H, int W, int E) {
  if (B < 1 || E < 4 || H < 2 || W < 2) return 0;
  const int ny = ml_slabs(B, E);
  return ny <= 1 ? 0 : static_cast<size_t>(ny) * H * W * E * sizeof(float);
}

}  // extern "C"
