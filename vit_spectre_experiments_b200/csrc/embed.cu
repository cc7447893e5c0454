// SpectralPatchEmbed around its one GEMM (spectre_vit/models/spectre/spectre.py:124-156).
//
// Re(rfft2(patch, ortho)), the per-frequency gains and the projection are all linear in the patch, so the host folds
// them into one matrix M [C*P*P, E] and the module is   tokens = patches @ M + b ;  out = dropout(cat(cls, tokens) + pos).
// The GEMM stays a library call; these kernels are what surrounds it, one pass each instead of the stock chain
// (unfold + permute copy + cast | cat copy + fp32 add + dropout + cast, and their autograd mirrors):
//   patch_gather      img fp32 [B, C, Hi, Wi]  ->  A [B * nH * nW, C * P * P] in the stream dtype   (spectre.py:127-131, 144)
//   embed_finish_fwd  out[b, 0] = cls + pos[0];  out[b, 1 + i] = tok[b, i] + pos[1 + i];  counter-based dropout  (:151-156)
//   embed_finish_bwd  dtok[b, i] = mask * dout[b, 1 + i];  dpos[t] = sum_b mask * dout[b, t]   (dcls = dpos[0] and
//                     dbias = sum_{t >= 1} dpos[t] follow on the host); per-slab partials summed in a fixed order.
#include "normact_common.cuh"

namespace sb200 {
namespace embed {

using na::Args;

template <class T> struct V4;
template <> struct V4<float> {
  static __device__ __forceinline__ void st(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <> struct V4<__nv_bfloat16> {
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float4 v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(na::bf_pack(v.x, v.y), na::bf_pack(v.z, v.w));
  }
};

// one thread = 4 consecutive pixels of one image row (P % 4 == 0): 16-byte reads that are contiguous across the
// threads of a warp; 4 (bf16: 8-byte) / (fp32: 16-byte) writes that fill whole 32-byte sectors of A
template <class T>
__global__ void __launch_bounds__(256) patch_gather_kernel(const float* __restrict__ img, T* __restrict__ out, int B, int C,
                                                           int Hi, int Wi, int P, int nH, int nW) {
  const int w4 = nW * P / 4;                          // 4-pixel groups per (cropped) image row
  const long long total = static_cast<long long>(B) * C * nH * P * w4;
  const int K = C * P * P;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int x4 = static_cast<int>(i % w4);
    long long r = i / w4;
    const int y = static_cast<int>(r % (nH * P));
    r /= nH * P;
    const int c = static_cast<int>(r % C), b = static_cast<int>(r / C);
    const float4 v = __ldcs(reinterpret_cast<const float4*>(img + ((static_cast<long long>(b) * C + c) * Hi + y) * Wi + x4 * 4));
    const int x = x4 * 4, ph = y / P, py = y % P, pw = x / P, px = x % P;
    T* dst = out + (static_cast<long long>(b) * nH * nW + ph * nW + pw) * K + (c * P + py) * P + px;
    V4<T>::st(dst, v);
  }
}

template <class T>
__device__ __forceinline__ void ld8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void ld8<float>(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void ld8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 q = *reinterpret_cast<const uint4*>(p);
  v[0] = na::bf_lo(q.x); v[1] = na::bf_hi(q.x); v[2] = na::bf_lo(q.y); v[3] = na::bf_hi(q.y);
  v[4] = na::bf_lo(q.z); v[5] = na::bf_hi(q.z); v[6] = na::bf_lo(q.w); v[7] = na::bf_hi(q.w);
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) = make_uint4(na::bf_pack(v[0], v[1]), na::bf_pack(v[2], v[3]), na::bf_pack(v[4], v[5]),
                                            na::bf_pack(v[6], v[7]));
}

// thread = (token t, 8 channels); images b = blockIdx.y, blockIdx.y + gridDim.y, ...
template <class T>
__global__ void __launch_bounds__(256) embed_finish_fwd_kernel(const T* __restrict__ tok, const float* __restrict__ cls,
                                                               const float* __restrict__ pos, T* __restrict__ out, int B,
                                                               int Tn, int E, const Args a) {
  const int nv = E / 8;
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= Tn * nv) return;
  const int t = idx / nv, v = idx % nv;
  float base[8];
  ld8<float>(pos + static_cast<size_t>(t) * E + v * 8, base);
  if (t == 0) {
    float c[8];
    ld8<float>(cls + v * 8, c);
#pragma unroll
    for (int e = 0; e < 8; ++e) base[e] += c[e];
  }
  const bool drop = a.drop_p > 0.f;
  const na::Seed sd = drop ? na::load_seed(a) : na::Seed{0u, 0u};
  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    float o[8];
    if (t == 0) {
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = base[e];
    } else {
      ld8<T>(tok + (static_cast<size_t>(b) * (Tn - 1) + (t - 1)) * E + v * 8, o);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] += base[e];
    }
    if (drop) {
      const unsigned dm = na::drop_mask8(a, na::row_seed(sd, static_cast<long long>(b) * Tn + t), static_cast<unsigned>(v));
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (dm >> e & 1u) ? 0.f : o[e] * a.keep_scale;
    }
    st8(out + (static_cast<size_t>(b) * Tn + t) * E + v * 8, o);
  }
}

template <class T>
__global__ void __launch_bounds__(256) embed_finish_bwd_kernel(const T* __restrict__ dout, T* __restrict__ dtok,
                                                               float* __restrict__ part, int B, int Tn, int E, const Args a) {
  const int nv = E / 8;
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= Tn * nv) return;
  const int t = idx / nv, v = idx % nv;
  const bool drop = a.drop_p > 0.f;
  const na::Seed sd = drop ? na::load_seed(a) : na::Seed{0u, 0u};
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    float d[8];
    ld8<T>(dout + (static_cast<size_t>(b) * Tn + t) * E + v * 8, d);
    if (drop) {
      const unsigned dm = na::drop_mask8(a, na::row_seed(sd, static_cast<long long>(b) * Tn + t), static_cast<unsigned>(v));
#pragma unroll
      for (int e = 0; e < 8; ++e) d[e] = (dm >> e & 1u) ? 0.f : d[e] * a.keep_scale;
    }
    if (t > 0) st8(dtok + (static_cast<size_t>(b) * (Tn - 1) + (t - 1)) * E + v * 8, d);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] += d[e];
  }
  st8(part + (static_cast<size_t>(blockIdx.y) * Tn + t) * E + v * 8, acc);
}

static int slabs_for(int B, int Tn, int E) {
  const int ctas_x = (Tn * (E / 8) + 255) / 256;
  int s = (device_sm_count() * 4 + ctas_x - 1) / ctas_x;
  if (s > B) s = B;
  if (s < 1) s = 1;
  return s;
}

}  // namespace embed
}  // namespace sb200

using namespace sb200;

extern "C" {

int spectre_patch_gather(const float* img, void* out, int B, int C, int Hi, int Wi, int P, int act_dtype, void* stream) {
  if (!img || !out) return fail(SPECTRE_E_NULL, "spectre_patch_gather: null pointer");
  if (B < 1 || C < 1 || P < 4 || (P % 4) || Hi < P || Wi < P || (Wi % 4))
    return fail(SPECTRE_E_SHAPE, "spectre_patch_gather: need P a multiple of 4, Wi a multiple of 4 (B=%d C=%d Hi=%d Wi=%d P=%d)", B, C, Hi, Wi, P);
  if (act_dtype != SPECTRE_DTYPE_F32 && act_dtype != SPECTRE_DTYPE_BF16) return fail(SPECTRE_E_DTYPE, "spectre_patch_gather: act_dtype=%d", act_dtype);
  if (!aligned16(img) || !aligned16(out)) return fail(SPECTRE_E_ALIGN, "spectre_patch_gather: pointers must be 16-byte aligned");
  const int nH = Hi / P, nW = Wi / P;
  const long long total = static_cast<long long>(B) * C * nH * P * (nW * P / 4);
  const long long want = (total + 255) / 256, cap = static_cast<long long>(device_sm_count()) * 16;
  const int grid = static_cast<int>(want < cap ? want : cap);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == SPECTRE_DTYPE_BF16)
    embed::patch_gather_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(img, static_cast<__nv_bfloat16*>(out), B, C, Hi, Wi, P, nH, nW);
  else
    embed::patch_gather_kernel<float><<<grid, 256, 0, st>>>(img, static_cast<float*>(out), B, C, Hi, Wi, P, nH, nW);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "patch_gather launch: %s", cudaGetErrorString(err));
  last_launch_count() = 1;
  last_error().clear();
  return SPECTRE_OK;
}

static int embed_check(const char* fn, int B, int T, int E, int act_dtype, float drop_p) {
  if (B < 1 || T < 2 || E < 8 || (E % 8)) return fail(SPECTRE_E_SHAPE, "%s: need B >= 1, T >= 2, E a multiple of 8 (B=%d T=%d E=%d)", fn, B, T, E);
  if (act_dtype != SPECTRE_DTYPE_F32 && act_dtype != SPECTRE_DTYPE_BF16) return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d", fn, act_dtype);
  if (drop_p < 0.f || drop_p >= 1.f) return fail(SPECTRE_E_SHAPE, "%s: drop_p=%g outside [0,1)", fn, drop_p);
  return SPECTRE_OK;
}

int spectre_embed_finish_fwd(const void* tok, const float* cls, const float* pos, void* out, int B, int T, int E,
                             int act_dtype, float drop_p, unsigned long long seed, const unsigned long long* seed_dev,
                             void* stream) {
  if (!tok || !cls || !pos || !out) return fail(SPECTRE_E_NULL, "spectre_embed_finish_fwd: null pointer");
  if (int rc = embed_check("spectre_embed_finish_fwd", B, T, E, act_dtype, drop_p)) return rc;
  if (!aligned16(tok) || !aligned16(cls) || !aligned16(pos) || !aligned16(out))
    return fail(SPECTRE_E_ALIGN, "spectre_embed_finish_fwd: pointers must be 16-byte aligned");
  na::Args a{};
  na::set_dropout(a, drop_p);
  a.seed_lo = static_cast<unsigned>(seed); a.seed_hi = static_cast<unsigned>(seed >> 32);
  a.seed_dev = seed_dev;
  const dim3 grid((T * (E / 8) + 255) / 256, embed::slabs_for(B, T, E));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == SPECTRE_DTYPE_BF16)
    embed::embed_finish_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(tok), cls, pos,
                                                                        static_cast<__nv_bfloat16*>(out), B, T, E, a);
  else
    embed::embed_finish_fwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(tok), cls, pos, static_cast<float*>(out), B, T, E, a);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "embed_finish_fwd launch: %s", cudaGetErrorString(err));
  last_launch_count() = 1;
  last_error().clear();
  return SPECTRE_OK;
}

size_t spectre_embed_finish_bwd_workspace_bytes(int B, int T, int E) {
  if (B < 1 || T < 2 || E < 8) return 0;
  return static_cast<size_t>(embed::slabs_for(B, T, E)) * T * E * sizeof(float);
}

int spectre_embed_finish_bwd(const void* dout, void* dtok, float* dpos, void* workspace, size_t workspace_bytes, int B,
                             int T, int E, int act_dtype, float drop_p, unsigned long long seed,
                             const unsigned long long* seed_dev, void* stream) {
  if (!dout || !dtok || !dpos || !workspace) return fail(SPECTRE_E_NULL, "spectre_embed_finish_bwd: null pointer");
  if (int rc = embed_check("spectre_embed_finish_bwd", B, T, E, act_dtype, drop_p)) return rc;
  if (!aligned16(dout) || !aligned16(dtok) || !aligned16(dpos) || !aligned16(workspace))
    return fail(SPECTRE_E_ALIGN, "spectre_embed_finish_bwd: pointers must be 16-byte aligned");
  const int slabs = embed::slabs_for(B, T, E);
  const size_t need = static_cast<size_t>(slabs) * T * E * sizeof(float);
  if (workspace_bytes < need) return fail(SPECTRE_E_WORKSPACE, "spectre_embed_finish_bwd: workspace %zu B < required %zu B", workspace_bytes, need);
  na::Args a{};
  na::set_dropout(a, drop_p);
  a.seed_lo = static_cast<unsigned>(seed); a.seed_hi = static_cast<unsigned>(seed >> 32);
  a.seed_dev = seed_dev;
  const dim3 grid((T * (E / 8) + 255) / 256, slabs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = static_cast<float*>(workspace);
  if (act_dtype == SPECTRE_DTYPE_BF16)
    embed::embed_finish_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dout),
                                                                        static_cast<__nv_bfloat16*>(dtok), part, B, T, E, a);
  else
    embed::embed_finish_bwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(dout), static_cast<float*>(dtok), part, B, T, E, a);
  launch_reduce_partials(part, dpos, static_cast<size_t>(T) * E, slabs, st);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "embed_finish_bwd launch: %s", cudaGetErrorString(err));
  last_launch_count() = 2;
  last_error().clear();
  return SPECTRE_OK;
}

}  // extern "C"
