// MHPermutMix gather (the reference's HEAD occupant of the mix_layer slot, spectre_vit/models/spectre/layers.py:68-70):
//
//   mixed[b, h, j] = x_flat[b, perms[h, j]] * signs[h, j]          x_flat = x.view(B, T*E),  j < F = T*E
//
// and its backward through the inverse permutations (perms[h] is a torch.randperm, layers.py:61-64):
//
//   dx_flat[b, i] = sum_h dmixed[b, h, inv[h, i]] * signs[h, inv[h, i]]        (heads summed in a fixed order)
//
// HBM-bound integer/byte work: int64 indices and signs are read once per (head, 4 columns) and reused for a slice of
// 8 images held in registers; the gathers hit a row of x (T*E elements, L2-resident), the heads-times larger output
// is written with 16-byte coalesced stores.  Sign flips are exact, so the forward is bit-identical to torch's.
#include "common.cuh"

namespace sb200 {
namespace pm {

constexpr int kImgs = 8;   // images per thread (index / sign loads amortised over them)

template <class T>
__device__ __forceinline__ float ldf(const T* p);
template <>
__device__ __forceinline__ float ldf<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(__ldg(p));
}
__device__ __forceinline__ void st4(float* p, const float (&v)[4]) {
  __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
}
__device__ __forceinline__ void st4(__nv_bfloat16* p, const float (&v)[4]) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  __stcs(reinterpret_cast<uint2*>(p), make_uint2(*reinterpret_cast<const unsigned*>(&a), *reinterpret_cast<const unsigned*>(&b)));
}

// grid: x = 4-column groups, y = head, z = image slice
template <class T>
__global__ void __launch_bounds__(256) permut_fwd_kernel(const T* __restrict__ x, const long long* __restrict__ perms,
                                                         const float* __restrict__ signs, T* __restrict__ out, int B,
                                                         long long F, int heads) {
  const long long j0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (j0 >= F) return;
  const int h = blockIdx.y;
  const longlong2 p01 = *reinterpret_cast<const longlong2*>(perms + h * F + j0);
  const longlong2 p23 = *reinterpret_cast<const longlong2*>(perms + h * F + j0 + 2);
  const float4 s = *reinterpret_cast<const float4*>(signs + h * F + j0);
  const long long idx[4] = {p01.x, p01.y, p23.x, p23.y};
  const float sg[4] = {s.x, s.y, s.z, s.w};
  const int b0 = blockIdx.z * kImgs;
#pragma unroll
  for (int i = 0; i < kImgs; ++i) {
    const int b = b0 + i;
    if (b < B) {
      float v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) v[q] = ldf(x + b * F + idx[q]) * sg[q];
      st4(out + (static_cast<long long>(b) * heads + h) * F + j0, v);
    }
  }
}

// grid: x = 4-column groups of dx, z = image slice
template <class T>
__global__ void __launch_bounds__(256) permut_bwd_kernel(const T* __restrict__ dout, const long long* __restrict__ inv,
                                                         const float* __restrict__ signs, T* __restrict__ dx, int B,
                                                         long long F, int heads) {
  const long long i0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (i0 >= F) return;
  const int b0 = blockIdx.z * kImgs;
  float acc[kImgs][4];
#pragma unroll
  for (int i = 0; i < kImgs; ++i)
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[i][q] = 0.f;
  for (int h = 0; h < heads; ++h) {
    const longlong2 p01 = *reinterpret_cast<const longlong2*>(inv + h * F + i0);
    const longlong2 p23 = *reinterpret_cast<const longlong2*>(inv + h * F + i0 + 2);
    const long long idx[4] = {p01.x, p01.y, p23.x, p23.y};
    float sg[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) sg[q] = __ldg(signs + h * F + idx[q]);
#pragma unroll
    for (int i = 0; i < kImgs; ++i) {
      const int b = b0 + i;
      if (b < B) {
        const T* row = dout + (static_cast<long long>(b) * heads + h) * F;
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[i][q] = fmaf(ldf(row + idx[q]), sg[q], acc[i][q]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kImgs; ++i) {
    const int b = b0 + i;
    if (b < B) st4(dx + b * F + i0, acc[i]);
  }
}

static int validate(const char* fn, const void* a, const void* b, const void* c, const void* d, int B, long long F,
                    int heads, int dtype) {
  if (!a || !b || !c || !d) return fail(SPECTRE_E_NULL, "%s: null pointer", fn);
  if (B < 1 || F < 4 || (F % 4) || heads < 1 || heads > 65535)
    return fail(SPECTRE_E_SHAPE, "%s: need B >= 1, T*E a positive multiple of 4, 1 <= heads <= 65535 (B=%d F=%lld heads=%d)",
                fn, B, F, heads);
  if (dtype != SPECTRE_DTYPE_F32 && dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, dtype);
  if (!aligned16(a) || !aligned16(b) || !aligned16(c) || !aligned16(d))
    return fail(SPECTRE_E_ALIGN, "%s: pointers must be 16-byte aligned", fn);
  if ((B + kImgs - 1) / kImgs > 65535) return fail(SPECTRE_E_SHAPE, "%s: batch %d too large", fn, B);
  return SPECTRE_OK;
}

}  // namespace pm
}  // namespace sb200

using namespace sb200;
using namespace sb200::pm;

extern "C" {

int spectre_permut_fwd(const void* x, const long long* perms, const float* signs, void* out, int B, long long F,
                       int heads, int act_dtype, void* stream) {
  if (int rc = validate("spectre_permut_fwd", x, perms, signs, out, B, F, heads, act_dtype)) return rc;
  const dim3 grid(static_cast<unsigned>((F / 4 + 255) / 256), heads, (B + kImgs - 1) / kImgs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == SPECTRE_DTYPE_F32)
    permut_fwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), perms, signs, static_cast<float*>(out), B, F, heads);
  else
    permut_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), perms, signs,
                                                           static_cast<__nv_bfloat16*>(out), B, F, heads);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "permut_fwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

int spectre_permut_bwd(const void* dout, const long long* inv_perms, const float* signs, void* dx, int B, long long F,
                       int heads, int act_dtype, void* stream) {
  if (int rc = validate("spectre_permut_bwd", dout, inv_perms, signs, dx, B, F, heads, act_dtype)) return rc;
  const dim3 grid(static_cast<unsigned>((F / 4 + 255) / 256), 1, (B + kImgs - 1) / kImgs);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == SPECTRE_DTYPE_F32)
    permut_bwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(dout), inv_perms, signs, static_cast<float*>(dx), B, F, heads);
  else
    permut_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dout), inv_perms, signs,
                                                           static_cast<__nv_bfloat16*>(dx), B, F, heads);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "permut_bwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

}  // extern "C"
