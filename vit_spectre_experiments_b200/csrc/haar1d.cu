// 1-D Haar mixers "dwt_token" / "dwt_embed", forward and backward (sm_100a).
//
// The reference names the two modes in the SpectreEncoderLayer docstring -- "dwt_embed - Wavelet transform in
// embedding dimension", "dwt_token - Wavelet transform in token dimension" (spectre_vit/models/spectre/spectre.py:33-34)
// -- and has no code for them, so the operator is defined here by analogy with the 2-D Haar mixer of dwtmix.cu
// (parity unpinned, like every Haar kernel of this repo):
//   single-level orthonormal Haar along ONE axis of x[:, cls:, :], a learned real weight per coefficient, inverse Haar;
//   CLS passes through.
//     axis 0 (token):  pairs of adjacent tokens (2p, 2p+1) per channel; weight float32 [2, ceil(N/2), E] = (low, high);
//                      an odd token count follows mode="zero": a zero token is appended, the result cropped
//     axis 1 (embed):  pairs of adjacent channels (2j, 2j+1) per token;  weight float32 [2, N, E/2]
// The pair transform is orthogonal and symmetric, so the backward is dx = H(w * H(dy)), dw = sum_b H(x_b) * H(dy_b).
//
// Memory-bound, local to a pair: a thread owns one (pair, 16-byte channel vector) [token axis] or one (token, 16-byte
// channel vector) [embed axis] for a strided set of images, keeps its weights (and weight-gradient accumulators) in
// registers and streams the activations with 16-byte coalesced accesses.  x once in, y once out.
#include "common.cuh"

namespace sb200 {
namespace {

constexpr float kR = 0.70710678118654752440f;

template <class T, int AXIS, bool BWD>
__global__ void __launch_bounds__(256) haar1d_kernel(const T* __restrict__ in, const T* __restrict__ xsave,
                                                     const float* __restrict__ wgt, T* __restrict__ out,
                                                     float* __restrict__ part, int B, int N, int E, int cls) {
  constexpr int L = VecIO<T>::kLanes;
  constexpr int WL = AXIS == 0 ? L : L / 2;       // weights per sub-band held by a thread
  const int nvec = E / L;
  const int np = AXIS == 0 ? (N + 1) / 2 : N;     // pairs (token axis) or tokens (embed axis) along the grid
  const int tokens = N + cls;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  const int nitems = np * nvec;
  if (item >= nitems + (cls ? nvec : 0)) return;

  if (item >= nitems) {   // CLS pass-through
    const int cv = item - nitems;
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
      const size_t off = static_cast<size_t>(b) * tokens * E + cv * L;
      float v[L];
      VecIO<T>::ld(in + off, v);
      VecIO<T>::st(out + off, v);
    }
    return;
  }

  const int cv = item % nvec, p = item / nvec;
  const int ch = cv * L;
  const size_t ta = static_cast<size_t>(cls + (AXIS == 0 ? 2 * p : p)) * E + ch;
  const bool vb = AXIS == 0 && 2 * p + 1 < N;     // second token of the pair inside the sequence?
  const int wrow = AXIS == 0 ? E : E / 2;         // weight row length
  const size_t band = static_cast<size_t>(np) * wrow;
  const size_t woff = static_cast<size_t>(p) * wrow + (AXIS == 0 ? ch : ch / 2);

  float w0[WL], w1[WL];
#pragma unroll
  for (int q = 0; q < WL; q += 2) {   // WL is 2, 4 or 8
    const float2 a = __ldg(reinterpret_cast<const float2*>(wgt + woff + q));
    const float2 b = __ldg(reinterpret_cast<const float2*>(wgt + band + woff + q));
    w0[q] = a.x; w0[q + 1] = a.y;
    w1[q] = b.x; w1[q + 1] = b.y;
  }
  float acc0[WL], acc1[WL];
  if (BWD) {
#pragma unroll
    for (int q = 0; q < WL; ++q) acc0[q] = acc1[q] = 0.f;
  }

  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    const size_t base = static_cast<size_t>(b) * tokens * E;
    float lo[WL], hi[WL];
    float va[L], vbv[L];
    VecIO<T>::ld(in + base + ta, va);
    if (AXIS == 0) {
      if (vb) {
        VecIO<T>::ld(in + base + ta + E, vbv);
      } else {
#pragma unroll
        for (int q = 0; q < L; ++q) vbv[q] = 0.f;
      }
#pragma unroll
      for (int q = 0; q < L; ++q) {
        lo[q] = (va[q] + vbv[q]) * kR;
        hi[q] = (va[q] - vbv[q]) * kR;
      }
    } else {
#pragma unroll
      for (int q = 0; q < WL; ++q) {
        lo[q] = (va[2 * q] + va[2 * q + 1]) * kR;
        hi[q] = (va[2 * q] - va[2 * q + 1]) * kR;
      }
    }
    if (BWD) {
      float xa[L], xb[L];
      VecIO<T>::ld(xsave + base + ta, xa);
      if (AXIS == 0) {
        if (vb) {
          VecIO<T>::ld(xsave + base + ta + E, xb);
        } else {
#pragma unroll
          for (int q = 0; q < L; ++q) xb[q] = 0.f;
        }
#pragma unroll
        for (int q = 0; q < L; ++q) {
          acc0[q] = fmaf((xa[q] + xb[q]) * kR, lo[q], acc0[q]);
          acc1[q] = fmaf((xa[q] - xb[q]) * kR, hi[q], acc1[q]);
        }
      } else {
#pragma unroll
        for (int q = 0; q < WL; ++q) {
          acc0[q] = fmaf((xa[2 * q] + xa[2 * q + 1]) * kR, lo[q], acc0[q]);
          acc1[q] = fmaf((xa[2 * q] - xa[2 * q + 1]) * kR, hi[q], acc1[q]);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < WL; ++q) {
      lo[q] *= w0[q];
      hi[q] *= w1[q];
    }
    if (AXIS == 0) {
#pragma unroll
      for (int q = 0; q < L; ++q) {
        va[q] = (lo[q] + hi[q]) * kR;
        vbv[q] = (lo[q] - hi[q]) * kR;
      }
      VecIO<T>::st(out + base + ta, va);
      if (vb) VecIO<T>::st(out + base + ta + E, vbv);
    } else {
#pragma unroll
      for (int q = 0; q < WL; ++q) {
        va[2 * q] = (lo[q] + hi[q]) * kR;
        va[2 * q + 1] = (lo[q] - hi[q]) * kR;
      }
      VecIO<T>::st(out + base + ta, va);
    }
  }

  if (BWD) {
    float* slab = part + static_cast<size_t>(blockIdx.y) * 2 * band;
#pragma unroll
    for (int q = 0; q < WL; q += 2) {
      *reinterpret_cast<float2*>(slab + woff + q) = make_float2(acc0[q], acc0[q + 1]);
      *reinterpret_cast<float2*>(slab + band + woff + q) = make_float2(acc1[q], acc1[q + 1]);
    }
  }
}

// image slabs (gridDim.y): forward ~8 CTAs of 256 threads per SM, backward ~2 (every slab costs one partial
// weight-gradient write + read); independent of has_cls and dtype so the workspace query and the launch agree
int h1_slabs(bool backward, int B, int N, int E, int axis) {
  const int np = axis == 0 ? (N + 1) / 2 : N;
  const int items = np * (E / 8);                       // bf16 vectors: the smaller item count of the two dtypes
  const int ctas_x = (items + 255) / 256;
  int ny = (device_sm_count() * (backward ? 2 : 8) + ctas_x - 1) / ctas_x;
  if (ny < 1) ny = 1;
  if (ny > B) ny = B;
  return ny;
}

size_t h1_param_floats(int N, int E, int axis) {
  return axis == 0 ? static_cast<size_t>(2) * ((N + 1) / 2) * E : static_cast<size_t>(N) * E;
}

template <class T, int AXIS>
int h1_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
              size_t ws_bytes, int B, int N, int E, int cls, cudaStream_t st) {
  constexpr int L = VecIO<T>::kLanes;
  const int np = AXIS == 0 ? (N + 1) / 2 : N;
  const int items = np * (E / L) + (cls ? E / L : 0);
  const int ny = h1_slabs(backward, B, N, E, AXIS);
  const dim3 grid((items + 255) / 256, ny);
  last_launch_count() = 1;
  if (!backward) {
    haar1d_kernel<T, AXIS, false><<<grid, 256, 0, st>>>(static_cast<const T*>(in), nullptr, w, static_cast<T*>(out),
                                                        nullptr, B, N, E, cls);
  } else {
    const size_t pf = h1_param_floats(N, E, AXIS);
    if (ny > 1 && ws_bytes < pf * sizeof(float) * ny)
      return fail(SPECTRE_E_WORKSPACE, "haar1d_bwd: workspace %zu B < required %zu B", ws_bytes, pf * sizeof(float) * ny);
    float* part = ny > 1 ? static_cast<float*>(ws) : dw;
    haar1d_kernel<T, AXIS, true><<<grid, 256, 0, st>>>(static_cast<const T*>(in), static_cast<const T*>(xsave), w,
                                                       static_cast<T*>(out), part, B, N, E, cls);
    if (ny > 1) {
      launch_reduce_partials(part, dw, pf, ny, st);
      last_launch_count() = 2;
    }
  }
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "haar1d launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

int h1_check(const char* fn, int B, int N, int E, int has_cls, int axis, int dtype) {
  if (B < 1 || N < 1 || E < 8 || (E & 7) || (has_cls != 0 && has_cls != 1))
    return fail(SPECTRE_E_SHAPE, "%s: bad shape B=%d N=%d E=%d has_cls=%d (need B, N >= 1, E a multiple of 8)", fn, B, N, E,
                has_cls);
  if (axis != 0 && axis != 1) return fail(SPECTRE_E_SHAPE, "%s: axis=%d (0 = token, 1 = embed)", fn, axis);
  if (dtype != SPECTRE_DTYPE_F32 && dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, dtype);
  return SPECTRE_OK;
}

int h1_dispatch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                size_t ws_bytes, int B, int N, int E, int cls, int axis, int dtype, cudaStream_t st) {
  if (dtype == SPECTRE_DTYPE_F32)
    return axis == 0 ? h1_launch<float, 0>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, N, E, cls, st)
                     : h1_launch<float, 1>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, N, E, cls, st);
  return axis == 0 ? h1_launch<__nv_bfloat16, 0>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, N, E, cls, st)
                   : h1_launch<__nv_bfloat16, 1>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, N, E, cls, st);
}

}  // namespace
}  // namespace sb200

using namespace sb200;

extern "C" {

int spectre_haar1d_fwd(const void* x, const float* w, void* y, int B, int N, int E, int has_cls, int axis, int act_dtype,
                       void* stream) {
  if (!x || !w || !y) return fail(SPECTRE_E_NULL, "spectre_haar1d_fwd: null pointer");
  if (int rc = h1_check("spectre_haar1d_fwd", B, N, E, has_cls, axis, act_dtype)) return rc;
  if (!aligned16(x) || !aligned16(w) || !aligned16(y))
    return fail(SPECTRE_E_ALIGN, "spectre_haar1d_fwd: pointers must be 16-byte aligned");
  const int rc = h1_dispatch(false, x, nullptr, w, y, nullptr, nullptr, 0, B, N, E, has_cls, axis, act_dtype,
                             static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

int spectre_haar1d_bwd(const void* dy, const void* x, const float* w, void* dx, float* dw, void* workspace,
                       size_t workspace_bytes, int B, int N, int E, int has_cls, int axis, int act_dtype, void* stream) {
  if (!dy || !x || !w || !dx || !dw) return fail(SPECTRE_E_NULL, "spectre_haar1d_bwd: null pointer");
  if (int rc = h1_check("spectre_haar1d_bwd", B, N, E, has_cls, axis, act_dtype)) return rc;
  if (!aligned16(dy) || !aligned16(x) || !aligned16(w) || !aligned16(dx) || !aligned16(dw) || !aligned16(workspace))
    return fail(SPECTRE_E_ALIGN, "spectre_haar1d_bwd: pointers must be 16-byte aligned");
  const int rc = h1_dispatch(true, dy, x, w, dx, dw, workspace, workspace ? workspace_bytes : 0, B, N, E, has_cls, axis,
                             act_dtype, static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

size_t spectre_haar1d_bwd_workspace_bytes(int B, int N, int E, int axis) {
  if (B < 1 || N < 1 || E < 8 || (axis != 0 && axis != 1)) return 0;
  const int ny = h1_slabs(true, B, N, E, axis);
  return ny <= 1 ? 0 : static_cast<size_t>(ny) * h1_param_floats(N, E, axis) * sizeof(float);
}

}  // extern "C"
