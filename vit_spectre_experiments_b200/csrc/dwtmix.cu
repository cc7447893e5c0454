// Haar DWT token mixer, forward and backward (sm_100a).
//
// Single-level orthonormal 2-D Haar over the token grid per channel, multiply by a
// learned real weight per (sub-band, position, channel), inverse Haar; CLS passes
// through.  Odd H or W follow mode="zero": one row / column of zeros is appended
// (pytorch_wavelets pads a length-2 filter at the end), the result is cropped back (Haar / mode="zero" convention of
// spectre_vit/repl/dwt_experiments.py:9,56; slot: spectre_vit/models/spectre/spectre.py:50,66).
//
// The transform is local to each 2x2 token block, so no shared memory is needed:
// a thread owns one (2x2 block, 16-byte channel vector) for a strided set of images,
// keeps the four weight vectors (and, in backward, the four weight-gradient
// accumulators) in registers, and streams the four tokens with 16-byte coalesced
// accesses (a warp reads 512 contiguous bytes of one token).  HBM traffic is the
// algorithmic minimum: x once in, y once out.
#include "common.cuh"

namespace sb200 {

template <int L>
__device__ __forceinline__ void haar_block(const float (&a)[L], const float (&b)[L], const float (&c)[L],
                                           const float (&d)[L], float (&ll)[L], float (&s1)[L], float (&s2)[L],
                                           float (&hh)[L]) {
#pragma unroll
  for (int i = 0; i < L; ++i) {
    const float p = a[i] + b[i], q = a[i] - b[i], r = c[i] + d[i], s = c[i] - d[i];
    ll[i] = 0.5f * (p + r);
    s1[i] = 0.5f * (p - r);  // low along W, high along H
    s2[i] = 0.5f * (q + s);  // high along W, low along H
    hh[i] = 0.5f * (q - s);
  }
}

template <class T>
__device__ __forceinline__ void ld_or_zero(const T* p, bool valid, float (&v)[VecIO<T>::kLanes]) {
  if (valid) {
    VecIO<T>::ld(p, v);
  } else {
#pragma unroll
    for (int i = 0; i < VecIO<T>::kLanes; ++i) v[i] = 0.f;
  }
}

// items: [0, nblk*nvec) = (2x2 block, channel vector); [nblk*nvec, nblk*nvec + nvec) = CLS row copy
template <class T, bool kBackward, bool kOdd>   // kOdd: H or W odd (zero-padded); even grids keep straight-line code
__global__ void __launch_bounds__(256) dwtmix_kernel(const T* __restrict__ in,      // x (fwd) or dy (bwd)
                                                     const T* __restrict__ xsave,   // x (bwd only)
                                                     const float* __restrict__ wgt, T* __restrict__ out,
                                                     float* __restrict__ part,      // bwd: [gridDim.y][4][H/2][W/2][E]
                                                     int B, int H, int W, int E, int cls) {
  constexpr int L = VecIO<T>::kLanes;
  const int nvec = E / L;
  const int h2 = (H + 1) / 2, w2 = (W + 1) / 2;   // odd sizes: zero-padded to even
  const int nblk = h2 * w2;
  const int tokens = H * W + cls;
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= nblk * nvec + (cls ? nvec : 0)) return;

  if (item >= nblk * nvec) {  // CLS passthrough
    const int cv = item - nblk * nvec;
    for (int b = blockIdx.y; b < B; b += gridDim.y) {
      const size_t off = static_cast<size_t>(b) * tokens * E + cv * L;
      float v[L];
      VecIO<T>::ld(in + off, v);
      VecIO<T>::st(out + off, v);
    }
    return;
  }

  const int cv = item % nvec, blk = item / nvec;
  const int i = blk / w2, j = blk % w2;
  const int ch = cv * L;
  // token offsets of the 2x2 block (a b / c d)
  const size_t ta = static_cast<size_t>(cls + (2 * i) * W + 2 * j) * E + ch;
  const size_t tb = ta + E, tc = ta + static_cast<size_t>(W) * E, td = tc + E;
  const bool vb = !kOdd || 2 * j + 1 < W, vc = !kOdd || 2 * i + 1 < H, vd = vb && vc;   // tokens b / c / d inside the grid?
  const size_t band = static_cast<size_t>(nblk) * E;  // stride between sub-bands in wgt

  float w[4][L];
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int q = 0; q < L; q += 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(wgt + s * band + static_cast<size_t>(blk) * E + ch + q));
      w[s][q] = t.x; w[s][q + 1] = t.y; w[s][q + 2] = t.z; w[s][q + 3] = t.w;
    }

  float acc[4][L];
  if (kBackward) {
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int q = 0; q < L; ++q) acc[s][q] = 0.f;
  }

  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    const size_t base = static_cast<size_t>(b) * tokens * E;
    float a[L], bb[L], c[L], d[L];
    VecIO<T>::ld(in + base + ta, a);
    ld_or_zero<T>(in + base + tb, vb, bb);
    ld_or_zero<T>(in + base + tc, vc, c);
    ld_or_zero<T>(in + base + td, vd, d);
    float s[4][L];
    haar_block<L>(a, bb, c, d, s[0], s[1], s[2], s[3]);
    if (kBackward) {
      float xa[L], xb[L], xc[L], xd[L], xs[4][L];
      VecIO<T>::ld(xsave + base + ta, xa);
      ld_or_zero<T>(xsave + base + tb, vb, xb);
      ld_or_zero<T>(xsave + base + tc, vc, xc);
      ld_or_zero<T>(xsave + base + td, vd, xd);
      haar_block<L>(xa, xb, xc, xd, xs[0], xs[1], xs[2], xs[3]);
#pragma unroll
      for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int q = 0; q < L; ++q) acc[k][q] = fmaf(xs[k][q], s[k][q], acc[k][q]);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < L; ++q) s[k][q] *= w[k][q];
    // the 4x4 Haar block matrix is orthogonal and symmetric: inverse = same butterflies
    haar_block<L>(s[0], s[1], s[2], s[3], a, bb, c, d);
    VecIO<T>::st(out + base + ta, a);
    if (vb) VecIO<T>::st(out + base + tb, bb);
    if (vc) VecIO<T>::st(out + base + tc, c);
    if (vd) VecIO<T>::st(out + base + td, d);
  }

  if (kBackward) {
    float* slab = part + static_cast<size_t>(blockIdx.y) * 4 * band;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int q = 0; q < L; q += 4)
        *reinterpret_cast<float4*>(slab + k * band + static_cast<size_t>(blk) * E + ch + q) =
            make_float4(acc[k][q], acc[k][q + 1], acc[k][q + 2], acc[k][q + 3]);
  }
}

// Image slabs (gridDim.y), independent of has_cls so the workspace query and the launch agree.
// Forward: ~8 CTAs of 256 threads per SM (4 x 16 B loads in flight per thread).  Backward: ~2 per
// SM (8 x 16 B loads in flight per thread) -- every slab costs one partial weight-gradient
// write + read, so fewer, longer-running threads keep that overhead to a few percent.
static int dwt_slabs(bool backward, int B, int H, int W, int E, int lanes) {
  const int items = ((H + 1) / 2) * ((W + 1) / 2) * (E / lanes);
  const int ctas_x = (items + 255) / 256;
  int ny = (device_sm_count() * (backward ? 2 : 8) + ctas_x - 1) / ctas_x;
  if (ny < 1) ny = 1;
  if (ny > B) ny = B;
  return ny;
}

template <class T>
static int dwt_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                      size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st) {
  constexpr int L = VecIO<T>::kLanes;
  if (E % L != 0) return fail(SPECTRE_E_SHAPE, "dwtmix: E=%d must be a multiple of %d for this dtype", E, L);
  const int items = ((H + 1) / 2) * ((W + 1) / 2) * (E / L) + (cls ? E / L : 0);
  const int ny = dwt_slabs(backward, B, H, W, E, L);
  const dim3 grid((items + 255) / 256, ny);
  last_launch_count() = 1;
  const bool odd = (H & 1) || (W & 1);
  if (!backward) {
    (odd ? dwtmix_kernel<T, false, true> : dwtmix_kernel<T, false, false>)<<<grid, 256, 0, st>>>(
        static_cast<const T*>(in), nullptr, w, static_cast<T*>(out), nullptr, B, H, W, E, cls);
  } else {
    const size_t slab_floats = static_cast<size_t>(4) * ((H + 1) / 2) * ((W + 1) / 2) * E;   // = H*W*E for even sizes
    if (ny > 1 && ws_bytes < slab_floats * sizeof(float) * ny)
      return fail(SPECTRE_E_WORKSPACE, "dwtmix_bwd: workspace %zu B < required %zu B", ws_bytes,
                  slab_floats * sizeof(float) * ny);
    float* part = ny > 1 ? static_cast<float*>(ws) : dw;
    (odd ? dwtmix_kernel<T, true, true> : dwtmix_kernel<T, true, false>)<<<grid, 256, 0, st>>>(
        static_cast<const T*>(in), static_cast<const T*>(xsave), w, static_cast<T*>(out), part, B, H, W, E, cls);
    if (ny > 1) {
      launch_reduce_partials(part, dw, slab_floats, ny, st);
      last_launch_count() = 2;
    }
  }
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "dwtmix launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

int dwtmix_dispatch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                    size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st) {
  return dtype == SPECTRE_DTYPE_F32
             ? dwt_launch<float>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, H, W, E, cls, st)
             : dwt_launch<__nv_bfloat16>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, H, W, E, cls, st);
}

size_t dwtmix_bwd_workspace(int B, int H, int W, int E) {
  const int a = dwt_slabs(true, B, H, W, E, 4), b = dwt_slabs(true, B, H, W, E, 8);
  const int ny = a > b ? a : b;
  if (ny <= 1) return 0;
  return static_cast<size_t>(ny) * 4 * ((H + 1) / 2) * ((W + 1) / 2) * E * sizeof(float);
}

}  // namespace sb200
