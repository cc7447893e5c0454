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
#include <cstdlib>

#include "common.cuh"
#include "sm100_ptx.cuh"

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

// ------------------------------------------------------------------ strip-staged variant (even grids)
// Same arithmetic, different data movement: the register-streaming kernel above holds 4 (backward: 8) 16-byte loads per
// thread and ~110 registers for bf16, i.e. ~32 KB in flight per SM -- 52-64 % of the HBM peak
// (profiles/r01_kernel_bench_v9.jsonl) where fp32 (twice the bytes per load) reaches 77-81 %.  Here the two token rows
// 2i, 2i+1 of one image ("strip": 2*W*E contiguous elements) arrive by ONE bulk async copy (cp.async.bulk + mbarrier) into
// a shared-memory ring, 3-6 strips in flight per CTA with no registers held for them.  A CTA owns strip index i
// (blockIdx.x) for a slab of images (blockIdx.y), so a thread's weights (and weight-gradient accumulators) stay in
// registers; results go straight from registers to HBM.  blockIdx.x == H/2 copies the CLS rows.
constexpr int kDwtMaxStages = 6;
constexpr int kDwtIpt = 2;   // at most this many (2x2 block, channel vector) items per thread

__device__ __forceinline__ void dwt_bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(ptx::smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(ptx::smem_u32(bar)) : "memory");
}
template <class T>
__device__ __forceinline__ void lds_vec(const T* p, float (&v)[VecIO<T>::kLanes]);
template <>
__device__ __forceinline__ void lds_vec<float>(const float* p, float (&v)[4]) {
  const float4 q = *reinterpret_cast<const float4*>(p);
  v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
template <>
__device__ __forceinline__ void lds_vec<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 q = *reinterpret_cast<const uint4*>(p);
  const unsigned u[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(u[i] << 16);
    v[2 * i + 1] = __uint_as_float(u[i] & 0xffff0000u);
  }
}

template <class T, bool kBackward, int IPT>
__global__ void __launch_bounds__(256) dwtmix_strip_kernel(const T* __restrict__ in, const T* __restrict__ xsave,
                                                           const float* __restrict__ wgt, T* __restrict__ out,
                                                           float* __restrict__ part, int B, int H, int W, int E, int cls,
                                                           int stages, int stage_bytes) {
  constexpr int L = VecIO<T>::kLanes;
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t full[kDwtMaxStages];
  pdl_launch_dependents();
  const int nvec = E / L, h2 = H / 2, w2 = W / 2;
  const int tokens = H * W + cls;
  const size_t img_stride = static_cast<size_t>(tokens) * E;
  if (static_cast<int>(blockIdx.x) == h2) {   // CLS rows: plain copy
    pdl_wait();
    for (int b = blockIdx.y; b < B; b += gridDim.y)
      for (int cv = threadIdx.x; cv < nvec; cv += blockDim.x) {
        float v[L];
        VecIO<T>::ld(in + b * img_stride + cv * L, v);
        VecIO<T>::st(out + b * img_stride + cv * L, v);
      }
    return;
  }
  const int strip = blockIdx.x;
  const int strip_elems = 2 * W * E;
  const size_t strip_off = static_cast<size_t>(cls + 2 * strip * W) * E;
  const uint32_t strip_bytes = static_cast<uint32_t>(strip_elems) * sizeof(T);
  const int n_it = static_cast<int>(blockIdx.y) < B ? (B - blockIdx.y + gridDim.y - 1) / gridDim.y : 0;
  const int S = stages;
  auto issue = [&](int it, int stg) {
    const size_t g = (static_cast<size_t>(blockIdx.y) + static_cast<size_t>(it) * gridDim.y) * img_stride + strip_off;
    uint8_t* st = ring + stg * stage_bytes;
    ptx::mbar_arrive_expect_tx(&full[stg], strip_bytes * (kBackward ? 2u : 1u));
    dwt_bulk_load(st, in + g, strip_bytes, &full[stg]);
    if (kBackward) dwt_bulk_load(st + strip_bytes, xsave + g, strip_bytes, &full[stg]);
  };
  if (threadIdx.x == 0) {
    for (int k = 0; k < S; ++k) ptx::mbar_init(&full[k], 1);
    ptx::mbar_fence_init();
  }
  pdl_wait();
  if (threadIdx.x == 0)
    for (int it = 0; it < S - 1 && it < n_it; ++it) issue(it, it);
  __syncthreads();

  // this thread's items: (column pair j, channel vector cv) of the strip
  const int items = w2 * nvec;
  const size_t band = static_cast<size_t>(h2) * w2 * E;
  int ta[IPT];       // element offset of token a inside the strip
  size_t woff[IPT];  // offset of (block, channel) inside one weight sub-band
  bool on[IPT];
  float w[IPT][4][L], acc[IPT][4][L];
#pragma unroll
  for (int k = 0; k < IPT; ++k) {
    const int item = threadIdx.x + k * 256;
    on[k] = item < items;
    const int cv = on[k] ? item % nvec : 0, j = on[k] ? item / nvec : 0;
    ta[k] = 2 * j * E + cv * L;
    woff[k] = static_cast<size_t>(strip * w2 + j) * E + cv * L;
#pragma unroll
    for (int sb = 0; sb < 4; ++sb)
#pragma unroll
      for (int q = 0; q < L; q += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(wgt + sb * band + woff[k] + q));
        w[k][sb][q] = t.x; w[k][sb][q + 1] = t.y; w[k][sb][q + 2] = t.z; w[k][sb][q + 3] = t.w;
        acc[k][sb][q] = acc[k][sb][q + 1] = acc[k][sb][q + 2] = acc[k][sb][q + 3] = 0.f;
      }
  }
  int stg = 0;
  uint32_t phase = 0;
  const int rowE = W * E;
  for (int it = 0; it < n_it; ++it) {
    if (threadIdx.x == 0 && it + S - 1 < n_it) issue(it + S - 1, stg == 0 ? S - 1 : stg - 1);
    ptx::mbar_wait(&full[stg], phase);
    const T* sIn = reinterpret_cast<const T*>(ring + stg * stage_bytes);
    const T* sX = sIn + strip_elems;
    T* o = out + (static_cast<size_t>(blockIdx.y) + static_cast<size_t>(it) * gridDim.y) * img_stride + strip_off;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      if (!on[k]) continue;
      float a[L], bb[L], c[L], d[L], s[4][L];
      lds_vec<T>(sIn + ta[k], a);
      lds_vec<T>(sIn + ta[k] + E, bb);
      lds_vec<T>(sIn + ta[k] + rowE, c);
      lds_vec<T>(sIn + ta[k] + rowE + E, d);
      haar_block<L>(a, bb, c, d, s[0], s[1], s[2], s[3]);
      if (kBackward) {
        float xs[4][L];
        lds_vec<T>(sX + ta[k], a);
        lds_vec<T>(sX + ta[k] + E, bb);
        lds_vec<T>(sX + ta[k] + rowE, c);
        lds_vec<T>(sX + ta[k] + rowE + E, d);
        haar_block<L>(a, bb, c, d, xs[0], xs[1], xs[2], xs[3]);
#pragma unroll
        for (int sb = 0; sb < 4; ++sb)
#pragma unroll
          for (int q = 0; q < L; ++q) acc[k][sb][q] = fmaf(xs[sb][q], s[sb][q], acc[k][sb][q]);
      }
#pragma unroll
      for (int sb = 0; sb < 4; ++sb)
#pragma unroll
        for (int q = 0; q < L; ++q) s[sb][q] *= w[k][sb][q];
      haar_block<L>(s[0], s[1], s[2], s[3], a, bb, c, d);
      VecIO<T>::st(o + ta[k], a);
      VecIO<T>::st(o + ta[k] + E, bb);
      VecIO<T>::st(o + ta[k] + rowE, c);
      VecIO<T>::st(o + ta[k] + rowE + E, d);
    }
    __syncthreads();   // stage consumed
    if (++stg == S) { stg = 0; phase ^= 1u; }
  }
  if (kBackward) {
    float* slab = part + static_cast<size_t>(blockIdx.y) * 4 * band;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      if (!on[k]) continue;
#pragma unroll
      for (int sb = 0; sb < 4; ++sb)
#pragma unroll
        for (int q = 0; q < L; q += 4)
          *reinterpret_cast<float4*>(slab + sb * band + woff[k] + q) =
              make_float4(acc[k][sb][q], acc[k][sb][q + 1], acc[k][sb][q + 2], acc[k][sb][q + 3]);
    }
  }
}

// launch plan of the strip kernel; stages == 0: shape not covered (odd grid, strip too large, too many items per thread)
struct StripPlan {
  int stages, stage_bytes, ny;
  size_t smem;
};
static StripPlan strip_plan(bool backward, int B, int H, int W, int E, int elem, int lanes) {
  StripPlan p{};
  if ((H & 1) || (W & 1) || E % lanes) return p;
  const int items = (W / 2) * (E / lanes);
  if (items > 256 * kDwtIpt) return p;
  const int strip_bytes = 2 * W * E * elem;
  if (strip_bytes % 16) return p;
  p.stage_bytes = strip_bytes * (backward ? 2 : 1);
  // CTAs per SM by registers (forward ~3 x 256 threads, backward 2) and a ~210 KB shared-memory budget
  const int per_sm = backward ? 2 : 3;
  int stages = (210 * 1024 / per_sm) / p.stage_bytes;
  if (stages > kDwtMaxStages) stages = kDwtMaxStages;
  if (stages < 2) return p;
  p.stages = stages;
  p.smem = static_cast<size_t>(stages) * p.stage_bytes;
  const int h2 = H / 2;
  int ny = (device_sm_count() * per_sm + h2 - 1) / h2;   // image slabs: fill the machine once
  if (ny > B) ny = B;
  if (ny < 1) ny = 1;
  p.ny = ny;
  return p;
}
static bool strip_enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_DWT_STRIP");
    return !(e && e[0] == '0');
  }();
  return on;
}

// Image slabs (gridDim.y), independent of has_cls so the workspace query and the launch agree.
// Forward: ~8 CTAs of 256 threads per SM (4 x 16 B loads in flight per thread).  Backward: ~2 per
// SM (8 x 16 B loads in flight per thread) -- every slab costs one partial weight-gradient
// write + read, so fewer, longer-running threads keep that overhead to a few percent.
static int dwt_slabs(bool backward, int B, int H, int W, int E, int lanes) {
  if (strip_enabled()) {
    const StripPlan sp = strip_plan(backward, B, H, W, E, lanes == 8 ? 2 : 4, lanes);
    if (sp.stages) return sp.ny;
  }
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
  const StripPlan sp = strip_enabled() ? strip_plan(backward, B, H, W, E, static_cast<int>(sizeof(T)), L) : StripPlan{};
  if (sp.stages) {
    const dim3 sgrid(H / 2 + (cls ? 1 : 0), sp.ny);
    float* part = nullptr;
    if (backward) {
      const size_t slab_floats = static_cast<size_t>(H) * W * E;
      if (sp.ny > 1 && ws_bytes < slab_floats * sizeof(float) * sp.ny)
        return fail(SPECTRE_E_WORKSPACE, "dwtmix_bwd: workspace %zu B < required %zu B", ws_bytes, slab_floats * sizeof(float) * sp.ny);
      part = sp.ny > 1 ? static_cast<float*>(ws) : dw;
    }
    const bool one = (W / 2) * (E / L) <= 256;   // one item per thread: half the registers
    auto kern = backward ? (one ? dwtmix_strip_kernel<T, true, 1> : dwtmix_strip_kernel<T, true, 2>)
                         : (one ? dwtmix_strip_kernel<T, false, 1> : dwtmix_strip_kernel<T, false, 2>);
    if (ensure_dynamic_smem(reinterpret_cast<const void*>(kern), 215 * 1024) != cudaSuccess)
      return fail(SPECTRE_E_CUDA, "dwtmix: cudaFuncSetAttribute failed");
    launch_pdl(kern, sgrid, dim3(256), sp.smem, st, static_cast<const T*>(in), static_cast<const T*>(xsave), w,
               static_cast<T*>(out), part, B, H, W, E, cls, sp.stages, sp.stage_bytes);
    if (backward && sp.ny > 1) {
      launch_reduce_partials(part, dw, static_cast<size_t>(H) * W * E, sp.ny, st);
      last_launch_count() = 2;
    }
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "dwtmix (strip) launch: %s", cudaGetErrorString(err));
    return SPECTRE_OK;
  }
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

// register-resident Haar kernel of dwtmix_ml_reg.cu (4x4 / 8x8 grids): at one level it is this mixer with the weights
// addressed sub-band major.  Its backward is faster than the kernels above (C3 shape bf16: 0.048 vs 0.056 ms; fp32
// 0.075 vs 0.079 ms), its forward the same -- so only the backward goes there.  SB200_DWT_REG_BWD=0 keeps it here.
int dwtml_reg_supported(int H, int W, int E, int dtype);
int dwtml_reg_launch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                     size_t ws_bytes, int B, int H, int E, int cls, int levels, int dtype, int max_slabs, int subband,
                     cudaStream_t st);
int dwtml_reg_bwd_slabs(int B, int H, int E, int dtype);
static bool reg_bwd_enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_DWT_REG_BWD");
    return !(e && e[0] == '0');
  }();
  return on;
}

// slabs the workspace query promises: the most any backward kernel of this shape (either dtype) asks for
static int bwd_slabs_for_workspace(int B, int H, int W, int E) {
  const int a = dwt_slabs(true, B, H, W, E, 4), b = dwt_slabs(true, B, H, W, E, 8);
  int ny = a > b ? a : b;
  if (reg_bwd_enabled())
    for (int dtype : {SPECTRE_DTYPE_F32, SPECTRE_DTYPE_BF16})
      if (dwtml_reg_supported(H, W, E, dtype)) {
        const int r = dwtml_reg_bwd_slabs(B, H, E, dtype);
        if (r > ny) ny = r;
      }
  return ny;
}

int dwtmix_dispatch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                    size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st) {
  if (backward && reg_bwd_enabled() && dwtml_reg_supported(H, W, E, dtype)) {
    return dwtml_reg_launch(true, in, xsave, w, out, dw, ws, ws_bytes, B, H, E, cls, 1, dtype, bwd_slabs_for_workspace(B, H, W, E), 1, st);
  }
  return dtype == SPECTRE_DTYPE_F32
             ? dwt_launch<float>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, H, W, E, cls, st)
             : dwt_launch<__nv_bfloat16>(backward, in, xsave, w, out, dw, ws, ws_bytes, B, H, W, E, cls, st);
}

size_t dwtmix_bwd_workspace(int B, int H, int W, int E) {
  const int ny = bwd_slabs_for_workspace(B, H, W, E);
  if (ny <= 1) return 0;
  return static_cast<size_t>(ny) * 4 * ((H + 1) / 2) * ((W + 1) / 2) * E * sizeof(float);
}

}  // namespace sb200
