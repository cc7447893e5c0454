// Fused row-wise tails around the mixer and the SpectreLinear GEMMs:
//
//   out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + pool(skip) )
//
// which is, with the right operands,
//   x1 = LN1(mix(x)) + x                      spectre_vit/models/spectre/spectre.py:66       (in0 = mix, skip = x)
//   a  = GELU(LN(Linear(x))) + avg_pool(x)    spectre_vit/models/spectre/layers.py:95-101    (in0 = Linear(x), skip = x)
//   x2 = LN2(x1 + ff)                         spectre.py:67                                  (in0 = x1, in1 = ff)
// and its backward (d_in0, d_in1, d_skip, dgamma, dbeta) in ONE pass over the rows each way, instead of
// the 6-10 elementwise / reduction launches of the stock composition (LayerNorm backward's gamma/beta
// reduction and the adaptive-pool kernels were 50 % of the ViT-S/16 step: profiles/r01_step_profile_c4.txt).
//
// Mapping: one CTA = one row at a time (grid-stride over rows); a thread owns VPT vectors of 8 consecutive
// columns (16-byte bf16 / 2x16-byte fp32 accesses, fully coalesced), so the per-column gamma/beta gradient
// accumulators live in registers across the whole row loop; row statistics are a warp shuffle + one small
// shared-memory exchange.  All arithmetic is fp32 (autocast runs layer_norm / gelu in fp32).
#include <cstdarg>

#include "common.cuh"

namespace sb200 {
namespace na {

enum : int { kSkipNone = 0, kSkipIdentity = 1, kSkipUp = 2, kSkipDown = 3, kSkipGeneral = 4 };

struct Args {
  const void* in0; const float* in1; const float* gamma; const float* beta; const float* skip;
  float* out; void* out_bf16; float* mean; float* rstd;
  const float* dout; const void* dout_bf16;
  void* d_in0; float* d_in1; float* d_skip; float* part;   // part: [grid][2][N] gamma/beta partials
  long long rows; int N; int Ns; int act; int skip_mode; int k;   // k = N/Ns (up) or Ns/N (down)
  float eps; float drop_p; unsigned long long seed;
};

// ------------------------------------------------------------------ small helpers
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 q = *reinterpret_cast<const uint4*>(p);
  const unsigned u[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(u[i] << 16);
    v[2 * i + 1] = __uint_as_float(u[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const float (&v)[8]) {
  unsigned u[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    u[i] = *reinterpret_cast<const unsigned*>(&h);
  }
  *reinterpret_cast<uint4*>(p) = make_uint4(u[0], u[1], u[2], u[3]);
}

// exact (erf) GELU, as nn.GELU() in layers.py:89
__device__ __forceinline__ float gelu(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad(float x) {
  return 0.5f * (1.0f + erff(x * 0.70710678118654752f)) + x * 0.39894228040143268f * __expf(-0.5f * x * x);
}

// counter-based dropout: keep-mask of element idx under (seed); one 64-bit mix per element
__device__ __forceinline__ bool keep(unsigned long long seed, unsigned long long idx, unsigned thresh) {
  unsigned long long z = seed + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return static_cast<unsigned>(z >> 32) >= thresh;
}

// sum over the CTA of up to two values; `slot` alternates so consecutive calls need one barrier each
template <int NV>
__device__ __forceinline__ void block_sum(float (&v)[NV], float (*red)[NV][8], int slot) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  const int nw = blockDim.x >> 5;
  if (nw > 1) {
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0)
#pragma unroll
      for (int i = 0; i < NV; ++i) red[slot][i][w] = v[i];
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float s = 0.f;
      for (int k = 0; k < nw; ++k) s += red[slot][i][k];
      v[i] = s;
    }
  }
}

// pooled skip value for output column j (adaptive average pool Ns -> N, torch semantics)
__device__ __forceinline__ float pool_general(const float* srow, int j, int N, int Ns) {
  const int s = static_cast<int>((static_cast<long long>(j) * Ns) / N);
  const int e = static_cast<int>((static_cast<long long>(j + 1) * Ns + N - 1) / N);
  float acc = 0.f;
  for (int i = s; i < e; ++i) acc += __ldg(srow + i);
  return acc / static_cast<float>(e - s);
}

// ------------------------------------------------------------------ forward
template <class T0, int VPT>
__global__ void __launch_bounds__(256) norm_act_fwd_kernel(const Args a) {
  __shared__ float red[3][1][8];
  const int t = threadIdx.x, nvec = a.N >> 3;
  const float inv_n = 1.0f / static_cast<float>(a.N);
  const bool drop = a.drop_p > 0.f;
  const unsigned thresh = drop ? static_cast<unsigned>(fminf(a.drop_p, 0.999999f) * 4294967296.0f) : 0u;
  const float keep_scale = drop ? 1.0f / (1.0f - a.drop_p) : 1.0f;
  for (long long row = blockIdx.x; row < a.rows; row += gridDim.x) {
    const T0* r0 = static_cast<const T0*>(a.in0) + row * a.N;
    float z[VPT][8];
    float s[1] = {0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * blockDim.x;
      if (v < nvec) {
        ld8(r0 + v * 8, z[i]);
        if (a.in1) {
          float w[8];
          ld8(a.in1 + row * a.N + v * 8, w);
#pragma unroll
          for (int e = 0; e < 8; ++e) z[i][e] += w[e];
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) s[0] += z[i][e];
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) z[i][e] = 0.f;
      }
    }
    block_sum<1>(s, red, 0);
    const float mean = s[0] * inv_n;
    float q[1] = {0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i)
      if (t + i * blockDim.x < nvec)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float d = z[i][e] - mean;
          q[0] += d * d;
        }
    block_sum<1>(q, red, 1);
    const float rstd = rsqrtf(q[0] * inv_n + a.eps);
    if (t == 0) {
      a.mean[row] = mean;
      a.rstd[row] = rstd;
    }
    const float* srow = a.skip ? a.skip + row * a.Ns : nullptr;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * blockDim.x;
      if (v >= nvec) continue;
      const int j0 = v * 8;
      float g[8], b[8], o[8];
      ld8(a.gamma + j0, g);
      ld8(a.beta + j0, b);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float u = (z[i][e] - mean) * rstd * g[e] + b[e];
        o[e] = a.act ? gelu(u) : u;
      }
      if (a.skip_mode == kSkipIdentity) {
        float sk[8];
        ld8(srow + j0, sk);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += sk[e];
      } else if (a.skip_mode == kSkipUp) {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += __ldg(srow + (j0 + e) / a.k);
      } else if (a.skip_mode == kSkipDown) {
        const float inv_k = 1.0f / static_cast<float>(a.k);
        const float* sp = srow + static_cast<long long>(j0) * a.k;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float acc = 0.f;
          if ((a.k & 3) == 0) {
            for (int r = 0; r < a.k; r += 4) {
              const float4 f = *reinterpret_cast<const float4*>(sp + e * a.k + r);
              acc += (f.x + f.y) + (f.z + f.w);
            }
          } else {
            for (int r = 0; r < a.k; ++r) acc += sp[e * a.k + r];
          }
          o[e] += acc * inv_k;
        }
      } else if (a.skip_mode == kSkipGeneral) {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += pool_general(srow, j0 + e, a.N, a.Ns);
      }
      if (drop) {
        const unsigned long long base = static_cast<unsigned long long>(row) * a.N + j0;
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = keep(a.seed, base + e, thresh) ? o[e] * keep_scale : 0.f;
      }
      st8(a.out + row * a.N + j0, o);
      if (a.out_bf16) st8(static_cast<__nv_bfloat16*>(a.out_bf16) + row * a.N + j0, o);
    }
  }
}

// ------------------------------------------------------------------ backward
// d = (dout + dout_bf16) * dropout_mask ; d_skip = pool^T(d) ; g = d * act'(u) ; LayerNorm backward of g.
template <class T0, int VPT>
__global__ void __launch_bounds__(256) norm_act_bwd_kernel(const Args a) {
  __shared__ float red[2][2][8];
  const int t = threadIdx.x, nvec = a.N >> 3;
  const float inv_n = 1.0f / static_cast<float>(a.N);
  const bool drop = a.drop_p > 0.f;
  const unsigned thresh = drop ? static_cast<unsigned>(fminf(a.drop_p, 0.999999f) * 4294967296.0f) : 0u;
  const float keep_scale = drop ? 1.0f / (1.0f - a.drop_p) : 1.0f;
  float dg[VPT][8], db[VPT][8], gam[VPT][8], bet[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = t + i * blockDim.x;
#pragma unroll
    for (int e = 0; e < 8; ++e) dg[i][e] = db[i][e] = gam[i][e] = bet[i][e] = 0.f;
    if (v < nvec) {
      ld8(a.gamma + v * 8, gam[i]);
      if (a.act) ld8(a.beta + v * 8, bet[i]);
    }
  }
  int parity = 0;
  for (long long row = blockIdx.x; row < a.rows; row += gridDim.x, parity ^= 1) {
    const T0* r0 = static_cast<const T0*>(a.in0) + row * a.N;
    const float mean = a.mean[row], rstd = a.rstd[row];
    float xh[VPT][8], gy[VPT][8];
    float s[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * blockDim.x;
      if (v >= nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) xh[i][e] = gy[i][e] = 0.f;
        continue;
      }
      const int j0 = v * 8;
      float d[8];
      ld8(a.dout + row * a.N + j0, d);
      if (a.dout_bf16) {
        float w[8];
        ld8(static_cast<const __nv_bfloat16*>(a.dout_bf16) + row * a.N + j0, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) d[e] += w[e];
      }
      if (drop) {
        const unsigned long long base = static_cast<unsigned long long>(row) * a.N + j0;
#pragma unroll
        for (int e = 0; e < 8; ++e) d[e] = keep(a.seed, base + e, thresh) ? d[e] * keep_scale : 0.f;
      }
      // ---- skip path: transpose of the pooling
      if (a.skip_mode == kSkipIdentity) {
        st8(a.d_skip + row * a.Ns + j0, d);
      } else if (a.skip_mode == kSkipUp) {         // N = k * Ns, k in {2,4,8}: 8/k skip columns per vector
        float* dp = a.d_skip + row * a.Ns + j0 / a.k;
        if (a.k == 2) {
          *reinterpret_cast<float4*>(dp) = make_float4(d[0] + d[1], d[2] + d[3], d[4] + d[5], d[6] + d[7]);
        } else if (a.k == 4) {
          *reinterpret_cast<float2*>(dp) = make_float2((d[0] + d[1]) + (d[2] + d[3]), (d[4] + d[5]) + (d[6] + d[7]));
        } else {
          *dp = ((d[0] + d[1]) + (d[2] + d[3])) + ((d[4] + d[5]) + (d[6] + d[7]));
        }
      } else if (a.skip_mode == kSkipDown) {       // Ns = k * N: every skip column of the window gets d / k
        const float inv_k = 1.0f / static_cast<float>(a.k);
        float* dp = a.d_skip + row * a.Ns + static_cast<long long>(j0) * a.k;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float val = d[e] * inv_k;
          if ((a.k & 3) == 0) {
            for (int r = 0; r < a.k; r += 4) *reinterpret_cast<float4*>(dp + e * a.k + r) = make_float4(val, val, val, val);
          } else {
            for (int r = 0; r < a.k; ++r) dp[e * a.k + r] = val;
          }
        }
      }
      // ---- main path
      float z[8];
      ld8(r0 + j0, z);
      if (a.in1) {
        float w[8];
        ld8(a.in1 + row * a.N + j0, w);
#pragma unroll
        for (int e = 0; e < 8; ++e) z[e] += w[e];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x = (z[e] - mean) * rstd;
        float g = d[e];
        if (a.act) g *= gelu_grad(x * gam[i][e] + bet[i][e]);
        db[i][e] += g;
        dg[i][e] += g * x;
        const float y = g * gam[i][e];
        xh[i][e] = x;
        gy[i][e] = y;
        s[0] += y;
        s[1] += y * x;
      }
    }
    block_sum<2>(s, red, parity);
    const float m1 = s[0] * inv_n, m2 = s[1] * inv_n;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = t + i * blockDim.x;
      if (v >= nvec) continue;
      float dz[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) dz[e] = rstd * (gy[i][e] - m1 - xh[i][e] * m2);
      if (a.d_in0) st8(static_cast<T0*>(a.d_in0) + row * a.N + v * 8, dz);
      if (a.d_in1) st8(a.d_in1 + row * a.N + v * 8, dz);
    }
  }
  // per-CTA partial gamma / beta gradients
  float* pg = a.part + static_cast<size_t>(blockIdx.x) * 2 * a.N;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = t + i * blockDim.x;
    if (v < nvec) {
      st8(pg + v * 8, dg[i]);
      st8(pg + a.N + v * 8, db[i]);
    }
  }
}

// general adaptive-pool transpose (only for widths that are not integer multiples): one thread per skip column
__global__ void pool_bwd_general_kernel(const Args a) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= a.rows * a.Ns) return;
  const long long row = i / a.Ns;
  const int c = static_cast<int>(i % a.Ns);
  const int jlo = static_cast<int>((static_cast<long long>(c) * a.N) / a.Ns);
  const int jhi = static_cast<int>((static_cast<long long>(c + 1) * a.N + a.Ns - 1) / a.Ns);   // exclusive
  const bool drop = a.drop_p > 0.f;
  const unsigned thresh = drop ? static_cast<unsigned>(fminf(a.drop_p, 0.999999f) * 4294967296.0f) : 0u;
  const float keep_scale = drop ? 1.0f / (1.0f - a.drop_p) : 1.0f;
  float acc = 0.f;
  for (int j = jlo; j < jhi; ++j) {
    const int s = static_cast<int>((static_cast<long long>(j) * a.Ns) / a.N);
    const int e = static_cast<int>((static_cast<long long>(j + 1) * a.Ns + a.N - 1) / a.N);
    if (c < s || c >= e) continue;
    float d = a.dout[row * a.N + j];
    if (a.dout_bf16) d += __bfloat162float(static_cast<const __nv_bfloat16*>(a.dout_bf16)[row * a.N + j]);
    if (drop) d = keep(a.seed, static_cast<unsigned long long>(row) * a.N + j, thresh) ? d * keep_scale : 0.f;
    acc += d / static_cast<float>(e - s);
  }
  a.d_skip[i] = acc;
}

// dgamma[c] = sum_s part[s][0][c], dbeta[c] = sum_s part[s][1][c]  (fixed order: deterministic)
__global__ void reduce_gamma_beta_kernel(const float* __restrict__ part, float* __restrict__ dgamma,
                                         float* __restrict__ dbeta, int N, int slabs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * N) return;
  float s = 0.f;
  for (int k = 0; k < slabs; ++k) s += part[static_cast<size_t>(k) * 2 * N + i];
  if (i < N) dgamma[i] = s;
  else dbeta[i - N] = s;
}

// ------------------------------------------------------------------ host side
static int pick_skip(const spectre_norm_act_t& d, int* mode, int* k) {
  *k = 1;
  if (!d.skip) { *mode = kSkipNone; return 0; }
  if (d.Ns == d.N) { *mode = kSkipIdentity; return 0; }
  if (d.N % d.Ns == 0 && (d.N / d.Ns == 2 || d.N / d.Ns == 4 || d.N / d.Ns == 8)) { *mode = kSkipUp; *k = d.N / d.Ns; return 0; }
  if (d.Ns % d.N == 0) { *mode = kSkipDown; *k = d.Ns / d.N; return 0; }
  *mode = kSkipGeneral;
  return 0;
}

static void launch_shape(int N, int* vpt, int* block) {
  const int nvec = N / 8;
  if (nvec <= 32) { *vpt = 1; *block = 32; return; }
  *vpt = 2;
  int b = ((nvec + 1) / 2 + 31) / 32 * 32;
  *block = b;
}

static int validate(const char* fn, const spectre_norm_act_t* d, bool bwd) {
  if (!d) return fail(SPECTRE_E_NULL, "%s: null descriptor", fn);
  if (!d->in0 || !d->gamma || !d->beta || !d->mean || !d->rstd) return fail(SPECTRE_E_NULL, "%s: null pointer", fn);
  if (d->rows < 1 || d->N < 8 || (d->N % 8) || d->N > 4096)
    return fail(SPECTRE_E_SHAPE, "%s: need rows >= 1 and N a multiple of 8 in [8, 4096] (rows=%lld N=%d)", fn, d->rows, d->N);
  if (d->skip && (d->Ns < 1 || (d->Ns % 4)))
    return fail(SPECTRE_E_SHAPE, "%s: skip width must be a positive multiple of 4 (Ns=%d)", fn, d->Ns);
  if (d->in0_dtype != SPECTRE_DTYPE_F32 && d->in0_dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: in0_dtype=%d (0=float32, 1=bfloat16)", fn, d->in0_dtype);
  if (d->drop_p < 0.f || d->drop_p >= 1.f) return fail(SPECTRE_E_SHAPE, "%s: drop_p=%g outside [0,1)", fn, d->drop_p);
  const void* ptrs[] = {d->in0, d->in1, d->gamma, d->beta, d->skip, d->out, d->out_bf16, d->dout, d->dout_bf16,
                        d->d_in0, d->d_in1, d->d_skip, d->dgamma, d->dbeta, d->workspace};
  for (const void* q : ptrs)
    if (!aligned16(q)) return fail(SPECTRE_E_ALIGN, "%s: pointers must be 16-byte aligned", fn);
  if (!bwd && !d->out) return fail(SPECTRE_E_NULL, "%s: null out", fn);
  if (bwd && (!d->dout || !d->dgamma || !d->dbeta || (d->skip && !d->d_skip) || (!d->d_in0 && !d->d_in1)))
    return fail(SPECTRE_E_NULL, "%s: null gradient pointer", fn);
  return SPECTRE_OK;
}

static Args to_args(const spectre_norm_act_t& d) {
  Args a{};
  a.in0 = d.in0; a.in1 = d.in1; a.gamma = d.gamma; a.beta = d.beta; a.skip = d.skip;
  a.out = d.out; a.out_bf16 = d.out_bf16; a.mean = d.mean; a.rstd = d.rstd;
  a.dout = d.dout; a.dout_bf16 = d.dout_bf16; a.d_in0 = d.d_in0; a.d_in1 = d.d_in1; a.d_skip = d.d_skip;
  a.part = static_cast<float*>(d.workspace);
  a.rows = d.rows; a.N = d.N; a.Ns = d.skip ? d.Ns : d.N; a.act = d.act; a.eps = d.eps; a.drop_p = d.drop_p; a.seed = d.seed;
  pick_skip(d, &a.skip_mode, &a.k);
  return a;
}

static int bwd_grid(long long rows) {
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  return static_cast<int>(rows < cap ? rows : cap);
}

}  // namespace na
}  // namespace sb200

using namespace sb200;
using namespace sb200::na;

extern "C" {

int spectre_norm_act_fwd(const spectre_norm_act_t* d, void* stream) {
  if (int rc = validate("spectre_norm_act_fwd", d, false)) return rc;
  const Args a = to_args(*d);
  int vpt, block;
  launch_shape(a.N, &vpt, &block);
  const long long cap = static_cast<long long>(device_sm_count()) * 16;
  const int grid = static_cast<int>(a.rows < cap ? a.rows : cap);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool bf = d->in0_dtype == SPECTRE_DTYPE_BF16;
  if (vpt == 1) {
    if (bf) norm_act_fwd_kernel<__nv_bfloat16, 1><<<grid, block, 0, st>>>(a);
    else norm_act_fwd_kernel<float, 1><<<grid, block, 0, st>>>(a);
  } else {
    if (bf) norm_act_fwd_kernel<__nv_bfloat16, 2><<<grid, block, 0, st>>>(a);
    else norm_act_fwd_kernel<float, 2><<<grid, block, 0, st>>>(a);
  }
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "norm_act_fwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

size_t spectre_norm_act_bwd_workspace_bytes(long long rows, int N) {
  if (rows < 1 || N < 8) return 0;
  return static_cast<size_t>(bwd_grid(rows)) * 2 * N * sizeof(float);
}

int spectre_norm_act_bwd(const spectre_norm_act_t* d, void* stream) {
  if (int rc = validate("spectre_norm_act_bwd", d, true)) return rc;
  Args a = to_args(*d);
  int vpt, block;
  launch_shape(a.N, &vpt, &block);
  const int grid = bwd_grid(a.rows);
  if (!d->workspace || d->workspace_bytes < static_cast<size_t>(grid) * 2 * a.N * sizeof(float))
    return fail(SPECTRE_E_WORKSPACE, "spectre_norm_act_bwd: workspace %zu B < required %zu B", d->workspace_bytes,
                static_cast<size_t>(grid) * 2 * a.N * sizeof(float));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool bf = d->in0_dtype == SPECTRE_DTYPE_BF16;
  if (vpt == 1) {
    if (bf) norm_act_bwd_kernel<__nv_bfloat16, 1><<<grid, block, 0, st>>>(a);
    else norm_act_bwd_kernel<float, 1><<<grid, block, 0, st>>>(a);
  } else {
    if (bf) norm_act_bwd_kernel<__nv_bfloat16, 2><<<grid, block, 0, st>>>(a);
    else norm_act_bwd_kernel<float, 2><<<grid, block, 0, st>>>(a);
  }
  int launches = 1;
  if (a.skip_mode == kSkipGeneral) {
    const long long n = a.rows * a.Ns;
    pool_bwd_general_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(a);
    ++launches;
  }
  reduce_gamma_beta_kernel<<<(2 * a.N + 255) / 256, 256, 0, st>>>(a.part, d->dgamma, d->dbeta, a.N, grid);
  ++launches;
  last_launch_count() = launches;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "norm_act_bwd launch: %s", cudaGetErrorString(err));
  last_error().clear();
  return SPECTRE_OK;
}

}  // extern "C"
