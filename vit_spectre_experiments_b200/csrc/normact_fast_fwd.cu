// Forward row-wise tail for the bf16 training stream at exact-fit widths (N = 8 * VPT * TPR):
//
//   out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + pool_up(skip) )      [+ out_pool = avg4(out)]
//
// spectre_vit/models/spectre/spectre.py:66-67 and spectre_vit/models/spectre/layers.py:95-101, as normact.cu (which
// keeps every other width / dtype / skip ratio).  Same data movement as there -- one bulk async copy per operand and
// row group into a shared-memory ring, 16-byte stores straight from registers -- but the arithmetic is rebuilt for
// issue slots, which is what bounded the generic kernels (46 SASS instructions per element at 70 % issue utilisation):
//   * statistics, LayerNorm affine, GELU and the skip add run on packed fp32x2 (FFMA2 / FMUL2 / FADD2);
//   * bf16 operands enter through FHADD.BF16 (unpack and pivot subtraction in one instruction);
//   * Phi(x) by the tanh-form fit on MUFU.TANH (normact_common.cuh) instead of the erf rational (2 MUFU + 9 FMA);
//   * dropout costs one hash + compare per 8 elements (drop_mask8), the 1/(1-p) scale one packed multiply per pair;
//   * N, the pooling ratio and the row split are compile-time: no column predicates, no idle lanes.
#include "normact_fast.cuh"

namespace sb200 {
namespace na {

// threads per CTA: 320 (two CTAs = 20 warps per SM within the register file) where the row split divides it, else 256
constexpr int fwd_threads(int tpr) { return 320 % tpr == 0 ? 320 : 256; }

template <int VPT, int TPR, int MODE, int ACT>
__global__ void __launch_bounds__(fwd_threads(TPR), 2) fast_fwd_kernel(const Args a) {
  constexpr int N = 8 * VPT * TPR, RPC = fwd_threads(TPR) / TPR, NV = N / 8;
  constexpr int NS = skip_width(MODE, N);
  static_assert(!contig_mode(MODE) || (MODE == kFastUp32 ? VPT == 3 : VPT == 2), "3:2 modes: 24 or 16 columns per thread");
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t full[kMaxStages], empty[kMaxStages];
  __shared__ float red[4][2][16];
  const int rg = threadIdx.x / TPR, t = threadIdx.x % TPR;
  const int S = a.stages;
  const bool has_in1 = a.in1 != nullptr;
  const bool drop = a.drop_p > 0.f;
  const long long groups = (a.rows + RPC - 1) / RPC;
  const int n_it = static_cast<int>((groups - blockIdx.x + gridDim.x - 1) / gridDim.x);
  // first row of the group of iteration `it`: first0 + it * gstep (ascending, or descending from the last group)
  const long long gstep = static_cast<long long>(gridDim.x) * RPC * (a.reverse ? -1 : 1);
  const long long first0 = (a.reverse ? groups - 1 - blockIdx.x : static_cast<long long>(blockIdx.x)) * RPC;
  const int off_in1 = RPC * N * 2;
  const int off_skip = off_in1 * (has_in1 ? 2 : 1);

  auto issue = [&](int it, int stg) {   // one thread: the group's rows are contiguous in every operand
    const long long row0 = first0 + it * gstep;
    const long long left = a.rows - row0;
    const uint32_t nr = static_cast<uint32_t>(left < RPC ? left : RPC);
    uint8_t* st = ring + stg * a.stage_bytes;
    uint64_t* bar = &full[stg];
    const uint32_t bmain = nr * N * 2u, bskip = nr * NS * 2u;
    ptx::mbar_arrive_expect_tx(bar, bmain * (has_in1 ? 2u : 1u) + bskip);
    bulk_load(st, static_cast<const __nv_bfloat16*>(a.in0) + row0 * N, bmain, bar);
    if (has_in1) bulk_load(st + off_in1, static_cast<const __nv_bfloat16*>(a.in1) + row0 * N, bmain, bar);
    if (MODE != kSkipNone) bulk_load(st + off_skip, static_cast<const __nv_bfloat16*>(a.skip) + row0 * NS, bskip, bar);
  };
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
    for (int k = 0; k < S; ++k) {
      ptx::mbar_init(&full[k], 1);
      ptx::mbar_init(&empty[k], blockDim.x / 32);   // one arrival per warp
    }
    ptx::mbar_fence_init();
  }
  RingProducer prod;
  pdl_wait();   // nothing above touches global memory
  const Seed sd = drop ? load_seed(a) : Seed{0u, 0u};
  // the ring is consumed in groups of kGroup row groups: one CTA-wide barrier and one refill round per group
  constexpr int kGroup = 2;
  if (threadIdx.x == 0) {
    if constexpr (kAsyncRing) prod.pump(0, n_it, S, empty, issue);
    else
      for (int it = 0; it < S - kGroup && it < n_it; ++it) issue(it, it);
  }
  __syncthreads();

  const u64 c0 = splat(kGeluC0), c1 = splat(kGeluC1), c2 = splat(kGeluC2), half2 = splat(0.5f);
  const u64 ks2 = splat(a.keep_scale);
  // gamma / beta of this thread's columns stay in registers for the whole kernel (loading them per row group kept the
  // LSU data pipe 72-86 % busy and bounded the kernel: profiles/r02_ncu_normact_fast_v1.txt)
  u64 gam[VPT][4], bet[VPT][4];
#pragma unroll
  for (int i = 0; i < VPT; ++i)
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      gam[i][k] = ldg2(a.gamma + vec_of<MODE, VPT, TPR>(t, i) * 8 + 2 * k);
      bet[i][k] = ldg2(a.beta + vec_of<MODE, VPT, TPR>(t, i) * 8 + 2 * k);
    }
  constexpr float inv_n = 1.0f / N;
  int stg = 0;
  uint32_t phase = 0;
  long long row = first0 + rg;
  const uint8_t* st = ring;
  for (int it = 0; it < n_it; ++it, row += gstep) {
    if (threadIdx.x == 0) {
      if constexpr (kAsyncRing) {
        prod.pump(it, n_it, S, empty, issue);
      } else if (it % kGroup == 0) {   // the stages of the previous group are free (barrier below)
#pragma unroll 1
        for (int k = 0; k < kGroup; ++k)
          if (it + S - kGroup + k < n_it) issue(it + S - kGroup + k, (it + S - kGroup + k) % S);
      }
    }
    ptx::mbar_wait(&full[stg], phase);
    const bool live = row < a.rows;
    const uint4* r0 = reinterpret_cast<const uint4*>(st) + rg * NV;
    const uint4* r1 = reinterpret_cast<const uint4*>(st + off_in1) + rg * NV;
    // ---- statistics in one pass about a pivot from the row itself (its first element): no cancellation in
    // E[d^2] - E[d]^2, and d = z - pivot is what stays in registers
    float pivot = bf_lo(reinterpret_cast<const unsigned*>(r0)[0]);
    if (has_in1) pivot += bf_lo(reinterpret_cast<const unsigned*>(r1)[0]);
    const float npiv = -pivot;
    u64 d[VPT][4], s0[VPT], s1[VPT];
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const uint4 q = r0[vec_of<MODE, VPT, TPR>(t, i)];
      const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) d[i][k] = bf2_add(w[k], npiv);
      if (has_in1) {
        const uint4 q1 = r1[vec_of<MODE, VPT, TPR>(t, i)];
        const unsigned w1[4] = {q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) d[i][k] = bf2_add(w1[k], d[i][k]);
      }
      s0[i] = add2(d[i][0], d[i][1]);
      s1[i] = mul2(d[i][0], d[i][0]);
      s1[i] = fma2(d[i][1], d[i][1], s1[i]);
      s0[i] = add2(s0[i], add2(d[i][2], d[i][3]));
      s1[i] = fma2(d[i][2], d[i][2], s1[i]);
      s1[i] = fma2(d[i][3], d[i][3], s1[i]);
    }
#pragma unroll
    for (int i = 1; i < VPT; ++i) {
      s0[0] = add2(s0[0], s0[i]);
      s1[0] = add2(s1[0], s1[i]);
    }
    float sa = hsum(s0[0]), sb = hsum(s1[0]);
    row_sum2<TPR>(sa, sb, red, it & 3, rg);
    const float dmean = sa * inv_n;
    const float var = fmaxf(fmaf(-dmean, dmean, sb * inv_n), 0.f);
    const float rstd = rsqrtf(var + a.eps);
    if (live && t == 0) {
      a.mean[row] = pivot + dmean;
      a.rstd[row] = rstd;
    }
    const u64 rs2 = splat(rstd), sh2 = splat(-dmean * rstd);
    const unsigned rseed = drop ? row_seed(sd, row) : 0u;
    __nv_bfloat16* orow = static_cast<__nv_bfloat16*>(a.out) + row * N;
    float sk[contig_mode(MODE) ? NS / TPR : 1];   // 3:2 modes: this thread's 16 / 24 skip columns
    if constexpr (contig_mode(MODE)) {
      const uint4* sp = reinterpret_cast<const uint4*>(st + off_skip) + rg * (NS / 8) + t * (NS / TPR / 8);
#pragma unroll
      for (int q = 0; q < NS / TPR / 8; ++q) unpack8(sp[q], sk + 8 * q);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = vec_of<MODE, VPT, TPR>(t, i), j0 = v * 8;
      u64 r[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const u64 u = fma2(fma2(d[i][k], rs2, sh2), gam[i][k], bet[i][k]);
        if constexpr (ACT) {
          const u64 u2 = mul2(u, u);
          const u64 th = tanh2(mul2(u, fma2(fma2(c2, u2, c1), u2, c0)));
          const u64 hu = mul2(u, half2);
          r[k] = fma2(hu, th, hu);
        } else {
          r[k] = u;
        }
      }
      if constexpr (MODE == kSkipIdentity) {
        const uint4 q = (reinterpret_cast<const uint4*>(st + off_skip) + rg * NV)[v];
        r[0] = bf2_add(q.x, r[0]); r[1] = bf2_add(q.y, r[1]); r[2] = bf2_add(q.z, r[2]); r[3] = bf2_add(q.w, r[3]);
      } else if constexpr (MODE == kSkipUp4) {   // 2 skip columns feed these 8 outputs
        const unsigned w = (reinterpret_cast<const unsigned*>(st + off_skip) + rg * (NS / 2))[v];
        const u64 sl = splat(bf_lo(w)), sh = splat(bf_hi(w));
        r[0] = add2(r[0], sl); r[1] = add2(r[1], sl); r[2] = add2(r[2], sh); r[3] = add2(r[3], sh);
      } else if constexpr (contig_mode(MODE)) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          r[k] = add2(r[k], pk(skip32_val<MODE>(sk, 8 * i + 2 * k), skip32_val<MODE>(sk, 8 * i + 2 * k + 1)));
      }
      if (drop) {
        const unsigned dm = drop_mask8(a, rseed, static_cast<unsigned>(v));
        if (dm) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float lo, hi;
            upk(r[k], lo, hi);
            r[k] = pk((dm >> (2 * k) & 1u) ? 0.f : lo, (dm >> (2 * k + 1) & 1u) ? 0.f : hi);
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = mul2(r[k], ks2);
      }
      if (live) {
        *reinterpret_cast<uint4*>(orow + j0) = make_uint4(bfpack2(r[0]), bfpack2(r[1]), bfpack2(r[2]), bfpack2(r[3]));
        if (a.out_pool) {   // pool_k == 4: two pooled columns per vector
          const float p0 = hsum(add2(r[0], r[1])) * 0.25f, p1 = hsum(add2(r[2], r[3])) * 0.25f;
          reinterpret_cast<unsigned*>(static_cast<__nv_bfloat16*>(a.out_pool) + row * (N / 4))[v] = bf_pack(p0, p1);
        }
      }
    }
    if constexpr (kAsyncRing) {
      __syncwarp();
      if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&empty[stg]);   // this warp is done with the stage
    } else {
      if (it % kGroup == kGroup - 1 || it + 1 == n_it) __syncthreads();   // the group is consumed: its stages may be refilled
    }
    st += a.stage_bytes;
    if (++stg == S) { stg = 0; phase ^= 1u; st = ring; }
  }
}

// ------------------------------------------------------------------ host side
typedef void (*FastFn)(const Args);
template <int VPT, int TPR>
static FastFn pick_fwd(int mode, int act) {
#define SB200_ROW(M)                                                                      \
  if (mode == M) return act ? fast_fwd_kernel<VPT, TPR, M, 1> : fast_fwd_kernel<VPT, TPR, M, 0>;
  SB200_ROW(kSkipNone)
  SB200_ROW(kSkipIdentity)
  SB200_ROW(kSkipUp4)
  if constexpr (VPT == 3) { SB200_ROW(kFastUp32) }
  if constexpr (VPT == 2) { SB200_ROW(kFastDown32) }
#undef SB200_ROW
  return nullptr;
}

bool fast_fwd_launch(const Args& a0, cudaStream_t st, int* rc) {
  FastShape fs;
  if (!fast_shape(a0.N, false, &fs)) return false;
  const int mode = fast_mode(a0.skip_mode, a0.N, a0.Ns);
  if (mode < 0) return false;
  if (a0.out_pool && (a0.pool_k != 4 || contig_mode(mode))) return false;
  FastFn fn = nullptr;
#define SB200_SHAPE(V, T) if (fs.vpt == V && fs.tpr == T) fn = pick_fwd<V, T>(mode, a0.act ? 1 : 0);
  SB200_SHAPE(3, 16) SB200_SHAPE(3, 32) SB200_SHAPE(3, 64) SB200_SHAPE(3, 128)
  SB200_SHAPE(2, 16) SB200_SHAPE(2, 32) SB200_SHAPE(2, 64) SB200_SHAPE(2, 128)
#undef SB200_SHAPE
  if (!fn) return false;
  Args a = a0;
  const int threads = fwd_threads(fs.tpr), rpc = threads / fs.tpr;
  const int ns = skip_width(mode, a.N);
  const int row_bytes = a.N * 2 * (a.in1 ? 2 : 1) + ns * 2;
  a.tpr = fs.tpr; a.rpc = rpc;
  a.stage_bytes = (rpc * row_bytes + 127) / 128 * 128;
  // two CTAs per SM: ring of up to ~100 KB each
  int stages = (100 * 1024) / a.stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 3) return false;   // 2 stages being read + at least 1 in flight
  a.stages = stages;
  const size_t smem = static_cast<size_t>(stages) * a.stage_bytes;
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(fn), kMaxSmem) != cudaSuccess) {
    *rc = fail(SPECTRE_E_CUDA, "norm_act_fwd: cudaFuncSetAttribute failed");
    return true;
  }
  const long long groups = (a.rows + rpc - 1) / rpc;
  const long long cap = static_cast<long long>(device_sm_count()) * 2;
  const int grid = static_cast<int>(groups < cap ? groups : cap);
  launch_pdl(fn, dim3(grid), dim3(threads), smem, st, a);
  const cudaError_t err = cudaGetLastError();
  *rc = err == cudaSuccess ? SPECTRE_OK : fail(SPECTRE_E_CUDA, "norm_act_fwd (fast) launch: %s", cudaGetErrorString(err));
  return true;
}

}  // namespace na
}  // namespace sb200
