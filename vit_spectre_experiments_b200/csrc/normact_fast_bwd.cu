// Backward row-wise tail for the bf16 training stream at exact-fit widths (N = 8 * VPT * TPR); see normact_fast_fwd.cu.
//
//   d      = (dout + dout2 + up(d_out_pool) / 4) * dropout_mask / (1 - p)
//   d_skip = pool_up^T(d)                                   (identity or 4 -> 1 column sums)
//   g      = d * act'(u),  u = xhat * gamma + beta,  xhat = (in0 + in1 - mean) * rstd
//   d_in0  = rstd * (g gamma - mean_row(g gamma) - xhat * mean_row(g gamma xhat))
//   dgamma += g xhat,  dbeta += g,  dcolsum += d_in0        (per column, over all rows)
//
// What differs from normact.cu: packed fp32x2 arithmetic; GELU' = Phi + x phi with Phi from the tanh-form fit
// (MUFU.TANH) and phi from MUFU.EX2; one hash + compare per 8 elements for the dropout mask; and the per-column
// accumulators live in REGISTERS for the whole kernel (a thread owns the same 8 * VPT columns in every row it visits) --
// the generic kernel's shared-memory accumulators cost 24 bytes of shared-memory traffic and 1.5 instructions per
// element.  They are folded once at the end: per row group through shared memory in a fixed order, one partial per CTA,
// summed across CTAs by reduce_gamma_beta_kernel (deterministic).
#include "normact_fast.cuh"

namespace sb200 {
namespace na {

template <int VPT, int TPR, int RPC, int MODE, int ACT, int NACC>
__global__ void __launch_bounds__(TPR * RPC, 1) fast_bwd_kernel(const Args a) {
  constexpr int N = 8 * VPT * TPR, NV = N / 8, NT = TPR * RPC;
  constexpr int NS = skip_width(MODE, N);
  static_assert(!contig_mode(MODE) || (MODE == kFastUp32 ? VPT == 3 : VPT == 2), "3:2 modes: 24 or 16 columns per thread");
  extern __shared__ __align__(128) uint8_t ring[];
  __shared__ uint64_t full[kMaxStages], empty[kMaxStages];
  __shared__ float red[4][2][16];   // one slot per row group in flight between two CTA-wide barriers (x2: see row_sum2)
  __shared__ __align__(16) float sgb[2 * N];   // gamma, beta
  const int rg = threadIdx.x / TPR, t = threadIdx.x % TPR;
  const int S = a.stages;
  const bool has_in1 = a.in1 != nullptr, has_d2 = a.dout2 != nullptr, has_dp = a.d_out_pool != nullptr;
  const bool drop = a.drop_p > 0.f;
  const long long groups = (a.rows + RPC - 1) / RPC;
  const int n_it = static_cast<int>((groups - blockIdx.x + gridDim.x - 1) / gridDim.x);
  const long long gstep = static_cast<long long>(gridDim.x) * RPC * (a.reverse ? -1 : 1);
  const long long first0 = (a.reverse ? groups - 1 - blockIdx.x : static_cast<long long>(blockIdx.x)) * RPC;
  // ring stage: dout | in0 | (in1) | (dout2) | (d_out_pool: N / 4 columns)
  constexpr int kOp = RPC * N * 2;
  const int off_in1 = 2 * kOp;
  const int off_d2 = off_in1 + (has_in1 ? kOp : 0);
  const int off_dp = off_d2 + (has_d2 ? kOp : 0);

  auto issue = [&](int it, int stg) {
    const long long row0 = first0 + it * gstep;
    const long long left = a.rows - row0;
    const uint32_t nr = static_cast<uint32_t>(left < RPC ? left : RPC);
    uint8_t* st = ring + stg * a.stage_bytes;
    uint64_t* bar = &full[stg];
    const uint32_t bmain = nr * N * 2u, bpool = nr * (N / 4) * 2u;
    ptx::mbar_arrive_expect_tx(bar, bmain * (2u + (has_in1 ? 1u : 0u) + (has_d2 ? 1u : 0u)) + (has_dp ? bpool : 0u));
    bulk_load(st, static_cast<const __nv_bfloat16*>(a.dout) + row0 * N, bmain, bar);
    bulk_load(st + kOp, static_cast<const __nv_bfloat16*>(a.in0) + row0 * N, bmain, bar);
    if (has_in1) bulk_load(st + off_in1, static_cast<const __nv_bfloat16*>(a.in1) + row0 * N, bmain, bar);
    if (has_d2) bulk_load(st + off_d2, static_cast<const __nv_bfloat16*>(a.dout2) + row0 * N, bmain, bar);
    if (has_dp) bulk_load(st + off_dp, static_cast<const __nv_bfloat16*>(a.d_out_pool) + row0 * (N / 4), bpool, bar);
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
  // The ring is consumed in GROUPS of kGroup row groups: one CTA-wide barrier (and one refill round by thread 0) per
  // group instead of per row group -- barrier stalls were 1.0 of the ~5 stall cycles per issued instruction (ncu v7).
  // kGroup stages are being read, S - kGroup are in flight.
  constexpr int kGroup = 2;
  if (threadIdx.x == 0) {
    if constexpr (kAsyncRing) prod.pump(0, n_it, S, empty, issue);
    else
      for (int it = 0; it < S - kGroup && it < n_it; ++it) issue(it, it);
  }
  for (int i = threadIdx.x; i < N; i += NT) {
    sgb[i] = a.gamma[i];
    sgb[N + i] = a.beta[i];
  }
  __syncthreads();

  // per-column accumulators of this thread's 8 * VPT columns: dgamma, dbeta (, column sums of d_in0)
  u64 acc[NACC][VPT][4];
#pragma unroll
  for (int n = 0; n < NACC; ++n)
#pragma unroll
    for (int i = 0; i < VPT; ++i)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[n][i][k] = 0ull;

  const u64 c0 = splat(kGeluC0), c1 = splat(kGeluC1), c2 = splat(kGeluC2), half2 = splat(0.5f);
  const u64 cphi = splat(kInvSqrt2Pi), cexp = splat(-kHalfLog2e), ks2 = splat(a.keep_scale), quarter2 = splat(0.25f);
  constexpr float inv_n = 1.0f / N;
  float mean_n = 0.f, rstd_n = 0.f;
  long long row = first0 + rg;
  if (n_it > 0 && row < a.rows) { mean_n = a.mean[row]; rstd_n = a.rstd[row]; }
  int stg = 0;
  uint32_t phase = 0;
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
    const bool live = row < a.rows;
    const float mean = mean_n, rstd = rstd_n;
    {
      const long long rown = row + gstep;
      if (it + 1 < n_it && rown < a.rows) { mean_n = a.mean[rown]; rstd_n = a.rstd[rown]; }
    }
    ptx::mbar_wait(&full[stg], phase);
    const u64 rs2 = splat(rstd), sh2 = splat(-mean * rstd);
    const unsigned rseed = drop ? row_seed(sd, row) : 0u;
    __nv_bfloat16* dsrow = static_cast<__nv_bfloat16*>(a.d_skip) + row * NS;
    u64 xh[VPT][4], gy[VPT][4];
    u64 s0 = 0ull, s1 = 0ull;
    if (live) {   // rows past the end (last group only) hold stale shared memory: they must not reach the accumulators
    float ds[contig_mode(MODE) ? NS / TPR : 1];   // 3:2 modes: gradient of this thread's 16 / 24 skip columns
    if constexpr (contig_mode(MODE)) {
#pragma unroll
      for (int q = 0; q < NS / TPR; ++q) ds[q] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = vec_of<MODE, VPT, TPR>(t, i), j0 = v * 8;
      const uint4 qd = (reinterpret_cast<const uint4*>(st) + rg * NV)[v];
      const uint4 qz = (reinterpret_cast<const uint4*>(st + kOp) + rg * NV)[v];
      const unsigned wd[4] = {qd.x, qd.y, qd.z, qd.w}, wz[4] = {qz.x, qz.y, qz.z, qz.w};
      u64 d[4], z[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        d[k] = bf2(wd[k]);
        z[k] = bf2(wz[k]);
      }
      if (has_in1) {
        const uint4 q = (reinterpret_cast<const uint4*>(st + off_in1) + rg * NV)[v];
        z[0] = bf2_add(q.x, z[0]); z[1] = bf2_add(q.y, z[1]); z[2] = bf2_add(q.z, z[2]); z[3] = bf2_add(q.w, z[3]);
      }
      if (has_d2) {
        const uint4 q = (reinterpret_cast<const uint4*>(st + off_d2) + rg * NV)[v];
        d[0] = bf2_add(q.x, d[0]); d[1] = bf2_add(q.y, d[1]); d[2] = bf2_add(q.z, d[2]); d[3] = bf2_add(q.w, d[3]);
      }
      if (has_dp) {   // gradient through the pooled copy: 2 pooled columns cover these 8
        const unsigned w = (reinterpret_cast<const unsigned*>(st + off_dp) + rg * (N / 8))[v];
        const u64 pl = splat(bf_lo(w)), ph = splat(bf_hi(w));
        d[0] = fma2(pl, quarter2, d[0]); d[1] = fma2(pl, quarter2, d[1]);
        d[2] = fma2(ph, quarter2, d[2]); d[3] = fma2(ph, quarter2, d[3]);
      }
      if (drop) {
        const unsigned dm = drop_mask8(a, rseed, static_cast<unsigned>(v));
        if (dm) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float lo, hi;
            upk(d[k], lo, hi);
            d[k] = pk((dm >> (2 * k) & 1u) ? 0.f : lo, (dm >> (2 * k + 1) & 1u) ? 0.f : hi);
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) d[k] = mul2(d[k], ks2);
      }
      // ---- skip path: transpose of the pooling
      if constexpr (MODE == kSkipIdentity) {
        *reinterpret_cast<uint4*>(dsrow + j0) = make_uint4(bfpack2(d[0]), bfpack2(d[1]), bfpack2(d[2]), bfpack2(d[3]));
      } else if constexpr (MODE == kSkipUp4) {
        reinterpret_cast<unsigned*>(dsrow)[v] = bf_pack(hsum(add2(d[0], d[1])), hsum(add2(d[2], d[3])));
      } else if constexpr (contig_mode(MODE)) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float lo, hi;
          upk(d[k], lo, hi);
          skip32_scatter<MODE>(ds, 8 * i + 2 * k, lo);
          skip32_scatter<MODE>(ds, 8 * i + 2 * k + 1, hi);
        }
      }
      // ---- main path
      const float4 ga = *reinterpret_cast<const float4*>(sgb + j0), gb = *reinterpret_cast<const float4*>(sgb + j0 + 4);
      const u64 gam[4] = {pk(ga.x, ga.y), pk(ga.z, ga.w), pk(gb.x, gb.y), pk(gb.z, gb.w)};
      u64 bet[4];
      if constexpr (ACT) {
        const float4 ba = *reinterpret_cast<const float4*>(sgb + N + j0), bb = *reinterpret_cast<const float4*>(sgb + N + j0 + 4);
        bet[0] = pk(ba.x, ba.y); bet[1] = pk(ba.z, ba.w); bet[2] = pk(bb.x, bb.y); bet[3] = pk(bb.z, bb.w);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const u64 x = fma2(z[k], rs2, sh2);
        u64 g = d[k];
        if constexpr (ACT) {
          const u64 u = fma2(x, gam[k], bet[k]);
          const u64 u2 = mul2(u, u);
          const u64 th = tanh2(mul2(u, fma2(fma2(c2, u2, c1), u2, c0)));
          const u64 e = ex2_2(mul2(u2, cexp));                       // exp(-u^2 / 2)
          const u64 cdf = fma2(th, half2, half2);
          g = mul2(g, fma2(mul2(u, e), cphi, cdf));                  // d * (Phi(u) + u phi(u))
        }
        acc[1][i][k] = add2(acc[1][i][k], g);
        acc[0][i][k] = fma2(g, x, acc[0][i][k]);
        const u64 y = mul2(g, gam[k]);
        xh[i][k] = x;
        gy[i][k] = y;
        s0 = add2(s0, y);
        s1 = fma2(y, x, s1);
      }
    }
    if constexpr (contig_mode(MODE)) {
      uint4* dp = reinterpret_cast<uint4*>(dsrow) + t * (NS / TPR / 8);
#pragma unroll
      for (int q = 0; q < NS / TPR / 8; ++q)
        dp[q] = make_uint4(bf_pack(ds[8 * q], ds[8 * q + 1]), bf_pack(ds[8 * q + 2], ds[8 * q + 3]),
                           bf_pack(ds[8 * q + 4], ds[8 * q + 5]), bf_pack(ds[8 * q + 6], ds[8 * q + 7]));
    }
    }
    if constexpr (kAsyncRing) {   // dout / in0 are in registers now: the stage can be refilled
      __syncwarp();
      if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&empty[stg]);
    }
    float sa = hsum(s0), sb = hsum(s1);
    row_sum2<TPR>(sa, sb, red, it & 3, rg);
    if (live) {
      const u64 cb = splat(-rstd * sa * inv_n), cx = splat(-rstd * sb * inv_n), cr = splat(rstd);
      __nv_bfloat16* drow = static_cast<__nv_bfloat16*>(a.d_in0) + row * N;
#pragma unroll
      for (int i = 0; i < VPT; ++i) {
        u64 dz[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          dz[k] = fma2(xh[i][k], cx, fma2(gy[i][k], cr, cb));
          if constexpr (NACC == 3) acc[2][i][k] = add2(acc[2][i][k], dz[k]);
        }
        *reinterpret_cast<uint4*>(drow + vec_of<MODE, VPT, TPR>(t, i) * 8) =
            make_uint4(bfpack2(dz[0]), bfpack2(dz[1]), bfpack2(dz[2]), bfpack2(dz[3]));
      }
    }
    if constexpr (!kAsyncRing) {
      if (it % kGroup == kGroup - 1 || it + 1 == n_it) __syncthreads();   // group consumed
    }
    st += a.stage_bytes;
    if (++stg == S) { stg = 0; phase ^= 1u; st = ring; }
  }
  // ---- fold the row groups of the CTA in a fixed order (through the now idle ring, one accumulator kind at a time)
  if constexpr (kAsyncRing) __syncthreads();   // every warp is past its last stage (and no bulk copy is in flight)
  float* fold = reinterpret_cast<float*>(ring);
  float* pg = a.part + static_cast<size_t>(blockIdx.x) * NACC * N;
#pragma unroll
  for (int n = 0; n < NACC; ++n) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      float* dst = fold + rg * N + vec_of<MODE, VPT, TPR>(t, i) * 8;
      float v[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) upk(acc[n][i][k], v[2 * k], v[2 * k + 1]);
      *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < N; c += NT) {
      float s = fold[c];
#pragma unroll
      for (int r = 1; r < RPC; ++r) s += fold[r * N + c];
      pg[n * N + c] = s;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ host side
typedef void (*FastFn)(const Args);
// (activation, accumulators) pairs that occur: LayerNorm alone -> (0, 2); GELU tail of a Linear -> (1, 3) with the
// bias gradient as column sums, (1, 2) without.  Anything else falls back to the generic kernels.
template <int VPT, int TPR, int RPC, int MODE, int ACT>
static FastFn pick_bwd_mode(int nacc) {
  if constexpr (ACT) {
    if (nacc == 3) return fast_bwd_kernel<VPT, TPR, RPC, MODE, 1, 3>;
    if (nacc == 2) return fast_bwd_kernel<VPT, TPR, RPC, MODE, 1, 2>;
  } else {
    if (nacc == 2) return fast_bwd_kernel<VPT, TPR, RPC, MODE, 0, 2>;
  }
  return nullptr;
}
template <int VPT, int TPR, int RPC, int ACT>
static FastFn pick_bwd(int mode, int nacc) {
  if (mode == kSkipNone) return pick_bwd_mode<VPT, TPR, RPC, kSkipNone, ACT>(nacc);
  if (mode == kSkipIdentity) return pick_bwd_mode<VPT, TPR, RPC, kSkipIdentity, ACT>(nacc);
  if (mode == kSkipUp4) return pick_bwd_mode<VPT, TPR, RPC, kSkipUp4, ACT>(nacc);
  if constexpr (VPT == 3 && ACT == 1) { if (mode == kFastUp32) return pick_bwd_mode<VPT, TPR, RPC, kFastUp32, ACT>(nacc); }
  if constexpr (VPT == 2 && ACT == 1) { if (mode == kFastDown32) return pick_bwd_mode<VPT, TPR, RPC, kFastDown32, ACT>(nacc); }
  return nullptr;
}

int fast_bwd_grid_bound() { return device_sm_count(); }

bool fast_bwd_launch(const Args& a0, size_t workspace_bytes, cudaStream_t st, int* rc, int* grid_out) {
  FastShape fs;
  if (!fast_shape(a0.N, true, &fs)) return false;
  const int mode = fast_mode(a0.skip_mode, a0.N, a0.Ns);
  if (mode < 0) return false;
  if (a0.d_out_pool && (a0.pool_k != 4 || contig_mode(mode))) return false;
  FastFn fn = nullptr;
  int threads = 0;
  // rows per CTA: the GELU kernels are issue-bound at 8-12 warps per SM (profiles/r02_ncu_normact_fast_v1.txt), so
  // they take the largest CTA their register count allows; the LayerNorm-only kernels run at the HBM roof as they are
#define SB200_SHAPE(V, T, R0, R1)                                              \
  if (fs.vpt == V && fs.tpr == T) {                                            \
    if (a0.act) {                                                              \
      fn = pick_bwd<V, T, R1, 1>(mode, a0.nacc);                          \
      threads = T * R1;                                                        \
    } else {                                                                   \
      fn = pick_bwd<V, T, R0, 0>(mode, a0.nacc);                          \
      threads = T * R0;                                                        \
    }                                                                          \
  }
  SB200_SHAPE(3, 16, 16, 24) SB200_SHAPE(3, 32, 8, 12) SB200_SHAPE(2, 96, 4, 5) SB200_SHAPE(2, 192, 2, 2)
  SB200_SHAPE(2, 16, 16, 24) SB200_SHAPE(2, 32, 8, 12) SB200_SHAPE(2, 64, 4, 6) SB200_SHAPE(2, 128, 2, 3)
  SB200_SHAPE(3, 64, 4, 6) SB200_SHAPE(3, 128, 2, 3)
#undef SB200_SHAPE
  if (!fn) return false;
  Args a = a0;
  const int rpc = threads / fs.tpr;
  const int row_bytes = a.N * 2 * (2 + (a.in1 ? 1 : 0) + (a.dout2 ? 1 : 0)) + (a.d_out_pool ? a.N / 4 * 2 : 0);
  a.tpr = fs.tpr; a.rpc = rpc;
  a.stage_bytes = (rpc * row_bytes + 127) / 128 * 128;
  // one CTA per SM (the accumulators bound the registers): ring of up to ~190 KB, at least the fold buffer
  int stages = (190 * 1024) / a.stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 3) return false;   // 2 stages being read + at least 1 in flight
  a.stages = stages;
  size_t smem = static_cast<size_t>(stages) * a.stage_bytes;
  const size_t fold = static_cast<size_t>(rpc) * a.N * sizeof(float);
  if (smem < fold) smem = fold;
  if (smem > static_cast<size_t>(kMaxSmem)) return false;
  const long long groups = (a.rows + rpc - 1) / rpc;
  const long long cap = device_sm_count();
  const int grid = static_cast<int>(groups < cap ? groups : cap);
  const size_t need = static_cast<size_t>(grid) * a.nacc * a.N * sizeof(float);
  if (!a.part || workspace_bytes < need) {
    *rc = fail(SPECTRE_E_WORKSPACE, "spectre_norm_act_bwd: workspace %zu B < required %zu B", workspace_bytes, need);
    return true;
  }
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(fn), kMaxSmem) != cudaSuccess) {
    *rc = fail(SPECTRE_E_CUDA, "norm_act_bwd: cudaFuncSetAttribute failed");
    return true;
  }
  launch_pdl(fn, dim3(grid), dim3(threads), smem, st, a);
  const cudaError_t err = cudaGetLastError();
  *rc = err == cudaSuccess ? SPECTRE_OK : fail(SPECTRE_E_CUDA, "norm_act_bwd (fast) launch: %s", cudaGetErrorString(err));
  *grid_out = grid;
  return true;
}

}  // namespace na
}  // namespace sb200
