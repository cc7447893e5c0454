// FFT token mixer for bf16 activations on the 5th-gen tensor cores (tcgen05 / TMEM / TMA).
//
// Same operator as fftmix.cu (rfft2 -> * filter -> irfft2 in the reference's mix_layer slot,
// spectre_vit/models/spectre/spectre.py:50,66; conventions of spectre.py:136), but the small
// DFTs are recast as dense real matmuls against constant DFT matrices (the escape clause of
// the north-star: the fp32 SIMT butterflies are issue-bound at 14x14 - profiles/r01_ncu_*).
//
//   tile            = one image x 128 channels.  M (the MMA row / TMEM lane) = channel, so
//                     every re-layout between the four separable stages is register renaming
//                     inside one thread (thread = TMEM lane = channel).
//   TMA             : [H*W tokens x 64 ch] boxes, 128B-swizzled, land as the MN-major A operand.
//   stage 1 (rows)  : per grid row h   D1[c, (kw,ri)]  = A_smem[c, w]        . B1   (SS, N=16, K=16)
//   stage 2 (cols)  : per kw           D2[c, (kh,ri)]  = A2_tmem[c, (h,ri)]  . B2   (TS, N=K=2H pad 16)
//   filter          : Z = D2 * filt (fp32, per-thread registers), packed to bf16 -> A3
//   stage 3 (cols^-1): per kw          D3[c, (h,ri)]   = A3_tmem[c, (kh,ri)] . B3
//   stage 4 (rows^-1): per grid row h  D4[c, w]        = A4_tmem[c, (kw,ri)] . B4   (C2R incl. weights)
//   output          : bf16 rows staged in a swizzled smem tile, written back by one TMA store.
// Accumulation is fp32 in TMEM; the operands of every stage are bf16 (three intermediate
// roundings + bf16 DFT matrices => <1e-2 of max|y|, inside the 3e-2 bf16 tolerance).
//
// Warp roles (384 threads, one CTA per SM, persistent over images of one channel tile):
//   warps 0-7  epilogue: warpgroup g handles half of the kw columns (stages 1-3) / half of
//              the grid rows (stage 4); its slice of the filter lives in registers.
//   warp  8    MMA issuer (one elected lane);  warp 9  TMA producer + CLS pass-through;  10-11 idle.
#include <cuda.h>

#include <type_traits>

#include "common.cuh"
#include "fft_codelets.cuh"
#include "sm100_ptx.cuh"

namespace sb200 {
namespace tc {

using namespace ptx;

// Debug timeline: CTA 0 records clock64 at phase boundaries; slot layout [who][tile][stamp].
//   who 0/1 = epilogue warpgroup 0/1 (first thread), 2 = MMA lane, 3 = TMA lane
constexpr int kProfStamps = 16, kProfTiles = 8;
#define SB200_STAMP(who, tile, idx)                                                              \
  do {                                                                                          \
    if (p.prof && blockIdx.x == 0 && (tile) < kProfTiles)                                       \
      p.prof[((who) * kProfTiles + (tile)) * kProfStamps + (idx)] = clock64();                  \
  } while (0)


// wall-clock stamps (debug, -DSB200_WALLCLOCK=1): %globaltimer + clock64 at entry / exit of CTA 0 and of the last CTA
#ifndef SB200_WALLCLOCK
#define SB200_WALLCLOCK 0
#endif
__device__ __forceinline__ long long global_ns() {
  long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define SB200_WALL(slot)                                                                         \
  do {                                                                                          \
    if (SB200_WALLCLOCK && p.prof && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == gridDim.x - 1)) { \
      long long* q = p.prof + ((4 * kProfTiles + (blockIdx.x == 0 ? 0 : 1)) * kProfStamps);     \
      q[(slot)] = global_ns();                                                                  \
      q[(slot) + 2] = clock64();                                                                \
    }                                                                                           \
    if (SB200_WALLCLOCK && p.prof && threadIdx.x == 0) p.prof[1024 + 2 * blockIdx.x + (slot)] = global_ns(); \
  } while (0)

constexpr int kEpiThreads = 256;
constexpr int kThreads = 384;
constexpr int kBarCount = kEpiThreads + 32;  // named barriers: epilogue arrives, MMA warp syncs
constexpr int kTileC = 128;
constexpr int kEpiRegs = 216, kAuxRegs = 72;  // setmaxnreg split: 8*32*216 + 4*32*72 = 384*168 (the launch allocation)

// K-major, no-swizzle B operand: byte offset of element (n, k) of an [N x K] bf16 matrix
__host__ __device__ constexpr uint32_t bmat_off(int n, int k, int K) {
  return (n % 8) * 16 + (n / 8) * (K / 8) * 128 + (k % 8) * 2 + (k / 8) * 128;
}

// round-to-nearest-even bf16 bit pattern of a double in the normal range (or 0), constexpr
constexpr uint16_t bf16_bits(double v) {
  if (v == 0.0) return 0;
  const uint16_t sign = v < 0 ? 0x8000u : 0u;
  double a = v < 0 ? -v : v;
  int e = 0;
  while (a >= 2.0) { a *= 0.5; ++e; }
  while (a < 1.0) { a *= 2.0; --e; }
  const double scaled = a * 128.0;                 // [128, 256)
  int m = static_cast<int>(scaled);
  const double frac = scaled - m;
  if (frac > 0.5 || (frac == 0.5 && (m & 1))) ++m;
  if (m == 256) { m = 128; ++e; }
  return static_cast<uint16_t>(sign | ((e + 127) << 7) | (m - 128));
}

template <int H_, int W_>
struct Cfg {
  static constexpr int H = H_, W = W_, HW = H * W, WH = W / 2 + 1;
  static_assert(W <= 14 && H <= 16 && WH <= 8, "tensor-core path: W <= 14, H <= 16");
  static constexpr int KWG = (WH + 1) / 2;             // kw columns per warpgroup (<= 4)
  static constexpr int HG = (H + 1) / 2;               // grid rows per warpgroup in stage 4
  static constexpr int K2 = (2 * H + 15) / 16 * 16;    // column-stage K and N
  static constexpr int A2C = K2 / 2;                   // TMEM columns per kw of a bf16 A operand
  static constexpr int kDCol = 0, kACol = 256;         // TMEM column bases
  static constexpr int kFCol = 384;                    // filter bins kh >= 8 (64 columns per warpgroup)
  static constexpr int HR = H < 8 ? H : 8;             // filter bins per kw kept in registers
  static constexpr int kHalfStride = ((HW + 16 - W) * 128 + 1023) / 1024 * 1024;  // 64-channel half
  static constexpr int kStageBytes = 2 * kHalfStride;
  static constexpr int kB1 = 16 * 16 * 2, kB2 = K2 * K2 * 2;
  static constexpr int kOffB1 = 0, kOffB4 = kB1, kOffB2 = 2 * kB1, kOffB3 = 2 * kB1 + kB2;
  static constexpr int kMatBytes = 2 * kB1 + 2 * kB2;
  // K / N slot of a half-spectrum column: warpgroup g owns slots [4g, 4g+4)
  __host__ __device__ static constexpr int slot(int kw) { return kw < KWG ? kw : 4 + (kw - KWG); }
  __host__ __device__ static constexpr int kw_of_slot(int s) {
    for (int q = 0; q < WH; ++q)
      if (slot(q) == s) return q;
    return -1;
  }
};

// The four DFT matrices, already in their shared-memory byte layout, built at compile time.
template <int H, int W>
struct MatImage {
  uint16_t w[Cfg<H, W>::kMatBytes / 2];
};

template <int H, int W>
constexpr MatImage<H, W> make_mats() {
  using C = Cfg<H, W>;
  MatImage<H, W> m{};
  for (int i = 0; i < C::kMatBytes / 2; ++i) m.w[i] = 0;
  for (int n = 0; n < 16; ++n)
    for (int k = 0; k < 16; ++k) {
      // B1[n = 2*slot(kw)+ri][k = w]: forward real DFT along W
      const int kw = C::kw_of_slot(n / 2);
      if (kw >= 0 && k < W) {
        const double c = cos2pi((kw * k) % W, W), s = sin2pi((kw * k) % W, W);
        m.w[(C::kOffB1 + bmat_off(n, k, 16)) / 2] = bf16_bits(n % 2 == 0 ? c : -s);
      }
      // B4[n = w][k = 2*slot(kw)+ri]: C2R along W with the Hermitian weights and 1/(H*W) folded in
      const int kw4 = C::kw_of_slot(k / 2);
      if (kw4 >= 0 && n < W) {
        const double c = cos2pi((kw4 * n) % W, W), s = sin2pi((kw4 * n) % W, W);
        // Hermitian weight and the ortho scale 1/(H*W) of the whole rfft2 -> irfft2 round trip
        const double wgt = ((kw4 == 0 || (W % 2 == 0 && kw4 == W / 2)) ? 1.0 : 2.0) / static_cast<double>(H * W);
        m.w[(C::kOffB4 + bmat_off(n, k, 16)) / 2] = bf16_bits(k % 2 == 0 ? wgt * c : -wgt * s);
      }
    }
  for (int n = 0; n < C::K2; ++n)
    for (int k = 0; k < C::K2; ++k) {
      // B2[n = 2kh+ri'][k = 2h+ri]: forward complex DFT along H (e^{-i phi});  B3: inverse (e^{+i phi})
      const int a = n / 2, ra = n % 2, b = k / 2, rb = k % 2;
      if (a < H && b < H) {
        const double c = cos2pi((a * b) % H, H), s = sin2pi((a * b) % H, H);
        m.w[(C::kOffB2 + bmat_off(n, k, C::K2)) / 2] = bf16_bits(ra == rb ? c : (ra == 0 ? s : -s));
        m.w[(C::kOffB3 + bmat_off(n, k, C::K2)) / 2] = bf16_bits(ra == rb ? c : (ra == 0 ? -s : s));
      }
    }
  return m;
}

template <int H, int W>
__device__ const MatImage<H, W> g_mats = make_mats<H, W>();

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ float f32(uint32_t u) { return __uint_as_float(u); }
__device__ __forceinline__ void st_shared_u16(uint32_t saddr, uint32_t v) {
  asm volatile("st.shared.u16 [%0], %1;" ::"r"(saddr), "h"(static_cast<unsigned short>(v)) : "memory");
}

// byte offset of (token row r, channel m of the 128-wide tile) in a [2 halves][rows][64 ch] 128B-swizzled tile
template <class C>
__device__ __forceinline__ uint32_t tile_off(int r, int m) {
  return (m >> 6) * C::kHalfStride + r * 128 + (((((m & 63) >> 3) ^ (r & 7))) << 4) + (m & 7) * 2;
}

template <int H_, int W_>
struct FwdSmem {
  using C = Cfg<H_, W_>;
  static constexpr int kBudget = 220 * 1024;
  static constexpr int kFit = (kBudget - C::kMatBytes) / C::kStageBytes - 1;   // input stages next to one output tile
  static constexpr int kStages = kFit > 4 ? 4 : kFit;
  static_assert(kStages >= 2, "forward: need at least two input stages");
  static constexpr int kOut = kStages * C::kStageBytes;
  static constexpr int kMat = kOut + C::kStageBytes;
  static constexpr int kBytes = kMat + C::kMatBytes + 1024 /*alignment slack*/;
};

struct FwdParams {
  const __nv_bfloat16* x;
  const float* filt;
  __nv_bfloat16* y;
  int B, E, cls, nct, img_stride;
  long long* prof;   // optional per-phase clock64 stamps of CTA 0 (debug; null in production)
};

template <int H, int W>
__global__ void __launch_bounds__(kThreads, 1)
fftmix_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_y,
                     const FwdParams p) {
  using C = Cfg<H, W>;
  using SM = FwdSmem<H, W>;
  constexpr int WH = C::WH, KWG = C::KWG, HG = C::HG, K2 = C::K2, A2C = C::A2C, S = SM::kStages;
  constexpr int WP = (W + 1) / 2;   // packed bf16 pairs per output row
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sOut = smem + SM::kOut;
  uint8_t* sMat = smem + SM::kMat;
  // dready_bar[stage][sub]: every stage's MMAs are committed in up to four sub-batches so the epilogue of the
  // first batch overlaps the issue / execution of the rest (the issuing thread needs ~50 clk per MMA)
  __shared__ uint64_t full_bar[S], empty_bar[S], dready_bar[4][4];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) uint8_t sDummy[8 * 16];   // sink for the padded rows of stmatrix

  const long long t_start = clock64();
  pdl_launch_dependents();
  SB200_WALL(0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ct = blockIdx.x % p.nct;
  const int img0 = blockIdx.x / p.nct;
  const int c0 = ct * kTileC;
  const int ntiles = img0 < p.B ? (p.B - img0 + p.img_stride - 1) / p.img_stride : 0;
  // images are swept from the last to the first: the producer of x (a row kernel sweeping upwards) left its most
  // recent ~100 MB in L2, and the consumer of y starts at image 0, which this sweep writes last
  const bool half1 = c0 + 64 < p.E;

  // ---- one-time setup
  if (warp == 8) tmem_alloc<512>(&tmem_base_s);
  if (tid == 9 * 32) {
    prefetch_tensormap(&tmap_x);
    prefetch_tensormap(&tmap_y);
    for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int k = 0; k < 16; ++k) mbar_init(&dready_bar[k >> 2][k & 3], 1);
    mbar_fence_init();
    // prologue loads: fill the input ring before the CTA-wide setup barrier (only the barriers are needed)
    pdl_wait();   // x is the preceding kernel's output; everything else in the setup is independent of it
    const uint32_t bytes0 = (half1 ? 2u : 1u) * C::HW * 128u;
    for (int it = 0; it < S && it < ntiles; ++it) {
      uint8_t* dst = smem + it * C::kStageBytes;
      const int b = img0 + (ntiles - 1 - it) * p.img_stride;
      mbar_arrive_expect_tx(&full_bar[it], bytes0);
      tma_load_3d(dst, &tmap_x, &full_bar[it], c0, 0, b);
      if (half1) tma_load_3d(dst + C::kHalfStride, &tmap_x, &full_bar[it], c0 + 64, 0, b);
    }
  }
  // epilogue threads: issue the loads of their filter slice now so the latency overlaps the setup
  float2 Fld[KWG][H];
  if (warp < 8) {
    const int g = warp >> 2, c = c0 + (warp & 3) * 32 + lane;
    const bool cvalid = c < p.E;
    const float2* fp = reinterpret_cast<const float2*>(p.filt) + (cvalid ? c : 0);
#pragma unroll
    for (int j = 0; j < KWG; ++j) {
      const int kw = g * KWG + j;
      const bool ok = kw < WH && cvalid;
#pragma unroll
      for (int kh = 0; kh < H; ++kh)   // branch-free, and no arithmetic on the values: the loads stay in flight
        Fld[j][kh] = __ldg(fp + (ok ? (static_cast<size_t>(kh) * WH + kw) * p.E : 0));
    }
  }
  for (int i = tid; i < C::kMatBytes / 16; i += kThreads)
    reinterpret_cast<uint4*>(sMat)[i] = reinterpret_cast<const uint4*>(g_mats<H, W>.w)[i];
  // rows past the image in every input half (read by the last grid row's K=16 window) must be finite
  {
    constexpr int per = (C::kHalfStride - C::HW * 128) / 16;
    for (int i = tid; i < S * 2 * per; i += kThreads)
      reinterpret_cast<uint4*>(smem + (i / per) * C::kHalfStride + C::HW * 128)[i % per] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  if (tid == 0 && p.prof && blockIdx.x == 0) { p.prof[14] = t_start; p.prof[15] = clock64(); }

  if (warp < 8) {
    // =========================================================== epilogue warpgroups
    setmaxnreg_inc<kEpiRegs>();
    const int g = warp >> 2;                                   // warpgroup
    const int m = (warp & 3) * 32 + lane;                      // TMEM lane = channel in tile
    const int c = c0 + m;
    const bool cvalid = c < p.E;
    const bool stamp = lane == 0 && (warp & 3) == 0;
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tD = tbase + lane_off + C::kDCol;
    const uint32_t tA = tbase + lane_off + C::kACol;

    // filter slice of this thread (the ortho 1/(H*W) lives in the C2R matrix): bins kh < 8 of kw = g*KWG + j
    // stay in registers, bins kh >= 8 are parked in spare TMEM columns of the thread's own lane
    // (16 columns per kw, read back next to the D2 loads) - 112 live filter registers would spill.
    // (The global loads were issued before the setup barrier; see `Fld` above.)
    float2 FR[KWG][C::HR];
    const uint32_t tF = tbase + lane_off + C::kFCol + g * 64;
#pragma unroll
    for (int j = 0; j < KWG; ++j) {
#pragma unroll
      for (int kh = 0; kh < C::HR; ++kh) FR[j][kh] = Fld[j][kh];
      if (H > 8) {
        uint32_t hi[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) hi[q] = 0u;
#pragma unroll
        for (int kh = 8; kh < H; ++kh) {
          hi[2 * (kh - 8)] = __float_as_uint(Fld[j][kh].x);
          hi[2 * (kh - 8) + 1] = __float_as_uint(Fld[j][kh].y);
        }
        tmem_st16(tF + j * 16, hi);
      }
    }
    if (H > 8) tmem_wait_st();
    if (tid == 0 && p.prof && blockIdx.x == 0) p.prof[13] = clock64();

    for (int it = 0; it < ntiles; ++it) {
      const uint32_t par = it & 1;
      const int b = img0 + (ntiles - 1 - it) * p.img_stride;
      // ---------------- E1: D1[h][slot,ri] -> A2[kw][h] (bf16 pairs)
      if (stamp) SB200_STAMP(g, it, 0);
      {
        // A2C is a multiple of 4: each batch of 4 grid rows becomes one 4-column store per kw; batch hb / 4 of
        // the row MMAs has its own commit barrier
#pragma unroll
        for (int hb = 0; hb < A2C; hb += 4) {
          if (hb < H) {
            mbar_wait(&dready_bar[0][hb >> 2], par);
            tc_fence_after_sync();
          }
          if (hb == 0 && stamp) SB200_STAMP(g, it, 1);
          uint32_t v[4][8];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (hb + i < H) tmem_ld8(tD + (hb + i) * 16 + 8 * g, v[i]);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < KWG; ++j) {
            uint32_t pk[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
              pk[i] = (hb + i < H) ? pack_bf16(f32(v[i][2 * j]), f32(v[i][2 * j + 1])) : 0u;
            const int kw = g * KWG + j;
            if (kw < WH) tmem_st4(tA + kw * A2C + hb, pk);
          }
        }
        tmem_wait_st();
        tc_fence_before_sync();
        named_bar_arrive(1, kBarCount);
      }
      if (stamp) SB200_STAMP(g, it, 2);
      // ---------------- E2: Z = D2 * filter -> A3[kw][kh]   (loads double-buffered)
      {
        // 16-column units (8 complex bins), double-buffered: unit u = (kw_local, half); the MMAs of local
        // column j (of both warpgroups) are committed on dready_bar[1][j]
        constexpr int UPK = K2 / 16, NU = KWG * UPK;
        uint32_t v[2][16];
        uint32_t ft[16];
        mbar_wait(&dready_bar[1][0], par);
        tc_fence_after_sync();
        if (stamp) SB200_STAMP(g, it, 3);
        if (g * KWG < WH) tmem_ld16(tD + (g * KWG) * K2, v[0]);
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const int j = u / UPK, hf = u % UPK;
          const int kw = g * KWG + j;
          tmem_wait_ld();
          if (u + 1 < NU) {
            const int jn = (u + 1) / UPK;
            const int kwn = g * KWG + jn;
            if (jn != j) {
              mbar_wait(&dready_bar[1][jn], par);
              tc_fence_after_sync();
            }
            if (kwn < WH) {
              tmem_ld16(tD + kwn * K2 + ((u + 1) % UPK) * 16, v[(u + 1) & 1]);
              if ((u + 1) % UPK == 1) tmem_ld16(tF + ((u + 1) / UPK) * 16, ft);
            }
          }
          if (kw < WH) {
            uint32_t z[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const int kh = hf * 8 + q;
              if (kh < H) {
                const float re = f32(v[u & 1][2 * q]), im = f32(v[u & 1][2 * q + 1]);
                const float2 f = hf == 0 ? FR[j][q < C::HR ? q : 0] : make_float2(f32(ft[2 * q]), f32(ft[2 * q + 1]));
                z[q] = pack_bf16(re * f.x - im * f.y, re * f.y + im * f.x);
              } else {
                z[q] = 0u;
              }
            }
            tmem_st8(tA + kw * A2C + hf * 8, z);
          }
        }
      }
      tmem_wait_st();
      tc_fence_before_sync();
      named_bar_arrive(2, kBarCount);
      if (stamp) SB200_STAMP(g, it, 4);
      // ---------------- E3: D3[kw][h,ri] -> A4[h][slot]
      {
        // two kw columns at a time: their (re, im) pairs go to adjacent K slots of every grid row.  No sub-batches
        // here: A4 [h][slot] overlays the A3 [kw] operands, so every inverse-column MMA must have completed.
        mbar_wait(&dready_bar[2][0], par);
        tc_fence_after_sync();
        if (stamp) SB200_STAMP(g, it, 5);
#pragma unroll
        for (int jp = 0; jp < 4; jp += 2) {
          const int kw0 = g * KWG + jp, kw1 = kw0 + 1;
          const bool live0 = jp < KWG && kw0 < WH, live1 = jp + 1 < KWG && kw1 < WH;
          uint32_t v0[K2], v1[K2];
          if (live0) tmem_ld<K2>(tD + kw0 * K2, v0);
          if (live1) tmem_ld<K2>(tD + kw1 * K2, v1);
          tmem_wait_ld();
#pragma unroll
          for (int h = 0; h < H; ++h)
            tmem_st2(tA + h * 8 + 4 * g + jp, live0 ? pack_bf16(f32(v0[2 * h]), f32(v0[2 * h + 1])) : 0u,
                     live1 ? pack_bf16(f32(v1[2 * h]), f32(v1[2 * h + 1])) : 0u);
        }
        tmem_wait_st();
        tc_fence_before_sync();
        named_bar_arrive(3, kBarCount);
      }
      if (stamp) SB200_STAMP(g, it, 6);
      // ---------------- E4: D4[h][w] -> bf16 -> swizzled smem tile -> TMA store
      if (tid == 0) tma_store_wait_read<0>();          // previous tile's store has drained sOut
      mbar_wait(&dready_bar[3][0], par);               // rows i of both warpgroups: sub-batch i / 2
      tc_fence_after_sync();
      if (stamp) SB200_STAMP(g, it, 7);
      {
        // TMEM is read in the mma C-fragment shape (thread = 2 adjacent tokens of one channel), packed to
        // bf16 pairs and transposed into the token-major output tile by stmatrix.trans: one instruction
        // moves a [8 tokens x 32 channels] block.  Lane T supplies the row address of matrix T/8, row T%8.
        const int q = warp & 3;
        const int mi = lane >> 3, rep = mi >> 1, up = mi & 1;
        const int wtok = rep * 8 + (lane & 7);                       // grid column this lane addresses
        const bool wvalid = wtok < W;
        const uint32_t obase = smem_u32(sOut) + (q >> 1) * C::kHalfStride;
        const uint32_t chunk0 = static_cast<uint32_t>(4 * (q & 1) + up);         // channels 32q + 8up + [0,8)
        const uint32_t dummy = smem_u32(sDummy) + (warp & 7) * 16;
        auto stage4 = [&](auto gc) {
          constexpr int G = decltype(gc)::value;
          constexpr int h0 = G * HG;
          uint32_t a[2][8], bq[2][8];
          if (h0 < H) {
            tmem_ld_16x256b_x2(tD + h0 * 16, a[0]);
            tmem_ld_16x256b_x2(tD + (16u << 16) + h0 * 16, bq[0]);
          }
          named_bar_sync(5, kEpiThreads);                        // sOut free (thread 0 waited above)
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            if (h0 + i < H) {
              tmem_wait_ld();
              if (i + 1 < HG && h0 + i + 1 < H) {
                if (((i + 1) & 1) == 0) {
                  mbar_wait(&dready_bar[3][(i + 1) >> 1], par);
                  tc_fence_after_sync();
                }
                tmem_ld_16x256b_x2(tD + (h0 + i + 1) * 16, a[(i + 1) & 1]);
                tmem_ld_16x256b_x2(tD + (16u << 16) + (h0 + i + 1) * 16, bq[(i + 1) & 1]);
              }
              const int t = (h0 + i) * W + wtok;
              const uint32_t row = obase + t * 128;
              const uint32_t x7 = static_cast<uint32_t>(t & 7);
              const uint32_t ad0 = wvalid ? row + ((chunk0 ^ x7) << 4) : dummy;
              const uint32_t ad1 = wvalid ? row + (((chunk0 + 2) ^ x7) << 4) : dummy;
              const uint32_t* u = a[i & 1];
              const uint32_t* w2 = bq[i & 1];
              stmatrix_x4_trans(ad0, pack_bf16(f32(u[0]), f32(u[1])), pack_bf16(f32(u[2]), f32(u[3])),
                                pack_bf16(f32(u[4]), f32(u[5])), pack_bf16(f32(u[6]), f32(u[7])));
              stmatrix_x4_trans(ad1, pack_bf16(f32(w2[0]), f32(w2[1])), pack_bf16(f32(w2[2]), f32(w2[3])),
                                pack_bf16(f32(w2[4]), f32(w2[5])), pack_bf16(f32(w2[6]), f32(w2[7])));
            }
          }
          tc_fence_before_sync();
          if (it + 1 < ntiles) named_bar_arrive(4, kBarCount);   // D region free for the next tile
          if (stamp) SB200_STAMP(G, it, 8);
        };
        if (g == 0) stage4(std::integral_constant<int, 0>{});
        else stage4(std::integral_constant<int, 1>{});
        fence_proxy_async_smem();
        named_bar_sync(6, kEpiThreads);
        if (tid == 0) {
          tma_store_3d(&tmap_y, sOut, c0, 0, b);
          if (half1) tma_store_3d(&tmap_y, sOut + C::kHalfStride, c0 + 64, 0, b);
          tma_store_commit();
          if (it + 1 == ntiles) tma_store_wait_all<0>();   // smem must outlive the last store
        }
      }
      if (stamp) SB200_STAMP(g, it, 9);
    }
  } else {
    setmaxnreg_dec<kAuxRegs>();
    if (warp == 8) {
      // ========================================================= MMA issuer
      const uint32_t tD = tbase + C::kDCol, tA = tbase + C::kACol;
      const uint32_t id_row = umma_idesc_bf16(128, 16, 1, 0);
      const uint32_t id_col = umma_idesc_bf16(128, K2, 0, 0);
      const uint32_t id_irow = umma_idesc_bf16(128, 16, 0, 0);
      const uint64_t b1 = umma_smem_desc(smem_u32(sMat + C::kOffB1), 128, 256, kSwzNone);
      const uint64_t b4 = umma_smem_desc(smem_u32(sMat + C::kOffB4), 128, 256, kSwzNone);
      const uint64_t b2 = umma_smem_desc(smem_u32(sMat + C::kOffB2), 128, (K2 / 8) * 128, kSwzNone);
      const uint64_t b3 = umma_smem_desc(smem_u32(sMat + C::kOffB3), 128, (K2 / 8) * 128, kSwzNone);
      for (int it = 0; it < ntiles; ++it) {
        const int s = it % S;
        if (lane == 0) SB200_STAMP(2, it, 0);
        mbar_wait(&full_bar[s], (it / S) & 1);
        if (it > 0) named_bar_sync(4, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 1);
        if (elect_one()) {
          const uint64_t a0 = umma_smem_desc(smem_u32(smem + s * C::kStageBytes), C::kHalfStride, 1024, kSwz128);
#pragma unroll
          for (int h = 0; h < H; ++h) {
            umma_ss(tD + h * 16, a0 + static_cast<uint64_t>(h * W * 8), b1, id_row, 0);
            if ((h & 3) == 3 || h == H - 1) umma_commit(&dready_bar[0][h >> 2]);
          }
          umma_commit(&empty_bar[s]);
        }
        __syncwarp();
        if (lane == 0) SB200_STAMP(2, it, 2);
        named_bar_sync(1, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 3);
        if (elect_one()) {
          // local column j of warpgroup 0 (kw = j) and of warpgroup 1 (kw = KWG + j), then one commit
#pragma unroll
          for (int j = 0; j < KWG; ++j) {
#pragma unroll
            for (int g2 = 0; g2 < 2; ++g2) {
              const int kw = g2 * KWG + j;
              if (kw < WH) {
#pragma unroll
                for (int ks = 0; ks < K2 / 16; ++ks)
                  umma_ts(tD + kw * K2, tA + kw * A2C + ks * 8, b2 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
              }
            }
            umma_commit(&dready_bar[1][j]);
          }
        }
        __syncwarp();
        if (lane == 0) SB200_STAMP(2, it, 4);
        named_bar_sync(2, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 5);
        if (elect_one()) {
#pragma unroll
          for (int kw = 0; kw < WH; ++kw)
#pragma unroll
            for (int ks = 0; ks < K2 / 16; ++ks)
              umma_ts(tD + kw * K2, tA + kw * A2C + ks * 8, b3 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
          umma_commit(&dready_bar[2][0]);
        }
        __syncwarp();
        if (lane == 0) SB200_STAMP(2, it, 6);
        named_bar_sync(3, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 7);
        if (elect_one()) {
          // grid rows i of both warpgroups (h = i and h = HG + i); one commit per two i
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            umma_ts(tD + i * 16, tA + i * 8, b4, id_irow, 0);
            if (HG + i < H) umma_ts(tD + (HG + i) * 16, tA + (HG + i) * 8, b4, id_irow, 0);
            if ((i & 1) == 1 || i == HG - 1) umma_commit(&dready_bar[3][i >> 1]);
          }
        }
        __syncwarp();
        if (lane == 0) SB200_STAMP(2, it, 8);
      }
    } else if (warp == 9) {
      // ========================================================= TMA producer + CLS pass-through
      const int T = C::HW + p.cls;
      const uint32_t bytes = (half1 ? 2u : 1u) * C::HW * 128u;
      for (int it = 0; it < ntiles; ++it) {
        const int s = it % S;
        const int b = img0 + (ntiles - 1 - it) * p.img_stride;
        if (it >= S) mbar_wait(&empty_bar[s], ((it / S) - 1) & 1);
        if (lane == 0 && it >= S) {
          SB200_STAMP(3, it, 0);
          uint8_t* dst = smem + s * C::kStageBytes;
          mbar_arrive_expect_tx(&full_bar[s], bytes);
          tma_load_3d(dst, &tmap_x, &full_bar[s], c0, 0, b);
          if (half1) tma_load_3d(dst + C::kHalfStride, &tmap_x, &full_bar[s], c0 + 64, 0, b);
        }
        if (p.cls && lane < 16 && c0 + lane * 8 < p.E) {
          const size_t off = static_cast<size_t>(b) * T * p.E + c0 + lane * 8;
          *reinterpret_cast<uint4*>(p.y + off) = __ldg(reinterpret_cast<const uint4*>(p.x + off));
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  SB200_WALL(1);
  if (warp == 8) tmem_dealloc<512>(tbase);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess) f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// [B][HW tokens][E] bf16 view of the patch grid (CLS skipped), 64-channel x HW-token boxes, 128B swizzle.
static int make_token_map(CUtensorMap* map, const void* base, int B, int HW, int T, int E, int cls) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
  bind_primary_context();
  const cuuint64_t dims[3] = {static_cast<cuuint64_t>(E), static_cast<cuuint64_t>(HW), static_cast<cuuint64_t>(B)};
  const cuuint64_t strides[2] = {static_cast<cuuint64_t>(E) * 2, static_cast<cuuint64_t>(T) * E * 2};
  const cuuint32_t box[3] = {64, static_cast<cuuint32_t>(HW), 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  void* gptr = const_cast<void*>(static_cast<const void*>(static_cast<const __nv_bfloat16*>(base) + cls * E));
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, gptr, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(SPECTRE_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", static_cast<int>(r));
  return SPECTRE_OK;
}

// grid = nct channel tiles x (images in flight); every CTA keeps one channel tile.
static void plan_grid(int B, int E, int* nct, int* img_stride) {
  *nct = (E + kTileC - 1) / kTileC;
  int per = device_sm_count() / *nct;
  if (per < 1) per = 1;
  if (per > B) per = B;
  *img_stride = per;
}

// backward: 64-channel blocks x image-pair groups
static void plan_bwd_grid(int B, int E, int* nblk, int* pair_stride) {
  *nblk = (E + 63) / 64;
  const int npairs = (B + 1) / 2;
  int per = device_sm_count() / *nblk;
  if (per < 1) per = 1;
  if (per > npairs) per = npairs;
  *pair_stride = per;
}

static long long* g_prof = nullptr;
void set_profile_buffer(long long* dev) { g_prof = dev; }

template <int H, int W>
static int fwd_launch(const void* x, const float* filt, void* y, int B, int E, int cls, cudaStream_t st) {
  using C = Cfg<H, W>;
  using SM = FwdSmem<H, W>;
  auto kern = fftmix_tc_fwd_kernel<H, W>;
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(kern), SM::kBytes) != cudaSuccess)
    return fail(SPECTRE_E_CUDA, "cudaFuncSetAttribute(tc fwd): %s", cudaGetErrorString(cudaGetLastError()));
  CUtensorMap tmx, tmy;
  if (int rc = make_token_map(&tmx, x, B, C::HW, C::HW + cls, E, cls)) return rc;
  if (int rc = make_token_map(&tmy, y, B, C::HW, C::HW + cls, E, cls)) return rc;
  FwdParams p;
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.filt = filt;
  p.y = static_cast<__nv_bfloat16*>(y);
  p.B = B; p.E = E; p.cls = cls;
  p.prof = g_prof;
  plan_grid(B, E, &p.nct, &p.img_stride);
  launch_pdl(kern, dim3(p.nct * p.img_stride), dim3(kThreads), SM::kBytes, st, tmx, tmy, p);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_tc_fwd launch: %s", cudaGetErrorString(err));
  return SPECTRE_OK;
}

// =====================================================================================================
// Backward:  dx = irfft2(conj(filt) * rfft2(dy)),  dfilt = sum_b conj(rfft2(x_b)) * (c * rfft2(dy_b))
// (SURVEY.md section 8 row A4).  Tile = TWO images x 64 channels (lanes 0-63 = image 2p, 64-127 = image
// 2p+1), so the 64-channel filter slice fits in shared memory (fp32) next to the dy / x / dx tiles and the
// filter-gradient accumulators fit in registers (each thread owns 4 of the 8 kw slots of its channel).
//   phases per tile (MMA -> epilogue):  rows(dy) -> A_dy | rows(x) -> A_x | cols(dy,x; slots 0-3) -> acc, Z |
//                                       cols(dy,x; slots 4-7) -> acc, Z | cols^-1(Z) -> A4 | rows^-1 -> dx
//   TMEM: D [0,256)  A_dy/Z [256,384)  A_x/A4 [384,512)
template <int H_, int W_>
struct BwdSmem {
  using C = Cfg<H_, W_>;
  static constexpr int kSlotF = (2 * H_ + 3) / 4 * 4;                // floats per slot (16-byte aligned)
  static constexpr int kFiltStride = 8 * kSlotF + 4;                 // floats per channel (8 slots, padded)
  static constexpr int kDy = 0, kX = C::kStageBytes, kOut = 2 * C::kStageBytes;
  static constexpr int kFilt = 3 * C::kStageBytes;
  static constexpr int kFiltBytes = 64 * kFiltStride * 4;
  static constexpr int kMat = kFilt + kFiltBytes;
  static constexpr int kBytes = kMat + C::kMatBytes + 1024;
  static_assert(kBytes <= 227 * 1024, "backward shared-memory budget");
};

struct BwdParams {
  const __nv_bfloat16* dy;
  const float* filt;
  __nv_bfloat16* dx;
  float* part;         // [slabs][H][WH][E][2] partial filter gradients (slab = image group of the CTA)
  int B, E, cls, nblk, pair_stride;
  long long* prof;
};

// E1 (shared by dy and x): D1[h][slot,ri] (fp32) -> bf16 A operand [slot][h] at tAdst, 4 grid rows per store;
// batch hb / 4 of the row MMAs is committed on bars[hb / 4]
template <class C>
__device__ __forceinline__ void epi_rows_to_cols(uint32_t tD, uint32_t tAdst, int g, uint64_t* bars, uint32_t par) {
  constexpr int H = C::H, A2C = C::A2C;
#pragma unroll
  for (int hb = 0; hb < A2C; hb += 4) {
    if (hb < H) {
      mbar_wait(&bars[hb >> 2], par);
      tc_fence_after_sync();
    }
    uint32_t v[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (hb + i < H) tmem_ld8(tD + (hb + i) * 16 + 8 * g, v[i]);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t pk[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pk[i] = (hb + i < H) ? pack_bf16(f32(v[i][2 * j]), f32(v[i][2 * j + 1])) : 0u;
      tmem_st4(tAdst + (4 * g + j) * A2C + hb, pk);
    }
  }
  tmem_wait_st();
}

template <int H, int W>
__global__ void __launch_bounds__(kThreads, 1)
fftmix_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap_dy, const __grid_constant__ CUtensorMap tmap_x,
                     const __grid_constant__ CUtensorMap tmap_dx, const BwdParams p) {
  using C = Cfg<H, W>;
  using SM = BwdSmem<H, W>;
  constexpr int WH = C::WH, KWG = C::KWG, HG = C::HG, K2 = C::K2, A2C = C::A2C, UPK = K2 / 16;
  constexpr int kADy = C::kACol, kAX = C::kACol + 128;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sDy = smem + SM::kDy;
  uint8_t* sX = smem + SM::kX;
  uint8_t* sOut = smem + SM::kOut;
  float* sF = reinterpret_cast<float*>(smem + SM::kFilt);
  uint8_t* sMat = smem + SM::kMat;
  __shared__ uint64_t full_bar[2], empty_bar[2], dready_bar[6][4];   // [phase][sub-batch], see the forward kernel
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) uint8_t sDummy[8 * 16];

  const long long t_start = clock64();
  pdl_launch_dependents();
  SB200_WALL(0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cb = blockIdx.x % p.nblk;
  const int pair0 = blockIdx.x / p.nblk;
  const int c0 = cb * 64;
  const int npairs = (p.B + 1) / 2;
  const int ntiles = pair0 < npairs ? (npairs - pair0 + p.pair_stride - 1) / p.pair_stride : 0;
  // slot -> kw (or -1): slots [0,KWG) and [4, 4+WH-KWG) are live
  auto kw_of = [](int s) { return s < 4 ? (s < KWG ? s : -1) : (s - 4 < WH - KWG ? KWG + s - 4 : -1); };

  // ---- one-time setup
  if (warp == 8) tmem_alloc<512>(&tmem_base_s);
  if (tid == 8 * 32) SB200_STAMP(5, 0, 0);
  if (tid == 9 * 32) {
    prefetch_tensormap(&tmap_dy);
    prefetch_tensormap(&tmap_x);
    prefetch_tensormap(&tmap_dx);
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int k = 0; k < 24; ++k) mbar_init(&dready_bar[k >> 2][k & 3], 1);
    mbar_fence_init();
    pdl_wait();   // dy is the preceding kernel's output; everything else in the setup is independent of it
    if (ntiles > 0) {   // prologue: first tile's dy and x
      const int b0 = 2 * (pair0 + (ntiles - 1) * p.pair_stride);   // first tile in the (descending) sweep
      const bool two = b0 + 1 < p.B;
      const uint32_t bytes = (two ? 2u : 1u) * C::HW * 128u;
      mbar_arrive_expect_tx(&full_bar[0], bytes);
      tma_load_3d(sDy, &tmap_dy, &full_bar[0], c0, 0, b0);
      if (two) tma_load_3d(sDy + C::kHalfStride, &tmap_dy, &full_bar[0], c0, 0, b0 + 1);
      mbar_arrive_expect_tx(&full_bar[1], bytes);
      tma_load_3d(sX, &tmap_x, &full_bar[1], c0, 0, b0);
      if (two) tma_load_3d(sX + C::kHalfStride, &tmap_x, &full_bar[1], c0, 0, b0 + 1);
    }
    SB200_STAMP(5, 0, 1);
  }
  {
    // conj(filt) of this CTA's 64 channels: sF[c][slot][kh][re,im], channel stride padded (the ortho 1/(H*W) lives
    // in the C2R matrix).  All global loads -- filter and DFT matrices -- are issued before the first shared-memory
    // store (memory-level parallelism: a load->store loop would serialise ~20 DRAM round trips per thread).
    constexpr int kIters = (64 * 8 * H + kThreads - 1) / kThreads;
    constexpr int kMatIters = (C::kMatBytes / 16 + kThreads - 1) / kThreads;
    float2 fv[kIters];
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
      const int i = tid + k * kThreads;
      const int cl = i & 63, rest = i >> 6, s = rest / H, kh = rest % H;
      const int kw = i < 64 * 8 * H ? kw_of(s) : -1;
      // branch-free (clamped address + select) so the loads issue back to back
      const bool ok = kw >= 0 && c0 + cl < p.E;
      const size_t off = ok ? (static_cast<size_t>(kh) * WH + kw) * p.E + c0 + cl : 0;
      const float2 f = __ldg(reinterpret_cast<const float2*>(p.filt) + off);
      fv[k] = ok ? f : make_float2(0.f, 0.f);
    }
    uint4 mv[kMatIters];
#pragma unroll
    for (int k = 0; k < kMatIters; ++k) {
      const int i = tid + k * kThreads;
      mv[k] = reinterpret_cast<const uint4*>(g_mats<H, W>.w)[i < C::kMatBytes / 16 ? i : 0];
    }
    if (tid == 0) SB200_STAMP(5, 0, 2);
#pragma unroll
    for (int k = 0; k < kMatIters; ++k) {
      const int i = tid + k * kThreads;
      if (i < C::kMatBytes / 16) reinterpret_cast<uint4*>(sMat)[i] = mv[k];
    }
    if (tid == 0) SB200_STAMP(5, 0, 3);
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
      const int i = tid + k * kThreads;
      const int cl = i & 63, rest = i >> 6, s = rest / H, kh = rest % H;
      if (i < 64 * 8 * H)
        *reinterpret_cast<float2*>(sF + cl * SM::kFiltStride + s * SM::kSlotF + kh * 2) = make_float2(fv[k].x, -fv[k].y);
    }
    if (tid == 0) SB200_STAMP(5, 0, 4);
    for (int i = tid; i < 64 * 4; i += kThreads) sF[(i >> 2) * SM::kFiltStride + 8 * SM::kSlotF + (i & 3)] = 0.f;
    if (SM::kSlotF > 2 * H)   // slot padding is read by the 16-byte filter loads: keep it finite
      for (int i = tid; i < 64 * 8 * (SM::kSlotF - 2 * H); i += kThreads) {
        constexpr int padn = SM::kSlotF - 2 * H > 0 ? SM::kSlotF - 2 * H : 1;
        const int cl = i & 63, rest = i >> 6;
        sF[cl * SM::kFiltStride + (rest / padn) * SM::kSlotF + 2 * H + rest % padn] = 0.f;
      }
    constexpr int per = (C::kHalfStride - C::HW * 128) / 16;
    for (int i = tid; i < 4 * per; i += kThreads)   // rows past the image in the dy / x halves
      reinterpret_cast<uint4*>(smem + (i / per) * C::kHalfStride + C::HW * 128)[i % per] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  if (tid == 0) SB200_STAMP(5, 0, 5);
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tmem_base_s;
  if (tid == 0 && p.prof && blockIdx.x == 0) { p.prof[14] = t_start; p.prof[15] = clock64(); }

  if (warp < 8) {
    // =========================================================== epilogue warpgroups
    setmaxnreg_inc<kEpiRegs>();
    const int g = warp >> 2;
    const int q = warp & 3;
    const int m = q * 32 + lane;                               // TMEM lane: image (m >> 6), channel (m & 63)
    const int c = c0 + (m & 63);
    const bool cvalid = c < p.E;
    const bool stamp = lane == 0 && q == 0;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t tD = tbase + lane_off + C::kDCol;
    const uint32_t tADy = tbase + lane_off + kADy;
    const uint32_t tAX = tbase + lane_off + kAX;
    const uint32_t fthr = smem_u32(sF) + static_cast<uint32_t>(m & 63) * SM::kFiltStride * 4;

    // filter-gradient accumulators: acc[hf][e][kh] for slot 4*hf + 2*g + e
    float2 acc[2][2][H];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int e = 0; e < 2; ++e)
#pragma unroll
        for (int kh = 0; kh < H; ++kh) acc[a][e][kh] = make_float2(0.f, 0.f);

    for (int it = 0; it < ntiles; ++it) {
      const uint32_t par = it & 1;
      const int b0 = 2 * (pair0 + (ntiles - 1 - it) * p.pair_stride);
      const bool two = b0 + 1 < p.B;
      const bool mvalid = m < 64 || two;
      // ---------------- rows(dy) -> A_dy ; rows(x) -> A_x
      if (stamp) SB200_STAMP(g, it, 0);
      epi_rows_to_cols<C>(tD, tADy, g, dready_bar[0], par);
      tc_fence_before_sync();
      named_bar_arrive(1, kBarCount);
      if (stamp) SB200_STAMP(g, it, 2);
      epi_rows_to_cols<C>(tD, tAX, g, dready_bar[1], par);
      tc_fence_before_sync();
      named_bar_arrive(2, kBarCount);
      if (stamp) SB200_STAMP(g, it, 4);
      // ---------------- cols: acc += conj(X) * DY ;  Z = conj(filt) * DY -> A_dy
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        constexpr int NU = 2 * UPK;                            // units: (e, half of the kh range)
        const bool live0 = kw_of(4 * hf + 2 * g) >= 0, live1 = kw_of(4 * hf + 2 * g + 1) >= 0;
#pragma unroll
        for (int u = 0; u < NU; ++u) {
          const int e = u / UPK, hk = u % UPK;
          if (hk == 0) {                                       // local slot e of both warpgroups: sub-batch e
            mbar_wait(&dready_bar[2 + hf][e], par);
            tc_fence_after_sync();
            if (e == 0 && stamp) SB200_STAMP(g, it, 5 + 2 * hf);
          }
          const bool live = e == 0 ? live0 : live1;
          if (live) {
            const int s = 4 * hf + 2 * g + e;
            uint32_t vd[16], vx[16];
            tmem_ld16(tD + (2 * g + e) * K2 + hk * 16, vd);
            tmem_ld16(tD + 128 + (2 * g + e) * K2 + hk * 16, vx);
            float4 f4[4];
#pragma unroll
            for (int qd = 0; qd < 4; ++qd)
              asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                           : "=f"(f4[qd].x), "=f"(f4[qd].y), "=f"(f4[qd].z), "=f"(f4[qd].w)
                           : "r"(fthr + static_cast<uint32_t>((s * SM::kSlotF + hk * 16 + qd * 4) * 4)));
            tmem_wait_ld();
            uint32_t z[8];
#pragma unroll
            for (int qq = 0; qq < 8; ++qq) {
              const int kh = hk * 8 + qq;
              if (kh < H) {
                const float dr = f32(vd[2 * qq]), di = f32(vd[2 * qq + 1]);
                const float xr = f32(vx[2 * qq]), xi = f32(vx[2 * qq + 1]);
                const float fr = (qq & 1) ? f4[qq >> 1].z : f4[qq >> 1].x;   // already conj(filt)
                const float fi = (qq & 1) ? f4[qq >> 1].w : f4[qq >> 1].y;
                if (mvalid) {
                  float2& a = acc[hf][e][kh < H ? kh : 0];
                  a.x += xr * dr + xi * di;
                  a.y += xr * di - xi * dr;
                }
                z[qq] = pack_bf16(fr * dr - fi * di, fr * di + fi * dr);
              } else {
                z[qq] = 0u;
              }
            }
            tmem_st8(tADy + s * A2C + hk * 8, z);
          }
        }
        tmem_wait_st();
        tc_fence_before_sync();
        named_bar_arrive(3 + hf, kBarCount);
        if (stamp) SB200_STAMP(g, it, 6 + 2 * hf);
      }
      // ---------------- cols^-1: D3[slot][h,ri] -> A4[h][slot] (in the A_x region)
#pragma unroll
      for (int jp = 0; jp < 4; jp += 2) {
        mbar_wait(&dready_bar[4][jp >> 1], par);               // slot pair jp of both warpgroups
        tc_fence_after_sync();
        if (jp == 0 && stamp) SB200_STAMP(g, it, 9);
        const int s0 = 4 * g + jp, s1 = s0 + 1;
        const bool live0 = kw_of(s0) >= 0, live1 = kw_of(s1) >= 0;
        uint32_t v0[K2], v1[K2];
        if (live0) tmem_ld<K2>(tD + s0 * K2, v0);
        if (live1) tmem_ld<K2>(tD + s1 * K2, v1);
        tmem_wait_ld();
#pragma unroll
        for (int h = 0; h < H; ++h)
          tmem_st2(tAX + h * 8 + s0, live0 ? pack_bf16(f32(v0[2 * h]), f32(v0[2 * h + 1])) : 0u,
                   live1 ? pack_bf16(f32(v1[2 * h]), f32(v1[2 * h + 1])) : 0u);
      }
      tmem_wait_st();
      tc_fence_before_sync();
      named_bar_arrive(5, kBarCount);
      if (stamp) SB200_STAMP(g, it, 10);
      // ---------------- rows^-1: D4[h][w] -> bf16 -> swizzled smem tile -> TMA store of dx
      if (tid == 0) tma_store_wait_read<0>();
      mbar_wait(&dready_bar[5][0], par);
      tc_fence_after_sync();
      if (stamp) SB200_STAMP(g, it, 11);
      {
        const int mi = lane >> 3, rep = mi >> 1, up = mi & 1;
        const int wtok = rep * 8 + (lane & 7);
        const bool wvalid = wtok < W;
        const uint32_t obase = smem_u32(sOut) + (q >> 1) * C::kHalfStride;
        const uint32_t chunk0 = static_cast<uint32_t>(4 * (q & 1) + up);
        const uint32_t dummy = smem_u32(sDummy) + (warp & 7) * 16;
        auto stage4 = [&](auto gc) {
          constexpr int G = decltype(gc)::value;
          constexpr int h0 = G * HG;
          uint32_t a[2][8], bq[2][8];
          if (h0 < H) {
            tmem_ld_16x256b_x2(tD + h0 * 16, a[0]);
            tmem_ld_16x256b_x2(tD + (16u << 16) + h0 * 16, bq[0]);
          }
          named_bar_sync(7, kEpiThreads);                      // sOut free (thread 0 waited above)
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            if (h0 + i < H) {
              tmem_wait_ld();
              if (i + 1 < HG && h0 + i + 1 < H) {
                if (((i + 1) & 1) == 0) {
                  mbar_wait(&dready_bar[5][(i + 1) >> 1], par);
                  tc_fence_after_sync();
                }
                tmem_ld_16x256b_x2(tD + (h0 + i + 1) * 16, a[(i + 1) & 1]);
                tmem_ld_16x256b_x2(tD + (16u << 16) + (h0 + i + 1) * 16, bq[(i + 1) & 1]);
              }
              const int t = (h0 + i) * W + wtok;
              const uint32_t row = obase + t * 128;
              const uint32_t x7 = static_cast<uint32_t>(t & 7);
              const uint32_t ad0 = wvalid ? row + ((chunk0 ^ x7) << 4) : dummy;
              const uint32_t ad1 = wvalid ? row + (((chunk0 + 2) ^ x7) << 4) : dummy;
              const uint32_t* u = a[i & 1];
              const uint32_t* w2 = bq[i & 1];
              stmatrix_x4_trans(ad0, pack_bf16(f32(u[0]), f32(u[1])), pack_bf16(f32(u[2]), f32(u[3])),
                                pack_bf16(f32(u[4]), f32(u[5])), pack_bf16(f32(u[6]), f32(u[7])));
              stmatrix_x4_trans(ad1, pack_bf16(f32(w2[0]), f32(w2[1])), pack_bf16(f32(w2[2]), f32(w2[3])),
                                pack_bf16(f32(w2[4]), f32(w2[5])), pack_bf16(f32(w2[6]), f32(w2[7])));
            }
          }
          tc_fence_before_sync();
          if (it + 1 < ntiles) named_bar_arrive(6, kBarCount);   // D region free for the next tile
          if (stamp) SB200_STAMP(G, it, 12);
        };
        if (g == 0) stage4(std::integral_constant<int, 0>{});
        else stage4(std::integral_constant<int, 1>{});
        fence_proxy_async_smem();
        named_bar_sync(8, kEpiThreads);
        if (tid == 0) {
          tma_store_3d(&tmap_dx, sOut, c0, 0, b0);
          if (two) tma_store_3d(&tmap_dx, sOut + C::kHalfStride, c0, 0, b0 + 1);
          tma_store_commit();
          if (it + 1 == ntiles) tma_store_wait_all<0>();
        }
      }
      if (it + 1 == ntiles) {
    // ---------------- all tiles done: add the two image halves of every channel, write this CTA's slab
      {
        float* scratch = reinterpret_cast<float*>(sDy);          // input tiles are dead by now
        constexpr int kStride = 4 * H * 2 + 1;
        float* mine = scratch + (g * 64 + (m & 63)) * kStride;
        if (m >= 64) {
#pragma unroll
          for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
              for (int kh = 0; kh < H; ++kh) {
                mine[((a * 2 + e) * H + kh) * 2] = acc[a][e][kh].x;
                mine[((a * 2 + e) * H + kh) * 2 + 1] = acc[a][e][kh].y;
              }
        }
        named_bar_sync(9, kEpiThreads);
        if (m < 64 && cvalid && ntiles > 0) {
          float2* slab = reinterpret_cast<float2*>(p.part) + static_cast<size_t>(pair0) * H * WH * p.E;
#pragma unroll
          for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int kw = kw_of(4 * a + 2 * g + e);
              if (kw >= 0) {
                const float wgt = ((kw == 0 || (W % 2 == 0 && kw == W / 2)) ? 1.0f : 2.0f) / static_cast<float>(H * W);
#pragma unroll
                for (int kh = 0; kh < H; ++kh) {
                  const float sx = acc[a][e][kh].x + mine[((a * 2 + e) * H + kh) * 2];
                  const float sy = acc[a][e][kh].y + mine[((a * 2 + e) * H + kh) * 2 + 1];
                  slab[(static_cast<size_t>(kh) * WH + kw) * p.E + c] = make_float2(sx * wgt, sy * wgt);
                }
              }
            }
        }
      }
      }
    }
  } else {
    setmaxnreg_dec<kAuxRegs>();
    if (warp == 8) {
      // ========================================================= MMA issuer
      const uint32_t tD0 = tbase + C::kDCol, tADy0 = tbase + kADy, tAX0 = tbase + kAX;
      const uint32_t id_row = umma_idesc_bf16(128, 16, 1, 0);
      const uint32_t id_col = umma_idesc_bf16(128, K2, 0, 0);
      const uint32_t id_irow = umma_idesc_bf16(128, 16, 0, 0);
      const uint64_t b1 = umma_smem_desc(smem_u32(sMat + C::kOffB1), 128, 256, kSwzNone);
      const uint64_t b4 = umma_smem_desc(smem_u32(sMat + C::kOffB4), 128, 256, kSwzNone);
      const uint64_t b2 = umma_smem_desc(smem_u32(sMat + C::kOffB2), 128, (K2 / 8) * 128, kSwzNone);
      const uint64_t b3 = umma_smem_desc(smem_u32(sMat + C::kOffB3), 128, (K2 / 8) * 128, kSwzNone);
      const uint64_t a_dy0 = umma_smem_desc(smem_u32(sDy), C::kHalfStride, 1024, kSwz128);
      const uint64_t a_x0 = umma_smem_desc(smem_u32(sX), C::kHalfStride, 1024, kSwz128);
      for (int it = 0; it < ntiles; ++it) {
        const uint32_t par = it & 1;
        // keep the operand bases opaque per iteration: otherwise every "base + constant" of the ~110 MMAs is
        // hoisted out of the loop and the issuing thread spills
        uint64_t a_dy = a_dy0, a_x = a_x0;
        uint32_t tD = tD0, tADy = tADy0, tAX = tAX0;
        asm volatile("" : "+l"(a_dy), "+l"(a_x), "+r"(tD), "+r"(tADy), "+r"(tAX));
        if (lane == 0) SB200_STAMP(2, it, 0);
        mbar_wait(&full_bar[0], par);
        if (it > 0) named_bar_sync(6, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 1);
        if (elect_one()) {
#pragma unroll
          for (int h = 0; h < H; ++h) {
            umma_ss(tD + h * 16, a_dy + static_cast<uint64_t>(h * W * 8), b1, id_row, 0);
            if ((h & 3) == 3 || h == H - 1) umma_commit(&dready_bar[0][h >> 2]);
          }
          umma_commit(&empty_bar[0]);
        }
        __syncwarp();
        mbar_wait(&full_bar[1], par);
        named_bar_sync(1, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 2);
        if (elect_one()) {
#pragma unroll
          for (int h = 0; h < H; ++h) {
            umma_ss(tD + h * 16, a_x + static_cast<uint64_t>(h * W * 8), b1, id_row, 0);
            if ((h & 3) == 3 || h == H - 1) umma_commit(&dready_bar[1][h >> 2]);
          }
          umma_commit(&empty_bar[1]);
        }
        __syncwarp();
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          named_bar_sync(2 + hf, kBarCount);
          tc_fence_after_sync();
          if (lane == 0) SB200_STAMP(2, it, 3 + hf);
          if (elect_one()) {
            // local slot e of warpgroup 0 (sl = e) and of warpgroup 1 (sl = 2 + e), then one commit
#pragma unroll
            for (int e = 0; e < 2; ++e) {
#pragma unroll
              for (int g2 = 0; g2 < 2; ++g2) {
                const int sl = 2 * g2 + e;
                if (C::kw_of_slot(4 * hf + sl) >= 0) {
#pragma unroll
                  for (int ks = 0; ks < K2 / 16; ++ks) {
                    umma_ts(tD + sl * K2, tADy + (4 * hf + sl) * A2C + ks * 8, b2 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
                    umma_ts(tD + 128 + sl * K2, tAX + (4 * hf + sl) * A2C + ks * 8, b2 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
                  }
                }
              }
              umma_commit(&dready_bar[2 + hf][e]);
            }
          }
          __syncwarp();
        }
        named_bar_sync(4, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 5);
        if (elect_one()) {
          // slot pair jp of warpgroup 0 (s = jp, jp+1) and of warpgroup 1 (s = 4+jp, 5+jp), then one commit
#pragma unroll
          for (int jp = 0; jp < 4; jp += 2) {
#pragma unroll
            for (int g2 = 0; g2 < 2; ++g2)
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int sl = 4 * g2 + jp + e;
                if (C::kw_of_slot(sl) >= 0) {
#pragma unroll
                  for (int ks = 0; ks < K2 / 16; ++ks)
                    umma_ts(tD + sl * K2, tADy + sl * A2C + ks * 8, b3 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
                }
              }
            umma_commit(&dready_bar[4][jp >> 1]);
          }
        }
        __syncwarp();
        named_bar_sync(5, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) SB200_STAMP(2, it, 6);
        if (elect_one()) {
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            umma_ts(tD + i * 16, tAX + i * 8, b4, id_irow, 0);
            if (HG + i < H) umma_ts(tD + (HG + i) * 16, tAX + (HG + i) * 8, b4, id_irow, 0);
            if ((i & 1) == 1 || i == HG - 1) umma_commit(&dready_bar[5][i >> 1]);
          }
        }
        __syncwarp();
        if (lane == 0) SB200_STAMP(2, it, 7);
      }
    } else if (warp == 9) {
      // ========================================================= TMA producer + CLS pass-through
      const int T = C::HW + p.cls;
      for (int it = 0; it < ntiles; ++it) {
        const int b0 = 2 * (pair0 + (ntiles - 1 - it) * p.pair_stride);
        const bool two = b0 + 1 < p.B;
        const uint32_t bytes = (two ? 2u : 1u) * C::HW * 128u;
        if (it > 0) {
          mbar_wait(&empty_bar[0], (it - 1) & 1);
          if (lane == 0) {
            SB200_STAMP(3, it, 0);
            mbar_arrive_expect_tx(&full_bar[0], bytes);
            tma_load_3d(sDy, &tmap_dy, &full_bar[0], c0, 0, b0);
            if (two) tma_load_3d(sDy + C::kHalfStride, &tmap_dy, &full_bar[0], c0, 0, b0 + 1);
          }
          mbar_wait(&empty_bar[1], (it - 1) & 1);
          if (lane == 0) {
            mbar_arrive_expect_tx(&full_bar[1], bytes);
            tma_load_3d(sX, &tmap_x, &full_bar[1], c0, 0, b0);
            if (two) tma_load_3d(sX + C::kHalfStride, &tmap_x, &full_bar[1], c0, 0, b0 + 1);
          }
        }
        if (p.cls && lane < 16) {
          const int img = b0 + (lane >> 3), cc = c0 + (lane & 7) * 8;
          if (img < p.B && cc < p.E) {
            const size_t off = static_cast<size_t>(img) * T * p.E + cc;
            *reinterpret_cast<uint4*>(p.dx + off) = __ldg(reinterpret_cast<const uint4*>(p.dy + off));
          }
        }
        __syncwarp();
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  SB200_WALL(1);
  if (warp == 8) tmem_dealloc<512>(tbase);
}

template <int H, int W>
static int bwd_launch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                      size_t ws_bytes, int B, int E, int cls, cudaStream_t st) {
  using C = Cfg<H, W>;
  using SM = BwdSmem<H, W>;
  auto kern = fftmix_tc_bwd_kernel<H, W>;
  if (ensure_dynamic_smem(reinterpret_cast<const void*>(kern), SM::kBytes) != cudaSuccess)
    return fail(SPECTRE_E_CUDA, "cudaFuncSetAttribute(tc bwd): %s", cudaGetErrorString(cudaGetLastError()));
  BwdParams p;
  plan_bwd_grid(B, E, &p.nblk, &p.pair_stride);
  const size_t slab_floats = static_cast<size_t>(H) * C::WH * E * 2;
  if (p.pair_stride > 1 && ws_bytes < slab_floats * sizeof(float) * p.pair_stride)
    return fail(SPECTRE_E_WORKSPACE, "fftmix_tc_bwd: workspace %zu B < required %zu B", ws_bytes,
                slab_floats * sizeof(float) * p.pair_stride);
  CUtensorMap tmdy, tmx, tmdx;
  if (int rc = make_token_map(&tmdy, dy, B, C::HW, C::HW + cls, E, cls)) return rc;
  if (int rc = make_token_map(&tmx, x, B, C::HW, C::HW + cls, E, cls)) return rc;
  if (int rc = make_token_map(&tmdx, dx, B, C::HW, C::HW + cls, E, cls)) return rc;
  p.dy = static_cast<const __nv_bfloat16*>(dy);
  p.filt = filt;
  p.dx = static_cast<__nv_bfloat16*>(dx);
  p.part = p.pair_stride > 1 ? static_cast<float*>(ws) : dfilt;
  p.B = B; p.E = E; p.cls = cls;
  p.prof = g_prof;
  launch_pdl(kern, dim3(p.nblk * p.pair_stride), dim3(kThreads), SM::kBytes, st, tmdy, tmx, tmdx, p);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_tc_bwd launch: %s", cudaGetErrorString(err));
  last_launch_count() = 1;
  if (p.pair_stride > 1) {
    launch_reduce_partials(p.part, dfilt, slab_floats, p.pair_stride, st);
    err = cudaGetLastError();
    if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_tc_bwd reduce launch: %s", cudaGetErrorString(err));
    last_launch_count() = 2;
  }
  return SPECTRE_OK;
}

}  // namespace tc

// (H, W) grids with a tensor-core kernel; E must be a multiple of 8 (16-byte TMA strides)
#define SB200_TC_GRIDS(F) F(7, 7) F(8, 8) F(14, 14)

int fftmix_tc_supported(int H, int W, int E) {
  if (E % 8) return 0;
#define F(h, w) if (H == h && W == w) return 1;
  SB200_TC_GRIDS(F)
#undef F
  return 0;
}

int fftmix_tc_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls,
                           cudaStream_t st) {
#define F(h, w) if (H == h && W == w) return tc::fwd_launch<h, w>(x, filt, y, B, E, cls, st);
  SB200_TC_GRIDS(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_tc: no tensor-core kernel for grid %dx%d", H, W);
}

int fftmix_tc_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                           size_t ws_bytes, int B, int H, int W, int E, int cls, cudaStream_t st) {
#define F(h, w) if (H == h && W == w) return tc::bwd_launch<h, w>(dy, x, filt, dx, dfilt, ws, ws_bytes, B, E, cls, st);
  SB200_TC_GRIDS(F)
#undef F
  return fail(SPECTRE_E_GRID, "fftmix_tc: no tensor-core kernel for grid %dx%d", H, W);
}

size_t fftmix_tc_bwd_workspace(int B, int H, int W, int E) {
  int nblk, stride;
  tc::plan_bwd_grid(B, E, &nblk, &stride);
  return stride > 1 ? static_cast<size_t>(stride) * H * (W / 2 + 1) * E * 2 * sizeof(float) : 0;
}

}  // namespace sb200

// Debug hook (not part of include/spectre_b200.h): device buffer of 6*8*16 int64 that CTA 0 of the
// tensor-core kernels fills with clock64 stamps; nullptr switches it off.
extern "C" void spectre_debug_tc_profile(long long* dev_buffer) { sb200::tc::set_profile_buffer(dev_buffer); }
