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
// Accumulation is fp32 in TMEM; the operands of every stage are bf16 (three intermediate
// roundings + bf16 DFT matrices => ~3e-3 of max|y|, inside the 3e-2 bf16 tolerance).
//
// Warp roles (384 threads, one CTA per SM, persistent over images of one channel tile):
//   warps 0-7  epilogue: warpgroup g handles half of the kw columns (stages 1-3) / half of
//              the grid rows (stage 4); its slice of the filter lives in registers.
//   warp  8    MMA issuer (one lane);  warp 9  TMA producer + CLS pass-through;  10-11 idle.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace sb200 {
namespace tc {

using namespace ptx;

constexpr int kEpiThreads = 256;
constexpr int kThreads = 384;
constexpr int kBarCount = kEpiThreads + 32;  // named barriers: epilogue arrives, MMA warp syncs
constexpr int kTileC = 128;

template <int H_, int W_>
struct Cfg {
  static constexpr int H = H_, W = W_, HW = H * W, WH = W / 2 + 1;
  static_assert(W <= 14 && H <= 16 && WH <= 8, "tensor-core path: W <= 14, H <= 16");
  static constexpr int KWG = (WH + 1) / 2;             // kw columns per warpgroup (<= 4)
  static constexpr int HG = (H + 1) / 2;               // grid rows per warpgroup in stage 4
  static constexpr int K2 = (2 * H + 15) / 16 * 16;    // column-stage K and N
  static constexpr int A2C = K2 / 2;                   // TMEM columns per kw of a bf16 A operand
  static constexpr int kDCol = 0, kACol = 256;         // TMEM column bases
  static constexpr int kHalfStride = ((HW + 16 - W) * 128 + 1023) / 1024 * 1024;  // 64-channel half
  static constexpr int kStageBytes = 2 * kHalfStride;
  static constexpr int kStages = (kStageBytes * 4 <= 200 * 1024) ? 4 : 3;
  static constexpr int kB1 = 16 * 16 * 2, kB2 = K2 * K2 * 2;
  static constexpr int kMatBytes = 2 * kB1 + 2 * kB2;
  static constexpr int kSmemBytes = kStages * kStageBytes + kMatBytes + 1024 /*align slack*/;
  // K / N slot of a half-spectrum column: warpgroup g owns slots [4g, 4g+4)
  __host__ __device__ static constexpr int slot(int kw) { return kw < KWG ? kw : 4 + (kw - KWG); }
};

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// K-major, no-swizzle B operand: element (n, k) of an [N x K] matrix
__device__ __forceinline__ uint32_t bmat_off(int n, int k, int K) {
  return (n % 8) * 16 + (n / 8) * (K / 8) * 128 + (k % 8) * 2 + (k / 8) * 128;
}

// cos/sin of 2*pi*num/den with exact argument reduction
__device__ __forceinline__ void cs2pi(int num, int den, float& c, float& s) {
  sincospif(2.0f * static_cast<float>(num % den) / static_cast<float>(den), &s, &c);
}

template <class C>
__device__ void build_matrices(uint8_t* sB1, uint8_t* sB2, uint8_t* sB3, uint8_t* sB4, bool conj_dir) {
  constexpr int H = C::H, W = C::W, WH = C::WH, K2 = C::K2;
  (void)conj_dir;
  auto put = [](uint8_t* base, uint32_t off, float v) { *reinterpret_cast<__nv_bfloat16*>(base + off) = __float2bfloat16(v); };
  // B1[n = 2*slot(kw)+ri][k = w]: forward real DFT along W
  for (int i = threadIdx.x; i < 16 * 16; i += blockDim.x) {
    const int n = i / 16, k = i % 16;
    float v = 0.f;
    const int sl = n / 2, ri = n % 2;
    int kw = -1;
    for (int q = 0; q < WH; ++q) if (C::slot(q) == sl) kw = q;
    if (kw >= 0 && k < W) {
      float c, s; cs2pi(kw * k, W, c, s);
      v = ri == 0 ? c : -s;
    }
    put(sB1, bmat_off(n, k, 16), v);
    // B4[n = w][k = 2*slot(kw)+ri]: C2R along W with the Hermitian weights folded in
    float u = 0.f;
    const int sl4 = k / 2, ri4 = k % 2;
    int kw4 = -1;
    for (int q = 0; q < WH; ++q) if (C::slot(q) == sl4) kw4 = q;
    if (kw4 >= 0 && n < W) {
      float c, s; cs2pi(kw4 * n, W, c, s);
      const float wgt = (kw4 == 0 || (W % 2 == 0 && kw4 == W / 2)) ? 1.f : 2.f;
      u = ri4 == 0 ? wgt * c : -wgt * s;
    }
    put(sB4, bmat_off(n, k, 16), u);
  }
  // B2[n = 2kh+ri'][k = 2h+ri]: forward complex DFT along H;  B3: inverse
  for (int i = threadIdx.x; i < K2 * K2; i += blockDim.x) {
    const int n = i / K2, k = i % K2;
    float f = 0.f, g = 0.f;
    const int a = n / 2, ra = n % 2, b = k / 2, rb = k % 2;
    if (a < H && b < H) {
      float c, s; cs2pi(a * b, H, c, s);
      // forward e^{-i phi}: re' = a cos + b sin ; im' = -a sin + b cos
      f = (ra == rb) ? c : (ra == 0 ? s : -s);
      // inverse e^{+i phi}: re' = a cos - b sin ; im' = a sin + b cos
      g = (ra == rb) ? c : (ra == 0 ? -s : s);
    }
    put(sB2, bmat_off(n, k, K2), f);
    put(sB3, bmat_off(n, k, K2), g);
  }
}

struct FwdParams {
  const __nv_bfloat16* x;
  const float* filt;
  __nv_bfloat16* y;
  int B, E, cls, nct, img_stride;
};

template <int H, int W>
__global__ void __launch_bounds__(kThreads, 1)
fftmix_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap_x, const FwdParams p) {
  using C = Cfg<H, W>;
  constexpr int WH = C::WH, KWG = C::KWG, HG = C::HG, K2 = C::K2, A2C = C::A2C, S = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB1 = smem + S * C::kStageBytes;
  uint8_t* sB4 = sB1 + C::kB1;
  uint8_t* sB2 = sB4 + C::kB1;
  uint8_t* sB3 = sB2 + C::kB2;
  __shared__ uint64_t full_bar[S], empty_bar[S], dready_bar[4];
  __shared__ uint32_t tmem_base_s;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ct = blockIdx.x % p.nct;
  const int img0 = blockIdx.x / p.nct;
  const int c0 = ct * kTileC;
  const int T = C::HW + p.cls;
  const int ntiles = img0 < p.B ? (p.B - img0 + p.img_stride - 1) / p.img_stride : 0;

  // ---- one-time setup
  if (warp == 8) tmem_alloc<512>(&tmem_base_s);
  if (tid == 9 * 32) {
    prefetch_tensormap(&tmap_x);
    for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int k = 0; k < 4; ++k) mbar_init(&dready_bar[k], 1);
    mbar_fence_init();
  }
  build_matrices<C>(sB1, sB2, sB3, sB4, false);
  // rows past the image in every half (read by the last row's K=16 window) must be finite
  for (int i = tid; i < S * 2 * ((C::kHalfStride - C::HW * 128) / 16); i += kThreads) {
    constexpr int per = (C::kHalfStride - C::HW * 128) / 16;
    const int half = i / per, j = i % per;
    reinterpret_cast<uint4*>(smem + half * C::kHalfStride + C::HW * 128)[j] = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tmem_base_s;

  if (warp < 8) {
    // =========================================================== epilogue warpgroups
    setmaxnreg_inc<224>();
    const int g = warp >> 2;                                   // warpgroup
    const int m = (warp & 3) * 32 + lane;                      // TMEM lane = channel in tile
    const int c = c0 + m;
    const bool cvalid = c < p.E;
    const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t tD = tbase + lane_off + C::kDCol;
    const uint32_t tA = tbase + lane_off + C::kACol;

    // filter slice of this thread: F[j][kh] for kw = g*KWG + j, pre-scaled by the ortho 1/(H*W)
    float2 F[KWG][H];
    {
      const float scale = 1.0f / static_cast<float>(H * W);
#pragma unroll
      for (int j = 0; j < KWG; ++j) {
        const int kw = g * KWG + j;
#pragma unroll
        for (int kh = 0; kh < H; ++kh) {
          float2 f = make_float2(0.f, 0.f);
          if (kw < WH && cvalid) {
            f = __ldg(reinterpret_cast<const float2*>(p.filt) + (static_cast<size_t>(kh) * WH + kw) * p.E + c);
            f.x *= scale; f.y *= scale;
          }
          F[j][kh] = f;
        }
      }
    }

    for (int it = 0; it < ntiles; ++it) {
      const uint32_t par = it & 1;
      const int b = img0 + it * p.img_stride;
      // ---------------- E1: D1[h][slot,ri] -> A2[kw][h] (bf16 pairs)
      mbar_wait(&dready_bar[0], par);
      tc_fence_after_sync();
      {
        uint32_t pk[KWG][A2C];
#pragma unroll
        for (int hb = 0; hb < H; hb += 4) {
          uint32_t v[4][8];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (hb + i < H) tmem_ld8(tD + (hb + i) * 16 + 8 * g, v[i]);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (hb + i < H) {
#pragma unroll
              for (int j = 0; j < KWG; ++j)
                pk[j][hb + i] = pack_bf16(__uint_as_float(v[i][2 * j]), __uint_as_float(v[i][2 * j + 1]));
            }
        }
#pragma unroll
        for (int j = 0; j < KWG; ++j) {
#pragma unroll
          for (int q = H; q < A2C; ++q) pk[j][q] = 0u;
          const int kw = g * KWG + j;
          if (kw < WH) tmem_st<A2C>(tA + kw * A2C, pk[j]);
        }
        tmem_wait_st();
        tc_fence_before_sync();
        named_bar_arrive(1, kBarCount);
      }
      // ---------------- E2: Z = D2 * filter -> A3[kw][kh]
      mbar_wait(&dready_bar[1], par);
      tc_fence_after_sync();
#pragma unroll
      for (int j = 0; j < KWG; ++j) {
        const int kw = g * KWG + j;
        if (kw < WH) {
          uint32_t v[K2];
          tmem_ld<K2>(tD + kw * K2, v);
          tmem_wait_ld();
          uint32_t z[A2C];
#pragma unroll
          for (int kh = 0; kh < H; ++kh) {
            const float re = __uint_as_float(v[2 * kh]), im = __uint_as_float(v[2 * kh + 1]);
            const float2 f = F[j][kh];
            z[kh] = pack_bf16(re * f.x - im * f.y, re * f.y + im * f.x);
          }
#pragma unroll
          for (int q = H; q < A2C; ++q) z[q] = 0u;
          tmem_st<A2C>(tA + kw * A2C, z);
        }
      }
      tmem_wait_st();
      tc_fence_before_sync();
      named_bar_arrive(2, kBarCount);
      // ---------------- E3: D3[kw][h,ri] -> A4[h][slot]
      mbar_wait(&dready_bar[2], par);
      tc_fence_after_sync();
      {
        uint32_t pk[H][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int kw = g * KWG + j;
          if (j < KWG && kw < WH) {
            uint32_t v[K2];
            tmem_ld<K2>(tD + kw * K2, v);
            tmem_wait_ld();
#pragma unroll
            for (int h = 0; h < H; ++h) pk[h][j] = pack_bf16(__uint_as_float(v[2 * h]), __uint_as_float(v[2 * h + 1]));
          } else {
#pragma unroll
            for (int h = 0; h < H; ++h) pk[h][j] = 0u;
          }
        }
#pragma unroll
        for (int h = 0; h < H; ++h) tmem_st4(tA + h * 8 + 4 * g, pk[h]);
        tmem_wait_st();
        tc_fence_before_sync();
        named_bar_arrive(3, kBarCount);
      }
      // ---------------- E4: D4[h][w] -> y
      mbar_wait(&dready_bar[3], par);
      tc_fence_after_sync();
      {
        uint32_t v[HG][16];
#pragma unroll
        for (int i = 0; i < HG; ++i)
          if (g * HG + i < H) tmem_ld16(tD + (g * HG + i) * 16, v[i]);
        tmem_wait_ld();
        tc_fence_before_sync();
        if (it + 1 < ntiles) named_bar_arrive(4, kBarCount);   // D region free for the next tile
        if (cvalid) {
          __nv_bfloat16* yb = p.y + (static_cast<size_t>(b) * T + p.cls) * p.E + c;
#pragma unroll
          for (int i = 0; i < HG; ++i) {
            const int h = g * HG + i;
            if (h < H) {
#pragma unroll
              for (int w = 0; w < W; ++w)
                __stcs(reinterpret_cast<unsigned short*>(yb + static_cast<size_t>(h * W + w) * p.E),
                       __bfloat16_as_ushort(__float2bfloat16(__uint_as_float(v[i][w]))));
            }
          }
        }
      }
    }
  } else {
    setmaxnreg_dec<56>();
    if (warp == 8) {
      // ========================================================= MMA issuer
      const uint32_t tD = tbase + C::kDCol, tA = tbase + C::kACol;
      const uint32_t id_row = umma_idesc_bf16(128, 16, 1, 0);
      const uint32_t id_col = umma_idesc_bf16(128, K2, 0, 0);
      const uint32_t id_irow = umma_idesc_bf16(128, 16, 0, 0);
      const uint64_t b1 = umma_smem_desc(smem_u32(sB1), 128, 256, kSwzNone);
      const uint64_t b4 = umma_smem_desc(smem_u32(sB4), 128, 256, kSwzNone);
      const uint64_t b2 = umma_smem_desc(smem_u32(sB2), 128, (K2 / 8) * 128, kSwzNone);
      const uint64_t b3 = umma_smem_desc(smem_u32(sB3), 128, (K2 / 8) * 128, kSwzNone);
      for (int it = 0; it < ntiles; ++it) {
        const int s = it % S;
        mbar_wait(&full_bar[s], (it / S) & 1);
        if (it > 0) named_bar_sync(4, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) {
          const uint64_t a0 = umma_smem_desc(smem_u32(smem + s * C::kStageBytes), C::kHalfStride, 1024, kSwz128);
#pragma unroll
          for (int h = 0; h < H; ++h) umma_ss(tD + h * 16, a0 + static_cast<uint64_t>(h * W * 8), b1, id_row, 0);
          umma_commit(&empty_bar[s]);
          umma_commit(&dready_bar[0]);
        }
        __syncwarp();
        named_bar_sync(1, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) {
#pragma unroll
          for (int kw = 0; kw < WH; ++kw)
#pragma unroll
            for (int ks = 0; ks < K2 / 16; ++ks)
              umma_ts(tD + kw * K2, tA + kw * A2C + ks * 8, b2 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
          umma_commit(&dready_bar[1]);
        }
        __syncwarp();
        named_bar_sync(2, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) {
#pragma unroll
          for (int kw = 0; kw < WH; ++kw)
#pragma unroll
            for (int ks = 0; ks < K2 / 16; ++ks)
              umma_ts(tD + kw * K2, tA + kw * A2C + ks * 8, b3 + static_cast<uint64_t>(ks * 16), id_col, ks > 0);
          umma_commit(&dready_bar[2]);
        }
        __syncwarp();
        named_bar_sync(3, kBarCount);
        tc_fence_after_sync();
        if (lane == 0) {
#pragma unroll
          for (int h = 0; h < H; ++h) umma_ts(tD + h * 16, tA + h * 8, b4, id_irow, 0);
          umma_commit(&dready_bar[3]);
        }
        __syncwarp();
      }
    } else if (warp == 9) {
      // ========================================================= TMA producer + CLS pass-through
      const bool half1 = c0 + 64 < p.E;
      const uint32_t bytes = (half1 ? 2u : 1u) * C::HW * 128u;
      for (int it = 0; it < ntiles; ++it) {
        const int s = it % S;
        const int b = img0 + it * p.img_stride;
        if (it >= S) mbar_wait(&empty_bar[s], ((it / S) - 1) & 1);
        if (lane == 0) {
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

template <int H, int W>
static int fwd_launch(const void* x, const float* filt, void* y, int B, int E, int cls, cudaStream_t st) {
  using C = Cfg<H, W>;
  auto kern = fftmix_tc_fwd_kernel<H, W>;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes) != cudaSuccess)
      return fail(SPECTRE_E_CUDA, "cudaFuncSetAttribute(tc fwd): %s", cudaGetErrorString(cudaGetLastError()));
    configured = true;
  }
  CUtensorMap tmap;
  if (int rc = make_token_map(&tmap, x, B, C::HW, C::HW + cls, E, cls)) return rc;
  FwdParams p;
  p.x = static_cast<const __nv_bfloat16*>(x);
  p.filt = filt;
  p.y = static_cast<__nv_bfloat16*>(y);
  p.B = B; p.E = E; p.cls = cls;
  plan_grid(B, E, &p.nct, &p.img_stride);
  kern<<<p.nct * p.img_stride, kThreads, C::kSmemBytes, st>>>(tmap, p);
  last_launch_count() = 1;
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "fftmix_tc_fwd launch: %s", cudaGetErrorString(err));
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

}  // namespace sb200
