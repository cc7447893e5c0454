// Probe for the tensor-core mixer design (profiles/microbench): measures what the guides
// leave open and validates the descriptor conventions of csrc/sm100_ptx.cuh on a real B200.
//   T1/T2  tcgen05.ld / tcgen05.st throughput per SM (4, 8, 16 warps)
//   T3     tcgen05.mma correctness: SS with K-major and MN-major(128B swizzle) A, TS (A in TMEM),
//          aligned and unaligned K offsets; issue rate of small-N MMAs
//   T4     TMA 4-D box load with out-of-bounds zero fill into the padded [h][16][64] layout
// build: nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -o tcgen05_probe tcgen05_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../vit_spectre_experiments_b200/csrc/sm100_ptx.cuh"

using namespace sb200::ptx;

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);           \
      exit(2);                                                                                  \
    }                                                                                           \
  } while (0)

// ------------------------------------------------------------------ T1/T2: TMEM <-> RF bandwidth
template <int MODE>  // 0 = ld x32, 1 = st x32, 2 = ld x16
__global__ void tmem_bw_kernel(long long* cycles, unsigned* sink, int iters) {
  __shared__ uint32_t tbase_s;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tbase_s);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tbase_s;
  const uint32_t lane_base = tbase + ((uint32_t)(32 * (warp & 3)) << 16);
  const int nw = blockDim.x >> 5;
  const int groups = nw / 4;           // warps sharing a lane quarter split the columns
  const int cols_per = 512 / groups;
  const int col0 = (warp >> 2) * cols_per;
  uint32_t acc = 0;
  uint32_t v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = threadIdx.x + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    for (int c = 0; c < cols_per; c += 32) {
      if (MODE == 0) {
        tmem_ld32(lane_base + col0 + c, v);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= v[i];
      } else if (MODE == 1) {
        tmem_st32(lane_base + col0 + c, v);
      } else {
        uint32_t a[16], b[16];
        tmem_ld16(lane_base + col0 + c, a);
        tmem_ld16(lane_base + col0 + c + 16, b);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) acc ^= a[i] ^ b[i];
      }
    }
    if (MODE == 1) tmem_wait_st();
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tbase);
}

// Pipelined variant: 4 loads in flight before one wait (what an epilogue would do).
__global__ void tmem_ld_pipelined_kernel(long long* cycles, unsigned* sink, int iters) {
  __shared__ uint32_t tbase_s;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tbase_s);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tbase_s;
  const uint32_t lane_base = tbase + ((uint32_t)(32 * (warp & 3)) << 16);
  const int nw = blockDim.x >> 5;
  const int groups = nw / 4;
  const int cols_per = 512 / groups;
  const int col0 = (warp >> 2) * cols_per;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    for (int c = 0; c < cols_per; c += 128) {
      uint32_t a[32], b[32], d[32], e[32];
      tmem_ld32(lane_base + col0 + c, a);
      tmem_ld32(lane_base + col0 + c + 32, b);
      tmem_ld32(lane_base + col0 + c + 64, d);
      tmem_ld32(lane_base + col0 + c + 96, e);
      tmem_wait_ld();
#pragma unroll
      for (int i = 0; i < 32; ++i) acc ^= a[i] ^ b[i] ^ d[i] ^ e[i];
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tbase);
}

// ------------------------------------------------------------------ T3: MMA correctness
// A: logical [128 x K] (m, k), B: logical [N x K] (n, k); D[m][n] = sum_k A[m][k] B[n][k].
// The host passes A and B in logical row-major form; the kernel places them in shared
// memory / TMEM in the layout under test.
struct MmaCase {
  int mode;      // 0: SS, A K-major no-swizzle   1: SS, A MN-major 128B swizzle (TMA-like rows)   2: TS (A in TMEM)
  int N, K;      // K multiple of 16
  int krow_off;  // mode 1: first K row inside the swizzled buffer (tests unaligned starts)
  int base_off_mode;  // mode 1: 0 = base_offset field 0, 1 = (start>>7)&7
  int repeat;    // >1: issue the whole MMA sequence `repeat` times for timing (result = last)
  int ndist;     // rate tests: round-robin over this many distinct D column ranges (0/1 = same D)
  int accmode;   // 0: accumulate = (ks > 0); 1: always accumulate
};

constexpr int kASmemBytes = 2 * 32768;  // two 64-channel halves of up to 256 K-rows of 128 B
constexpr int kBSmemBytes = 256 * 64 * 2;

__global__ void __launch_bounds__(128, 1)
mma_probe_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ Bm, float* __restrict__ D,
                 long long* cyc, MmaCase cs) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + kASmemBytes;
  __shared__ uint64_t bar;
  __shared__ uint32_t tbase_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = cs.N, K = cs.K;

  if (warp == 0) tmem_alloc<512>(&tbase_s);
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  // zero A region so padded / unread rows are finite
  for (int i = tid; i < kASmemBytes / 16; i += 128) reinterpret_cast<uint4*>(sA)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();

  // ---- B: K-major, no swizzle: (n%8)*16 + (n/8)*SBO + (k%8)*2 + (k/8)*LBO, LBO=128, SBO=(K/8)*128
  const uint32_t b_lbo = 128, b_sbo = (uint32_t)(K / 8) * 128;
  for (int i = tid; i < N * K; i += 128) {
    const int n = i / K, k = i % K;
    const uint32_t off = (n % 8) * 16 + (n / 8) * b_sbo + (k % 8) * 2 + (k / 8) * b_lbo;
    *reinterpret_cast<__nv_bfloat16*>(sB + off) = Bm[i];
  }
  // ---- A
  uint32_t a_lbo = 0, a_sbo = 0;
  if (cs.mode == 0) {
    a_lbo = 128;
    a_sbo = (uint32_t)(K / 8) * 128;
    for (int i = tid; i < 128 * K; i += 128) {
      const int m = i / K, k = i % K;
      const uint32_t off = (m % 8) * 16 + (m / 8) * a_sbo + (k % 8) * 2 + (k / 8) * a_lbo;
      *reinterpret_cast<__nv_bfloat16*>(sA + off) = A[i];
    }
  } else if (cs.mode == 1) {
    // rows r = krow_off + k, 128 B per row = 64 m; chunk c16 = (m%64)/8 XOR (r%8); halves at 32768
    for (int i = tid; i < 128 * K; i += 128) {
      const int m = i / K, k = i % K;
      const int r = cs.krow_off + k;
      const uint32_t off = (m / 64) * 32768 + r * 128 + ((((m % 64) / 8) ^ (r % 8)) * 16) + (m % 8) * 2;
      *reinterpret_cast<__nv_bfloat16*>(sA + off) = A[i];
    }
    a_lbo = 32768;  // between 64-wide MN blocks
    a_sbo = 1024;   // between 8-row K groups
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tbase = tbase_s;
  const uint32_t d_tmem = tbase;             // columns [0, N)
  const uint32_t a_tmem = tbase + 256;       // columns [256, 256 + K/2)
  if (cs.mode == 2) {
    // thread = lane = row m; column j holds (k=2j, k=2j+1) as (lo, hi) bf16
    const int m = tid;
    for (int j0 = 0; j0 < K / 2; j0 += 8) {
      uint32_t v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const __nv_bfloat16 lo = A[m * K + 2 * (j0 + j)], hi = A[m * K + 2 * (j0 + j) + 1];
        v[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
      }
      tmem_st8(a_tmem + ((uint32_t)(32 * warp) << 16) + j0, v);
    }
    tmem_wait_st();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
  }

  long long t0 = 0, t1 = 0;
  if (warp == 0) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, N, cs.mode == 1 ? 1 : 0, 0);
      t0 = clock64();
      const int ndist = cs.ndist > 1 ? cs.ndist : 1;
      for (int rep = 0; rep < cs.repeat; ++rep) {
        const uint32_t d_tmem = tbase + (uint32_t)((rep % ndist) * N);
        for (int ks = 0; ks < K / 16; ++ks) {
          const uint64_t bdesc = umma_smem_desc(smem_u32(sB) + ks * 2 * b_lbo, b_lbo, b_sbo, kSwzNone);
          if (cs.mode == 0) {
            const uint64_t adesc = umma_smem_desc(smem_u32(sA) + ks * 2 * a_lbo, a_lbo, a_sbo, kSwzNone);
            umma_ss(d_tmem, adesc, bdesc, idesc, cs.accmode ? 1 : (ks > 0));
          } else if (cs.mode == 1) {
            const uint32_t start = smem_u32(sA) + (cs.krow_off + ks * 16) * 128;
            const uint32_t bo = cs.base_off_mode ? ((start >> 7) & 7) : 0;
            const uint64_t adesc = umma_smem_desc(start, a_lbo, a_sbo, kSwz128, bo);
            umma_ss(d_tmem, adesc, bdesc, idesc, cs.accmode ? 1 : (ks > 0));
          } else {
            umma_ts(d_tmem, a_tmem + ks * 8, bdesc, idesc, cs.accmode ? 1 : (ks > 0));
          }
        }
      }
      const long long ti = clock64();
      umma_commit(&bar);
      mbar_wait(&bar, 0);
      t1 = clock64();
      cyc[0] = t1 - t0;
      cyc[1] = ti - t0;
    }
    __syncwarp();
  }
  __syncthreads();
  tc_fence_after_sync();
  // read D: thread = row
  for (int n0 = 0; n0 < N; n0 += 8) {
    uint32_t v[8];
    tmem_ld8(d_tmem + ((uint32_t)(32 * warp) << 16) + n0, v);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 8; ++j) D[tid * N + n0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tbase);
}


// ------------------------------------------------------------------ T5: unrolled issue rate (what the real kernel does)
// One tile's worth of MMAs of the 14x14 mixer: 14 SS (N=16) + 16 TS (N=32) + 16 TS + 14 TS (N=16),
// all descriptors compile-time offsets from registers, fully unrolled.  Garbage operands; timing only.
__global__ void __launch_bounds__(128, 1) mma_unrolled_kernel(long long* cyc, int reps, int row_stride16) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tbase_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (warp == 0) tmem_alloc<512>(&tbase_s);
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  for (int i = tid; i < 65536 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tb = tbase_s;
  if (warp == 0 && elect_one()) {
    const uint32_t sa = smem_u32(smem);
    const uint64_t a0 = umma_smem_desc(sa, 28672, 1024, kSwz128);
    const uint64_t b1 = umma_smem_desc(sa + 60000, 128, 256, kSwzNone);
    const uint64_t b2 = umma_smem_desc(sa + 61000, 128, 512, kSwzNone);
    const uint32_t id_row = umma_idesc_bf16(128, 16, 1, 0);
    const uint32_t id_c32 = umma_idesc_bf16(128, 32, 0, 0);
    const uint32_t id_c16 = umma_idesc_bf16(128, 16, 0, 0);
    uint32_t phase = 0;
    long long t_issue = 0, t_total = 0;
    for (int r = 0; r < reps; ++r) {
      const long long t0 = clock64();
#pragma unroll
      for (int h = 0; h < 14; ++h) umma_ss(tb + h * 16, a0 + (uint64_t)(h * row_stride16), b1, id_row, 0);
      const long long t1 = clock64();
#pragma unroll
      for (int kw = 0; kw < 8; ++kw) {
        umma_ts(tb + kw * 32, tb + 256 + kw * 16, b2, id_c32, 0);
        umma_ts(tb + kw * 32, tb + 256 + kw * 16 + 8, b2 + 16, id_c32, 1);
      }
      const long long t2 = clock64();
#pragma unroll
      for (int h = 0; h < 14; ++h) umma_ts(tb + h * 16, tb + 256 + h * 8, b1, id_c16, 0);
      const long long t3 = clock64();
      umma_commit(&bar);
      mbar_wait(&bar, phase);
      phase ^= 1;
      const long long t4 = clock64();
      if (r == reps - 1) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t3 - t2; cyc[3] = t4 - t3; }
      t_issue += t3 - t0; t_total += t4 - t0;
    }
    cyc[4] = t_issue; cyc[5] = t_total;
  }
  __syncthreads();
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tb);
}

// ------------------------------------------------------------------ T4: TMA load with OOB zero fill
__global__ void __launch_bounds__(128, 1)
tma_probe_kernel(const __grid_constant__ CUtensorMap tmap, uint8_t* out, int out_bytes, int b, int c0) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < out_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0xdeadbeefu;
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, out_bytes);
    tma_load_4d(smem, &tmap, &bar, c0, 0, 0, b);
  }
  mbar_wait(&bar, 0);
  for (int i = threadIdx.x; i < out_bytes / 4; i += 128)
    reinterpret_cast<uint32_t*>(out)[i] = reinterpret_cast<uint32_t*>(smem)[i];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int run_mma_case(const MmaCase& cs, const char* name) {
  const int N = cs.N, K = cs.K;
  std::vector<__nv_bfloat16> hA(128 * K), hB(N * K);
  std::vector<float> fA(128 * K), fB(N * K);
  srand(1234 + N * 7 + K);
  for (int i = 0; i < 128 * K; ++i) { fA[i] = (float)((rand() % 17) - 8); hA[i] = __float2bfloat16(fA[i]); }
  for (int i = 0; i < N * K; ++i) { fB[i] = (float)((rand() % 9) - 4) * 0.5f; hB[i] = __float2bfloat16(fB[i]); }
  __nv_bfloat16 *dA, *dB; float* dD; long long* dc;
  CK(cudaMalloc(&dA, hA.size() * 2)); CK(cudaMalloc(&dB, hB.size() * 2));
  CK(cudaMalloc(&dD, 128 * N * 4)); CK(cudaMalloc(&dc, 16));
  CK(cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0xff, 128 * N * 4));
  const int smem = kASmemBytes + kBSmemBytes;
  CK(cudaFuncSetAttribute(mma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  mma_probe_kernel<<<1, 128, smem>>>(dA, dB, dD, dc, cs);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("[T3] %-44s CUDA ERROR %s\n", name, cudaGetErrorString(e)); return -1; }
  std::vector<float> hD(128 * N); long long cyc = 0, cyc_issue = 0;
  CK(cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&cyc, dc, 8, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&cyc_issue, dc + 1, 8, cudaMemcpyDeviceToHost));
  int bad = 0; double maxerr = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fA[m * K + k] * fB[n * K + k];
      const double err = fabs(ref - hD[m * N + n]);
      if (!(err <= 1e-3)) ++bad;
      if (err > maxerr || err != err) maxerr = err;
    }
  const int nmma = cs.repeat * (K / 16);
  printf("[T3] %-44s N=%3d K=%3d  mismatches=%5d/%d maxerr=%.3g  cycles=%lld issue=%lld (%d mma, %.1f cyc/mma) ndist=%d acc=%d\n", name, N, K, bad,
         128 * N, maxerr, cyc, cyc_issue, nmma, (double)cyc / nmma, cs.ndist, cs.accmode);
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dc);
  return bad;
}

int main(int argc, char** argv) {
  const char* which = argc > 1 ? argv[1] : "all";
  const int sel = argc > 2 ? atoi(argv[2]) : -1;
  auto want = [&](const char* t) { return !strcmp(which, "all") || !strcmp(which, t); };
  int dev = 0; cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  printf("device: %s sm_%d%d SMs=%d\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount);
  const int nsm = prop.multiProcessorCount;

  // ---------------- T1/T2
  if (want("t1")) {
  long long* dcyc; unsigned* dsink;
  CK(cudaMalloc(&dcyc, nsm * 8)); CK(cudaMalloc(&dsink, nsm * 512 * 4));
  const int iters = 200;
  for (int mode = 0; mode < 4; ++mode) {
    for (int nw : {4, 8, 16}) {
      for (int grid : {1, nsm}) {
        for (int rep = 0; rep < 2; ++rep) {
          if (mode == 0) tmem_bw_kernel<0><<<grid, nw * 32>>>(dcyc, dsink, iters);
          if (mode == 1) tmem_bw_kernel<1><<<grid, nw * 32>>>(dcyc, dsink, iters);
          if (mode == 2) tmem_bw_kernel<2><<<grid, nw * 32>>>(dcyc, dsink, iters);
          if (mode == 3) tmem_ld_pipelined_kernel<<<grid, nw * 32>>>(dcyc, dsink, iters);
          CK(cudaDeviceSynchronize());
        }
        std::vector<long long> h(grid);
        CK(cudaMemcpy(h.data(), dcyc, grid * 8, cudaMemcpyDeviceToHost));
        long long mx = 0; for (auto c : h) mx = c > mx ? c : mx;
        const double bytes = (double)iters * 512 * 128 * 4;  // whole TMEM per iteration
        const char* nm[4] = {"ld.x32 wait-each", "st.x32", "ld.x16x2 wait-each", "ld.x32 x4 pipelined"};
        printf("[T1] %-20s warps=%2d grid=%3d  %.1f B/clk/SM (cycles=%lld)\n", nm[mode], nw, grid, bytes / mx, mx);
      }
    }
  }

  }
  // ---------------- T3
  if (want("t3")) {
    int idx = 0;
#define CASE(...) do { if (sel < 0 || sel == idx) run_mma_case(__VA_ARGS__); ++idx; } while (0)
    CASE(MmaCase{0, 16, 16, 0, 0, 1}, "SS A K-major noswz");
    CASE(MmaCase{0, 32, 32, 0, 0, 1}, "SS A K-major noswz (2 k-steps)");
    CASE(MmaCase{1, 16, 16, 0, 0, 1}, "SS A MN-major sw128 row0");
    CASE(MmaCase{1, 16, 16, 16, 0, 1}, "SS A MN-major sw128 row16");
    CASE(MmaCase{1, 16, 32, 32, 0, 1}, "SS A MN-major sw128 row32 K=32");
    CASE(MmaCase{1, 16, 16, 14, 0, 1}, "SS A MN-major sw128 row14 (unaligned) bo=0");
    CASE(MmaCase{1, 16, 16, 14, 1, 1}, "SS A MN-major sw128 row14 (unaligned) bo=1");
    CASE(MmaCase{1, 16, 16, 4, 0, 1}, "SS A MN-major sw128 row4 (unaligned) bo=0");
    CASE(MmaCase{1, 16, 16, 4, 1, 1}, "SS A MN-major sw128 row4 (unaligned) bo=1");
    CASE(MmaCase{2, 16, 16, 0, 0, 1}, "TS A in TMEM");
    CASE(MmaCase{2, 32, 32, 0, 0, 1}, "TS A in TMEM (2 k-steps)");
    CASE(MmaCase{2, 32, 64, 0, 0, 1}, "TS A in TMEM (4 k-steps)");
  // issue-rate
    CASE(MmaCase{1, 16, 16, 0, 0, 512}, "rate: SS MN sw128 N=16 x512");
    CASE(MmaCase{1, 32, 16, 0, 0, 512}, "rate: SS MN sw128 N=32 x512");
    CASE(MmaCase{2, 16, 16, 0, 0, 512}, "rate: TS N=16 x512");
    CASE(MmaCase{2, 32, 32, 0, 0, 256}, "rate: TS N=32 K=32 x256");
    CASE(MmaCase{2, 64, 32, 0, 0, 256}, "rate: TS N=64 K=32 x256");
    CASE(MmaCase{2, 128, 32, 0, 0, 256}, "rate: TS N=128 K=32 x256");

    for (int len : {8, 32, 256}) {
      CASE(MmaCase{1, 16, 16, 0, 0, len, 1, 1}, "rate2: SS N=16 same D acc=1");
      CASE(MmaCase{1, 16, 16, 0, 0, len, 14, 0}, "rate2: SS N=16 14 D regions acc=0");
      CASE(MmaCase{1, 16, 16, 0, 0, len, 14, 1}, "rate2: SS N=16 14 D regions acc=1");
      CASE(MmaCase{0, 16, 16, 0, 0, len, 14, 0}, "rate2: SS(Kmajor noswz) N=16 14 D acc=0");
      CASE(MmaCase{2, 16, 16, 0, 0, len, 14, 0}, "rate2: TS N=16 14 D regions acc=0");
      CASE(MmaCase{2, 32, 32, 0, 0, len, 8, 0}, "rate2: TS N=32 K=32 8 D regions");
      CASE(MmaCase{2, 32, 32, 0, 0, len, 1, 1}, "rate2: TS N=32 K=32 same D acc=1");
      CASE(MmaCase{2, 128, 32, 0, 0, len, 2, 0}, "rate2: TS N=128 K=32 2 D regions");
      CASE(MmaCase{2, 128, 64, 0, 0, len, 1, 1}, "rate2: TS N=128 K=64 same D acc=1");
      CASE(MmaCase{2, 256, 64, 0, 0, len, 1, 1}, "rate2: TS N=256 K=64 same D acc=1");
    }

  }
  // ---------------- T4
  if (want("t4")) {
    void* fn = nullptr; cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn) { printf("[T4] cuTensorMapEncodeTiled not found\n"); return 1; }
    EncodeTiledFn encode = (EncodeTiledFn)fn;
    const int B = 3, H = 14, W = 14, E = 384, T = 1 + H * W;
    std::vector<__nv_bfloat16> hx((size_t)B * T * E);
    for (size_t i = 0; i < hx.size(); ++i) hx[i] = __float2bfloat16((float)((i * 37) % 251) - 125.f);
    __nv_bfloat16* dx; CK(cudaMalloc(&dx, hx.size() * 2));
    CK(cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice));
    CUtensorMap tmap;
    cuuint64_t dims[4] = {(cuuint64_t)E, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {(cuuint64_t)E * 2, (cuuint64_t)W * E * 2, (cuuint64_t)T * E * 2};
    cuuint32_t box[4] = {64, 16, 14, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)(dx + E), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("[T4] cuTensorMapEncodeTiled -> %d\n", (int)r);
    if (r == CUDA_SUCCESS) {
      const int bytes = 64 * 16 * 14 * 2;
      uint8_t* dout; CK(cudaMalloc(&dout, bytes));
      CK(cudaFuncSetAttribute(tma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
      const int b = 2, c0 = 128;
      tma_probe_kernel<<<1, 128, bytes>>>(tmap, dout, bytes, b, c0);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("[T4] CUDA ERROR %s\n", cudaGetErrorString(e)); return 1; }
      std::vector<uint8_t> ho(bytes);
      CK(cudaMemcpy(ho.data(), dout, bytes, cudaMemcpyDeviceToHost));
      int bad = 0, badpad = 0;
      for (int h = 0; h < 14; ++h)
        for (int w = 0; w < 16; ++w)
          for (int c = 0; c < 64; ++c) {
            const int rrow = h * 16 + w;
            const int off = rrow * 128 + (((c / 8) ^ (rrow % 8)) * 16) + (c % 8) * 2;
            const float got = __bfloat162float(*reinterpret_cast<__nv_bfloat16*>(&ho[off]));
            if (w < 14) {
              const float ref = __bfloat162float(hx[((size_t)b * T + 1 + h * W + w) * E + c0 + c]);
              if (got != ref) ++bad;
            } else if (got != 0.f) ++badpad;
          }
      printf("[T4] TMA padded-box load: mismatches=%d pad-nonzero=%d (of %d)\n", bad, badpad, 64 * 16 * 14);
    }
  }

  if (want("t5")) {
    long long* dc; CK(cudaMalloc(&dc, 64));
    CK(cudaFuncSetAttribute(mma_unrolled_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    for (int rs : {128, 112}) for (int reps : {1, 16}) {
      printf("row stride %d B: ", rs * 16);
      mma_unrolled_kernel<<<1, 128, 65536>>>(dc, reps, rs);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("[T5] CUDA ERROR %s\n", cudaGetErrorString(e)); return 1; }
      long long h[8]; CK(cudaMemcpy(h, dc, 48, cudaMemcpyDeviceToHost));
      printf("[T5] reps=%2d last: row14(SS N16)=%lld col16(TS N32)=%lld irow14(TS N16)=%lld drain=%lld | per rep: issue=%.0f total=%.0f (44 mma)\n",
             reps, h[0], h[1], h[2], h[3], (double)h[4] / reps, (double)h[5] / reps);
    }
  }
  printf("probe done\n");
  return 0;
}
