// Pipe-rate microbenchmark used to choose the mixer kernels' arithmetic shape
// (scalar FFMA vs packed FFMA2, SHFL, LDS, legacy mma.sync).  Not part of the
// product; results are recorded in profiles/r01_pipes.txt and DESIGN.md.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

#define ITERS 2048

template <int MODE>
__global__ void __launch_bounds__(1024) k(float* out, long long* cyc, int dummy) {
  extern __shared__ float sm[];
  const int tid = threadIdx.x;
  float2 a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = make_float2(tid * 0.001f + i, tid * 0.002f - i);
  float2 m = make_float2(1.0001f + dummy, 0.9999f + dummy), c = make_float2(0.001f * dummy + 1e-3f, 1e-4f);
  for (int i = tid; i < 8192; i += blockDim.x) sm[i] = i;
  unsigned ra[4] = {1u + tid, 2u, 3u, 4u}, rb[2] = {5u + tid, 6u};
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {  // 16 scalar FFMA (3-reg)
#pragma unroll
      for (int i = 0; i < 8; ++i) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }
    } else if (MODE == 1) {  // 8 packed FFMA2
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = __ffma2_rn(a[i], m, c);
    } else if (MODE == 2) {  // 16 scalar FADD
#pragma unroll
      for (int i = 0; i < 8; ++i) { a[i].x = a[i].x + c.x; a[i].y = a[i].y + c.y; }
    } else if (MODE == 3) {  // 8 packed FADD2
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = __fadd2_rn(a[i], c);
    } else if (MODE == 4) {  // 8 SHFL
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i].x = __shfl_xor_sync(0xffffffffu, a[i].x, 1 + (i & 3));
    } else if (MODE == 5) {  // 8 LDS.64 conflict-free
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float2 v = *reinterpret_cast<float2*>(&sm[(((int)a[i].x & 63) * 64 + (tid & 31) * 2) & 8191]);
        a[i].x = v.x; a[i].y += v.y;
      }
    } else if (MODE == 6) {  // 8 LDS.128 conflict-free
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float4 v = *reinterpret_cast<float4*>(&sm[(((int)a[i].x & 31) * 128 + (tid & 31) * 4) & 8191]);
        a[i].x = v.x; a[i].y += v.y + v.z + v.w;
      }
    } else if (MODE == 7) {  // 8 mma.sync m16n8k16 bf16
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                     : "r"(ra[0]), "r"(ra[1]), "r"(ra[2]), "r"(ra[3]), "r"(rb[0]), "r"(rb[1]));
    } else if (MODE == 8) {  // 8 mma.sync m16n8k8 tf32
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                     : "r"(ra[0]), "r"(ra[1]), "r"(ra[2]), "r"(ra[3]), "r"(rb[0]), "r"(rb[1]));
    } else if (MODE == 9) {  // mixed: 8 FFMA2 + 8 FADD2 (butterfly-like)
#pragma unroll
      for (int i = 0; i < 8; ++i) { a[i] = __ffma2_rn(a[i], m, c); a[i] = __fadd2_rn(a[i], c); }
    } else if (MODE == 10) {  // 8 packed FMUL2
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = __fmul2_rn(a[i], m);
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y + acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
  out[blockIdx.x * blockDim.x + tid] = s;
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int instr_per_iter, int threads) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  k<MODE><<<148, threads, 32768>>>(out, cyc, 0);
  k<MODE><<<148, threads, 32768>>>(out, cyc, 0);
  cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
  double warp_instr = (double)(threads / 32) * ITERS * instr_per_iter;
  printf("%-34s threads/SM=%4d  cycles=%9.0f  warp-instr/clk/SM=%.3f  (err=%s)\n", name, threads, avg,
         warp_instr / avg, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int threads : {256, 512, 1024}) {
    run<0>("FFMA scalar 3-reg (16/iter)", 16, threads);
    run<1>("FFMA2 packed (8/iter)", 8, threads);
    run<2>("FADD scalar (16/iter)", 16, threads);
    run<3>("FADD2 packed (8/iter)", 8, threads);
    run<10>("FMUL2 packed (8/iter)", 8, threads);
    run<9>("FFMA2+FADD2 dependent (16/iter)", 16, threads);
    run<4>("SHFL.BFLY (8/iter)", 8, threads);
    run<5>("LDS.64 (8/iter)", 8, threads);
    run<6>("LDS.128 (8/iter)", 8, threads);
    run<7>("mma.sync m16n8k16 bf16 (8/iter)", 8, threads);
    run<8>("mma.sync m16n8k8 tf32 (8/iter)", 8, threads);
  }
  return 0;
}
