// Pipe-rate microbenchmark, round 2: the instructions the bf16 row-wise tails are rebuilt from.
//   FHADD.BF16 / FHFMA.BF16 (mixed-precision add / fma: bf16 operands straight out of a packed register half,
//   fp32 result -- PTX add.rn.f32.bf16 / fma.rn.f32.bf16), MUFU.TANH / EX2, F2FP pack, SHF/LOP3 unpack, and mixes
//   of a packed-FMA stream with ALU / MUFU streams (do they overlap?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes2 pipes2.cu && ./pipes2
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#define ITERS 2048
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fhadd(unsigned short h, float c) { float r; asm volatile("add.rn.f32.bf16 %0, %1, %2;" : "=f"(r) : "h"(h), "f"(c)); return r; }
__device__ __forceinline__ float fhfma(unsigned short a, unsigned short b, float c) { float r; asm volatile("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(r) : "h"(a), "h"(b), "f"(c)); return r; }
__device__ __forceinline__ float tanh_a(float x) { float r; asm volatile("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ex2_a(float x) { float r; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

template <int MODE>
__global__ void __launch_bounds__(1024) k(float* out, long long* cyc, int dummy) {
  const int tid = threadIdx.x;
  float a[16];
  u64 p[8];
  unsigned w[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = tid * 0.001f + i * 0.01f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { p[i] = (u64)__float_as_uint(a[i]) << 32 | __float_as_uint(a[i + 8]); w[i] = 0x3f803f80u + tid + i + dummy; }
  const u64 m2 = (u64)__float_as_uint(0.9999f) << 32 | __float_as_uint(1.0001f + dummy), c2 = (u64)__float_as_uint(1e-3f) << 32 | __float_as_uint(1e-4f);
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
    if (MODE == 0) {          // 16 FHADD.BF16
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fhadd(static_cast<unsigned short>(i & 1 ? w[i >> 1] >> 16 : w[i >> 1] & 0xffff), a[i]);
    } else if (MODE == 1) {   // 16 FHFMA.BF16
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fhfma(static_cast<unsigned short>(w[i >> 1] & 0xffff), static_cast<unsigned short>(w[i >> 1] >> 16), a[i]);
    } else if (MODE == 2) {   // 16 MUFU.TANH
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = tanh_a(a[i]);
    } else if (MODE == 3) {   // 16 MUFU.EX2
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = ex2_a(a[i]);
    } else if (MODE == 4) {   // 8 F2FP packs + 8 unpacks (SHL) to keep a dependency
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
        unsigned u = *reinterpret_cast<unsigned*>(&h);
        a[2 * i] = __uint_as_float(u << 16);
      }
    } else if (MODE == 5) {   // 16 unpack ops (8 SHL + 8 LOP3), xor back to keep them live
#pragma unroll
      for (int i = 0; i < 8; ++i) { unsigned lo = w[i] << 16, hi = w[i] & 0xffff0000u; w[i] = lo ^ (hi >> 3); }
    } else if (MODE == 6) {   // 8 FFMA2 + 16 ALU ops (independent streams)
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], m2, c2); unsigned lo = w[i] << 16, hi = w[i] & 0xffff0000u; w[i] = lo ^ (hi >> 3); }
    } else if (MODE == 7) {   // 8 FFMA2 + 8 MUFU.TANH (independent streams)
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], m2, c2); a[i] = tanh_a(a[i]); }
    } else if (MODE == 8) {   // 8 FFMA2 + 8 FHADD
#pragma unroll
      for (int i = 0; i < 8; ++i) { p[i] = fma2(p[i], m2, c2); a[i] = fhadd(static_cast<unsigned short>(w[i] & 0xffff), a[i]); }
    } else if (MODE == 9) {   // 8 FFMA2 alone (reference)
#pragma unroll
      for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], m2, c2);
    } else if (MODE == 10) {  // 16 FHADD + 16 scalar FFMA
#pragma unroll
      for (int i = 0; i < 16; ++i) { a[i] = fhadd(static_cast<unsigned short>(w[i >> 1] & 0xffff), a[i]); a[i] = fmaf(a[i], 0.999f, 1e-3f); }
    }
  }
  long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s += __uint_as_float(static_cast<unsigned>(p[i])) + __uint_as_float(static_cast<unsigned>(p[i] >> 32)) + __uint_as_float(w[i]);
  out[blockIdx.x * blockDim.x + tid] = s;
  if (tid == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int per_iter, int threads) {
  float* out; long long* cyc;
  int sms = 148;
  cudaMalloc(&out, sizeof(float) * sms * threads);
  cudaMalloc(&cyc, sizeof(long long) * sms);
  k<MODE><<<sms, threads>>>(out, cyc, 0);
  k<MODE><<<sms, threads>>>(out, cyc, 0);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < sms; ++i) mx = h[i] > mx ? h[i] : mx;
  const double winstr = (double)ITERS * per_iter * (threads / 32);
  printf("%-44s threads/SM=%4d cycles=%9lld warp-instr/clk/SM=%.3f (%s)\n", name, threads, mx, winstr / mx, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int th : {256, 512, 1024}) {
    run<0>("FHADD.BF16 (16/iter)", 16, th);
    run<1>("FHFMA.BF16 (16/iter)", 16, th);
    run<2>("MUFU.TANH (16/iter)", 16, th);
    run<3>("MUFU.EX2 (16/iter)", 16, th);
    run<4>("F2FP pack + SHL (8+8/iter)", 16, th);
    run<5>("SHL + LOP3 unpack (+xor,shr) (32 ALU/iter)", 32, th);
    run<9>("FFMA2 alone (8/iter)", 8, th);
    run<6>("FFMA2 + 4 ALU each (8+32/iter)", 40, th);
    run<7>("FFMA2 + MUFU.TANH (8+8/iter)", 16, th);
    run<8>("FFMA2 + FHADD (8+8/iter)", 16, th);
    run<10>("FHADD + FFMA scalar dependent (16+16/iter)", 32, th);
  }
  return 0;
}
