// Issue-rate probe: scalar FFMA vs packed FFMA2 (fma.rn.f32x2), sm_100a.  Prints lane-FMAs per clock per SM.
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
template <int PACKED>
__global__ void __launch_bounds__(256) probe(float* out, float s, int iters) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  const float b = s, c = s * 0.5f;
  for (int it = 0; it < iters; ++it) {
    if (PACKED) {
      unsigned long long bb, cc;
      asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
      asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        unsigned long long v;
        asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[i]), "f"(a[i + 1]));
        v = fma2(v, bb, cc);
        asm("mov.b64 {%0, %1}, %2;" : "=f"(a[i]), "=f"(a[i + 1]) : "l"(v));
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(b), "f"(c));
    }
  }
  float t = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) t += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  for (int packed = 0; packed < 2; ++packed) {
    const int iters = 20000;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (packed) probe<1><<<148 * 8, 256>>>(out, 0.999f, iters); else probe<0><<<148 * 8, 256>>>(out, 0.999f, iters);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
    }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fmas = 148.0 * 8 * 256 * 16 * iters;
    printf("%s: %.3f ms, %.2f lane-FMA/ns (%.1f per SM per clock at %d MHz nominal)\n", packed ? "FFMA2" : "FFMA ", ms,
           fmas / (ms * 1e6), fmas / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
  }
  return 0;
}
