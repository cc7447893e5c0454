// Probe: register <-> (lane, column) mapping of tcgen05.ld.16x256b, to plan a TMEM -> stmatrix.trans
// output path.  TMEM is filled through 32x32b stores with value = lane*1000 + column.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include "../../vit_spectre_experiments_b200/csrc/sm100_ptx.cuh"
using namespace sb200::ptx;

__global__ void k(uint32_t* out) {
  __shared__ uint32_t tb_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc<64>(&tb_s);
  tc_fence_before_sync(); __syncthreads(); tc_fence_after_sync();
  const uint32_t tb = tb_s;
  const uint32_t q = (uint32_t)(32 * warp) << 16;
  uint32_t v[32];
  for (int i = 0; i < 32; ++i) v[i] = (32 * warp + lane) * 1000 + i;
  tmem_st32(tb + q, v);
  tmem_wait_st();
  tc_fence_before_sync(); __syncthreads(); tc_fence_after_sync();
  // 16x256b.x2: 8 registers; lane base +0 and +16 inside the warp's quarter
  for (int half = 0; half < 2; ++half) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(tb + q + ((uint32_t)(16 * half) << 16) + 4 /*column offset 4*/));
    tmem_wait_ld();
    for (int i = 0; i < 8; ++i) out[((warp * 2 + half) * 32 + lane) * 8 + i] = r[i];
  }
  tc_fence_before_sync(); __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tb);
}
int main() {
  uint32_t* d; cudaMalloc(&d, 4 * 2 * 32 * 8 * 4);
  k<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CUDA ERROR %s\n", cudaGetErrorString(e)); return 1; }
  static uint32_t h[4 * 2 * 32 * 8];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  for (int w = 0; w < 2; ++w) for (int half = 0; half < 2; ++half) {
    printf("warp %d half %d (values = lane*1000+col; column base 4)\n", w, half);
    for (int l = 0; l < 32; ++l) { printf(" T%02d:", l); for (int i = 0; i < 8; ++i) printf(" %6u", h[((w*2+half)*32+l)*8+i]); printf("\n"); }
  }
  // check the hypothesis: r[2j+e] (j<2: rows T/4, T/4+8 ; repeat rep) 
  int bad = 0;
  for (int w = 0; w < 4; ++w) for (int half = 0; half < 2; ++half) for (int l = 0; l < 32; ++l)
    for (int rep = 0; rep < 2; ++rep) for (int up = 0; up < 2; ++up) for (int e2 = 0; e2 < 2; ++e2) {
      const uint32_t got = h[((w*2+half)*32+l)*8 + rep*4 + up*2 + e2];
      const uint32_t exp = (32*w + 16*half + l/4 + 8*up) * 1000 + 4 + rep*8 + 2*(l%4) + e2;
      if (got != exp) ++bad;
    }
  printf("hypothesis r[rep*4+up*2+e] = (lane T/4+8up, col 8rep+2(T%%4)+e): mismatches=%d\n", bad);
  return 0;
}
