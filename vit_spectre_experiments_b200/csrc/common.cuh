// Shared helpers for the mixer kernels: streaming global I/O of channel pairs /
// channel vectors, the deterministic partial-sum reduction, launch bookkeeping.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/spectre_b200.h"

namespace sb200 {

// ------------------------------------------------------------------ error / launch bookkeeping
std::string& last_error();
int& last_launch_count();
int fail(int code, const char* fmt, ...);
int device_sm_count();
bool aligned16(const void* p);
// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE function attribute: set it on the first use of
// (kernel, current device) -- a process driving several GPUs must not rely on a once-per-process flag.
cudaError_t ensure_dynamic_smem(const void* kernel, int bytes);
// The driver-API tensor-map encoder (cuTensorMapEncodeTiled) wants a context current on the calling thread; a thread that
// has made no runtime-API call yet (an autograd worker) has none and gets CUDA_ERROR_INVALID_CONTEXT.  Binds the
// primary context of the current device to this thread (cudaSetDevice; once per thread and device).
void bind_primary_context();
// cudaOccupancyMaxActiveBlocksPerMultiprocessor, remembered per (kernel, device, threads, dynamic smem): the launchers
// size their grids from it on every call, and the query costs a few microseconds of host time in front of kernels that
// run for 20-80 us.
int cached_blocks_per_sm(const void* kernel, int threads, size_t smem);

// ------------------------------------------------------------------ programmatic dependent launch
// Every kernel here is enqueued with cudaLaunchAttributeProgrammaticStreamSerialization (launch_pdl) and runs
//   pdl_launch_dependents();   as its first statement: the next kernel of the stream may be scheduled as soon as SMs
//                              free up, so its launch latency and its prologue (barrier init, TMEM allocation, filter /
//                              DFT-matrix staging) overlap this kernel's tail;
//   pdl_wait();                before its first access to memory a predecessor may still be writing (or reading, for
//                              buffers it overwrites): blocks until the preceding grid has completed and flushed.
// Launched without the attribute (SB200_PDL=0, or by a caller using <<<>>>) both are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool pdl_enabled();
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ------------------------------------------------------------------ channel-pair I/O
// One thread moves two adjacent channels: a 64-bit access for fp32, a 32-bit
// access for bf16.  Activations are touched exactly once, so loads/stores use the
// streaming (evict-first) cache policy and leave L2 to the filter / partials.
template <class T>
struct PairIO;

template <>
struct PairIO<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return __ldcs(reinterpret_cast<const float2*>(p)); }
  static __device__ __forceinline__ void st(float* p, float2 v) { __stcs(reinterpret_cast<float2*>(p), v); }
};

template <>
struct PairIO<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p) {
    const unsigned u = __ldcs(reinterpret_cast<const unsigned*>(p));
    return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float2 v) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
    __stcs(reinterpret_cast<unsigned*>(p), *reinterpret_cast<const unsigned*>(&h));
  }
};

// ------------------------------------------------------------------ 16-byte channel-vector I/O (DWT kernels)
template <class T>
struct VecIO;

template <>
struct VecIO<float> {
  static constexpr int kLanes = 4;
  static __device__ __forceinline__ void ld(const float* p, float (&v)[4]) {
    const float4 q = __ldcs(reinterpret_cast<const float4*>(p));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  }
  static __device__ __forceinline__ void st(float* p, const float (&v)[4]) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
  }
};

template <>
struct VecIO<__nv_bfloat16> {
  static constexpr int kLanes = 8;
  static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 q = __ldcs(reinterpret_cast<const uint4*>(p));
    const unsigned u[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(u[i] << 16);
      v[2 * i + 1] = __uint_as_float(u[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, const float (&v)[8]) {
    unsigned u[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      u[i] = *reinterpret_cast<const unsigned*>(&h);
    }
    __stcs(reinterpret_cast<uint4*>(p), make_uint4(u[0], u[1], u[2], u[3]));
  }
};

// out[i] = sum_{s<slabs} part[s*n4 + i]  (float4 granularity, fixed order => deterministic)
void launch_reduce_partials(const float* part, float* out, size_t n_floats, int slabs, cudaStream_t stream);

}  // namespace sb200
