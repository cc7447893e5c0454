// extern "C" entry points declared in include/spectre_b200.h: argument validation,
// error bookkeeping, dispatch to the templated launchers.  No torch types here.
#include <cstdarg>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace sb200 {

int fftmix_supported(int H, int W);
int fftmix_fwd_dispatch(const void* x, const float* filt, void* y, int B, int H, int W, int E, int cls, int dtype,
                        cudaStream_t st);
int fftmix_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                        size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st);
size_t fftmix_bwd_workspace(int B, int H, int W, int E);
int dwtmix_dispatch(bool backward, const void* in, const void* xsave, const float* w, void* out, float* dw, void* ws,
                    size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st);
size_t dwtmix_bwd_workspace(int B, int H, int W, int E);

std::string& last_error() {
  thread_local std::string msg;
  return msg;
}
int& last_launch_count() {
  thread_local int n = 0;
  return n;
}
int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error() = buf;
  last_launch_count() = 0;
  return code;
}
int device_sm_count() {
  thread_local int cached_dev = -1, cached_sms = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceGetAttribute(&cached_sms, cudaDevAttrMultiProcessorCount, dev);
    cached_dev = dev;
  }
  return cached_sms > 0 ? cached_sms : 148;
}
void bind_primary_context() {
  thread_local int bound = -1;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  if (dev != bound) {
    cudaSetDevice(dev);
    bound = dev;
  }
}
bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("SB200_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

cudaError_t ensure_dynamic_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::unordered_map<const void*, unsigned long long> done[2];   // bit d (of word d / 64) = device d configured
  int dev = 0;
  if (cudaError_t e = cudaGetDevice(&dev); e != cudaSuccess) return e;
  if (dev < 0 || dev >= 128) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  const unsigned long long bit = 1ull << (dev & 63);
  std::lock_guard<std::mutex> lock(mu);
  unsigned long long& mask = done[dev >> 6][kernel];
  if (mask & bit) return cudaSuccess;
  const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) mask |= bit;
  return e;
}

int cached_blocks_per_sm(const void* kernel, int threads, size_t smem) {
  struct Key {
    const void* k;
    int dev, threads;
    size_t smem;
    bool operator==(const Key& o) const { return k == o.k && dev == o.dev && threads == o.threads && smem == o.smem; }
  };
  struct Hash {
    size_t operator()(const Key& a) const {
      return std::hash<const void*>()(a.k) ^ (static_cast<size_t>(a.dev) * 0x9E3779B97F4A7C15ull) ^ (a.smem << 7) ^
             static_cast<size_t>(a.threads);
    }
  };
  static std::mutex mu;
  static std::unordered_map<Key, int, Hash> memo;
  Key key{kernel, 0, threads, smem};
  cudaGetDevice(&key.dev);
  {
    std::lock_guard<std::mutex> lock(mu);
    auto it = memo.find(key);
    if (it != memo.end()) return it->second;
  }
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem) != cudaSuccess || per_sm < 1) {
    cudaGetLastError();
    return 1;   // not remembered: the attribute may not be set yet
  }
  std::lock_guard<std::mutex> lock(mu);
  memo[key] = per_sm;
  return per_sm;
}

// out[i] = sum_k part[k][i], float4 granularity.  CTA = 64 float4 columns x 4 slab lanes; a lane strides over the
// slabs with four independent accumulators (16 loads in flight per column), the lanes are combined through shared
// memory in a fixed order => deterministic, and ~1 us instead of a serial chain of `slabs` L2 round trips.
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float4* __restrict__ part, float4* __restrict__ out,
                                                              size_t n4, int slabs) {
  __shared__ float4 sm[4][64];
  pdl_launch_dependents();
  pdl_wait();
  const int col = threadIdx.x & 63, ln = threadIdx.x >> 6;
  const size_t i = static_cast<size_t>(blockIdx.x) * 64 + col;
  float4 a[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n4) {
    int k = ln;
    for (; k + 12 < slabs; k += 16) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 t = part[static_cast<size_t>(k + 4 * u) * n4 + i];
        a[u].x += t.x; a[u].y += t.y; a[u].z += t.z; a[u].w += t.w;
      }
    }
    for (; k < slabs; k += 4) {
      const float4 t = part[static_cast<size_t>(k) * n4 + i];
      a[0].x += t.x; a[0].y += t.y; a[0].z += t.z; a[0].w += t.w;
    }
  }
  sm[ln][col] = make_float4((a[0].x + a[1].x) + (a[2].x + a[3].x), (a[0].y + a[1].y) + (a[2].y + a[3].y),
                            (a[0].z + a[1].z) + (a[2].z + a[3].z), (a[0].w + a[1].w) + (a[2].w + a[3].w));
  __syncthreads();
  if (ln == 0 && i < n4) {
    const float4 p0 = sm[0][col], p1 = sm[1][col], p2 = sm[2][col], p3 = sm[3][col];
    out[i] = make_float4((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y),
                         (p0.z + p1.z) + (p2.z + p3.z), (p0.w + p1.w) + (p2.w + p3.w));
  }
}

// measurement hook: an event to record between a backward mixer's main kernel and its partial-gradient reduction
static thread_local cudaEvent_t g_main_end_event = nullptr;

void launch_reduce_partials(const float* part, float* out, size_t n_floats, int slabs, cudaStream_t stream) {
  if (g_main_end_event) {
    cudaEventRecord(g_main_end_event, stream);
    g_main_end_event = nullptr;
  }
  const size_t n4 = n_floats / 4;
  const unsigned blocks = static_cast<unsigned>((n4 + 63) / 64);
  launch_pdl(reduce_partials_kernel, dim3(blocks), dim3(256), 0, stream, reinterpret_cast<const float4*>(part),
             reinterpret_cast<float4*>(out), n4, slabs);
}

static int check_common(const char* fn, int B, int H, int W, int E, int has_cls, int dtype) {
  if (B < 1 || H < 1 || W < 1 || E < 2 || (E & 1) || (has_cls != 0 && has_cls != 1))
    return fail(SPECTRE_E_SHAPE, "%s: bad shape B=%d H=%d W=%d E=%d has_cls=%d (need B>=1, E even)", fn, B, H, W, E,
                has_cls);
  if (dtype != SPECTRE_DTYPE_F32 && dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, dtype);
  return SPECTRE_OK;
}

}  // namespace sb200

using namespace sb200;

extern "C" {

void spectre_debug_mark_main_kernel_end(void* event) { g_main_end_event = static_cast<cudaEvent_t>(event); }
int spectre_debug_mark_pending(void) {
  const int pending = g_main_end_event != nullptr;
  g_main_end_event = nullptr;
  return pending;
}

int spectre_fftmix_fwd(const void* x, const float* filt, void* y, int B, int H, int W, int E, int has_cls, int act_dtype,
                       void* stream) {
  if (!x || !filt || !y) return fail(SPECTRE_E_NULL, "spectre_fftmix_fwd: null pointer");
  if (int rc = check_common("spectre_fftmix_fwd", B, H, W, E, has_cls, act_dtype)) return rc;
  if (!aligned16(x) || !aligned16(filt) || !aligned16(y))
    return fail(SPECTRE_E_ALIGN, "spectre_fftmix_fwd: pointers must be 16-byte aligned");
  if (act_dtype == SPECTRE_DTYPE_BF16 && ((static_cast<long long>(H) * W + has_cls) * E) % 2)
    return fail(SPECTRE_E_SHAPE, "spectre_fftmix_fwd: odd element count");
  const int rc = fftmix_fwd_dispatch(x, filt, y, B, H, W, E, has_cls, act_dtype, static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

int spectre_fftmix_bwd(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* workspace,
                       size_t workspace_bytes, int B, int H, int W, int E, int has_cls, int act_dtype, void* stream) {
  if (!dy || !x || !filt || !dx || !dfilt) return fail(SPECTRE_E_NULL, "spectre_fftmix_bwd: null pointer");
  if (int rc = check_common("spectre_fftmix_bwd", B, H, W, E, has_cls, act_dtype)) return rc;
  if (!aligned16(dy) || !aligned16(x) || !aligned16(filt) || !aligned16(dx) || !aligned16(dfilt) ||
      !aligned16(workspace))
    return fail(SPECTRE_E_ALIGN, "spectre_fftmix_bwd: pointers must be 16-byte aligned");
  const int rc = fftmix_bwd_dispatch(dy, x, filt, dx, dfilt, workspace, workspace ? workspace_bytes : 0, B, H, W, E,
                                     has_cls, act_dtype, static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

size_t spectre_fftmix_bwd_workspace_bytes(int B, int H, int W, int E) {
  if (B < 1 || E < 2 || !fftmix_supported(H, W)) return 0;
  return fftmix_bwd_workspace(B, H, W, E);
}

int spectre_dwtmix_fwd(const void* x, const float* w, void* y, int B, int H, int W, int E, int has_cls, int act_dtype,
                       void* stream) {
  if (!x || !w || !y) return fail(SPECTRE_E_NULL, "spectre_dwtmix_fwd: null pointer");
  if (int rc = check_common("spectre_dwtmix_fwd", B, H, W, E, has_cls, act_dtype)) return rc;
  if (!aligned16(x) || !aligned16(w) || !aligned16(y))
    return fail(SPECTRE_E_ALIGN, "spectre_dwtmix_fwd: pointers must be 16-byte aligned");
  const int rc = dwtmix_dispatch(false, x, nullptr, w, y, nullptr, nullptr, 0, B, H, W, E, has_cls, act_dtype,
                                 static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

int spectre_dwtmix_bwd(const void* dy, const void* x, const float* w, void* dx, float* dw, void* workspace,
                       size_t workspace_bytes, int B, int H, int W, int E, int has_cls, int act_dtype, void* stream) {
  if (!dy || !x || !w || !dx || !dw) return fail(SPECTRE_E_NULL, "spectre_dwtmix_bwd: null pointer");
  if (int rc = check_common("spectre_dwtmix_bwd", B, H, W, E, has_cls, act_dtype)) return rc;
  if (!aligned16(dy) || !aligned16(x) || !aligned16(w) || !aligned16(dx) || !aligned16(dw) || !aligned16(workspace))
    return fail(SPECTRE_E_ALIGN, "spectre_dwtmix_bwd: pointers must be 16-byte aligned");
  const int rc = dwtmix_dispatch(true, dy, x, w, dx, dw, workspace, workspace ? workspace_bytes : 0, B, H, W, E, has_cls,
                                 act_dtype, static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}

size_t spectre_dwtmix_bwd_workspace_bytes(int B, int H, int W, int E) {
  if (B < 1 || E < 2 || H < 1 || W < 1) return 0;
  return dwtmix_bwd_workspace(B, H, W, E);
}

int spectre_fftmix_supported(int H, int W) { return fftmix_supported(H, W); }
int spectre_last_launch_count(void) { return last_launch_count(); }
const char* spectre_last_error(void) { return last_error().c_str(); }
const char* spectre_version(void) { return "spectre_b200 0.1.0 sm_100a"; }

}  // extern "C"
