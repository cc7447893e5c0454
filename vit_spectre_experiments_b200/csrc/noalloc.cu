// spectre_fftmix_bwd_noalloc: the FFT mixer backward for callers that cannot allocate device memory
// (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  Same operator, kernels and
// results as spectre_fftmix_bwd (capi.cu) -- the only difference is where the scratch for the per-CTA partial
// filter gradients lives: a file-scope __device__ array of SPECTRE_NOALLOC_SCRATCH_BYTES instead of a
// caller-provided workspace.  Every GPU that loads the library holds that array until the process exits.
//
// Rules this file keeps:
//   * host code never passes `g_scratch` itself to a kernel (in host code the symbol is a host-side stand-in,
//     not a device address): its device address is taken ONCE PER DEVICE, with that device current, by
//     cudaGetSymbolAddress, the return code is checked, and the resulting pointer is what the launchers get;
//   * no cudaMalloc / cudaFree / allocating library routine on the call path;
//   * the array is shared by every call and keeps the last call's contents.  Nothing in it has to start at
//     zero: each launch fully overwrites the slabs its own reduction reads before reading them, so there is
//     nothing to clear -- but two calls that overlap on the GPU (different streams, or graph branches without a
//     dependency) race on it.  Calls on one stream are ordered and safe.
#include <mutex>

#include "common.cuh"

namespace sb200 {

int fftmix_supported(int H, int W);
int fftmix_bwd_dispatch(const void* dy, const void* x, const float* filt, void* dx, float* dfilt, void* ws,
                        size_t ws_bytes, int B, int H, int W, int E, int cls, int dtype, cudaStream_t st);
size_t fftmix_bwd_workspace(int B, int H, int W, int E);

namespace {

static __device__ __align__(256) unsigned char g_scratch[SPECTRE_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;

// Device address of g_scratch on the CURRENT device (each device has its own instance of the array).
int scratch_address(void** out) {
  static std::mutex mu;
  static void* addr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return fail(SPECTRE_E_CUDA, "spectre_fftmix_bwd_noalloc: cudaGetDevice: %s", cudaGetErrorString(err));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(SPECTRE_E_CUDA, "spectre_fftmix_bwd_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    err = cudaGetSymbolAddress(&p, g_scratch);
    if (err != cudaSuccess || !p)
      return fail(SPECTRE_E_CUDA, "spectre_fftmix_bwd_noalloc: cudaGetSymbolAddress(device %d): %s", dev,
                  cudaGetErrorString(err));
    addr[dev] = p;
  }
  *out = addr[dev];
  return SPECTRE_OK;
}

}  // namespace
}  // namespace sb200

using namespace sb200;

extern "C" int spectre_fftmix_bwd_noalloc(const void* dy, const void* x, const float* filt, void* dx, float* dfilt,
                                          int B, int H, int W, int E, int has_cls, int act_dtype, void* stream) {
  const char* fn = "spectre_fftmix_bwd_noalloc";
  if (!dy || !x || !filt || !dx || !dfilt) return fail(SPECTRE_E_NULL, "%s: null pointer", fn);
  if (B < 1 || H < 1 || W < 1 || E < 2 || (E & 1) || (has_cls != 0 && has_cls != 1))
    return fail(SPECTRE_E_SHAPE, "%s: bad shape B=%d H=%d W=%d E=%d has_cls=%d (need B>=1, E even)", fn, B, H, W, E,
                has_cls);
  if (act_dtype != SPECTRE_DTYPE_F32 && act_dtype != SPECTRE_DTYPE_BF16)
    return fail(SPECTRE_E_DTYPE, "%s: act_dtype=%d (0=float32, 1=bfloat16)", fn, act_dtype);
  if (!aligned16(dy) || !aligned16(x) || !aligned16(filt) || !aligned16(dx) || !aligned16(dfilt))
    return fail(SPECTRE_E_ALIGN, "%s: pointers must be 16-byte aligned", fn);
  if (!fftmix_supported(H, W)) return fail(SPECTRE_E_GRID, "%s: no kernel compiled for grid %dx%d", fn, H, W);
  const size_t need = fftmix_bwd_workspace(B, H, W, E);
  if (need > static_cast<size_t>(SPECTRE_NOALLOC_SCRATCH_BYTES))
    return fail(SPECTRE_E_WORKSPACE,
                "%s: B=%d grid %dx%d E=%d needs %zu bytes of scratch, the built-in array holds %zu "
                "(SPECTRE_NOALLOC_SCRATCH_BYTES); use spectre_fftmix_bwd with a caller-provided workspace",
                fn, B, H, W, E, need, static_cast<size_t>(SPECTRE_NOALLOC_SCRATCH_BYTES));
  void* scratch = nullptr;
  if (int rc = scratch_address(&scratch)) return rc;
  const int rc = fftmix_bwd_dispatch(dy, x, filt, dx, dfilt, scratch, SPECTRE_NOALLOC_SCRATCH_BYTES, B, H, W, E, has_cls,
                                     act_dtype, static_cast<cudaStream_t>(stream));
  if (rc == SPECTRE_OK) last_error().clear();
  return rc;
}
