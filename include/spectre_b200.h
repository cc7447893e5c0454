/* spectre_b200.h -- C ABI of the B200 (sm_100a) spectral token-mixer kernels.
 *
 * This is the drop-in boundary for the reference's `mix_layer` slot
 *   spectre_vit/models/spectre/spectre.py:50   self.mix_layer = MHPermutMix(d_model, seq_length, nhead, d_model)
 *   spectre_vit/models/spectre/spectre.py:66   x = self.norm1(self.mix_layer(x)) + x
 * i.e. the arithmetic a `mix_layer.forward(x:[B,T,E]) -> [B,T,E]` occupant runs in
 * forward and in autograd backward.  Plain pointers and sizes only; no torch types.
 *
 * Common conventions
 *   - activations x, y, dy, dx: contiguous [B, T, E] with T = has_cls + H*W, E innermost,
 *     token 0 = CLS when has_cls=1 (spectre.py:151-152,182,198); CLS is passed through
 *     (y[:,0] = x[:,0], dx[:,0] = dy[:,0]).
 *   - act_dtype: 0 = float32, 1 = bfloat16.  All arithmetic is fp32; bf16 is I/O only
 *     (torch.fft up-casts bf16 under autocast: SURVEY.md section 0.4).
 *   - learned parameters and their gradients are always float32.
 *   - every pointer is DEVICE memory on the current CUDA device, 16-byte aligned;
 *     E must be even.  The caller owns every buffer (outputs and workspace included);
 *     the library never allocates, frees or synchronises, and only enqueues work on
 *     `stream` (a cudaStream_t passed as void*; NULL = legacy default stream), so the
 *     calls are CUDA-graph capturable.
 *   - return value: 0 on success; negative SPECTRE_E_* otherwise, with a message
 *     retrievable through spectre_last_error() (thread-local).  There is no CPU
 *     fallback and no fallback to torch.fft: unsupported grids are an error.
 */
#ifndef SPECTRE_B200_H_
#define SPECTRE_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPECTRE_DTYPE_F32 0
#define SPECTRE_DTYPE_BF16 1

#define SPECTRE_OK 0
#define SPECTRE_E_NULL (-1)        /* a required pointer is NULL */
#define SPECTRE_E_SHAPE (-2)       /* B/E/H/W out of range, E odd, H or W odd for DWT */
#define SPECTRE_E_GRID (-3)        /* (H, W) has no compiled kernel */
#define SPECTRE_E_DTYPE (-4)       /* act_dtype not 0/1 */
#define SPECTRE_E_ALIGN (-5)       /* pointer not 16-byte aligned */
#define SPECTRE_E_WORKSPACE (-6)   /* workspace_bytes too small */
#define SPECTRE_E_CUDA (-7)        /* launch failed; message holds cudaGetErrorString */

/* FFT mixer forward  (replaces: rfft2 -> * filter -> irfft2 in the mix_layer slot;
 * conventions of torch.fft.rfft2(..., norm="ortho") at spectre.py:136 and the
 * 2-D FFT token mixer at spectre_vit/repl/orthogonal_permut.py:23-28).
 *   y[:,1:] = irfft2( rfft2(x[:,1:] as [B,H,W,E], dims (1,2), ortho) * filt, s=(H,W), ortho )
 * filt: float32 [H, W/2+1, E, 2] (re, im interleaved). */
int spectre_fftmix_fwd(const void* x, const float* filt, void* y,
                       int B, int H, int W, int E, int has_cls, int act_dtype, void* stream);

/* FFT mixer backward (autograd of the above):
 *   dx    = irfft2( conj(filt) * rfft2(dy) )
 *   dfilt = sum_b conj(rfft2(x_b)) * (c * rfft2(dy_b)),  c[kw] = 1 for kw=0 and (W even) kw=W/2, else 2
 * dfilt: float32 [H, W/2+1, E, 2], fully overwritten (not accumulated).
 * workspace: >= spectre_fftmix_bwd_workspace_bytes(B,H,W,E) bytes, holds the per-CTA
 * partial filter gradients that a second tiny kernel sums in a fixed order
 * (deterministic; no atomics). */
int spectre_fftmix_bwd(const void* dy, const void* x, const float* filt,
                       void* dx, float* dfilt, void* workspace, size_t workspace_bytes,
                       int B, int H, int W, int E, int has_cls, int act_dtype, void* stream);
size_t spectre_fftmix_bwd_workspace_bytes(int B, int H, int W, int E);

/* FFT mixer backward that allocates nothing.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernels and results as spectre_fftmix_bwd; the scratch for the per-CTA partial filter
 * gradients is a file-scope __device__ array of SPECTRE_NOALLOC_SCRATCH_BYTES inside the library (held by every
 * GPU that loads it until the process exits) instead of a caller workspace.  Calls no cudaMalloc / cudaFree.
 * Returns SPECTRE_E_WORKSPACE when spectre_fftmix_bwd_workspace_bytes(B,H,W,E) exceeds the array (12 MiB covers
 * every compiled grid on a B200, e.g. 8.3 MB for the 14x14 ViT grids).  All calls share the array: calls on one
 * stream are ordered, calls that overlap on the GPU (other streams, independent graph branches) RACE. */
#define SPECTRE_NOALLOC_SCRATCH_BYTES (12u << 20)
int spectre_fftmix_bwd_noalloc(const void* dy, const void* x, const float* filt,
                               void* dx, float* dfilt,
                               int B, int H, int W, int E, int has_cls, int act_dtype, void* stream);

/* Haar DWT mixer forward (single-level orthonormal 2-D Haar over the token grid;
 * convention of DWTForward(wave="haar", mode="zero") at spectre_vit/repl/dwt_experiments.py:9,56;
 * mode names dwt_token/dwt_embed at spectre.py:33-34):
 *   y[:,1:] = IDWT( w * DWT(x[:,1:] as [B,H,W,E]) )
 * Odd H or W follow mode="zero": a row / column of zeros is appended before the transform and the result is
 * cropped back to H x W.  With h2 = ceil(H/2), w2 = ceil(W/2):
 * w: float32 [4, h2, w2, E], sub-band order (LL, W-low/H-high, W-high/H-low, HH). */
int spectre_dwtmix_fwd(const void* x, const float* w, void* y,
                       int B, int H, int W, int E, int has_cls, int act_dtype, void* stream);

/* Haar DWT mixer backward (row A6): dx = IDWT(w * DWT(dy)); dw = sum_b DWT(x_b) * DWT(dy_b).
 * dw: float32 [4, ceil(H/2), ceil(W/2), E], fully overwritten. */
int spectre_dwtmix_bwd(const void* dy, const void* x, const float* w,
                       void* dx, float* dw, void* workspace, size_t workspace_bytes,
                       int B, int H, int W, int E, int has_cls, int act_dtype, void* stream);
size_t spectre_dwtmix_bwd_workspace_bytes(int B, int H, int W, int E);


/* ---------------------------------------------------------------------------------------------------
 * Fused row-wise tail of the layers around the mixer:
 *
 *     out = dropout( act( LayerNorm(in0 + in1) * gamma + beta ) + adaptive_avg_pool(skip) )
 *
 * replaces, in one pass per direction,
 *   spectre_vit/models/spectre/spectre.py:66      x = self.norm1(self.mix_layer(x)) + x     (in0 = mix, skip = x)
 *   spectre_vit/models/spectre/layers.py:95-101   local_head(x) + avg_pool(x)               (in0 = Linear(x), act = GELU, skip = x)
 *   spectre_vit/models/spectre/spectre.py:67      self.norm2(x + self._ff_block(x))         (in0 = x, in1 = ff)
 * rows x N row-major.  Every activation tensor of a call (in0, in1, skip, out, dout, d_in0, d_skip) has the
 * element type act_dtype (float32, or bfloat16 for the bf16 training stream); gamma / beta, the row
 * statistics and all arithmetic are float32.
 * skip is [rows, Ns] and is pooled to N with torch.nn.AdaptiveAvgPool1d semantics (Ns == N: identity).
 * Dropout is counter-based: element (row, col) is kept iff hash(seed, row, col) >= drop_p * 2^32, so
 * the backward recomputes the mask from (seed, drop_p); drop_p = 0 disables it.
 * Backward: dout -> d_in0 (also the gradient of in1: the same values), d_skip [rows, Ns] (non-null iff the
 * forward had a skip; `skip` itself is not read), dgamma, dbeta (fully overwritten).
 * mean / rstd [rows] are written by forward and read by backward.  workspace >=
 * spectre_norm_act_bwd_workspace_bytes(rows, N) bytes (per-CTA gamma/beta partials, summed in a fixed order). */
typedef struct spectre_norm_act_t {
  const void* in0; const void* in1; const float* gamma; const float* beta; const void* skip;
  void* out; float* mean; float* rstd;
  const void* dout; const void* dout2 /* optional second addend of the upstream gradient */;
  void* d_in0; void* d_skip; float* dgamma; float* dbeta;
  float* dcolsum /* optional [N]: column sums of d_in0 = bias gradient of the Linear that produced in0 */;
  void* workspace; size_t workspace_bytes;
  long long rows; int N; int Ns; int act_dtype; int act /* 0 identity, 1 GELU(erf) */;
  float eps; float drop_p; unsigned long long seed;
  int reverse /* 0: rows swept first to last; 1: last to first (same result; lets a caller that knows which end of
                 the operands its producer left hot in L2 start there) */;
  /* Optional pooled copy of the result: out_pool[r, i] = mean(out[r, i*pool_k .. (i+1)*pool_k)), pool_k in {2, 4, 8},
   * N a multiple of 8*pool_k -- what AdaptiveAvgPool1d(N / pool_k) of `out` would give, i.e. the skip operand of the
   * NEXT SpectreLinear when it narrows N -> N / pool_k (layers.py:92,101), emitted here so that consumer reads
   * N / pool_k columns per row instead of N.  Backward: d_out_pool [rows, N / pool_k] is the gradient that arrived
   * through out_pool; it is added as d_out_pool[r, j / pool_k] / pool_k to dout[r, j].  NULL = unused. */
  void* out_pool; const void* d_out_pool; int pool_k;
  /* Optional device word added (times an odd constant) to `seed` inside the kernel.  A step captured in a CUDA graph
   * freezes the host-side seed; the graph bumps this word between replays so every replay draws a fresh mask, and the
   * backward of the same replay reads the same value.  NULL = unused. */
  const unsigned long long* seed_dev;
} spectre_norm_act_t;

int spectre_norm_act_fwd(const spectre_norm_act_t* desc, void* stream);
int spectre_norm_act_bwd(const spectre_norm_act_t* desc, void* stream);
size_t spectre_norm_act_bwd_workspace_bytes(long long rows, int N);
/* 1 if rows of width N with a skip of width Ns (pass Ns = N for none) fit the kernels in both directions
 * (N, Ns multiples of 8, N <= 4096, Ns <= 16384, shared-memory budget), else 0: callers keep the stock
 * torch composition for such widths. */
int spectre_norm_act_supported(int N, int Ns, int act_dtype);

/* Multi-level Haar DWT mixer (J = levels, the reference call site uses J=3: dwt_experiments.py:56; H and W multiples
 * of 2^levels, H*W <= 256, E a multiple of 4):  y[:,1:] = T^-1( w * T(x[:,1:]) ), T = J-level orthonormal 2-D Haar.
 * Coefficients are addressed IN PLACE on the token lattice: level l works on the 2x2 blocks (a b / c d, b next along W,
 * c next along H) of the points whose coordinates are multiples of 2^(l-1) and leaves LL at a, low-W/high-H at c,
 * high-W/low-H at b, HH at d; the next level recurses on the LL points.  w and dw: float32 [H, W, E] in that layout
 * (for levels = 1: w[0::2,0::2] = LL, w[1::2,0::2], w[0::2,1::2], w[1::2,1::2] = the three detail bands of
 * spectre_dwtmix_fwd in its order).  Backward: dx = T^-1(w * T dy), dw = sum_b T(x_b) * T(dy_b), fully overwritten. */
int spectre_dwtmix_ml_fwd(const void* x, const float* w, void* y,
                          int B, int H, int W, int E, int has_cls, int levels, int act_dtype, void* stream);
int spectre_dwtmix_ml_bwd(const void* dy, const void* x, const float* w,
                          void* dx, float* dw, void* workspace, size_t workspace_bytes,
                          int B, int H, int W, int E, int has_cls, int levels, int act_dtype, void* stream);
size_t spectre_dwtmix_ml_bwd_workspace_bytes(int B, int H, int W, int E);

/* 1-D Haar mixers "dwt_token" (axis 0) / "dwt_embed" (axis 1): the two modes the reference's SpectreEncoderLayer
 * docstring names -- "Wavelet transform in token dimension" / "in embedding dimension",
 * spectre_vit/models/spectre/spectre.py:33-34 -- without code; defined here by analogy with the 2-D Haar mixer above:
 * single-level orthonormal Haar along one axis of x[:, has_cls:, :] ([B, has_cls + N, E]), a learned real weight per
 * coefficient, inverse Haar; the CLS row passes through.
 *   axis 0: pairs of adjacent tokens per channel, w float32 [2, ceil(N/2), E] = (low, high); odd N: a zero token is
 *           appended (mode="zero") and the result cropped
 *   axis 1: pairs of adjacent channels per token, w float32 [2, N, E/2]
 * Backward: dx = Haar(w * Haar(dy)), dw = sum_b Haar(x_b) * Haar(dy_b) (fully overwritten; partial slabs in the
 * workspace, summed in a fixed order).  E a multiple of 8. */
int spectre_haar1d_fwd(const void* x, const float* w, void* y,
                       int B, int N, int E, int has_cls, int axis, int act_dtype, void* stream);
int spectre_haar1d_bwd(const void* dy, const void* x, const float* w,
                       void* dx, float* dw, void* workspace, size_t workspace_bytes,
                       int B, int N, int E, int has_cls, int axis, int act_dtype, void* stream);
size_t spectre_haar1d_bwd_workspace_bytes(int B, int N, int E, int axis);

/* ---------------------------------------------------------------------------------------------------
 * MHPermutMix gather -- the reference's HEAD occupant of the slot (spectre_vit/models/spectre/layers.py:68-70):
 *   mixed[b, h, j] = x.view(B, T*E)[b, perms[h, j]] * signs[h, j]            -> [B, heads, F], F = T*E
 * perms: int64 [heads, F] (each row a permutation, layers.py:61-64); signs: float32 [heads, F] (+-1).
 * Backward takes the inverse permutations inv[h, perms[h, j]] = j:
 *   dx[b, i] = sum_h dmixed[b, h, inv[h, i]] * signs[h, inv[h, i]]            (heads summed in a fixed order)
 * x / out / dout / dx share act_dtype; F must be a multiple of 4; all pointers 16-byte aligned. */
int spectre_permut_fwd(const void* x, const long long* perms, const float* signs, void* out,
                       int B, long long F, int heads, int act_dtype, void* stream);
int spectre_permut_bwd(const void* dout, const long long* inv_perms, const float* signs, void* dx,
                       int B, long long F, int heads, int act_dtype, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * SpectralPatchEmbed around its GEMM (spectre_vit/models/spectre/spectre.py:124-156).
 * Re(rfft2(patch, norm="ortho")) (:136), the frequency gains (:139-141) and the projection (:144) are linear in the
 * patch, so the host folds them into one matrix and the module becomes  tokens = patches @ M + b;
 * out = dropout(cat(cls, tokens) + pos).  The GEMM is the caller's; these are the passes around it.
 *
 * spectre_patch_gather: the unfold of :127-131 + the (c, i, j) flattening of :144 + the cast to the stream dtype.
 *   img float32 [B, C, Hi, Wi] -> out [B * (Hi/P) * (Wi/P), C * P * P] (act_dtype); P and Wi multiples of 4.
 * spectre_embed_finish_fwd (:151-156): out[b, 0] = cls + pos[0]; out[b, 1 + i] = tok[b, i] + pos[1 + i]; dropout
 *   (counter-based, as in spectre_norm_act_fwd).  tok [B, T-1, E] and out [B, T, E] in act_dtype, cls [E], pos [T, E] fp32.
 * spectre_embed_finish_bwd: dtok[b, i] = mask * dout[b, 1 + i] / (1 - p);  dpos[t] = sum_b mask * dout[b, t] / (1 - p)
 *   (fully overwritten; d cls = dpos[0], d bias of the projection = sum_{t >= 1} dpos[t]).  workspace >=
 *   spectre_embed_finish_bwd_workspace_bytes(B, T, E): per-slab partials, summed in a fixed order.
 *   seed_dev: optional device word added to the seed (see spectre_norm_act_t.seed_dev); NULL = unused. */
int spectre_patch_gather(const float* img, void* out, int B, int C, int Hi, int Wi, int P, int act_dtype, void* stream);
int spectre_embed_finish_fwd(const void* tok, const float* cls, const float* pos, void* out, int B, int T, int E,
                             int act_dtype, float drop_p, unsigned long long seed, const unsigned long long* seed_dev,
                             void* stream);
int spectre_embed_finish_bwd(const void* dout, void* dtok, float* dpos, void* workspace, size_t workspace_bytes,
                             int B, int T, int E, int act_dtype, float drop_p, unsigned long long seed,
                             const unsigned long long* seed_dev, void* stream);
size_t spectre_embed_finish_bwd_workspace_bytes(int B, int T, int E);

/* Measurement hook (bench.py's per-kernel roofline): the next backward mixer call on this thread records `event`
 * (a cudaEvent_t) on its stream right after its main kernel, i.e. before the small partial-gradient reduction kernel.
 * spectre_debug_mark_pending() returns 1 and disarms the hook if no call consumed it (single-slab launches have no
 * reduction kernel). */
void spectre_debug_mark_main_kernel_end(void* event);
int spectre_debug_mark_pending(void);

/* 1 if (H, W) has a compiled FFT-mixer kernel, else 0. */
int spectre_fftmix_supported(int H, int W);

/* Number of kernels the last successful call on this thread enqueued (for launch accounting). */
int spectre_last_launch_count(void);

/* Thread-local message for the last non-zero return on this thread ("" if none). */
const char* spectre_last_error(void);

/* "spectre_b200 <version> sm_100a" */
const char* spectre_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SPECTRE_B200_H_ */
