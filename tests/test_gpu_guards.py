"""Out-of-bounds write check without compute-sanitizer (closed on this pool: profiles/r02_sanitizer_unavailable.txt).

Every output and workspace of a C-ABI call is carved out of a larger allocation; the guard zones on both sides hold a
sentinel pattern and must be bit-identical after the call.  Ragged shapes (last row group / image group / channel tile
partially filled) are the ones where a stray store would land in a guard.  `-m gpu`."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096          # bytes on each side
SENT = 0x5A


class Guarded:
    """Device buffers with sentinel guard zones; ``check()`` asserts no guard byte changed."""

    def __init__(self, dev):
        self.dev, self.bufs = dev, []

    def alloc(self, shape, dtype):
        n = 1
        for s in shape:
            n *= s
        nbytes = max(n, 1) * torch.empty((), dtype=dtype).element_size()
        pad = (-nbytes) % 256
        raw = torch.full((GUARD + nbytes + pad + GUARD,), SENT, dtype=torch.uint8, device=self.dev)
        self.bufs.append((raw, nbytes))
        return raw[GUARD:GUARD + nbytes].view(dtype).view(*shape)

    def check(self):
        torch.cuda.synchronize()
        for raw, nbytes in self.bufs:
            assert bool((raw[:GUARD] == SENT).all()), "write before the buffer"
            assert bool((raw[GUARD + nbytes:] == SENT).all()), "write past the buffer"


def _st(dev):
    return torch.cuda.current_stream(dev).cuda_stream


@pytest.mark.parametrize("B,H,W,E,dt", [(5, 14, 14, 384, torch.bfloat16), (3, 14, 14, 200, torch.bfloat16),
                                        (7, 8, 8, 512, torch.bfloat16), (9, 7, 7, 48, torch.float32),
                                        (3, 8, 8, 24, torch.float32), (301, 14, 14, 64, torch.bfloat16),
                                        # TMA-fed fp32 kernels: ragged image groups (images past B are zero-filled
                                        # on load and must not be stored), one- and two-channel lanes
                                        (33, 8, 8, 96, torch.float32), (5, 14, 14, 384, torch.float32),
                                        (7, 4, 4, 32, torch.float32), (3, 16, 16, 16, torch.float32)])
def test_fftmix_guards(cuda_device, B, H, W, E, dt):
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    T, code = H * W + 1, 0 if dt == torch.float32 else 1
    x = torch.randn(B, T, E, device=cuda_device).to(dt)
    dy = torch.randn(B, T, E, device=cuda_device).to(dt)
    f = torch.randn(H, W // 2 + 1, E, 2, device=cuda_device)
    y, dx = g.alloc((B, T, E), dt), g.alloc((B, T, E), dt)
    df = g.alloc((H, W // 2 + 1, E, 2), torch.float32)
    nws = lib.spectre_fftmix_bwd_workspace_bytes(B, H, W, E)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    _lib.check(lib.spectre_fftmix_fwd(x.data_ptr(), f.data_ptr(), y.data_ptr(), B, H, W, E, 1, code, _st(cuda_device)), "fwd")
    _lib.check(lib.spectre_fftmix_bwd(dy.data_ptr(), x.data_ptr(), f.data_ptr(), dx.data_ptr(), df.data_ptr(), ws.data_ptr(),
                                      nws, B, H, W, E, 1, code, _st(cuda_device)), "bwd")
    g.check()
    assert torch.isfinite(y.float()).all() and torch.isfinite(dx.float()).all() and torch.isfinite(df).all()


@pytest.mark.parametrize("B,H,W,E,dt", [(33, 8, 8, 512, torch.bfloat16), (5, 14, 14, 384, torch.bfloat16),
                                        (9, 7, 7, 48, torch.float32), (1, 2, 2, 8, torch.bfloat16), (4, 8, 8, 512, torch.float32)])
def test_dwtmix_guards(cuda_device, B, H, W, E, dt):
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    T, code = H * W + 1, 0 if dt == torch.float32 else 1
    x = torch.randn(B, T, E, device=cuda_device).to(dt)
    dy = torch.randn(B, T, E, device=cuda_device).to(dt)
    w = torch.randn(4, (H + 1) // 2, (W + 1) // 2, E, device=cuda_device)
    y, dx = g.alloc((B, T, E), dt), g.alloc((B, T, E), dt)
    dw = g.alloc(tuple(w.shape), torch.float32)
    nws = lib.spectre_dwtmix_bwd_workspace_bytes(B, H, W, E)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    _lib.check(lib.spectre_dwtmix_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, H, W, E, 1, code, _st(cuda_device)), "fwd")
    _lib.check(lib.spectre_dwtmix_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(), ws.data_ptr(),
                                      nws, B, H, W, E, 1, code, _st(cuda_device)), "bwd")
    g.check()
    assert torch.isfinite(y.float()).all() and torch.isfinite(dx.float()).all() and torch.isfinite(dw).all()


@pytest.mark.parametrize("B,H,E,levels,dt", [(33, 8, 128, 3, torch.bfloat16), (5, 8, 64, 2, torch.float32),
                                             (7, 4, 64, 2, torch.bfloat16), (3, 4, 32, 1, torch.float32),
                                             (3, 8, 48, 3, torch.float32)])
def test_dwtmix_ml_guards(cuda_device, B, H, E, levels, dt):
    """Multi-level Haar: register / TMA kernel (E a multiple of the 128-byte tile) and the shared-memory kernel (E = 48)."""
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    W, T, code = H, H * H + 1, 0 if dt == torch.float32 else 1
    x = torch.randn(B, T, E, device=cuda_device).to(dt)
    dy = torch.randn(B, T, E, device=cuda_device).to(dt)
    w = torch.randn(H, W, E, device=cuda_device)
    y, dx = g.alloc((B, T, E), dt), g.alloc((B, T, E), dt)
    dw = g.alloc((H, W, E), torch.float32)
    nws = lib.spectre_dwtmix_ml_bwd_workspace_bytes(B, H, W, E)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    _lib.check(lib.spectre_dwtmix_ml_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, H, W, E, 1, levels, code,
                                         _st(cuda_device)), "fwd")
    _lib.check(lib.spectre_dwtmix_ml_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(), ws.data_ptr(),
                                         nws, B, H, W, E, 1, levels, code, _st(cuda_device)), "bwd")
    g.check()
    assert torch.isfinite(y.float()).all() and torch.isfinite(dx.float()).all() and torch.isfinite(dw).all()


@pytest.mark.parametrize("B,N,E,dt", [(33, 49, 48, torch.float32), (5, 196, 384, torch.bfloat16), (300, 7, 8, torch.bfloat16)])
@pytest.mark.parametrize("axis", [0, 1])
def test_haar1d_guards(cuda_device, B, N, E, dt, axis):
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    T, code = N + 1, 0 if dt == torch.float32 else 1
    x = torch.randn(B, T, E, device=cuda_device).to(dt)
    dy = torch.randn(B, T, E, device=cuda_device).to(dt)
    wshape = (2, (N + 1) // 2, E) if axis == 0 else (2, N, E // 2)
    w = torch.randn(*wshape, device=cuda_device)
    y, dx = g.alloc((B, T, E), dt), g.alloc((B, T, E), dt)
    dw = g.alloc(wshape, torch.float32)
    nws = lib.spectre_haar1d_bwd_workspace_bytes(B, N, E, axis)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    _lib.check(lib.spectre_haar1d_fwd(x.data_ptr(), w.data_ptr(), y.data_ptr(), B, N, E, 1, axis, code, _st(cuda_device)), "fwd")
    _lib.check(lib.spectre_haar1d_bwd(dy.data_ptr(), x.data_ptr(), w.data_ptr(), dx.data_ptr(), dw.data_ptr(), ws.data_ptr(),
                                      nws, B, N, E, 1, axis, code, _st(cuda_device)), "bwd")
    g.check()
    assert torch.isfinite(y.float()).all() and torch.isfinite(dx.float()).all() and torch.isfinite(dw).all()


@pytest.mark.parametrize("rows,N,Ns,dt,pool_k", [(1003, 1536, 384, torch.bfloat16, 4), (1003, 384, 384, torch.bfloat16, 0),
                                                 (77, 768, 512, torch.bfloat16, 0), (50, 48, 256, torch.float32, 0),
                                                 (403, 3072, 768, torch.bfloat16, 4), (19, 512, 512, torch.float32, 0)])
def test_norm_act_guards(cuda_device, rows, N, Ns, dt, pool_k):
    """Ragged last row group (rows is not a multiple of the rows per CTA) on the fast and the generic kernels."""
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    code = 0 if dt == torch.float32 else 1
    rnd = lambda *s: torch.randn(*s, device=cuda_device)   # noqa: E731
    in0, skip, dout = rnd(rows, N).to(dt), rnd(rows, Ns).to(dt), rnd(rows, N).to(dt)
    gamma, beta = 1 + 0.1 * rnd(N), 0.1 * rnd(N)
    out, d_in0, d_skip = g.alloc((rows, N), dt), g.alloc((rows, N), dt), g.alloc((rows, Ns), dt)
    mean, rstd = g.alloc((rows,), torch.float32), g.alloc((rows,), torch.float32)
    dg, db, dc = g.alloc((N,), torch.float32), g.alloc((N,), torch.float32), g.alloc((N,), torch.float32)
    pooled = g.alloc((rows, N // pool_k), dt) if pool_k else None
    dpool = rnd(rows, N // pool_k).to(dt) if pool_k else None
    nws = lib.spectre_norm_act_bwd_workspace_bytes(rows, N)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    d = _lib.NormActDesc()
    d.in0, d.gamma, d.beta, d.skip, d.out, d.mean, d.rstd = (in0.data_ptr(), gamma.data_ptr(), beta.data_ptr(), skip.data_ptr(),
                                                            out.data_ptr(), mean.data_ptr(), rstd.data_ptr())
    d.rows, d.N, d.Ns, d.act_dtype, d.act, d.eps, d.drop_p, d.seed = rows, N, Ns, code, 1, 1e-5, 0.01, 5
    if pool_k:
        d.out_pool, d.pool_k = pooled.data_ptr(), pool_k
    _lib.check(lib.spectre_norm_act_fwd(ctypes.byref(d), _st(cuda_device)), "fwd")
    d.dout, d.d_in0, d.d_skip, d.dgamma, d.dbeta, d.dcolsum = (dout.data_ptr(), d_in0.data_ptr(), d_skip.data_ptr(), dg.data_ptr(),
                                                               db.data_ptr(), dc.data_ptr())
    d.workspace, d.workspace_bytes = ws.data_ptr(), nws
    if pool_k:
        d.d_out_pool = dpool.data_ptr()
    _lib.check(lib.spectre_norm_act_bwd(ctypes.byref(d), _st(cuda_device)), "bwd")
    g.check()
    for t in (out, d_in0, d_skip, mean, rstd, dg, db, dc):
        assert torch.isfinite(t.float()).all()


def test_embed_guards(cuda_device):
    from vit_spectre_experiments_b200 import _lib

    lib, g = _lib.load(), Guarded(cuda_device)
    B, C, P, E, n = 5, 3, 4, 48, 64
    img = torch.randn(B, C, 32, 32, device=cuda_device)     # 32x32 -> 8x8 patches of 4x4
    A = g.alloc((B * n, C * P * P), torch.bfloat16)
    _lib.check(lib.spectre_patch_gather(img.data_ptr(), A.data_ptr(), B, C, 32, 32, P, 1, _st(cuda_device)), "gather")
    tok = torch.randn(B * n, E, device=cuda_device).bfloat16()
    cls, pos = torch.randn(E, device=cuda_device), torch.randn(n + 1, E, device=cuda_device)
    out = g.alloc((B, n + 1, E), torch.bfloat16)
    _lib.check(lib.spectre_embed_finish_fwd(tok.data_ptr(), cls.data_ptr(), pos.data_ptr(), out.data_ptr(), B, n + 1, E, 1, 0.1, 3,
                                            None, _st(cuda_device)), "finish_fwd")
    dout = torch.randn(B, n + 1, E, device=cuda_device).bfloat16()
    dtok, dpos = g.alloc((B * n, E), torch.bfloat16), g.alloc((n + 1, E), torch.float32)
    nws = lib.spectre_embed_finish_bwd_workspace_bytes(B, n + 1, E)
    ws = g.alloc((max(nws, 16),), torch.uint8)
    _lib.check(lib.spectre_embed_finish_bwd(dout.data_ptr(), dtok.data_ptr(), dpos.data_ptr(), ws.data_ptr(), nws, B, n + 1, E, 1,
                                            0.1, 3, None, _st(cuda_device)), "finish_bwd")
    g.check()
    ref = img.reshape(B, C, 8, P, 8, P).permute(0, 2, 4, 1, 3, 5).reshape(B * n, C * P * P).bfloat16()
    assert torch.equal(A, ref)


def test_tma_kernels_from_a_thread_without_a_cuda_context(cuda_device):
    """The TMA-fed kernels encode their tensor maps with the driver API, which needs a context current on the calling
    thread; a thread that has made no runtime-API call yet (an autograd worker, a serving thread) has none.  The C ABI
    binds the primary context itself (`bind_primary_context`): calls made from a brand-new thread must succeed even
    when everything the launcher would query is already cached."""
    import threading

    from vit_spectre_experiments_b200 import _lib

    lib = _lib.load()
    dev = cuda_device
    B, E = 6, 64
    jobs = []
    for (H, dt) in ((8, torch.float32), (14, torch.bfloat16), (8, torch.bfloat16)):
        T, code = H * H + 1, 0 if dt == torch.float32 else 1
        x = torch.randn(B, T, E, device=dev).to(dt)
        f = torch.randn(H, H // 2 + 1, E, 2, device=dev)
        w = torch.randn(H, H, E, device=dev)
        y, y2 = torch.empty_like(x), torch.empty_like(x)
        jobs.append((H, code, x, f, w, y, y2))
    st = _st(dev)

    def run(results):
        for (H, code, x, f, w, y, y2) in jobs:
            results.append(lib.spectre_fftmix_fwd(x.data_ptr(), f.data_ptr(), y.data_ptr(), B, H, H, E, 1, code, st))
            if H == 8:
                results.append(lib.spectre_dwtmix_ml_fwd(x.data_ptr(), w.data_ptr(), y2.data_ptr(), B, H, H, E, 1, 3, code, st))

    warm = []
    run(warm)                       # main thread: fills the per-kernel caches
    torch.cuda.synchronize()
    ref = [(j[5].clone(), j[6].clone()) for j in jobs]
    for j in jobs:
        j[5].zero_(); j[6].zero_()
    torch.cuda.synchronize()
    res = []
    t = threading.Thread(target=run, args=(res,))
    t.start(); t.join()
    torch.cuda.synchronize()
    assert warm and all(rc == 0 for rc in warm), warm
    assert res and all(rc == 0 for rc in res), (res, lib.spectre_last_error().decode())
    for j, (ya, yb) in zip(jobs, ref):
        assert torch.equal(j[5], ya)
        if j[0] == 8:
            assert torch.equal(j[6], yb)
