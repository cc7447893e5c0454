"""Oracle parity AT THE SHAPES bench.py TIMES (`-m gpu`).  The small cases of test_gpu_parity.py run 1-3 tiles per
CTA; at the bench batch the tensor-core mixers run 21-43 tiles per CTA (TMA ring wrap, mbarrier parity flips,
filter-gradient accumulation over the whole batch, reversed image sweep) and the row-wise tails run ~900 row groups per
CTA.  The oracle (``oracle/mixers_ref.py``: plain ``torch.fft`` calls / Haar butterflies in fp32) is evaluated on the
GPU here so the full batch is compared element by element, with a random non-Hermitian filter / random weights.

Bar: y, dx <= 3e-2 (bf16) / 1e-5 (fp32) of max|ref|; dfilt / dw <= 2e-2 (bf16) / 2e-5 (fp32).
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import mixers_ref as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    """max|a-b| / max|b| on the device (the tensors are up to 155 M elements)."""
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _chunked(fn, B, chunk):
    """Run the oracle over image chunks (keeps its complex intermediates small); concatenates / sums the results."""
    outs = [fn(slice(i, min(i + chunk, B))) for i in range(0, B, chunk)]
    return outs


FFT_BENCH = [
    # B, H, W, E, dtype                 where it is timed
    (2048, 14, 14, 384, torch.bfloat16),   # bench.py default (C4, 42-43 tiles per CTA in the backward)
    (1024, 14, 14, 384, torch.bfloat16),   # profiles/*b1024*
    (256, 14, 14, 384, torch.bfloat16),    # SURVEY's batch (configs/vit_s16_fft.py:batch_size)
    (512, 14, 14, 768, torch.bfloat16),    # C5 bench batch
    (4096, 7, 7, 48, torch.float32),       # C2 bandwidth shape
    (1024, 8, 8, 512, torch.float32),      # C1 widths at the kernel-bench batch
    (1024, 8, 8, 512, torch.bfloat16),     # 8x8 tensor-core path
]


@pytest.mark.parametrize("B,H,W,E,dtype", FFT_BENCH)
def test_fftmix_bench_shape(cuda_device, B, H, W, E, dtype):
    from vit_spectre_experiments_b200 import fft_mix

    g = torch.Generator(device="cpu").manual_seed(B + E)
    T = 1 + H * W
    filt = torch.randn(H, W // 2 + 1, E, 2, generator=g).to(cuda_device)
    gd = torch.Generator(device=cuda_device).manual_seed(B * 3 + E)
    x = torch.randn(B, T, E, device=cuda_device, generator=gd).to(dtype)
    dy = torch.randn(B, T, E, device=cuda_device, generator=gd).to(dtype)
    xd, fd = x.clone().requires_grad_(True), filt.clone().requires_grad_(True)
    y = fft_mix(xd, fd, (H, W))
    y.backward(dy)
    bf = dtype == torch.bfloat16
    tol, ptol = (3e-2, 2e-2) if bf else (1e-5, 2e-5)

    chunk = 256
    y_ref = torch.cat(_chunked(lambda s: O.fft_mix_ref(x[s].float(), filt, H, W).to(dtype), B, chunk))
    assert rel_err(y, y_ref) <= tol
    del y_ref
    parts = _chunked(lambda s: O.fft_mix_bwd_ref(dy[s].float(), x[s].float(), filt, H, W), B, chunk)
    dx_ref = torch.cat([p[0].to(dtype) for p in parts])
    df_ref = torch.stack([p[1].double() for p in parts]).sum(0)
    assert rel_err(xd.grad, dx_ref) <= tol
    assert (fd.grad.double() - df_ref).abs().max().item() <= ptol * df_ref.abs().max().item()
    assert torch.equal(y[:, 0], x[:, 0]) and torch.equal(xd.grad[:, 0], dy[:, 0])
    # determinism of the fused filter-gradient reduction at this batch
    xd2, fd2 = x.clone().requires_grad_(True), filt.clone().requires_grad_(True)
    fft_mix(xd2, fd2, (H, W)).backward(dy)
    assert torch.equal(fd.grad, fd2.grad) and torch.equal(xd.grad, xd2.grad)


@pytest.mark.parametrize("B,H,W,E,dtype", [
    (1024, 8, 8, 512, torch.bfloat16),     # C3 at its full shape, RANDOM weights
    (1024, 8, 8, 512, torch.float32),
    (2048, 14, 14, 384, torch.bfloat16),
])
def test_dwtmix_bench_shape(cuda_device, B, H, W, E, dtype):
    from vit_spectre_experiments_b200 import dwt_mix

    gd = torch.Generator(device=cuda_device).manual_seed(B + E + 1)
    T = 1 + H * W
    x = torch.randn(B, T, E, device=cuda_device, generator=gd).to(dtype)
    dy = torch.randn(B, T, E, device=cuda_device, generator=gd).to(dtype)
    w = 1.0 + 0.5 * torch.randn(4, H // 2, W // 2, E, device=cuda_device, generator=gd)
    xd, wd = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y = dwt_mix(xd, wd, (H, W))
    y.backward(dy)
    bf = dtype == torch.bfloat16
    tol, ptol = (3e-2, 2e-3) if bf else (1e-5, 2e-5)
    assert rel_err(y, O.dwt_mix_ref(x.float(), w, H, W).to(dtype)) <= tol
    dx_ref, dw_ref = O.dwt_mix_bwd_ref(dy.double(), x.double(), w.double(), H, W)
    assert rel_err(xd.grad, dx_ref.float().to(dtype)) <= tol
    assert (wd.grad.double() - dw_ref).abs().max().item() <= ptol * dw_ref.abs().max().item()


def _ref_tail(in0, gamma, beta, in1, skip, act):
    z = in0 if in1 is None else in0 + in1
    u = F.layer_norm(z, (z.shape[-1],), gamma, beta, 1e-5)
    if act == "gelu":
        u = F.gelu(u)
    if skip is not None:
        N = z.shape[-1]
        u = u + (skip if skip.shape[-1] == N else F.adaptive_avg_pool1d(skip, N))
    return u


@pytest.mark.parametrize("rows,N,Ns,has_in1,act", [
    (2048 * 197, 1536, 384, False, "gelu"),   # GELU(LN(h1)) + pool_up(x1) at the bench batch: rows = 403,456
    (2048 * 197, 384, 1536, False, "gelu"),   # GELU(LN(h2)) + pool_down(a1)
    (2048 * 197, 384, 384, False, "none"),    # LN1(mix) + x
    (2048 * 197, 384, None, True, "none"),    # LN2(x1 + ff)
    (1024 * 65, 768, 512, False, "gelu"),     # C3 general widths at its batch
    (1024 * 65, 512, 768, False, "gelu"),
])
def test_norm_act_bench_rows(cuda_device, rows, N, Ns, has_in1, act):
    """The row-wise tails at the row counts of the timed steps (bf16 stream), vs the fp32 torch composition on the GPU."""
    from vit_spectre_experiments_b200.fused import norm_act

    dt = torch.bfloat16
    gd = torch.Generator(device=cuda_device).manual_seed(rows % 1000 + N)

    def rnd(*s):
        return torch.randn(*s, device=cuda_device, generator=gd)

    in0 = rnd(rows, N).to(dt)
    in1 = rnd(rows, N).to(dt) if has_in1 else None
    skip = rnd(rows, Ns).to(dt) if Ns else None
    gamma, beta = 1 + 0.3 * rnd(N), 0.2 * rnd(N)
    dout = rnd(rows, N).to(dt)
    leaves = [t.clone().requires_grad_(True) if t is not None else None for t in (in0, gamma, beta, in1, skip)]
    y = norm_act(leaves[0], leaves[1], leaves[2], in1=leaves[3], skip=leaves[4], act=act)
    y.backward(dout)

    # reference in row chunks (fp32 on the bf16-rounded inputs); parameter gradients accumulated in float64
    dg = torch.zeros(N, dtype=torch.float64, device=cuda_device)
    db = torch.zeros_like(dg)
    worst = {"y": 0.0, "d_in0": 0.0, "d_skip": 0.0}
    scale = {"y": 0.0, "d_in0": 0.0, "d_skip": 0.0}
    chunk = 32768
    for i in range(0, rows, chunk):
        s = slice(i, min(i + chunk, rows))
        r = [None if t is None else t[s].float().requires_grad_(True) for t in (in0, in1, skip)]
        gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        yr = _ref_tail(r[0], gr, br, r[1], r[2], act)
        (yr * dout[s].float()).sum().backward()
        dg += gr.grad.double()
        db += br.grad.double()
        for name, a, b in (("y", y[s], yr.detach()), ("d_in0", leaves[0].grad[s], r[0].grad),
                           ("d_skip", None if skip is None else leaves[4].grad[s], None if skip is None else r[2].grad)):
            if a is None:
                continue
            worst[name] = max(worst[name], (a.float() - b).abs().max().item())
            scale[name] = max(scale[name], b.abs().max().item())
    for name in worst:
        if scale[name] > 0:
            assert worst[name] <= 1e-2 * scale[name], (name, worst[name], scale[name])
    if has_in1:
        assert torch.equal(leaves[3].grad, leaves[0].grad)
    # fp32 parameter gradients summed over 4e5 bf16 rows: the rounding of d_in0's inputs is shared with the reference,
    # what is left is fp32 summation order
    ptol = 2e-3 if act == "gelu" else 2e-4     # bf16 stream: Phi through MUFU.TANH (relative error 2^-11)
    assert (leaves[1].grad.double() - dg).abs().max().item() <= ptol * dg.abs().max().item()
    assert (leaves[2].grad.double() - db).abs().max().item() <= ptol * db.abs().max().item()
