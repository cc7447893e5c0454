"""Parity of the fused row-wise tail kernels (C ABI ``spectre_norm_act_*``)
against the stock torch composition they replace (``spectre.py:65-73``, ``layers.py:95-101``).  `-m gpu` only."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def ref_norm_act(in0, gamma, beta, in1, skip, act, eps=1e-5):
    z = in0 if in1 is None else in0 + in1
    u = F.layer_norm(z, (z.shape[-1],), gamma, beta, eps)
    if act == "gelu":
        u = F.gelu(u)
    if skip is not None:
        N = z.shape[-1]
        u = u + (skip if skip.shape[-1] == N else F.adaptive_avg_pool1d(skip, N))
    return u


CASES = [
    # rows-shape, N, Ns, dtype, in1, act      which line of the reference layer
    ((6, 197), 384, 384, torch.bfloat16, False, "none"),    # x1 = LN1(mix) + x            (spectre.py:66)
    ((6, 197), 1536, 384, torch.bfloat16, False, "gelu"),   # GELU(LN(h1)) + pool_up(x1)    (layers.py:95-101)
    ((6, 197), 384, 1536, torch.bfloat16, False, "gelu"),   # GELU(LN(h2)) + pool_down(a1)
    ((6, 197), 384, None, torch.bfloat16, True, "none"),    # x2 = LN2(x1 + ff)            (spectre.py:67)
    ((6, 197), 384, 384, torch.float32, False, "none"),     # the same four in float32
    ((6, 197), 1536, 384, torch.float32, False, "gelu"),
    ((6, 197), 384, 1536, torch.float32, False, "gelu"),
    ((6, 197), 384, None, torch.float32, True, "none"),
    ((3, 65), 768, 512, torch.float32, False, "gelu"),      # C1 widths: general adaptive pool 512 -> 768
    ((3, 65), 512, 768, torch.float32, False, "gelu"),      #            and 768 -> 512
    ((3, 65), 768, 512, torch.bfloat16, False, "gelu"),     # C3 (bf16 stream): the fast kernels' 3:2 modes
    ((3, 65), 512, 768, torch.bfloat16, False, "gelu"),
    ((41, 7), 768, 512, torch.bfloat16, False, "none"),     # 3:2 up without activation (forward fast, backward generic)
    ((41, 7), 1536, 1024, torch.bfloat16, False, "gelu"),   # other widths in ratio 3:2
    ((41, 7), 1024, 1536, torch.bfloat16, True, "gelu"),
    ((5, 50), 48, 48, torch.float32, False, "none"),        # C2 width 48 (single partial warp)
    ((5, 50), 256, 48, torch.float32, False, "gelu"),       # 48 -> 256 (general)
    ((5, 50), 48, 256, torch.float32, False, "gelu"),       # 256 -> 48 (general)
    ((2, 197), 3072, 768, torch.bfloat16, False, "gelu"),   # ViT-B widths
    ((2, 197), 768, 3072, torch.bfloat16, False, "gelu"),
    ((1, 37), 4096, 512, torch.float32, False, "gelu"),     # widest row (one CTA per row), ratio 8
    ((1, 37), 512, 4096, torch.bfloat16, False, "none"),
    ((1, 1), 128, 256, torch.float32, True, "gelu"),        # single row, ratio 2 down, with in1
    ((7, 11), 256, 128, torch.bfloat16, True, "none"),      # ratio 2 up, with in1
]


@pytest.mark.parametrize("shape,N,Ns,dtype,has_in1,act", CASES)
def test_norm_act_fwd_bwd(cuda_device, shape, N, Ns, dtype, has_in1, act):
    """Tolerances (relative to the max magnitude of the reference tensor): float32 5e-6 forward / 2e-5 backward
    (erf by the A&S 7.1.26 rational, |err| < 2e-7); bfloat16 1e-2 (output rounding only)."""
    from vit_spectre_experiments_b200.fused import norm_act

    g = torch.Generator().manual_seed(N * 7 + (Ns or 0))
    in0 = torch.randn(*shape, N, generator=g).to(dtype)
    in1 = torch.randn(*shape, N, generator=g).to(dtype) if has_in1 else None
    skip = torch.randn(*shape, Ns, generator=g).to(dtype) if Ns else None
    gamma = 1 + 0.3 * torch.randn(N, generator=g)
    beta = 0.2 * torch.randn(N, generator=g)
    dout = torch.randn(*shape, N, generator=g).to(dtype)

    def leaf(t):
        return None if t is None else t.clone().requires_grad_(True)

    # reference (fp32 torch on the same, dtype-rounded, inputs)
    r = [leaf(None if t is None else t.float()) for t in (in0, gamma, beta, in1, skip)]
    y_ref = ref_norm_act(r[0], r[1], r[2], r[3], r[4], act)
    (y_ref * dout.float()).sum().backward()

    d = [leaf(None if t is None else t.to(cuda_device)) for t in (in0, gamma, beta, in1, skip)]
    y = norm_act(d[0], d[1], d[2], in1=d[3], skip=d[4], act=act)
    bf = dtype == torch.bfloat16
    assert y.dtype == dtype and rel_err(y, y_ref) <= (1e-2 if bf else 5e-6)
    y.backward(dout.to(cuda_device))
    tol = 1e-2 if bf else 2e-5
    assert d[0].grad.dtype == dtype and rel_err(d[0].grad, r[0].grad) <= tol
    # fp32 parameter gradients; the bf16 stream evaluates Phi through MUFU.TANH (relative error 2^-11)
    ptol = 2e-3 if (bf and act == "gelu") else 3e-5
    assert rel_err(d[1].grad, r[1].grad) <= ptol and rel_err(d[2].grad, r[2].grad) <= ptol
    if has_in1:
        assert rel_err(d[3].grad, r[3].grad) <= tol
    if Ns:
        assert rel_err(d[4].grad, r[4].grad) <= tol


def test_norm_act_large_rows_deterministic(cuda_device):
    """Many rows (several row groups per CTA, partial last group): gamma/beta gradients are bit-identical
    between runs (fixed-order partial sums, no atomics) and match torch."""
    from vit_spectre_experiments_b200.fused import norm_act

    torch.manual_seed(5)
    x = torch.randn(20011, 384, device=cuda_device)
    gamma = (1 + 0.1 * torch.randn(384, device=cuda_device)).requires_grad_(True)
    beta = torch.zeros(384, device=cuda_device, requires_grad=True)
    w = torch.randn_like(x)
    grads = []
    for _ in range(2):
        gamma.grad = beta.grad = None
        (norm_act(x, gamma, beta, act="gelu") * w).sum().backward()
        grads.append((gamma.grad.clone(), beta.grad.clone()))
    assert torch.equal(grads[0][0], grads[1][0]) and torch.equal(grads[0][1], grads[1][1])
    g2, b2 = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    (F.gelu(F.layer_norm(x, (384,), g2, b2)) * w).sum().backward()
    assert rel_err(grads[0][0], g2.grad) <= 3e-5 and rel_err(grads[0][1], b2.grad) <= 3e-5


def test_norm_act_dropout_is_consistent(cuda_device):
    """Counter-based dropout: forward and backward use the same mask; keep-rate ~ 1 - p; scaling 1/(1-p)."""
    from vit_spectre_experiments_b200.fused import norm_act

    x = torch.randn(64, 197, 384, device=cuda_device).requires_grad_(True)
    gamma, beta = torch.ones(384, device=cuda_device), torch.zeros(384, device=cuda_device)
    y0 = norm_act(x, gamma, beta)
    y = norm_act(x, gamma, beta, drop_p=0.25, seed=1234)
    kept = y != 0
    assert abs(kept.float().mean().item() - 0.75) < 5e-3
    assert torch.allclose(y[kept], y0[kept] / 0.75, rtol=1e-6, atol=1e-6)
    y2 = norm_act(x, gamma, beta, drop_p=0.25, seed=1234)
    assert torch.equal(y, y2)
    y3 = norm_act(x, gamma, beta, drop_p=0.25, seed=1235)
    assert not torch.equal(y, y3)
    # small p (the configs use 0.001): keep-rate still right
    y4 = norm_act(x, gamma, beta, drop_p=0.001, seed=7)
    assert abs((y4 == 0).float().mean().item() - 0.001) < 2e-4
    # gradient flows only through kept elements: d/dx of sum(y * w) with w = 1 on dropped positions is 0
    (y * (~kept).float()).sum().backward()
    assert x.grad.abs().max().item() == 0.0


def test_norm_act_bwd_second_addend_and_colsum(cuda_device):
    """Backward extras used by the one-node encoder layer: ``dout2`` is added to ``dout`` inside the kernel and
    ``dcolsum`` = column sums of ``d_in0`` (the bias gradient of the Linear that produced in0)."""
    from vit_spectre_experiments_b200.fused import na_bwd, na_fwd

    torch.manual_seed(9)
    for dt, tol in ((torch.float32, 2e-5), (torch.bfloat16, 1e-2)):
        rows, N = 1000, 384
        in0 = torch.randn(rows, N, device=cuda_device).to(dt)
        skip = torch.randn(rows, N, device=cuda_device).to(dt)
        gamma = 1 + 0.2 * torch.randn(N, device=cuda_device)
        beta = 0.1 * torch.randn(N, device=cuda_device)
        da, db = torch.randn(rows, N, device=cuda_device).to(dt), torch.randn(rows, N, device=cuda_device).to(dt)
        out, mean, rstd = na_fwd(in0, gamma, beta, skip=skip, act=1)
        d_in0, d_skip, dg, dbt, dcol = na_bwd(da, in0, gamma, beta, mean, rstd, skip_shape=(rows, N), act=1, dout2=db,
                                              colsum=True)
        x = in0.float().requires_grad_(True)
        sk = skip.float().requires_grad_(True)
        g2, b2 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        y = F.gelu(F.layer_norm(x, (N,), g2, b2)) + sk
        y.backward(da.float() + db.float())
        assert rel_err(d_in0, x.grad) <= tol and rel_err(d_skip, sk.grad) <= tol
        ptol = 3e-5 if dt == torch.float32 else 2e-3   # bf16 stream: Phi through MUFU.TANH (relative error 2^-11)
        assert rel_err(dg, g2.grad) <= ptol and rel_err(dbt, b2.grad) <= ptol
        assert rel_err(dcol, x.grad.sum(0)) <= ptol   # fp32 sums of the unrounded d_in0


@pytest.mark.parametrize("rows,N,k,dt", [(1003, 1536, 4, torch.bfloat16), (1003, 1536, 4, torch.float32),
                                         (77, 256, 2, torch.bfloat16), (77, 512, 8, torch.float32),
                                         (403, 3072, 4, torch.bfloat16)])
def test_norm_act_pooled_second_output(cuda_device, rows, N, k, dt):
    """``out_pool`` = AdaptiveAvgPool1d(N / k) of the tail's result (what the next SpectreLinear's skip reads,
    layers.py:92,101) and its transpose ``d_out_pool`` in the backward, against torch autograd of the pair."""
    from vit_spectre_experiments_b200.fused import na_bwd, na_fwd

    g = torch.Generator(device=cuda_device).manual_seed(rows + N)
    rnd = lambda *s: torch.randn(*s, device=cuda_device, generator=g)   # noqa: E731
    Ns = N // 4
    in0, skip = rnd(rows, N).to(dt), rnd(rows, Ns).to(dt)
    gamma, beta = 1 + 0.2 * rnd(N), 0.1 * rnd(N)
    dout, dpool = rnd(rows, N).to(dt), rnd(rows, N // k).to(dt)
    out, mean, rstd, pooled = na_fwd(in0, gamma, beta, skip=skip, act=1, pool_k=k)
    d_in0, d_skip, dg, dbt, _ = na_bwd(dout, in0, gamma, beta, mean, rstd, skip_shape=(rows, Ns), act=1,
                                       d_out_pool=dpool, pool_k=k)
    x, sk = in0.float().requires_grad_(True), skip.float().requires_grad_(True)
    g2, b2 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y = F.gelu(F.layer_norm(x, (N,), g2, b2)) + F.adaptive_avg_pool1d(sk, N)
    yp = F.adaptive_avg_pool1d(y, N // k)
    (y * dout.float()).sum().add((yp * dpool.float()).sum()).backward()
    bf = dt == torch.bfloat16
    tol = 1e-2 if bf else 2e-5
    assert rel_err(out, y) <= (1e-2 if bf else 5e-6) and rel_err(pooled, yp) <= (1e-2 if bf else 5e-6)
    assert rel_err(d_in0, x.grad) <= tol and rel_err(d_skip, sk.grad) <= tol
    assert rel_err(dg, g2.grad) <= (2e-3 if bf else 3e-5) and rel_err(dbt, b2.grad) <= (2e-3 if bf else 3e-5)


def test_norm_act_sweep_direction_is_invisible(cuda_device):
    """``reverse`` only changes the order in which row groups are visited (to start where the producer left L2 hot):
    per-row results are bit-identical, the column reductions agree to rounding."""
    from vit_spectre_experiments_b200.fused import na_bwd, na_fwd

    torch.manual_seed(5)
    rows, N, Ns = 30001, 384, 1536          # ragged last row group
    for dt in (torch.float32, torch.bfloat16):
        in0 = torch.randn(rows, N, device=cuda_device).to(dt)
        skip = torch.randn(rows, Ns, device=cuda_device).to(dt)
        gamma, beta = 1 + 0.2 * torch.randn(N, device=cuda_device), 0.1 * torch.randn(N, device=cuda_device)
        dout = torch.randn(rows, N, device=cuda_device).to(dt)
        kw = dict(act=1, drop_p=0.05, seed=123)
        fa, fb = na_fwd(in0, gamma, beta, skip=skip, **kw), na_fwd(in0, gamma, beta, skip=skip, reverse=True, **kw)
        for u, v in zip(fa, fb):
            assert torch.equal(u, v)
        ba = na_bwd(dout, in0, gamma, beta, fa[1], fa[2], skip_shape=(rows, Ns), colsum=True, **kw)
        bb = na_bwd(dout, in0, gamma, beta, fa[1], fa[2], skip_shape=(rows, Ns), colsum=True, reverse=True, **kw)
        assert torch.equal(ba[0], bb[0]) and torch.equal(ba[1], bb[1])
        for u, v in zip(ba[2:], bb[2:]):
            assert rel_err(u, v) <= 1e-5


@pytest.mark.parametrize("mixer,grid", [("fft", (8, 8)), ("dwt", (8, 8)), ("permut", None)])
@pytest.mark.parametrize("autocast", [False, True])
def test_encoder_layer_fused_matches_torch(cuda_device, autocast, mixer, grid):
    """Whole SpectreEncoderLayer: fused path == stock torch composition (same parameters, same input).  With a CUDA
    mixer in the slot the fused path is the one-node ``fused.encoder_layer``; with MHPermutMix it is the
    per-kernel autograd composition."""
    from vit_spectre_experiments_b200 import SpectreEncoderLayer

    torch.manual_seed(3)
    kw = dict(seq_length=65, d_model=128, nhead=4, dim_feedforward=256, dropout=0.0, activation="gelu", mixer=mixer,
              grid=grid)
    a = SpectreEncoderLayer(**kw, fused=True).to(cuda_device)
    b = SpectreEncoderLayer(**kw, fused=False).to(cuda_device)
    b.load_state_dict(a.state_dict())
    x = torch.randn(16, 65, 128, device=cuda_device)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        ya, yb = a(xa), b(xb)
    assert a.use_fused(x) and not b.use_fused(x)
    w = torch.randn_like(ya)
    (ya * w).sum().backward()
    (yb * w).sum().backward()
    tol = 3e-2 if autocast else 2e-5   # autocast: bf16 residual stream vs torch's fp32 LayerNorm outputs
    assert rel_err(ya, yb) <= tol and rel_err(xa.grad, xb.grad) <= tol
    pa, pb = dict(a.named_parameters()), dict(b.named_parameters())
    for k in pa:
        assert rel_err(pa[k].grad, pb[k].grad) <= (5e-2 if autocast else 1e-4), k


@pytest.mark.parametrize("autocast", [False, True])
def test_encoder_layer_c4_widths_pooled_handover(cuda_device, autocast):
    """ViT-S/16 widths (E=384, hidden=1536, 14x14 grid): the first tail emits the pooled copy of a1 that the second
    tail's skip reads, and the skip gradient returns through ``d_out_pool`` -- same layer output and gradients as the
    stock torch composition, with dropout off."""
    from vit_spectre_experiments_b200 import SpectreEncoderLayer

    torch.manual_seed(4)
    kw = dict(seq_length=197, d_model=384, nhead=6, dim_feedforward=1536, dropout=0.0, activation="gelu", mixer="fft",
              grid=(14, 14))
    a = SpectreEncoderLayer(**kw, fused=True).to(cuda_device)
    b = SpectreEncoderLayer(**kw, fused=False).to(cuda_device)
    b.load_state_dict(a.state_dict())
    x = torch.randn(5, 197, 384, device=cuda_device)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        ya, yb = a(xa), b(xb)
    w = torch.randn_like(ya)
    (ya * w).sum().backward()
    (yb * w).sum().backward()
    tol = 3e-2 if autocast else 2e-5
    assert rel_err(ya, yb) <= tol and rel_err(xa.grad, xb.grad) <= tol
    pa, pb = dict(a.named_parameters()), dict(b.named_parameters())
    for k in pa:
        assert rel_err(pa[k].grad, pb[k].grad) <= (5e-2 if autocast else 1e-4), k


@pytest.mark.parametrize("autocast", [False, True])
def test_stacked_layers_hand_the_gradient_over_in_two_parts(cuda_device, autocast):
    """SpectreEncoder chains the one-node layers through (y, y_twin): the two addends of a layer's input gradient
    travel separately and are summed inside the LN2 backward of the layer below.  Same loss and gradients as
    calling the layers one by one (where each backward adds the parts itself)."""
    from vit_spectre_experiments_b200 import SpectreEncoder, SpectreEncoderLayer

    torch.manual_seed(11)
    layer = SpectreEncoderLayer(seq_length=65, d_model=128, nhead=4, dim_feedforward=256, dropout=0.0,
                                activation="gelu", mixer="fft", grid=(8, 8), fused=True)
    enc = SpectreEncoder(layer, 3).to(cuda_device)
    with torch.no_grad():
        for p in enc.parameters():
            p.add_(0.05 * torch.randn_like(p))
    x = torch.randn(8, 65, 128, device=cuda_device)
    w = torch.randn(8, 65, 128, device=cuda_device)

    def run(chained):
        xa = x.clone().requires_grad_(True)
        enc.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            if chained:
                y = enc(xa)
            else:
                out = xa
                for l in enc.layers:
                    out = l.forward_fused(out)
                y = out.to(xa.dtype) + xa
        (y.float() * w).sum().backward()
        return y.detach(), xa.grad.clone(), [p.grad.clone() for p in enc.parameters()]

    ya, ga, pa = run(True)
    yb, gb, pb = run(False)
    tol = 2e-2 if autocast else 2e-5
    assert rel_err(ya, yb) == 0.0
    assert rel_err(ga, gb) <= tol
    for u, v in zip(pa, pb):
        assert rel_err(u, v) <= tol


def test_host_batch_prefetcher_feeds_the_same_batches(cuda_device):
    """train.HostBatchPrefetcher (bench.py's end-to-end feed): double-buffered pinned-host -> device copies on a side
    stream deliver exactly the staged batches, in order."""
    from vit_spectre_experiments_b200.train import HostBatchPrefetcher

    like_img = torch.empty(4, 3, 8, 8, device=cuda_device)
    like_lab = torch.empty(4, dtype=torch.long, device=cuda_device)
    feed = HostBatchPrefetcher(cuda_device, like_img, like_lab)
    host = [(torch.full((4, 3, 8, 8), float(i)).pin_memory(), torch.full((4,), i, dtype=torch.long).pin_memory())
            for i in range(5)]
    feed.stage(*host[0])
    for i in range(5):
        img, lab = feed.next()
        if i + 1 < 5:
            feed.stage(*host[i + 1])
        s = img.sum() + lab.sum()          # consumer work enqueued on the current stream
        feed.release()
        assert s.item() == i * 4 * 3 * 8 * 8 + i * 4
    with pytest.raises(RuntimeError):
        feed.next()


@pytest.mark.parametrize("B,T,E,heads,dtype", [(8, 65, 512, 16, torch.float32), (3, 50, 48, 8, torch.float32),
                                               (37, 17, 24, 3, torch.bfloat16), (1, 5, 4, 1, torch.float32)])
def test_permut_mix_gather(cuda_device, B, T, E, heads, dtype):
    """MHPermutMix gather kernels (C ABI ``spectre_permut_*``; layers.py:68-70): forward bit-identical to
    ``x.view(B,-1)[:, perms] * signs``; backward (inverse-permutation gather, heads summed in a fixed order) matches
    torch's index backward to fp32 rounding and is bit-reproducible."""
    from vit_spectre_experiments_b200.functional import inverse_permutations, permut_mix

    g = torch.Generator().manual_seed(B * 100 + heads)
    F_ = T * E
    perms = torch.stack([torch.randperm(F_, generator=g) for _ in range(heads)]).to(cuda_device)
    signs = (torch.randint(0, 2, (1, heads, F_), generator=g, dtype=torch.float32) * 2 - 1).to(cuda_device)
    x = torch.randn(B, T, E, generator=g).to(dtype).to(cuda_device)
    inv = inverse_permutations(perms)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya = permut_mix(xa, perms, signs, inv)
    yb = (xb.reshape(B, -1)[:, perms] * signs).to(dtype)
    assert ya.shape == (B, heads, F_) and ya.dtype == dtype and torch.equal(ya, yb)
    w = torch.randn(B, heads, F_, generator=g).to(dtype).to(cuda_device)
    ya.backward(w)
    yb.backward(w)
    assert rel_err(xa.grad, xb.grad) <= (2e-6 if dtype == torch.float32 else 1e-2)
    xc = x.clone().requires_grad_(True)
    permut_mix(xc, perms, signs, inv).backward(w)
    assert torch.equal(xc.grad, xa.grad)
    with pytest.raises(RuntimeError):
        inverse_permutations(torch.zeros(2, 8, dtype=torch.int64, device=cuda_device))



@pytest.mark.parametrize("img_size,P,E,C,dt", [(28, 4, 48, 3, torch.float32), (32, 4, 512, 3, torch.float32),
                                               (224, 16, 384, 3, torch.bfloat16), (224, 16, 64, 3, torch.float32),
                                               (30, 4, 16, 1, torch.float32)])
def test_patch_embed_kernels_match_reference_module(cuda_device, img_size, P, E, C, dt):
    """SpectralPatchEmbed through csrc/embed.cu (patch gather + cast | GEMM | cls / position / dropout pass) against
    the module's stock composition (unfold -> rfft2(ortho).real -> gains -> Linear -> cat -> + pos; spectre.py:124-156)
    on the CPU: output and the gradients of all six parameters.  30x30 with P=4 exercises the crop to 28x28."""
    from vit_spectre_experiments_b200.model import SpectralPatchEmbed

    torch.manual_seed(img_size + E)
    n = (img_size // P) ** 2
    ref = SpectralPatchEmbed(E, P, n, 0.0, C)
    with torch.no_grad():
        ref.freq_weight_h.add_(0.3 * torch.randn(P))
        ref.freq_weight_w.add_(0.3 * torch.randn(P // 2 + 1))
    dev = SpectralPatchEmbed(E, P, n, 0.0, C)
    dev.load_state_dict(ref.state_dict())
    dev.to(cuda_device)
    img = torch.randn(3, C, img_size, img_size)
    w = torch.randn(3, n + 1, E)
    y_ref = ref(img)                                   # CPU: the stock path
    (y_ref * w).sum().backward()
    bf = dt == torch.bfloat16
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=bf):
        y = dev(img.to(cuda_device))
    assert y.dtype == dt
    (y.float() * w.to(cuda_device)).sum().backward()
    tol = 2e-2 if bf else 2e-5
    assert rel_err(y, y_ref) <= tol
    for (k, pr), (_, pd) in zip(ref.named_parameters(), dev.named_parameters()):
        assert pd.grad is not None and rel_err(pd.grad, pr.grad) <= (3e-2 if bf else 1e-4), k


def test_patch_embed_dropout_consistent(cuda_device):
    """Counter-based dropout of the embedding pass: keep-rate ~ 1 - p, kept values scaled by 1 / (1 - p), and the
    backward recomputes the same mask (no gradient through dropped positions)."""
    from vit_spectre_experiments_b200.functional import patch_embed

    torch.manual_seed(2)
    B, C, S, P, E = 16, 3, 32, 4, 64
    n = (S // P) ** 2
    img = torch.randn(B, C, S, S, device=cuda_device)
    M = torch.randn(C * P * P, E, device=cuda_device, requires_grad=True)
    bias = torch.randn(E, device=cuda_device, requires_grad=True)
    cls = torch.randn(1, 1, E, device=cuda_device, requires_grad=True)
    pos = torch.randn(1, n + 1, E, device=cuda_device, requires_grad=True)
    y0 = patch_embed(img, M, bias, cls, pos, P, 0.0, 0)
    y = patch_embed(img, M, bias, cls, pos, P, 0.2, 99)
    kept = y != 0
    assert abs(kept.float().mean().item() - 0.8) < 1e-2
    assert torch.allclose(y[kept], y0[kept] / 0.8, rtol=1e-5, atol=1e-5)
    assert torch.equal(y, patch_embed(img, M, bias, cls, pos, P, 0.2, 99))
    (y * (~kept).float()).sum().backward()
    assert pos.grad.abs().max().item() == 0.0 and M.grad.abs().max().item() == 0.0


@pytest.mark.parametrize("mixer,precision", [("fft", "bf16"), ("dwt", "bf16"), ("fft", "fp32")])
def test_training_step_is_cuda_graph_capturable(cuda_device, mixer, precision):
    """include/spectre_b200.h promises that every entry point only enqueues work on the given stream: the WHOLE step
    (forward, loss, backward with the fused parameter-gradient reductions, AdamW) is captured in one CUDA graph by
    ``Trainer(cuda_graph=True)`` and replayed.  A replay from the same state is bit-identical; against the eager step
    it agrees to rounding; with dropout on, every replay draws a fresh mask through the device-resident seed word."""
    from types import SimpleNamespace

    from vit_spectre_experiments_b200.train import Trainer

    def cfg(p):
        return SimpleNamespace(img_size=32, patch_size=4, in_channels=3, num_classes=10, embed_dim=64, num_encoders=2,
                               num_heads=4, hidden_dim=256, dropout=p, activation="gelu", mixer=mixer, precision=precision,
                               random_seed=42, learning_rate=1e-3, adam_betas=(0.9, 0.999), adam_weight_decay=0.01)

    def snapshot(tr):
        """Parameters and AdamW state, to be restored IN PLACE (the captured graph holds the tensors' addresses)."""
        return ([p.detach().clone() for p in tr.model.parameters()],
                [{k: v.clone() for k, v in tr.opt.state[p].items() if torch.is_tensor(v)} for p in tr.model.parameters()])

    def restore(tr, snap):
        with torch.no_grad():
            for p, q in zip(tr.model.parameters(), snap[0]):
                p.copy_(q)
            for p, st in zip(tr.model.parameters(), snap[1]):
                for k, v in st.items():
                    tr.opt.state[p][k].copy_(v)

    tg = Trainer(cfg(0.0), cuda_device, cuda_graph=True)
    img, lab = tg.synthetic_batch(32)
    tg.step(img, lab)                                   # 3 eager warm-up steps + capture + first replay
    state = snapshot(tg)

    def replay_from_state():
        restore(tg, state)
        loss = tg.step(img, lab).clone()
        return loss, [p.detach().clone() for p in tg.model.parameters()]

    la, pa = replay_from_state()
    lb, pb = replay_from_state()
    assert torch.equal(la, lb) and all(torch.equal(u, v) for u, v in zip(pa, pb))
    # the same step, eagerly, from the same state
    restore(tg, state)
    le = tg._eager_step(img, lab)
    assert abs(le.item() - la.item()) <= 1e-5 * abs(la.item()) + 1e-6
    for u, v in zip(pa, tg.model.parameters()):
        assert rel_err(u, v) <= (2e-2 if precision == "bf16" else 1e-4)
    losses = [tg.step(img, lab).item() for _ in range(8)]
    assert losses[-1] < losses[0]
    # dropout inside a graph: fresh masks per replay
    td = Trainer(cfg(0.3), cuda_device, cuda_graph=True)
    td.step(img, lab)
    st = snapshot(td)
    ls = []
    for _ in range(3):
        restore(td, st)
        ls.append(td.step(img, lab).item())
    assert len(set(ls)) == 3
    from vit_spectre_experiments_b200 import _lib
    _lib.graph_seed["tensor"] = None


def test_patch_embed_folded_matrix_cache(cuda_device):
    """Without autograd the folded matrix ``M`` is cached until a parameter changes (in-place update or
    ``load_state_dict``); under autograd it is rebuilt so gradients reach the parameters."""
    from vit_spectre_experiments_b200.model import SpectralPatchEmbed

    torch.manual_seed(1)
    m = SpectralPatchEmbed(32, 4, 64, 0.0, 3).to(cuda_device).eval()
    img = torch.randn(2, 3, 32, 32, device=cuda_device)
    with torch.no_grad():
        y0 = m(img)
        M0 = m._M
        assert m(img) is not None and m._M is M0                    # second call: cache hit
        m.freq_weight_h.mul_(1.5)                                    # in-place parameter update
        y1 = m(img)
        assert m._M is not M0 and not torch.equal(y0, y1)
    y2 = m(img)                                                      # autograd on: fresh M with a graph
    y2.sum().backward()
    assert m.freq_weight_h.grad is not None and torch.allclose(y2, y1, rtol=1e-5, atol=1e-5)
