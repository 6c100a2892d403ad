"""Numerics of the sm_100a kernels against the plain-PyTorch fp32 oracle of the same op
(``fdt.ops.reference``).  Everything here needs a GPU."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _ops():
    from fdt import ops
    return ops


def _close(a, b, rtol, atol, what=""):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item()
    ref = b.abs().max().item()
    assert torch.allclose(a, b, rtol=rtol, atol=atol), f"{what}: max err {err:.4g} (ref max {ref:.4g})"


def _relnorm(a, b, tol, what=""):
    """Relative Frobenius error.  Used for gradients that pass through a ReLU/CELU mask: a bf16
    conv output that lands on the other side of zero than the fp32 oracle flips single mask
    entries, which is a legitimate O(1) difference in a few elements but tiny in norm."""
    a, b = a.float(), b.float()
    rel = ((a - b).norm() / b.norm().clamp_min(1e-12)).item()
    assert rel < tol, f"{what}: relative Frobenius error {rel:.4g} >= {tol}"


@pytest.mark.parametrize("cin,cout,k,hw,act,res", [(64, 64, 3, 16, "relu", False), (64, 256, 1, 8, "relu", True),
                                                    (16, 32, 3, 8, "celu", True), (128, 128, 3, 4, "none", False)])
def test_conv_bnstat_act_fwd_bwd(cin, cout, k, hw, act, res):
    ops = _ops()
    torch.manual_seed(0)
    N = 8
    x = torch.randn(N, cin, hw, hw, device=DEV)
    w = torch.randn(cout, cin, k, k, device=DEV) / math.sqrt(cin * k * k)
    r = torch.randn(N, cout, hw, hw, device=DEV) if res else None
    gout = torch.randn(N, cout, hw, hw, device=DEV)
    # oracle in fp32 on bf16-rounded inputs
    xr = x.bfloat16().float().requires_grad_(True)
    wr = w.bfloat16().float().requires_grad_(True)
    rr = r.bfloat16().float().requires_grad_(True) if res else None
    out_ref = ops.ref.conv_bnstat_act(xr, wr, 1, k // 2, 1e-3, act, rr)
    out_ref.backward(gout.bfloat16().float())
    xn = x.bfloat16().to(memory_format=torch.channels_last).requires_grad_(True)
    wn = w.clone().requires_grad_(True)
    rn = r.bfloat16().to(memory_format=torch.channels_last).requires_grad_(True) if res else None
    out = ops.conv_bnstat_act(xn, wn, 1, k // 2, 1e-3, act, rn)
    assert out.dtype == torch.bfloat16
    out.backward(gout.bfloat16().to(memory_format=torch.channels_last))
    _close(out, out_ref, 3e-2, 3e-2, "out")
    _relnorm(xn.grad, xr.grad, 3e-2, "dx")
    _relnorm(wn.grad, wr.grad, 3e-2, "dw")
    if res:
        _relnorm(rn.grad, rr.grad, 3e-2, "dres")


@pytest.mark.parametrize("affine", [False, True])
def test_bn_backward_mask_from_y_equals_mask_from_output(affine, monkeypatch):
    """Backward of a ReLU unit without a residual input: the kernels that re-derive the mask from the conv output
    (scale*y + shift > 0, the forward's expression) give the gradients of the ones that read it from the activated output."""
    ops = _ops()
    from fdt.ops import cuda_ops
    torch.manual_seed(5)
    N, cin, cout, hw = 8, 64, 128, 8
    x = torch.randn(N, cin, hw, hw, device=DEV)
    w = torch.randn(cout, cin, 3, 3, device=DEV) / math.sqrt(cin * 9)
    gamma = torch.rand(cout, device=DEV) + 0.5
    beta = torch.randn(cout, device=DEV) * 0.3
    gout = torch.randn(N, cout, hw, hw, device=DEV).bfloat16().to(memory_format=torch.channels_last)
    res = {}
    for flag in (False, True):
        monkeypatch.setattr(cuda_ops, "_BWD_MASKY", flag)
        before = _lib_launches("fdt_bn_bwd_apply_masky")
        xi = x.bfloat16().to(memory_format=torch.channels_last).requires_grad_(True)
        wi = w.clone().requires_grad_(True)
        if affine:
            g, b = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
            rm, rv, nb = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV), torch.zeros((), dtype=torch.long, device=DEV)
            o = ops.conv_bnaffine_act(xi, wi, g, b, rm, rv, nb, 1, 1, 1e-5, 0.1, True, "relu", None)
        else:
            o = ops.conv_bnstat_act(xi, wi, 1, 1, 1e-3, "relu", None)
        o.backward(gout)
        assert (_lib_launches("fdt_bn_bwd_apply_masky") > before) == flag
        res[flag] = [xi.grad.float(), wi.grad.float()] + ([g.grad, b.grad] if affine else [])
    for a, b_ in zip(res[False], res[True]):
        assert torch.isfinite(b_).all()
        rel = ((a - b_).norm() / a.norm().clamp_min(1e-12)).item()
        assert rel < 1e-3, rel          # (the channel sums are atomic accumulations: the two runs are not bit-identical)


def _lib_launches(name):
    from fdt.ops import _lib
    return _lib.LAUNCHES.by_name.get(name, 0)


@pytest.mark.parametrize("training", [True, False])
def test_conv_bnaffine_act(training):
    ops = _ops()
    torch.manual_seed(1)
    N, cin, cout, hw = 8, 32, 64, 8
    x = torch.randn(N, cin, hw, hw, device=DEV)
    w = torch.randn(cout, cin, 3, 3, device=DEV) / math.sqrt(cin * 9)
    gamma = torch.rand(cout, device=DEV) + 0.5
    beta = torch.randn(cout, device=DEV) * 0.1
    gout = torch.randn(N, cout, hw // 2, hw // 2, device=DEV)

    def run(native):
        rm, rv, nb = torch.zeros(cout, device=DEV), torch.ones(cout, device=DEV), torch.zeros((), dtype=torch.long, device=DEV)
        g, b = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        if native:
            xi = x.bfloat16().to(memory_format=torch.channels_last).requires_grad_(True)
            wi = w.clone().requires_grad_(True)
            o = ops.conv_bnaffine_act(xi, wi, g, b, rm, rv, nb, 2, 1, 1e-5, 0.1, training, "relu", None)
            o.backward(gout.bfloat16().to(memory_format=torch.channels_last))
        else:
            xi = x.bfloat16().float().requires_grad_(True)
            wi = w.bfloat16().float().requires_grad_(True)
            o = ops.ref.conv_bnaffine_act(xi, wi, g, b, rm, rv, nb, 2, 1, 1e-5, 0.1, training, "relu", None)
            o.backward(gout.bfloat16().float())
        return o, xi.grad, wi.grad, g.grad, b.grad, rm, rv

    a, b_ = run(True), run(False)
    names = ["out", "dx", "dw", "dgamma", "dbeta", "running_mean", "running_var"]
    for n, u, v in zip(names, a, b_):
        if u is None and v is None:
            continue
        _relnorm(u, v, 3e-2, n)


def test_layernorm_fwd_bwd():
    ops = _ops()
    torch.manual_seed(2)
    x = torch.randn(6, 37, 512, device=DEV) * 2 + 0.3
    a = torch.rand(512, device=DEV) + 0.5
    b = torch.randn(512, device=DEV) * 0.1
    g = torch.randn(6, 37, 512, device=DEV)
    xr, ar, br = x.clone().requires_grad_(True), a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ops.ref.layernorm_ref(xr, ar, br, 1e-6).backward(g.bfloat16().float())
    xn, an, bn = x.clone().requires_grad_(True), a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = ops.layernorm_ref(xn, an, bn, 1e-6)
    assert y.dtype == torch.bfloat16
    y.backward(g.bfloat16())
    _close(y, ops.ref.layernorm_ref(x, a, b, 1e-6), 1e-2, 2e-2, "y")
    _close(xn.grad, xr.grad, 1e-3, 1e-3, "dx")
    _close(an.grad, ar.grad, 1e-3, 1e-2, "da")
    _close(bn.grad, br.grad, 1e-3, 1e-2, "db")


@pytest.mark.parametrize("per_sample", [False, True])
def test_mixup_ce(per_sample):
    ops = _ops()
    torch.manual_seed(3)
    B, C = 77, 10
    logits = torch.randn(B, C, device=DEV) * 3
    ya, yb = torch.randint(0, C, (B,), device=DEV), torch.randint(0, C, (B,), device=DEV)
    lam = torch.rand(B, device=DEV) if per_sample else torch.rand((), device=DEV)
    lr_ = logits.clone().requires_grad_(True)
    loss_ref = ops.ref.mixup_cross_entropy(lr_, ya, yb, lam)
    loss_ref.backward()
    ln = logits.clone().requires_grad_(True)
    loss = ops.mixup_cross_entropy(ln, ya, yb, lam)
    loss.backward()
    _close(loss, loss_ref, 1e-5, 1e-5, "loss")
    _close(ln.grad, lr_.grad, 1e-4, 1e-6, "dlogits")


def test_embed_fwd_bwd():
    ops = _ops()
    torch.manual_seed(4)
    B, L, D, V = 5, 33, 512, 1000
    tok = torch.randint(0, V, (B, L), device=DEV)
    typ = torch.randint(0, 3, (B, L), device=DEV)
    ws = [torch.randn(V, D, device=DEV), torch.randn(512, D, device=DEV), torch.randn(3, D, device=DEV)]
    g = torch.randn(B, L, D, device=DEV)
    wr = [w.clone().requires_grad_(True) for w in ws]
    wn = [w.clone().requires_grad_(True) for w in ws]
    o_ref = ops.ref.embed(tok, typ, *wr, math.sqrt(D))
    o_ref.backward(g)
    o = ops.embed(tok, typ, *wn, math.sqrt(D))
    o.backward(g)
    _close(o, o_ref, 1e-6, 1e-5, "out")
    for n, a, b in zip(["dtok", "dpos", "dseg"], wn, wr):
        _close(a.grad, b.grad, 1e-4, 1e-3, n)


def test_dropout_add():
    ops = _ops()
    torch.manual_seed(5)
    x = torch.randn(64, 128, 512, device=DEV).bfloat16()
    r = torch.randn(64, 128, 512, device=DEV)
    o = ops.dropout_add(x, r, 0.1, False)
    _close(o, r + x.float(), 1e-6, 1e-6, "eval")
    xn = x.clone().requires_grad_(True)
    o = ops.dropout_add(xn, r, 0.1, True)
    kept = ((o - r).abs() > 0) | (x.float() == 0)
    frac = 1.0 - kept.float().mean().item()
    assert abs(frac - 0.1) < 5e-3, frac
    o.backward(torch.ones_like(o))
    mask_b = xn.grad.float() != 0
    assert torch.equal(mask_b, (o - r).abs() > 0) or (mask_b ^ ((o - r).abs() > 0)).float().mean() < 1e-3
    _close(xn.grad[mask_b], torch.full_like(xn.grad[mask_b], 1 / 0.9), 1e-2, 1e-2, "scale")


@pytest.mark.parametrize("n,N,D", [(3, 200, 64), (2, 100, 160), (4, 576, 64), (2, 27, 3), (1, 7, 5)])
def test_ngd_eig_update(n, N, D):
    ops = _ops()
    from fdt.optim import ngd as M
    torch.manual_seed(6)
    fb_a = M.FactorBatch(n, N, D, 4.0, -1, 4, 0.1, torch.float32, DEV)
    fb_b = M.FactorBatch(n, N, D, 4.0, -1, 4, 0.1, torch.float32, DEV)
    for it in range(5):
        X = torch.randn(n, N, D, device=DEV) * (1 + it)
        ops.force_torch(True)
        out_b = M.precondition(fb_b, X, left=False)
        ops.force_torch(False)
        out_a = M.precondition(fb_a, X, left=False)
        _close(out_a, out_b, 2e-3, 2e-3 * out_b.abs().max().item(), f"precond it{it}")
        _close(fb_a.d, fb_b.d, 5e-3, 5e-3 * fb_b.d.abs().max().item(), f"d it{it}")
        _close(fb_a.rho, fb_b.rho, 5e-3, 1e-6, f"rho it{it}")
        WtWa = torch.bmm(fb_a.W.transpose(1, 2), fb_a.W)
        WtWb = torch.bmm(fb_b.W.transpose(1, 2), fb_b.W)
        _close(WtWa, WtWb, 1e-2, 1e-2 * WtWb.abs().max().item(), f"WtW it{it}")


def test_flat_kernels():
    ops = _ops()
    cu = ops._cu()
    torch.manual_seed(7)
    n = 1_000_003
    g = torch.randn(n, device=DEV)
    sq = cu.sqnorm(g)
    _close(sq, (g.double() ** 2).sum().float(), 1e-4, 0, "sqnorm")
    g2 = g.clone()
    sk = torch.zeros((), device=DEV)
    cu.clip_scale_(g2, sq, 10.0, sk)
    _close(g2, g * (10.0 / (sq.sqrt() + 1e-6)), 1e-5, 1e-7, "clip")
    p, buf = torch.randn(n, device=DEV), torch.zeros(n, device=DEV)
    p_ref = p.clone()
    lr = torch.tensor(0.1, device=DEV)
    cu.sgd_momentum_(p, g, buf, lr, 0.9, 0.0, False, True)
    cu.sgd_momentum_(p, g, buf, lr, 0.9, 0.0, False, False)
    b = g.clone(); p_ref -= 0.1 * b; b = 0.9 * b + g; p_ref -= 0.1 * b
    _close(p, p_ref, 1e-5, 1e-6, "sgd")
    y = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    cu.cast_bf16_(g, y)
    assert torch.equal(y, g.bfloat16())
    bad = torch.full((8,), float("nan"), device=DEV)
    cu.clip_scale_(bad, cu.sqnorm(bad), 10.0, sk)
    assert sk.item() == 1 and (bad == 0).all()


def test_pool_fc_fwd_bwd():
    ops = _ops()
    torch.manual_seed(8)
    x = torch.randn(32, 2048, 4, 4, device=DEV).bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(10, 2048, device=DEV) / 45).requires_grad_(True)
    b = torch.randn(10, device=DEV).requires_grad_(True)
    g = torch.randn(32, 10, device=DEV)
    xn = x.clone().requires_grad_(True)
    out = ops.pool_fc(xn, w, b)
    out.backward(g)
    xr, wr, br = x.float().requires_grad_(True), w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    ref = ops.ref.pool_fc(xr, wr, br)
    ref.backward(g)
    _close(out, ref, 1e-4, 1e-4, "logits")
    _relnorm(xn.grad, xr.grad, 1e-2, "dx")
    _relnorm(w.grad, wr.grad, 1e-4, "dW")
    _relnorm(b.grad, br.grad, 1e-5, "db")


def test_input_aug_mixup_kernel():
    ops = _ops()
    from fdt.data import cifar
    cu = ops._cu()
    torch.manual_seed(9)
    imgs = torch.randint(0, 256, (16, 32, 32, 3), dtype=torch.uint8, device=DEV)
    # eval mode: plain normalise, exact against the torch definition
    out = cu.input_aug_mixup(imgs, False, None, None, cifar.MEAN, cifar.STD)
    ref = cifar.gpu_transform(imgs, train=False)
    assert out.shape == (16, 3, 32, 32) and out.is_contiguous(memory_format=torch.channels_last)
    _close(out, ref, 1e-2, 2e-2, "normalise")
    # train + mixup: lam = 1 must return an augmented version of each image (values from the image or padding)
    perm = torch.randperm(16, device=DEV)
    one = torch.ones((), device=DEV)
    aug = cu.input_aug_mixup(imgs, True, perm, one, cifar.MEAN, cifar.STD).float()
    pad_val = torch.tensor([(0 - m) / s for m, s in zip(cifar.MEAN, cifar.STD)], device=DEV).view(1, 3, 1, 1)
    frac_pad = ((aug - pad_val).abs() < 2e-2).float().mean().item()
    assert 0.0 < frac_pad < 0.4
    # blend linearity: lam=0.3 equals 0.3*T(x) + 0.7*T(x[perm]) when the same augmentation draw is used
    tab = cu._seed_state[imgs.device]               # rewind the seed table so the three calls take the same row
    at = tab.idx
    a = cu.input_aug_mixup(imgs, True, perm, one, cifar.MEAN, cifar.STD).float()
    tab.idx = at
    b = cu.input_aug_mixup(imgs, True, perm, torch.zeros((), device=DEV), cifar.MEAN, cifar.STD).float()
    tab.idx = at
    c = cu.input_aug_mixup(imgs, True, perm, torch.full((), 0.3, device=DEV), cifar.MEAN, cifar.STD).float()
    _close(c, 0.3 * a + 0.7 * b, 2e-2, 3e-2, "blend")


@pytest.mark.parametrize("n,N,D,R,left", [(2, 256, 64, 32, False), (3, 200, 64, 32, True), (2, 100, 160, 80, False),
                                           (1, 4608, 512, 80, True), (1, 10, 2048, 80, False), (2, 36, 128, 64, True),
                                           (1, 30, 24, 12, False)])
def test_ngd_tcgen05_gemms(n, N, D, R, left):
    """Batched 3xTF32 tcgen05 kernels for H = XWᵀ, X̂ = X − HW (+norms) and J = HᵀX: fp32-level accuracy, operands
    used in place (no transposed copies)."""
    ops = _ops()
    cu = ops._cu()
    assert cu.ngd_tf32_ok(n, N, D, R, left) or R < 16
    torch.manual_seed(11)
    X = torch.randn(n, N, D, device=DEV)
    W = torch.randn(n, R, D, device=DEV) * 0.1
    Xin = X.transpose(1, 2).contiguous() if left else X
    H, out, p0 = cu.ngd_project(Xin, W, left)
    Xd, Wd = X.double(), W.double()
    Href = torch.bmm(Xd, Wd.transpose(1, 2))
    Xhref = Xd - torch.bmm(Href, Wd)
    scale = torch.sqrt((Xd * Xd).sum((1, 2)) / (Xhref * Xhref).sum((1, 2)))
    want = Xhref * scale[:, None, None]
    got = out.transpose(1, 2) if left else out
    # 3xTF32 error is dominated by the tensor core's truncating fp32 accumulation: ~7e-9 * K relative (measured:
    # 8e-7 at K = 64, 1.5e-5 at K = 2048; cuBLAS fp32 SIMT: 1e-7 .. 8e-7; plain tf32: ~1e-3)
    _relnorm(H.double(), Href, 3e-5, "H")
    _relnorm(got.double(), want, 3e-5, "rescaled Xhat")
    _relnorm(p0.double(), (Xd * Xd).sum((1, 2)), 1e-5, "p0")
    J = torch.full((n, R, D), 3.0, device=DEV)
    Hf = Href.float().contiguous()
    cu.ngd_stats_j(Hf, Xin, J, left)
    _relnorm(J.double(), torch.bmm(Hf.double().transpose(1, 2), Xd), 3e-5, "J")
    if R >= 16:
        K = torch.full((n, R, R), 5.0, device=DEV)
        L = torch.full((n, R, R), 5.0, device=DEV)
        Jf = J.clone()
        cu.ngd_stats_kl(Jf, W, K, L)
        _relnorm(K.double(), torch.bmm(Jf.double(), Jf.double().transpose(1, 2)), 3e-5, "K")
        _relnorm(L.double(), torch.bmm(Jf.double(), Wd.transpose(1, 2)), 3e-5, "L")


@pytest.mark.parametrize("n,A,D,B", [(3, 64, 3, 192), (2, 5, 3, 7), (1, 1, 10, 2048), (4, 576, 3, 1), (2, 9, 4, 36), (1, 3, 12, 20)])
@pytest.mark.parametrize("with_j", [False, True])
def test_ngd_small_axis(n, A, D, B, with_j):
    """Short-axis NGD pass on the [n, A, D, B] view vs the two-GEMM formulation on the [n, N, D] matrix."""
    from fdt.ops import cuda_ops as cu
    torch.manual_seed(0)
    R = (D + 1) // 2
    X = torch.randn(n, A, D, B, device=DEV)
    W = torch.randn(n, R, D, device=DEV) * 0.3
    J = torch.full((n, R, D), 7.0, device=DEV) if with_j else None
    p0, out = cu.ngd_small_axis(X, W, A, D, B, J)
    Xm = X.double().movedim(2, -1).reshape(n, A * B, D)
    Wd = W.double()
    H = Xm @ Wd.transpose(1, 2)
    Xh = Xm - H @ Wd
    p0r = (Xm * Xm).sum((1, 2))
    p1r = (Xh * Xh).sum((1, 2))
    ref = (Xh * torch.sqrt(p0r / p1r)[:, None, None]).reshape(n, A, B, D).movedim(-1, 2)
    _close(p0, p0r, 1e-5, 1e-5, "p0")
    _close(out, ref, 1e-4, 1e-5, "out")
    if with_j:
        _close(J, H.transpose(1, 2) @ Xm, 1e-4, 1e-3, "J")


def test_ngd_small_axis_matches_general_path():
    """One NGD trajectory of a conv-shaped tensor with the short-axis kernel on and off."""
    from fdt import ops
    from fdt.optim import NGD
    from fdt.ops import cuda_ops as cu
    outs = []
    for enabled in (True, False):
        torch.manual_seed(0)
        p = torch.nn.Parameter(torch.randn(16, 8, 3, 3, device=DEV))
        opt = NGD([p], lr=0.1, momentum=0.9)
        saved = cu.ngd_small_axis_ok
        if not enabled:
            cu.ngd_small_axis_ok = lambda D, R: False
        try:
            g = torch.Generator(device=DEV).manual_seed(1)
            for _ in range(14):
                p.grad = torch.randn(p.shape, device=DEV, generator=g)
                opt.step()
        finally:
            cu.ngd_small_axis_ok = saved
        outs.append(p.detach().clone())
    _relnorm(outs[0], outs[1], 2e-3, "small-axis vs general NGD trajectory")


def test_dgrad_weight_shadow_layouts():
    """All data-gradient operand layouts produced by one launch (csrc/weight_prep.cu) vs the torch transposes."""
    from fdt.ops import cuda_ops as cu
    from fdt.ops import gemm_sm100 as g
    from fdt.parallel.flat import FlatParameters
    torch.manual_seed(0)
    convs = [torch.nn.Parameter(torch.randn(64, 32, 3, 3)), torch.nn.Parameter(torch.randn(72, 40, 1, 1)),
             torch.nn.Parameter(torch.randn(128, 64, 3, 3)), torch.nn.Parameter(torch.randn(16, 3, 3, 3))]
    lins = [torch.nn.Parameter(torch.randn(96, 200)), torch.nn.Parameter(torch.randn(10, 64))]
    modes = [0, 0, 1, 0, 0, 0]
    for p, m in zip(convs + lins, modes):
        p._fdt_dgrad_mode = m
    flat = FlatParameters(convs + lins, device=torch.device(DEV))
    flat.refresh_shadow()
    torch.cuda.synchronize()
    for p, m in zip(convs + lins, modes):
        w16 = cu._w16(p)
        got = cu._wdgrad(p, m)
        if p.shape[0] % 8 or p.shape[1] % 8:
            assert got is None
            continue
        assert got is not None and cu._wdgrad(p, 1 - m) is None
        if p.dim() == 2:
            want = w16.t()
        elif m == 0:
            want = g.dgrad_weight(w16)
            assert got.is_contiguous(memory_format=torch.channels_last)
        else:
            O, I, kh, kw = p.shape
            want = g._weight_2d(w16, kh * kw * I, kh * kw * I).t()
        assert got.shape == want.shape
        assert torch.equal(got.float(), want.float())


def test_linear_gelu_dropout_fused():
    """Linear → GELU → dropout as GEMM + one elementwise pass; the backward regenerates the mask from (seed, offset).
    Check: keep rate, kept values = gelu(z)/(1-p), and gradients equal to autograd through the recovered mask."""
    ops = _ops()
    torch.manual_seed(3)
    T, K, N, p = 512, 128, 256, 0.25
    x = torch.randn(T, K, device=DEV).bfloat16().requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(N, K, device=DEV) * 0.1)
    b = torch.nn.Parameter(torch.randn(N, device=DEV) * 0.1)
    out = ops.linear(x, w, b, act="gelu", dropout_p=p, training=True)
    z = torch.nn.functional.linear(x.float(), w.bfloat16().float(), b)
    full = torch.nn.functional.gelu(z)
    mask = (out != 0)
    keep = mask.float().mean().item()
    assert abs(keep - (1 - p)) < 0.02, keep
    _close(out.float()[mask], (full / (1 - p))[mask], 3e-2, 3e-2, "kept values")
    gout = torch.randn_like(out)
    out.backward(gout)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().bfloat16().float().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True)
    ref = torch.nn.functional.gelu(torch.nn.functional.linear(xr, wr, br)) * mask.float() / (1 - p)
    ref.backward(gout.float())
    _relnorm(x.grad, xr.grad, 2e-2, "dx")
    _relnorm(w.grad, wr.grad, 2e-2, "dw")
    _relnorm(b.grad, br.grad, 2e-2, "db")
    # eval / p = 0: plain gelu
    out0 = ops.linear(x.detach(), w, b, act="gelu", dropout_p=p, training=False)
    _close(out0, full, 3e-2, 3e-2, "eval")


def test_meta_mixup_kernel_forward_and_dtheta():
    """Fused meta-mixup (crop + flip + normalise + sigmoid(theta) blend): the blend is linear in lam, so with a fixed
    (seed, offset) out(theta) = lam*A + (1-lam)*Bp for partner images A, Bp recovered from theta = ±inf; dtheta from the
    backward kernel must equal sum g*(A - Bp) * sigmoid'(theta)."""
    import ctypes
    from fdt.ops._lib import call
    torch.manual_seed(3)
    B, H, W = 16, 32, 32
    dev = "cuda"
    x = torch.randint(0, 256, (B, H, W, 3), dtype=torch.uint8, device=dev)
    perm = torch.randperm(B, device=dev)
    so = torch.tensor([12345, 7], dtype=torch.int64, device=dev)
    m3, s3 = (ctypes.c_float * 3)(0.49, 0.48, 0.45), (ctypes.c_float * 3)(0.25, 0.24, 0.26)

    def fwd(theta):
        out = torch.empty((B, 3, H, W), dtype=torch.bfloat16, device=dev, memory_format=torch.channels_last)
        call("fdt_input_aug_mixup_ex", x, perm, theta.contiguous(), B, 1, so, out, B, H, W, 4, 1, m3, s3)
        return out.permute(0, 2, 3, 1).float()          # [B, H, W, 3]

    A = fwd(torch.full((B,), 40.0, device=dev))         # lam = 1
    Bp = fwd(torch.full((B,), -40.0, device=dev))       # lam = 0
    theta = torch.randn(B, device=dev)
    lam = torch.sigmoid(theta)
    out = fwd(theta)
    ref = lam.view(B, 1, 1, 1) * A + (1 - lam.view(B, 1, 1, 1)) * Bp
    assert (out - ref).abs().max() < 0.05
    assert (A - Bp).abs().mean() > 0.1                  # the partners really differ (perm + independent crops)
    g = torch.randn(B, 3, H, W, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    dth = torch.zeros(B, device=dev)
    call("fdt_input_aug_mixup_bwd", x, perm, theta.contiguous(), 1, so, g, dth, B, H, W, 4, 1, m3, s3)
    want = (g.permute(0, 2, 3, 1).float() * (A - Bp)).sum((1, 2, 3)) * lam * (1 - lam)
    assert torch.allclose(dth, want, rtol=2e-2, atol=2e-2 * want.abs().max().item()), (dth - want).abs().max()


def test_mixup_ce_per_sample_lambda_gradient():
    from fdt import ops
    torch.manual_seed(4)
    B, C = 64, 10
    logits = torch.randn(B, C, device="cuda", requires_grad=True)
    ya, yb = torch.randint(0, C, (B,), device="cuda"), torch.randint(0, C, (B,), device="cuda")
    lam = torch.rand(B, device="cuda", requires_grad=True)
    loss = ops.mixup_cross_entropy(logits, ya, yb, lam)
    loss.backward()
    l2, lam2 = logits.detach().clone().requires_grad_(True), lam.detach().clone().requires_grad_(True)
    ref = ops.ref.mixup_cross_entropy(l2, ya, yb, lam2)
    ref.backward()
    assert torch.allclose(loss, ref, rtol=1e-4, atol=1e-5)
    assert torch.allclose(logits.grad, l2.grad, rtol=1e-3, atol=1e-6)
    assert torch.allclose(lam.grad, lam2.grad, rtol=1e-3, atol=1e-6)


def test_head_kernels_match_torch():
    """cls gather / scatter and the dropout + manifold-mixup kernel of the transformer head."""
    from fdt import ops
    torch.manual_seed(5)
    B, L, D = 16, 9, 512
    x = torch.randn(B, L, D, device="cuda", requires_grad=True)
    cls = ops.cls_gather(x)
    assert cls.dtype == torch.bfloat16 and torch.equal(cls, x[:, 0, :].bfloat16())
    g = torch.randn(B, D, device="cuda").bfloat16()
    cls.backward(g)
    want = torch.zeros(B, L, D, device="cuda")
    want[:, 0, :] = g.float()
    assert torch.equal(x.grad, want)
    z = torch.randn(B, D, device="cuda").bfloat16().requires_grad_(True)
    perm = torch.randperm(B, device="cuda")
    lam = torch.tensor(0.3, device="cuda")
    out = ops.head_mix(z, perm, lam, 0.0, True)
    ref = 0.3 * z.float() + 0.7 * z.float()[perm]
    assert torch.allclose(out.float(), ref, rtol=2e-2, atol=2e-2)
    go = torch.randn(B, D, device="cuda").bfloat16()
    out.backward(go)
    zr = z.detach().float().requires_grad_(True)
    (0.3 * zr + 0.7 * zr[perm]).backward(go.float())
    assert torch.allclose(z.grad.float(), zr.grad, rtol=2e-2, atol=2e-2)
    # with dropout: same mask in forward and backward (d out / d z is the mask pattern), mean preserved
    z2 = torch.ones(B, D, device="cuda").bfloat16().requires_grad_(True)
    o2 = ops.head_mix(z2, None, None, 0.1, True)
    o2.float().sum().backward()
    assert torch.allclose(o2.float(), z2.grad.float(), rtol=1e-2, atol=1e-2)
    assert abs(o2.float().mean().item() - 1.0) < 0.05


def test_ngd_gpu_trajectory_matches_reference_and_uses_no_library_gemm():
    """8 NGD steps on the GPU (native kernels: tcgen05 3xTF32 products incl. the padded-pitch and split-K forms, short-axis
    and whole-factor-in-shared-memory kernels, Jacobi eigensolver) against the UNMODIFIED reference NGD on the same
    gradients; every factor has N >= R (rank-deficient factors amplify eigensolver rounding, see tests/test_ngd.py).
    Under strict mode: no cuBLAS / ATen fallback anywhere in the optimizer, warm start included."""
    from _refmods import load
    from fdt.ops import _lib
    from fdt.optim.ngd import NGD
    from fdt.parallel.flat import FlatParameters
    ref = load("ngd_optimizer")
    # conv (tc + short axes), a stem-like [48,5,3,3] (shared-memory kernel: rows of 45 floats), Linear, an "embedding" whose row
    # count is not a multiple of 4 (padded pitch, split-K projection), a tensor with only short axes (10, 7, 3, 3)
    shapes = [(64, 32, 3, 3), (64, 32, 3, 3), (48, 5, 3, 3), (128, 64), (1002, 128), (2050, 96), (10, 7, 3, 3)]
    g = torch.Generator().manual_seed(0)
    p_ref = [torch.nn.Parameter((torch.randn(*s, generator=g) * 0.1).cuda()) for s in shapes]
    p_new = [torch.nn.Parameter(p.detach().clone()) for p in p_ref]
    FlatParameters(p_new, torch.device("cuda"))
    o_ref = ref.NGD(p_ref, lr=0.01, momentum=0.9, weight_decay=1e-4)
    o_new = NGD(p_new, lr=0.01, momentum=0.9, weight_decay=1e-4)
    gg = torch.Generator(device="cuda").manual_seed(1)
    _lib.set_strict(True)
    try:
        for step in range(8):
            for a, b in zip(p_ref, p_new):
                grad = torch.randn(a.shape, device="cuda", generator=gg)
                a.grad = grad.clone()
                b.grad.copy_(grad)
            o_ref.step()
            o_new.step()
    finally:
        _lib.set_strict(False)
    torch.cuda.synchronize()
    for s, a, b in zip(shapes, p_ref, p_new):
        rel = ((a - b).norm() / a.norm()).item()
        print(s, f"rel {rel:.2e}")
        assert rel < 1e-3, (s, rel)


def test_ngd_norm_chain_equals_per_axis_rescale(monkeypatch):
    """On steps without a Fisher refresh the per-axis norm restorations are folded into ONE rescale at the end of a
    tensor's chain of axes (fdt.ops.cuda_ops.NormChain): same parameters as rescaling after every axis
    (ngd_optimizer.py:168), to fp32 rounding, over 16 steps (refresh and plain steps mixed)."""
    from fdt.ops import _lib
    from fdt.optim.ngd import NGD
    from fdt.parallel.flat import FlatParameters
    shapes = [(64, 32, 3, 3), (64, 32, 3, 3), (128, 64), (256, 128, 1, 1), (10, 7, 3, 3)]
    g = torch.Generator().manual_seed(0)
    init = [(torch.randn(*s, generator=g) * 0.1).cuda() for s in shapes]
    outs = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("FDT_NGD_NORM_CHAIN", mode)
        ps = [torch.nn.Parameter(p.clone()) for p in init]
        FlatParameters(ps, torch.device("cuda"))
        opt = NGD(ps, lr=0.01, momentum=0.9, weight_decay=1e-4)
        gg = torch.Generator(device="cuda").manual_seed(1)
        before = _lib.LAUNCHES.by_name.get("fdt_ngd_rescale2", 0)
        _lib.set_strict(True)
        try:
            for step in range(16):
                for p in ps:
                    p.grad.copy_(torch.randn(p.shape, device="cuda", generator=gg))
                opt.step()
        finally:
            _lib.set_strict(False)
        torch.cuda.synchronize()
        used = _lib.LAUNCHES.by_name.get("fdt_ngd_rescale2", 0) - before
        assert (used > 0) == (mode == "1"), (mode, used)
        outs[mode] = [p.detach().clone() for p in ps]
    for s, a, b in zip(shapes, outs["0"], outs["1"]):
        rel = ((a - b).norm() / a.norm()).item()
        # fp32 rounding (1e-7 per step) grows ~100x over 16 NGD steps (atomics make even two identical runs differ by ~1e-5);
        # a chain that restored the wrong norm would be off by tens of per cent
        assert rel < 5e-4, (s, rel)


def test_batch_stat_normalisation_form_is_the_references():
    """The FusedConvBN form — UNBIASED variance, eps added to the standard deviation (resnet.py:85-100) — checked where it
    matters: 32 values per channel (biased vs unbiased differ by 1.6 %) and a conv output whose std is comparable to eps
    (eps inside vs outside the sqrt differ by tens of per cent).  Runs under strict mode: native kernels only."""
    from fdt.ops import _lib
    ops = _ops()
    torch.manual_seed(11)
    N, C, hw = 2, 64, 4
    x = torch.randn(N, C, hw, hw, device=DEV)
    w = torch.randn(C, C, 1, 1, device=DEV) / math.sqrt(C) * 2e-3          # conv output std ~ 2e-3, eps = 1e-3
    xr, wr = x.bfloat16().float(), w.bfloat16().float()
    y = torch.nn.functional.conv2d(xr, wr)
    mean = y.mean((0, 2, 3), keepdim=True)
    want = (y - mean) / (y.var((0, 2, 3), unbiased=True, keepdim=True).sqrt() + 1e-3)
    biased = (y - mean) / (y.var((0, 2, 3), unbiased=False, keepdim=True).sqrt() + 1e-3)
    inside = (y - mean) / (y.var((0, 2, 3), unbiased=True, keepdim=True) + 1e-3).sqrt()
    _lib.set_strict(True)
    try:
        out = ops.conv_bnstat_act(x.bfloat16().to(memory_format=torch.channels_last), w.clone(), 1, 0, 1e-3, "none", None)
    finally:
        _lib.set_strict(False)
    rel = lambda a, b: ((a.float() - b).norm() / b.norm()).item()
    assert rel(out, want) < 6e-3, rel(out, want)
    assert rel(out, want) < 0.5 * rel(out, biased) and rel(out, want) < 0.1 * rel(out, inside)


def test_mixup_sample_kernel_draws():
    """Device-side Beta(alpha, alpha) + random permutation (one launch): permutations are valid and differ between calls,
    lambda has the Beta(0.99, 0.99) mean / variance (1/2, 1/(4(2a+1)) = 0.0839), alpha <= 0 passes through."""
    from fdt.ops import cuda_ops as cu
    dev = torch.device("cuda")
    lams, perms = [], []
    for _ in range(400):
        lam, perm = cu.mixup_sample(0.99, 128, dev)
        lams.append(lam)
        perms.append(perm)
    P = torch.stack(perms)
    assert torch.equal(P.sort(dim=1).values, torch.arange(128, device=dev).expand(400, 128))
    assert len({tuple(p.tolist()) for p in perms[:50]}) == 50
    pos0 = (P == 0).float().argmax(1).float()             # where element 0 lands: uniform on 0..127
    assert abs(pos0.mean().item() - 63.5) < 8.0
    L = torch.stack(lams)
    assert (L > 0).all() and (L < 1).all()
    assert abs(L.mean().item() - 0.5) < 0.06 and abs(L.var().item() - 0.0839) < 0.025
    lam0, _ = cu.mixup_sample(0.0, 4, dev)
    assert float(lam0) == 0.0


def test_linear_dropout_add_sublayer_tail():
    """residual + dropout(x Wᵀ + b): forward and all four gradients vs fp32 (p = 0), and with p > 0 the backward mask equals
    the forward mask (d out / d bias is the column count of kept elements / (1 - p))."""
    from fdt import ops
    torch.manual_seed(6)
    T, K, N = 512, 1024, 512
    x = (torch.randn(4, T // 4, K, device=DEV) / math.sqrt(K)).bfloat16().requires_grad_(True)
    w = torch.randn(N, K, device=DEV, requires_grad=True)
    b = torch.randn(N, device=DEV, requires_grad=True)
    res = torch.randn(4, T // 4, N, device=DEV, requires_grad=True)
    g = torch.randn(4, T // 4, N, device=DEV)
    out = ops.linear_dropout_add(x, w, b, res, 0.0, True)
    assert out.dtype == torch.float32
    out.backward(g)
    xr = x.detach().float().requires_grad_(True)
    wr = w.detach().bfloat16().float().requires_grad_(True)
    br, rr = b.detach().clone().requires_grad_(True), res.detach().clone().requires_grad_(True)
    ref = rr + torch.nn.functional.linear(xr, wr, br)
    ref.backward(g)
    rel = lambda a, c: ((a.float() - c.float()).norm() / c.float().norm()).item()
    assert rel(out, ref) < 5e-3
    assert rel(x.grad, xr.grad) < 1e-2 and rel(w.grad, wr.grad) < 1e-2 and rel(b.grad, br.grad) < 1e-2
    assert torch.equal(res.grad, g)
    # dropout: forward mask == backward mask
    b2 = torch.zeros(N, device=DEV, requires_grad=True)
    x0 = torch.zeros(4, T // 4, K, device=DEV).bfloat16()
    ones = torch.ones(N, device=DEV)
    o2 = ops.linear_dropout_add(x0, w.detach(), (b2 + ones), torch.zeros_like(res), 0.1, True)   # = mask / (1 - p)
    o2.sum().backward()
    assert torch.allclose(b2.grad, o2.sum((0, 1)), rtol=2e-2, atol=1.0)
    assert abs(o2.mean().item() - 1.0) < 0.02
