"""tcgen05 GEMM / implicit-GEMM convolution against fp32 PyTorch references."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 64, 128), (1000, 256, 512), (8192, 512, 1024), (77, 1536, 512)])
@pytest.mark.parametrize("act,bias", [("none", False), ("relu", True), ("gelu", True)])
def test_gemm_nt(M, N, K, act, bias):
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(0)
    a = (torch.randn(M, K, device="cuda") / math.sqrt(K)).bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bi = torch.randn(N, device="cuda") if bias else None
    out = g.gemm_nt(a, b, bi, act)
    assert out is not None and out.dtype == torch.bfloat16
    ref = a.float() @ b.float().t()
    if bias:
        ref = ref + bi
    ref = {"none": lambda t: t, "relu": F.relu, "gelu": F.gelu}[act](ref)
    assert _rel(out, ref) < 6e-3, _rel(out, ref)
    assert (out.float() - ref).abs().max() < 0.05 * max(ref.abs().max().item(), 1.0)


def test_gemm_nt_fp32_out_and_stats():
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(1)
    M, N, K = 4096 + 64, 256, 192
    a = (torch.randn(M, K, device="cuda") / math.sqrt(K)).bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    sums = torch.zeros(2 * N, device="cuda")
    out = g.gemm_nt(a, b, None, "none", torch.float32, sums)
    ref = a.float() @ b.float().t()
    assert _rel(out, ref) < 1e-4
    assert _rel(sums[:N], ref.sum(0)) < 1e-3 and _rel(sums[N:], (ref * ref).sum(0)) < 1e-3


@pytest.mark.parametrize("N,C,O,HW,k", [(4, 64, 64, 32, 3), (8, 128, 128, 16, 3), (16, 256, 256, 8, 3), (16, 512, 512, 4, 3),
                                          (2, 64, 128, 32, 3), (4, 64, 256, 32, 1), (9, 128, 64, 16, 3)])
def test_conv_fwd(N, C, O, HW, k):
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(2)
    x = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    sums = torch.zeros(2 * O, device="cuda")
    y = g.conv_fwd(x, w, 1, k // 2, sums)
    assert y is not None and y.shape == (N, O, HW, HW) and y.is_contiguous(memory_format=torch.channels_last)
    ref = F.conv2d(x.float(), w.float(), padding=k // 2)
    assert _rel(y, ref) < 6e-3, _rel(y, ref)
    assert _rel(sums[:O], ref.sum((0, 2, 3))) < 5e-3 or (sums[:O] - ref.sum((0, 2, 3))).abs().max() < 0.5
    assert _rel(sums[O:], (ref * ref).sum((0, 2, 3))) < 5e-3


@pytest.mark.parametrize("N,C,O,HW,k", [(4, 64, 64, 32, 3), (8, 256, 128, 8, 3), (4, 128, 256, 16, 1)])
def test_conv_dgrad(N, C, O, HW, k):
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(3)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, O, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dx = g.conv_dgrad(dy, w, (N, C, HW, HW), 1, k // 2)
    assert dx is not None
    ref = F.conv_transpose2d(dy.float(), w.float(), padding=k // 2)
    assert _rel(dx, ref) < 6e-3, _rel(dx, ref)


@pytest.mark.parametrize("N,C,O,HW,k", [(4, 64, 64, 32, 3), (8, 128, 128, 16, 3), (16, 256, 256, 8, 3), (16, 512, 512, 4, 3),
                                          (4, 64, 256, 32, 1), (8, 256, 64, 16, 1), (8, 128, 192, 8, 3),
                                          # 4x4 maps with a batch that is not a multiple of the 4 images of a 64-pixel k-block
                                          (10, 128, 64, 4, 3), (17, 64, 128, 4, 1)])
def test_conv_wgrad(N, C, O, HW, k):
    from fdt.ops import wgrad_sm100 as wg
    torch.manual_seed(4)
    x = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, O, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dw = wg.conv_wgrad(dy, x, (O, C, k, k), 1, k // 2)
    assert dw is not None and dw.shape == (O, C, k, k) and dw.dtype == torch.float32
    w = torch.zeros(O, C, k, k, device="cuda", requires_grad=True)
    F.conv2d(x.float(), w, padding=k // 2).backward(dy.float())
    assert _rel(dw, w.grad) < 2e-3, _rel(dw, w.grad)
    # accumulate-into form adds on top of existing contents
    slot = torch.ones(O, k, k, C, device="cuda")
    assert wg.conv_wgrad(dy, x, (O, C, k, k), 1, k // 2, slot) is True
    assert _rel(slot.permute(0, 3, 1, 2) - 1.0, w.grad) < 2e-3


@pytest.mark.parametrize("M,N,K", [(8192, 512, 512), (8192, 1024, 512), (64, 8, 64), (4096, 1536, 192)])
def test_linear_wgrad(M, N, K):
    from fdt.ops import wgrad_sm100 as wg
    torch.manual_seed(5)
    gz = torch.randn(M, N, device="cuda").bfloat16()
    x = torch.randn(M, K, device="cuda").bfloat16()
    dw = wg.linear_wgrad(gz, x)
    assert dw is not None
    ref = gz.float().t() @ x.float()
    assert _rel(dw, ref) < 2e-3, _rel(dw, ref)


def test_linear_autograd_native():
    from fdt import ops
    torch.manual_seed(6)
    x = torch.randn(4, 128, 512, device="cuda", requires_grad=True)
    w = (torch.randn(1024, 512, device="cuda") / 22).requires_grad_(True)
    b = torch.randn(1024, device="cuda").requires_grad_(True)
    g = torch.randn(4, 128, 1024, device="cuda")
    out = ops.linear(x, w, b, "gelu")
    out.backward(g.bfloat16())
    xr, wr, br = [t.detach().clone().requires_grad_(True) for t in (x, w, b)]
    ops.ref.linear(xr.bfloat16().float(), wr.bfloat16().float(), br, "gelu").backward(g.bfloat16().float())
    assert _rel(out, ops.ref.linear(x.bfloat16().float(), w.bfloat16().float(), b, "gelu")) < 1e-2
    assert _rel(w.grad, wr.grad) < 2e-2 and _rel(b.grad, br.grad) < 2e-2 and _rel(x.grad, xr.grad) < 2e-2


@pytest.mark.parametrize("B,H,L,masked,compat", [(4, 8, 128, True, False), (3, 8, 77, True, False), (2, 4, 128, False, False),
                                                  (2, 8, 64, True, True)])
def test_attention_fwd_bwd(B, H, L, masked, compat):
    from fdt import ops
    from fdt.ops import attention_sm100 as at
    torch.manual_seed(7)
    D = 64

    def mk():
        return torch.randn(B, L, H * D, device="cuda").bfloat16().view(B, L, H, D).transpose(1, 2).requires_grad_(True)

    q, k, v = mk(), mk(), mk()
    km = None
    if masked:
        lens = torch.randint(L // 3, L + 1, (B,), device="cuda")
        km = (torch.arange(L, device="cuda")[None, :] < lens[:, None]).long()
    g = torch.randn(B, H, L, D, device="cuda").bfloat16()
    out = at.attention(q, k, v, km, 0.0, compat)
    assert out is not None and out.shape == (B, H, L, D)
    out.backward(g)
    qr, kr, vr = [t.detach().float().requires_grad_(True) for t in (q, k, v)]
    ref = ops.ref.attention(qr, kr, vr, km, 0.0, False, compat)
    ref.backward(g.float())
    assert _rel(out, ref) < 1e-2, _rel(out, ref)
    for name, a, b in (("dq", q.grad, qr.grad), ("dk", k.grad, kr.grad), ("dv", v.grad, vr.grad)):
        assert _rel(a, b) < 2e-2, (name, _rel(a, b))


def test_attention_dropout_is_unbiased_and_consistent():
    from fdt import ops
    from fdt.ops import attention_sm100 as at
    torch.manual_seed(8)
    B, H, L, D = 2, 8, 128, 64
    mk = lambda: torch.randn(B, L, H * D, device="cuda").bfloat16().view(B, L, H, D).transpose(1, 2)
    q, k, v = mk(), mk(), mk()
    base = at.attention(q, k, v, None, 0.0, False).float()
    acc = torch.zeros_like(base)
    n = 48
    for _ in range(n):
        acc += at.attention(q, k, v, None, 0.1, False).float()
    assert _rel(acc / n, base) < 0.08
    # backward regenerates the same mask: gradient of sum(out) w.r.t. v equals Pᵀ·1 with the dropped P
    vv = v.clone().requires_grad_(True)
    o = at.attention(q, k, vv, None, 0.1, False)
    o.float().sum().backward()
    assert torch.isfinite(vv.grad.float()).all()
    col = vv.grad.float()[..., 0]                       # Σ_q P_drop[q, key], identical for every d
    assert torch.allclose(vv.grad.float(), col.unsqueeze(-1).expand_as(vv.grad), rtol=2e-2, atol=2e-2)
    assert abs(col.sum().item() / (B * H * L) - 1.0) < 0.05      # rows of P_drop sum to ~1 in expectation


@pytest.mark.parametrize("N,C,O,HW,k,stride,pad", [(8, 128, 128, 32, 3, 2, 1), (8, 256, 512, 32, 1, 2, 0), (16, 3, 64, 32, 3, 1, 1),
                                                    (4, 512, 512, 8, 3, 2, 1), (8, 64, 64, 12, 3, 1, 1)])
def test_conv_im2col_path(N, C, O, HW, k, stride, pad):
    """Strided / stem / odd-width convolutions: explicit im2col + tcgen05 GEMM, all three directions."""
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(9)
    x = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    y = g.conv_fwd_im2col(x, w, stride, pad)
    xr, wr = x.float().requires_grad_(True), w.float().requires_grad_(True)
    ref = F.conv2d(xr, wr, stride=stride, padding=pad)
    assert y is not None and y.shape == ref.shape
    assert _rel(y, ref) < 6e-3
    dy = torch.randn_like(ref).bfloat16().contiguous(memory_format=torch.channels_last)
    ref.backward(dy.float())
    dx = g.conv_dgrad_im2col(dy, w, x.shape, stride, pad)
    assert dx is not None and _rel(dx, xr.grad) < 1e-2, _rel(dx, xr.grad)
    dw = g.conv_wgrad_im2col(dy, x, w.shape, stride, pad)
    assert dw is not None and dw.shape == w.shape and _rel(dw, wr.grad) < 3e-3, _rel(dw, wr.grad)


# ----------------------------------------------------------------------------- stride 2 without im2col / col2im
_S2 = [(8, 128, 128, 32, 3), (8, 256, 256, 16, 3), (16, 512, 512, 8, 3), (8, 256, 512, 32, 1), (8, 512, 1024, 16, 1),
       (16, 1024, 2048, 8, 1)]


@pytest.mark.parametrize("N,C,O,HW,k", _S2)
def test_conv_fwd_stride2(N, C, O, HW, k):
    """Strided convolution on the implicit-GEMM kernel (TMA element strides) vs F.conv2d."""
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(12)
    x = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    sums = torch.zeros(2 * O, device="cuda")
    y = g.conv_fwd(x, w, 2, k // 2, sums)
    assert y is not None and y.shape == (N, O, HW // 2, HW // 2)
    ref = F.conv2d(x.float(), w.float(), stride=2, padding=k // 2)
    assert _rel(y, ref) < 6e-3, _rel(y, ref)
    assert _rel(sums[O:], (ref * ref).sum((0, 2, 3))) < 5e-3


@pytest.mark.parametrize("N,C,O,HW,k", _S2)
@pytest.mark.parametrize("with_addend", [False, True])
def test_conv_dgrad_stride2(N, C, O, HW, k, with_addend):
    """Stride-2 data gradient as parity-class implicit GEMMs, optionally accumulating into the other branch's gradient."""
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(13)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, O, HW // 2, HW // 2, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    add = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last) if with_addend else None
    add_ref = add.float().clone() if with_addend else 0.0
    dx = g.conv_dgrad(dy, w, (N, C, HW, HW), 2, k // 2, None, add)
    assert dx is not None and dx.shape == (N, C, HW, HW)
    ref = F.conv_transpose2d(dy.float(), w.float(), stride=2, padding=k // 2, output_padding=1) + add_ref
    assert _rel(dx, ref) < 8e-3, _rel(dx, ref)


@pytest.mark.parametrize("N,C,O,HW,k", _S2)
def test_conv_wgrad_stride2(N, C, O, HW, k):
    from fdt.ops import wgrad_sm100 as wg
    torch.manual_seed(14)
    x = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, O, HW // 2, HW // 2, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dw = wg.conv_wgrad(dy, x, (O, C, k, k), 2, k // 2)
    assert dw is not None
    ref = torch.ops.aten.convolution_backward(dy.float(), x.float(), torch.zeros(O, C, k, k, device="cuda"), None, [2, 2],
                                              [k // 2, k // 2], [1, 1], False, [0, 0], 1, [False, True, False])[1]
    assert _rel(dw, ref) < 6e-3, _rel(dw, ref)


@pytest.mark.parametrize("N,C,O,HW,k", [(4, 64, 64, 32, 3), (4, 128, 256, 16, 1), (16, 512, 2048, 4, 1)])
def test_conv_dgrad_with_addend(N, C, O, HW, k):
    """Residual-fork gradient sum folded into the stride-1 dgrad epilogue."""
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(15)
    w = (torch.randn(O, C, k, k, device="cuda") / math.sqrt(C * k * k)).bfloat16().contiguous(memory_format=torch.channels_last)
    dy = torch.randn(N, O, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    add = torch.randn(N, C, HW, HW, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    dx = g.conv_dgrad(dy, w, (N, C, HW, HW), 1, k // 2, None, add)
    assert dx is not None
    ref = F.conv_transpose2d(dy.float(), w.float(), padding=k // 2) + add.float()
    assert _rel(dx, ref) < 6e-3, _rel(dx, ref)


def test_resnet50_step_takes_no_library_fallback():
    """A ResNet-50 training step at 32x32 under FDT_STRICT semantics: no cuDNN / cuBLAS / ATen-add fallback is taken
    (stride-2 convolutions, shortcut projections and the residual-gradient sums all run on the native kernels)."""
    from fdt.models.resnet import resnet50
    from fdt.ops import _lib
    from fdt.parallel.ddp import DataParallelEngine
    torch.manual_seed(0)
    eng = DataParallelEngine(resnet50(10).cuda())
    x = torch.randn(64, 3, 32, 32, device="cuda")
    before = dict(_lib.FALLBACKS.by_name)
    _lib.set_strict(True)
    try:
        eng.zero_grad()
        loss = eng(x).float().square().mean()
        loss.backward()
    finally:
        _lib.set_strict(False)
    torch.cuda.synchronize()
    assert torch.isfinite(loss)
    assert dict(_lib.FALLBACKS.by_name) == before
    assert torch.isfinite(eng.flat.grad_flat).all()


# ----------------------------------------------------------------------------- attention beyond one tile, packed QKV
@pytest.mark.parametrize("B,H,L", [(3, 8, 77), (2, 8, 128), (3, 4, 200), (2, 8, 384), (2, 8, 512), (5, 2, 130)])
@pytest.mark.parametrize("masked", [False, True])
def test_attention_long_sequences(B, H, L, masked):
    """KV-tile loop with online softmax (forward) and per-key-tile dK/dV + accumulated dQ (backward) vs the fp32 oracle,
    at the reference's sequence range (maxlen 512, padding='longest': transformer.py:35, transformer_test.py:97)."""
    from fdt import ops
    from fdt.ops import attention_sm100 as at
    torch.manual_seed(17)
    D = 64

    def mk():
        return torch.randn(B, L, H * D, device="cuda").bfloat16().view(B, L, H, D).transpose(1, 2).requires_grad_(True)

    q, k, v = mk(), mk(), mk()
    km = None
    if masked:
        lens = torch.randint(L // 3, L + 1, (B,), device="cuda")
        km = (torch.arange(L, device="cuda")[None, :] < lens[:, None]).long()
    g = torch.randn(B, H, L, D, device="cuda").bfloat16()
    out = at.attention(q, k, v, km, 0.0, False)
    assert out is not None and out.shape == (B, H, L, D)
    out.backward(g)
    qr, kr, vr = [t.detach().float().requires_grad_(True) for t in (q, k, v)]
    ref = ops.ref.attention(qr, kr, vr, km, 0.0, False, False)
    ref.backward(g.float())
    assert _rel(out, ref) < 1e-2, _rel(out, ref)
    for name, a, b in (("dq", q.grad, qr.grad), ("dk", k.grad, kr.grad), ("dv", v.grad, vr.grad)):
        assert _rel(a, b) < 2e-2, (name, _rel(a, b))


@pytest.mark.parametrize("B,L", [(4, 128), (3, 200), (2, 512)])
def test_attention_packed_qkv(B, L):
    """The packed [B, L, 3*h*d] form (output of the fused Q/K/V GEMM) equals attention on the three slices."""
    from fdt import ops
    from fdt.ops import attention_sm100 as at
    torch.manual_seed(18)
    H, D = 8, 64
    qkv = torch.randn(B, L, 3 * H * D, device="cuda").bfloat16().requires_grad_(True)
    lens = torch.randint(L // 2, L + 1, (B,), device="cuda")
    km = (torch.arange(L, device="cuda")[None, :] < lens[:, None]).long()
    g = torch.randn(B, L, H * D, device="cuda").bfloat16()
    o = at.attention_packed(qkv, H, km, 0.0, False)
    assert o is not None and o.shape == (B, L, H * D)
    o.backward(g)
    ref_in = qkv.detach().float().requires_grad_(True)
    q, k, v = [t.reshape(B, L, H, D).transpose(1, 2) for t in ref_in.split(H * D, dim=-1)]
    ref = ops.ref.attention(q, k, v, km, 0.0, False, False).transpose(1, 2).reshape(B, L, H * D)
    ref.backward(g.float())
    assert _rel(o, ref) < 1e-2, _rel(o, ref)
    assert _rel(qkv.grad, ref_in.grad) < 2e-2, _rel(qkv.grad, ref_in.grad)


def test_attention_dropout_long_sequence_consistent():
    """Dropout masks are regenerated identically in the backward pass for multi-tile sequences."""
    from fdt.ops import attention_sm100 as at
    torch.manual_seed(19)
    B, H, L, D = 2, 4, 300, 64
    mk = lambda: torch.randn(B, L, H * D, device="cuda").bfloat16().view(B, L, H, D).transpose(1, 2)
    q, k = mk(), mk()
    vv = mk().clone().requires_grad_(True)
    o = at.attention(q, k, vv, None, 0.1, False)
    o.float().sum().backward()
    col = vv.grad.float()[..., 0]
    assert torch.isfinite(vv.grad.float()).all()
    assert torch.allclose(vv.grad.float(), col.unsqueeze(-1).expand_as(vv.grad), rtol=2e-2, atol=2e-2)
    assert abs(col.sum().item() / (B * H * L) - 1.0) < 0.05


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 256, 128), (1000, 512, 512), (8192, 1536, 512), (333, 256, 192), (4096, 1024, 1024)])
@pytest.mark.parametrize("act,bias,addend", [("none", False, False), ("gelu", True, False), ("none", False, True)])
def test_gemm2_two_cta(M, N, K, act, bias, addend):
    """2-CTA kernel (thread-block clusters, tcgen05.mma.cta_group::2, 256 x 256 tiles) against fp32 matmul."""
    from fdt.ops import gemm_sm100 as g
    torch.manual_seed(21)
    a = (torch.randn(M, K, device="cuda") / math.sqrt(K)).bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    bi = torch.randn(N, device="cuda") if bias else None
    ad = torch.randn(M, N, device="cuda").bfloat16() if addend else None
    out = g.gemm2_nt(a, b, bi, act, ad)
    assert out is not None and out.dtype == torch.bfloat16
    ref = a.float() @ b.float().t()
    if bias:
        ref = ref + bi
    ref = {"none": lambda t: t, "gelu": F.gelu}[act](ref)
    if addend:
        ref = ref + ad.float()
    assert _rel(out, ref) < 6e-3, _rel(out, ref)
