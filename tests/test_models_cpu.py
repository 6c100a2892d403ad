"""CPU tier: model parity with the reference modules (state-dict keys, parameter counts, forward and
gradient values with the reference's weights loaded), op-level oracles against the reference's own
autograd Functions."""
import pytest
import torch

from fdt import ops
from fdt.models.resnet import resnet18, resnet50
from fdt.models.transformer import Transformer

from _refmods import load


def test_resnet50_keys_and_param_count_match_reference():
    ref = load("resnet")
    a, b = resnet50(10), ref.resnet50(10)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys())
    assert all(sa[k].shape == sb[k].shape for k in sa)
    assert sum(p.numel() for p in a.parameters()) == 23_477_194
    assert len(list(a.parameters())) == 69


@pytest.mark.parametrize("factory", ["resnet18", "resnet50"])
def test_resnet_forward_backward_matches_reference(factory):
    ref = load("resnet")
    import fdt.models.resnet as ours
    torch.manual_seed(0)
    a, b = getattr(ours, factory)(10), getattr(ref, factory)(10)
    a.load_state_dict(b.state_dict())
    x = torch.randn(4, 3, 32, 32)
    ya, yb = a(x), b(x)
    assert torch.allclose(ya, yb, rtol=1e-4, atol=1e-4), (ya - yb).abs().max()
    ya.square().mean().backward()
    yb.square().mean().backward()
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(p.grad, q.grad, rtol=2e-3, atol=1e-5), n
    a.eval(), b.eval()
    with torch.no_grad():
        assert torch.allclose(a(x), b(x), rtol=1e-4, atol=1e-4)


def test_conv_bnstat_matches_reference_function_and_gradcheck():
    ref = load("resnet")
    torch.manual_seed(1)
    X = torch.rand(2, 3, 12, 12, dtype=torch.double, requires_grad=True)
    W = torch.rand(5, 3, 3, 3, dtype=torch.double, requires_grad=True)
    ours = ops.ref.conv_bnstat_act(X, W, 1, 1, 1e-3, "none", None)
    theirs = ref.FusedConvBN2DFunction.apply(X, W, 1, 1, 1e-3)
    assert torch.allclose(ours, theirs, rtol=1e-10, atol=1e-12)
    g = torch.rand_like(ours)
    ga = torch.autograd.grad(ours, (X, W), g)
    gb = torch.autograd.grad(theirs, (X, W), g)
    for u, v in zip(ga, gb):
        assert torch.allclose(u, v, rtol=1e-8, atol=1e-10)
    assert torch.autograd.gradcheck(lambda x, w: ops.ref.conv_bnstat_act(x, w, 1, 1, 1e-3, "celu", None), (X, W))


def test_transformer_keys_and_forward_match_reference():
    ref = load("transformer")
    ref.device = "cpu"
    torch.manual_seed(0)
    a = Transformer(4, 1000, n_layers=2, mask_compat=True).eval()
    b = ref.Transformer(4, 1000, n_layers=2, alpha=0.0).eval()     # alpha=0 → lam=0 in the reference's forward mixup
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys()) and all(sa[k].shape == sb[k].shape for k in sa)
    a.load_state_dict(sb)
    tok = torch.randint(0, 1000, (3, 17))
    typ = torch.zeros(3, 17, dtype=torch.long)
    mask = torch.ones(3, 17, dtype=torch.long)
    mask[:, 12:] = 0
    with torch.no_grad():
        ya = a(tok, typ, None, mask)
        enc = a.encode(tok, typ, mask)
        # reference: pooled → (lam=0 mix picks x[index]) → classifier; compare the un-mixed path
        emb = b.input_embeddings(tok, typ, torch.arange(512))
        x = emb + b.input_encodings(emb)
        for i in range(2):
            x = b.sublayer_ffn[i](b.sublayer_attention[i](x, mask.view(3, 1, 1, 17)))
        pooled = b.pooler(x)
        yb = b.classifier(pooled)
    assert torch.allclose(enc, pooled, rtol=1e-4, atol=1e-5), (enc - pooled).abs().max()
    assert torch.allclose(ya, yb, rtol=1e-4, atol=1e-5)
    assert sum(p.numel() for p in Transformer(4, 30522).parameters()) == 29_299_716


def test_layernorm_and_mlp_head_match_reference():
    ref = load("transformer")
    torch.manual_seed(2)
    ln = ref.LayerNorm(32)
    x = torch.randn(5, 7, 32, dtype=torch.float64)
    ln = ln.double()
    assert torch.allclose(ops.ref.layernorm_ref(x, ln.a_2, ln.b_2, ln.eps), ln(x), rtol=1e-12, atol=1e-12)
    W1, b1 = torch.randn(6, 4, requires_grad=True), torch.randn(1, 6, requires_grad=True)
    W2, b2 = torch.randn(3, 6, requires_grad=True), torch.randn(1, 3, requires_grad=True)
    X = torch.randn(5, 4, requires_grad=True)
    g = torch.randn(5, 3)
    ours = torch.autograd.grad(ops.ref.mlp_head(X, W1, b1, W2, b2, True), (X, W1, b1, W2, b2), g)
    theirs = torch.autograd.grad(ref.MLPScratch.apply(X, W1, b1, W2, b2), (X, W1, b1, W2, b2), g)
    for u, v in zip(ours, theirs):
        assert torch.allclose(u, v, rtol=1e-5, atol=1e-6)          # compat mode reproduces the mean-bias quirk
    fixed = torch.autograd.grad(ops.ref.mlp_head(X, W1, b1, W2, b2, False), (b1, b2), g)
    assert torch.allclose(fixed[0], theirs[2] * 5) and torch.allclose(fixed[1], theirs[4] * 5)


def test_mixup_and_loss_semantics():
    from fdt.data import mixup as mx
    torch.manual_seed(3)
    x, y = torch.randn(16, 3, 8, 8), torch.randint(0, 10, (16,))
    mixed, ya, yb, lam = mx.mixup_data(x, y, 0.99)
    assert mixed.shape == x.shape and lam.dim() == 0 and 0 <= float(lam) <= 1
    logits = torch.randn(16, 10)
    ce = torch.nn.CrossEntropyLoss()
    expect = float(lam) * ce(logits, ya) + (1 - float(lam)) * ce(logits, yb)        # resnet50_test.py:451-452
    assert torch.allclose(mx.mixup_criterion(logits, ya, yb, lam), expect, rtol=1e-5, atol=1e-6)
    meta = mx.MetaMixup(16)
    mixed, ya, yb, lam_s = meta(x, y)
    cen = torch.nn.CrossEntropyLoss(reduction="none")
    l = lam_s.reshape(-1)
    expect = torch.mean(l * cen(logits, ya) + (1 - l) * cen(logits, yb))           # resnet50_test.py:455-457
    assert torch.allclose(mx.mixup_criterion_meta(logits, ya, yb, l), expect, rtol=1e-5, atol=1e-6)
    assert meta.lam.requires_grad and mixed.requires_grad                          # λ is learnable (Q3)
    same = mx.mixup_data(x, torch.zeros(16, dtype=torch.long), 0.5, intra_only=True)[0]
    assert torch.allclose(same, x)                                                 # same-class pairs stay unmixed
