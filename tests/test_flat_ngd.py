"""NGD on flat-buffer-backed parameters (zero-copy stacked views, channels_last conv weights,
fused momentum/update range) must follow the same trajectory as NGD on loose tensors."""
import torch

from fdt.optim.ngd import NGD
from fdt.parallel.flat import FlatParameters


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    shapes = [(16, 8, 3, 3), (12, 16, 1, 1), (16, 8, 3, 3), (10, 12), (12,), (10, 12)]
    return [torch.nn.Parameter(torch.randn(*s, generator=g, dtype=torch.float32) * 0.1) for s in shapes]


def test_flat_matches_loose():
    loose = _params(0)
    flatp = [torch.nn.Parameter(p.detach().clone()) for p in loose]
    flat = FlatParameters(flatp, "cpu")
    assert not flatp[0].data.is_contiguous() and flatp[0].shape == (16, 8, 3, 3)
    for a, b in zip(loose, flatp):
        assert torch.equal(a.data, b.data)
    o1 = NGD(loose, lr=0.05, momentum=0.9, weight_decay=1e-4)
    o2 = NGD(flatp, lr=0.05, momentum=0.9, weight_decay=1e-4)
    g = torch.Generator().manual_seed(1)
    for step in range(6):
        flat.zero_grad()
        for a, b in zip(loose, flatp):
            grad = torch.randn(a.shape, generator=g)
            a.grad = grad.clone()
            b.grad.copy_(grad)
        o1.step()
        o2.step()
        for i, (a, b) in enumerate(zip(loose, flatp)):
            if a.dim() == 1:
                continue  # rank-deficient factors are noise-sensitive (see test_ngd.py)
            assert torch.allclose(a, b, rtol=2e-3, atol=2e-5), f"step {step} param {i}: {(a - b).abs().max()}"
    # the conv-weight shape group really used the zero-copy physical view
    sg = [s for s in o2._groups[0] if s.shape == (16, 8, 3, 3)][0]
    assert sg._flat is not None and sg.axis_map == [0, 3, 1, 2]
    assert o2._flat_span(0, o2._groups[0]) is not None


def test_sharded_layout_regions():
    ps = _params(2)
    flat = FlatParameters(ps, "cpu", world_size=4, shard=True)
    sizes = {hi - lo for lo, hi in flat.region}
    assert len(sizes) == 1 and flat.total == 4 * next(iter(sizes))
    owned = [flat.owned_params(r) for r in range(4)]
    assert sum(len(o) for o in owned) == len(ps)
    for r, (lo, hi) in enumerate(flat.region):
        for g in flat.groups:
            if g.owner == r:
                assert lo <= g.offset and g.offset + g.size <= hi


def test_sharded_layout_cuts_oversized_shape_groups():
    """FSDP layout: a shape group larger than a rank's fair share is cut into whole-tensor parts, every parameter has
    exactly one owner, and the padded regions stay close to total / world."""
    import torch
    from fdt.models.resnet import resnet50
    from fdt.parallel.flat import FlatParameters
    params = list(resnet50(10).parameters())
    real = sum(p.numel() for p in params)
    ref = [p.detach().clone() for p in params]
    flat = FlatParameters(params, device="cpu", world_size=8, shard=True)
    assert flat.total / real < 1.25, flat.total / real                  # was 2.4x with whole shape groups
    owned = [flat.owned_params(r) for r in range(8)]
    ids = [id(p) for ps in owned for p in ps]
    assert len(ids) == len(set(ids)) == len(params)
    for r, (lo, hi) in enumerate(flat.region):
        assert hi - lo == flat.total // 8
    loads = [sum(p.numel() for p in ps) for ps in owned]
    assert max(loads) < 1.25 * real / 8 + max(p.numel() for p in params)
    for p, q in zip(params, ref):                                          # values survived the move into the flat buffer
        assert torch.equal(p.detach(), q)


def test_linear_dropout_dispatch_cpu():
    import torch
    from fdt import ops
    torch.manual_seed(0)
    x = torch.randn(64, 32)
    w, b = torch.randn(48, 32), torch.randn(48)
    full = torch.nn.functional.gelu(torch.nn.functional.linear(x, w, b))
    assert torch.allclose(ops.linear(x, w, b, act="gelu", dropout_p=0.5, training=False), full)
    out = ops.linear(x, w, b, act="gelu", dropout_p=0.5, training=True)
    kept = out != 0
    assert 0.35 < kept.float().mean().item() < 0.65
    assert torch.allclose(out[kept], 2.0 * full[kept], atol=1e-6)
