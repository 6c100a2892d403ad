"""CIFAR-variant ResNet family (18/34/50/101/152) with reference-compatible state-dict keys.

Capability parity with ``resnet.py:147-313`` of the reference: 3x3 stride-1 stem with
CELU(0.075), no max-pool, BasicBlock / BottleNeck stages, projection shortcuts,
global average pool + Linear.  The module tree reproduces the reference's key names
(``conv1.0.conv_weight``, ``convN_x.B.residual_function.{0,2,4}.conv_weight``,
``...shortcut.1.running_mean``, ``fc.weight`` — SURVEY.md §5.4) so checkpoints are
interchangeable, but execution is block-level: each block's ``forward`` issues *fused units*
(``conv → normalise → [+residual] → activation``) through :mod:`fdt.ops`, which on a B200
are sm_100a kernels working on NHWC bf16 tensors, and on CPU are the PyTorch oracle.

Semantics kept from the reference:
* ``ConvBNStat`` (the reference's ``FusedConvBN``, ``resnet.py:116-144``) has a single
  parameter ``conv_weight``, normalises with *batch* statistics in train and eval,
  ``(y - mean) / (sqrt(var_unbiased) + 1e-3)``, no affine, no running stats, stride 1 only.
* stride-2 3x3 convs and projection shortcuts use ``Conv2d`` + ``BatchNorm2d`` maths.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .. import ops


class ConvBNStat(nn.Module):
    """Parameter holder for a conv followed by batch-stat normalisation (no affine)."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=0, eps=1e-3):
        super().__init__()
        if stride != 1:
            raise ValueError("ConvBNStat supports stride 1 only (reference resnet.py:120)")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride, self.padding, self.eps = kernel_size, stride, padding, eps
        self.conv_weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size, kernel_size))
        # eligible for the tcgen05 wgrad kernel → its gradient can be all-reduced inside that kernel
        self.conv_weight._fdt_wgrad_fused = in_channels % 64 == 0 and out_channels % 8 == 0
        self.reset_parameters()

    def reset_parameters(self):
        bound = 1.0 / math.sqrt(self.in_channels * self.kernel_size * self.kernel_size)
        with torch.no_grad():
            self.conv_weight.uniform_(-bound, bound)

    def forward(self, x, act="none", residual=None, passthrough=False):
        return ops.conv_bnstat_act(x, self.conv_weight, self.stride, self.padding, self.eps, act, residual,
                                   getattr(self, "_fdt_arena", None), passthrough)

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, k={self.kernel_size}, pad={self.padding}, eps={self.eps}"


class Act(nn.Module):
    """Placeholder that keeps the reference's Sequential indices; the activation itself is
    fused into the preceding unit."""

    def __init__(self, kind):
        super().__init__()
        self.kind = kind

    def forward(self, x):
        return ops.ref.activation(x, self.kind)

    def extra_repr(self):
        return self.kind


def _conv_bn(conv: nn.Conv2d, bn: nn.BatchNorm2d, x, act="none", residual=None, passthrough=False):
    # num_batches_tracked of every BatchNorm is advanced by ONE multi-tensor add per forward pass (ResNet.forward) when the
    # model runs on the native path; the per-unit increment stays for the CPU oracle and for blocks used on their own
    nbt = None if getattr(bn, "_fdt_nbt_batched", False) else bn.num_batches_tracked
    return ops.conv_bnaffine_act(
        x, conv.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var, nbt,
        conv.stride[0], conv.padding[0], bn.eps, bn.momentum, bn.training, act, residual,
        getattr(bn, "_fdt_arena", None), passthrough)


class _Block(nn.Module):
    """A residual block's input feeds two branches, so its gradient is a sum.  With ``passthrough=True`` a fused unit also
    returns its input as a second output; the OTHER branch consumes that alias, so both gradients reach the unit's
    backward together and the sum happens in the epilogue of its data-gradient GEMM instead of in a separate add kernel
    (16 of them per ResNet-50 step)."""

    def _fork(self, x, first, act):
        """Returns (first unit's output, shortcut tensor)."""
        if len(self.shortcut) == 0:
            o, xp = first(x, act, True)
            return o, xp
        sc, xp = _conv_bn(self.shortcut[0], self.shortcut[1], x, passthrough=True)
        return first(xp, act, False), sc


class BasicBlock(_Block):
    """Two 3x3 units, CELU activations (resnet.py:147-190)."""
    expansion = 1

    def __init__(self, in_channels, out_channels, stride=1):
        super().__init__()
        self.stride = stride
        if stride != 1:
            self.residual_function = nn.Sequential(
                nn.Conv2d(in_channels, out_channels, 3, stride=stride, padding=1, bias=False),
                nn.BatchNorm2d(out_channels),
                Act("celu"),
                ConvBNStat(out_channels, out_channels * self.expansion, 3, padding=1))
        else:
            self.residual_function = nn.Sequential(
                ConvBNStat(in_channels, out_channels, 3, padding=1),
                Act("celu"),
                ConvBNStat(out_channels, out_channels * self.expansion, 3, padding=1))
        self.shortcut = nn.Sequential()
        if stride != 1 or in_channels != out_channels * self.expansion:
            self.shortcut = nn.Sequential(
                nn.Conv2d(in_channels, out_channels * self.expansion, 1, stride=stride, bias=False),
                nn.BatchNorm2d(out_channels * self.expansion))

    def forward(self, x):
        rf = self.residual_function
        if self.stride != 1:
            o, sc = self._fork(x, lambda t, a, pt: _conv_bn(rf[0], rf[1], t, act=a, passthrough=pt), "celu")
            last = rf[3]
        else:
            o, sc = self._fork(x, lambda t, a, pt: rf[0](t, act=a, passthrough=pt), "celu")
            last = rf[2]
        return last(o, act="celu", residual=sc)


class BottleNeck(_Block):
    """1x1 → 3x3 → 1x1 (x4) units, ReLU activations (resnet.py:193-227)."""
    expansion = 4

    def __init__(self, in_channels, out_channels, stride=1):
        super().__init__()
        self.stride = stride
        if stride != 1:
            self.residual_function = nn.Sequential(
                ConvBNStat(in_channels, out_channels, 1),
                Act("relu"),
                nn.Conv2d(out_channels, out_channels, 3, stride=stride, padding=1, bias=False),
                nn.BatchNorm2d(out_channels),
                Act("relu"),
                ConvBNStat(out_channels, out_channels * self.expansion, 1))
        else:
            self.residual_function = nn.Sequential(
                ConvBNStat(in_channels, out_channels, 1),
                Act("relu"),
                ConvBNStat(out_channels, out_channels, 3, padding=1),
                Act("relu"),
                ConvBNStat(out_channels, out_channels * self.expansion, 1))
        self.shortcut = nn.Sequential()
        if stride != 1 or in_channels != out_channels * self.expansion:
            self.shortcut = nn.Sequential(
                nn.Conv2d(in_channels, out_channels * self.expansion, 1, stride=stride, bias=False),
                nn.BatchNorm2d(out_channels * self.expansion))

    def forward(self, x):
        rf = self.residual_function
        o, sc = self._fork(x, lambda t, a, pt: rf[0](t, act=a, passthrough=pt), "relu")
        if self.stride != 1:
            o = _conv_bn(rf[2], rf[3], o, act="relu")
            last = rf[5]
        else:
            o = rf[2](o, act="relu")
            last = rf[4]
        return last(o, act="relu", residual=sc)


class ResNet(nn.Module):
    def __init__(self, block, num_block, num_classes=100):
        super().__init__()
        self.in_channels = 64
        self.conv1 = nn.Sequential(ConvBNStat(3, 64, 3, padding=1), Act("celu"))
        self.conv2_x = self._make_layer(block, 64, num_block[0], 1)
        self.conv3_x = self._make_layer(block, 128, num_block[1], 2)
        self.conv4_x = self._make_layer(block, 256, num_block[2], 2)
        self.conv5_x = self._make_layer(block, 512, num_block[3], 2)
        self.avg_pool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        # layout of the transposed filter the dX GEMM wants (FlatParameters.refresh_dgrad_shadow makes them all in
        # one launch per step): 0 = tap-rotated [Cin][kh][kw][Cout] (implicit GEMM), 1 = [K, Cout] (col2im path)
        # batch-statistics slots of all conv+BN units in one arena, cleared by a single fill per forward pass
        self._bn_total = 0
        self._bn_arena = None
        for m in self.modules():
            C = m.out_channels if isinstance(m, ConvBNStat) else (m.num_features if isinstance(m, nn.BatchNorm2d) else 0)
            if C:
                m._fdt_arena_off = self._bn_total
                self._bn_total += 4 * C
        for m in self.modules():
            if isinstance(m, ConvBNStat):
                m.conv_weight._fdt_dgrad_mode = 0
            elif isinstance(m, nn.Conv2d):
                # stride 1 and stride 2 "same"-padded convs: tap-rotated filter (implicit-GEMM dgrad); anything else: [K, Cout]
                native = (m.stride[0] in (1, 2) and 2 * m.padding[0] == m.kernel_size[0] - 1
                          and m.in_channels % 64 == 0 and m.out_channels % 64 == 0)
                m.weight._fdt_dgrad_mode = 0 if native else 1

    def _make_layer(self, block, out_channels, num_blocks, stride):
        layers = []
        for s in [stride] + [1] * (num_blocks - 1):
            layers.append(block(self.in_channels, out_channels, s))
            self.in_channels = out_channels * block.expansion
        return nn.Sequential(*layers)

    def forward(self, x):
        """``x``: [N, 3, H, W] (any float dtype / memory format).  On CUDA the activations
        are bf16 NHWC end to end; the logits come back in fp32."""
        if ops.use_native(x):
            x = ops._cu().to_nhwc_bf16(x)
            if self._bn_arena is None:
                self._bn_arena = ops._cu().BNArena()
                for m in self.modules():
                    if hasattr(m, "_fdt_arena_off"):
                        m._fdt_arena = (self._bn_arena, m._fdt_arena_off)
            self._bn_arena.begin(self._bn_total, x.device)     # also in eval: the FusedConvBN units always use batch statistics
            bns = [m for m in self.modules() if isinstance(m, nn.BatchNorm2d)]
            for m in bns:
                m._fdt_nbt_batched = True
            if self.training and bns:
                torch._foreach_add_([m.num_batches_tracked for m in bns], 1)
        else:
            for m in self.modules():
                if isinstance(m, nn.BatchNorm2d):
                    m._fdt_nbt_batched = False
        # Stage boundaries for the parallel engines: forward "stage s" = the s-th entry of STAGES (its parameters may still
        # be arriving — all-gather overlapped with the previous stage's compute); gradient "bucket k" = the stage whose
        # gradients are final once backward crosses the boundary in front of it (reduce-scatter overlapped with the rest
        # of backward).  Both calls are no-ops without an engine that asks for them.
        ops.stage_begin(self, 0)
        o = self.conv1[0](x, act="celu")
        o = self.conv2_x(o)
        ops.stage_begin(self, 1)
        o = ops.bucket_boundary(o, self, 2)
        o = self.conv3_x(o)
        ops.stage_begin(self, 2)
        o = ops.bucket_boundary(o, self, 1)
        o = self.conv4_x(o)
        ops.stage_begin(self, 3)
        o = ops.bucket_boundary(o, self, 0)
        o = self.conv5_x(o)
        return ops.pool_fc(o, self.fc.weight, self.fc.bias)

    # forward stages, in order; backward finishes them in reverse (gradient bucket k = stage NUM_STAGES-1-k)
    STAGES = (("conv1", "conv2_x"), ("conv3_x",), ("conv4_x",), ("conv5_x", "fc"))

    def stage_of(self, p) -> int:
        """Index of the forward stage that owns parameter ``p``."""
        if not hasattr(self, "_stage_index"):
            self._stage_index = {}
            for si, names in enumerate(self.STAGES):
                for n in names:
                    for q in getattr(self, n).parameters():
                        self._stage_index[id(q)] = si
        return self._stage_index[id(p)]

    def grad_bucket_of(self, p) -> int:
        return len(self.STAGES) - 1 - self.stage_of(p)


def resnet18(num_classes=10):
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes)


def resnet34(num_classes=10):
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes)


def resnet50(num_classes=10):
    return ResNet(BottleNeck, [3, 4, 6, 3], num_classes)


def resnet101(num_classes=10):
    return ResNet(BottleNeck, [3, 4, 23, 3], num_classes)


def resnet152(num_classes=10):
    return ResNet(BottleNeck, [3, 8, 36, 3], num_classes)
