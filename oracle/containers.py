"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Pure-torch module trees with the reference's structure and ``state_dict`` keys, so that the CPU arm of
``bench.py`` (``--impl reference`` / ``cpu_baseline``) can walk a network with ``oracle.net_restated`` WITHOUT
importing the product package (round-1 verdict: the reference arm listed ``libsoda_b200.so`` as loaded).

Restates the construction side of the reference only - no arithmetic lives here:

* ``BlockGen``      models/generator.py:35-143   (branch lists, Residual / Dense / forward merge, state flags)
* layer generators  models/modules/layer_gen.py:96-254 (Conv -> nn.Conv2d(bias=False, padding=k//2), Norm ->
                    nn.BatchNorm2d with bias=None, LIF / LI -> cell markers the oracle executor dispatches on by
                    class name, Pool("S") -> SumPool2d marker, Pass -> nn.Identity, Return -> Storage)
* ``TinyYolo``      models/tiny_yolo.py:10-89, ``SODa.__init__`` models/soda.py:77-96, ``Head`` / ``HeadGen``
                    models/generator.py:359-538 (anchor sizes / ratios as frozen parameters, utils/anchors.py:26-27)
* weight init       models/generator.py:245-256 (kaiming_normal_ fan_out for convs, BatchNorm weight 1)
"""

from __future__ import annotations

from typing import List

import torch
from torch import nn


# ---- list markers and layer generators (names as in the reference) -----------------------------------------
class Residual(list):
    pass


class Dense(list):
    pass


class LIFCell(nn.Module):          # dispatched on by class name in oracle.net_restated.block_forward
    def forward(self, x, state=None):
        raise RuntimeError("marker module: the oracle executor supplies the arithmetic")


class LICell(LIFCell):
    pass


class SumPool2d(nn.Module):
    def __init__(self, kernel_size: int, stride: int):
        super().__init__()
        self.kernel_size, self.stride = kernel_size, stride


class Storage(nn.Module):
    def __init__(self):
        super().__init__()
        self.storage = None

    def forward(self, X):
        self.storage = X
        return X

    def get_storage(self):
        t, self.storage = self.storage, None
        return t


class Conv:
    def __init__(self, out_channels=None, kernel_size=3, stride=1):
        self.out_channels, self.kernel_size, self.stride = out_channels, kernel_size, stride

    def get(self, cin):
        out = cin if self.out_channels is None else self.out_channels
        return nn.Conv2d(cin, out, self.kernel_size, padding=int(self.kernel_size / 2), stride=self.stride, bias=False), out


class Norm:
    def get(self, cin):
        m = nn.BatchNorm2d(cin)
        m.bias = None
        return m, cin


class LIF:
    def get(self, cin):
        return LIFCell(), cin


class LI:
    def get(self, cin):
        return LICell(), cin


class Tanh:
    def get(self, cin):
        return nn.Tanh(), cin


class Pass:
    def get(self, cin):
        return nn.Identity(), cin


class Pool:
    def __init__(self, type: str, kernel_size: int = 2, stride=None):
        assert type == "S", "container: only the sum pool of the generator nets is needed"
        self.kernel_size, self.stride = kernel_size, kernel_size if stride is None else stride

    def get(self, cin):
        return SumPool2d(self.kernel_size, self.stride), cin


class Return:
    def __init__(self):
        self.out_channels = 0

    def get(self, cin):
        self.out_channels = cin
        return Storage(), cin


class BlockGen(nn.Module):
    def __init__(self, in_channels: int, cfgs):
        super().__init__()
        self.merge = "residual" if isinstance(cfgs, Residual) else "dense" if isinstance(cfgs, Dense) else "forward"
        if self.merge == "forward":
            cfgs = [cfgs]
        branches, self.branch_state, self.out_channels = [], [], 0
        for cfg in cfgs:
            layers, flags, ch = [], [], in_channels
            for g in cfg:
                if isinstance(g, list):
                    layer = BlockGen(ch, g)
                    ch = layer.out_channels
                    flags.append(True)
                else:
                    layer, ch = g.get(ch)
                    flags.append(isinstance(layer, LIFCell))
                layers.append(layer)
            branches.append(nn.ModuleList(layers))
            self.branch_state.append(flags)
            self.out_channels = self.out_channels + ch if self.merge == "dense" else ch
        self.net = nn.ModuleList(branches)


class _Gen(nn.Module):
    def __init__(self, cfg, in_channels):
        super().__init__()
        self.net = BlockGen(in_channels, cfg)
        self.out_channels = self.net.out_channels


class _AnchorGen(nn.Module):
    def __init__(self, sizes, ratios):
        super().__init__()
        self.sizes = nn.Parameter(sizes, requires_grad=False)
        self.ratios = nn.Parameter(ratios, requires_grad=False)


class _HeadGen(nn.Module):
    def __init__(self, cfg, in_channels):
        super().__init__()
        self.base_net = BlockGen(in_channels, [cfg[0]])
        self.box_net = BlockGen(self.base_net.out_channels, [cfg[1]])
        self.cls_net = BlockGen(self.base_net.out_channels, [cfg[2]])


class _Head(nn.Module):
    def __init__(self, cfg_fn, num_classes: int, in_shape: List[int]):
        super().__init__()
        self.num_classes = num_classes
        hi, lo, spp = 0.75, 0.08, 3
        sizes = torch.arange(lo, hi, (hi - lo) / (len(in_shape) * spp), dtype=torch.float32).reshape((-1, spp))
        ratios = torch.tensor((0.5, 1.0, 2), dtype=torch.float32)
        n_anchor = spp * len(ratios)
        for idx, ch in enumerate(in_shape):
            setattr(self, f"anchor_gen_{idx}", _AnchorGen(sizes[idx], ratios))
            setattr(self, f"model_{idx}", _HeadGen(cfg_fn(n_anchor * 4, n_anchor * (num_classes + 1)), ch))


def _init(module: nn.Module) -> None:
    for m in module.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.constant_(m.weight, 1)


class TinyYolo(nn.Module):
    """models/tiny_yolo.py:10-89 as a container: base_net / neck_net / head_net with the reference's keys."""

    def __init__(self, num_classes: int = 2):
        super().__init__()
        self.base_net = _Gen([*self._conv(64, 3, 2), *self._c2f(64, 2), *self._conv(128, 3, 2), *self._c2f(128, 3)], 2)
        _init(self.base_net)
        returns, neck = [], []
        for depth in (4, 3, 2):
            r = Return()
            returns.append(r)
            neck += [*self._conv(256, 3, 2), *self._c2f(256, depth), r]
        self.neck_net = _Gen(neck, self.base_net.out_channels)
        _init(self.neck_net)
        self.head_net = _Head(lambda box, cls: [[Conv(kernel_size=1), Norm(), LI(), Tanh()], [Conv(box, 1)], [Conv(cls, 1)]],
                              num_classes, [r.out_channels for r in returns])
        _init(self.head_net)

    @staticmethod
    def _conv(out=None, k=3, s=1):
        return Conv(out, stride=s, kernel_size=k), Norm(), LIF()

    def _rec(self, n):
        if n == 0:
            return []
        return (Dense([[Residual([[*self._conv()], [Pass()]]), *self._rec(n - 1)], [Pass()]]),)

    def _c2f(self, out, n):
        half = out // 2
        return Conv(out, 1), Dense([[Conv(half, 1), *self._rec(n)], [Conv(half, 1)]]), Conv(out, 1)


def vgg_net() -> BlockGen:
    """vgg_block(8)/Pool/vgg_block(32)/Pool/vgg_block(64) (reference README.md:57-64, BASELINE config 1)."""
    blk = lambda c: (Conv(c, 3), Norm(), LIF())  # noqa: E731
    net = BlockGen(2, [*blk(8), Pool("S"), *blk(32), Pool("S"), *blk(64)])
    _init(net)
    return net
