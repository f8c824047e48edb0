"""YOLOv8-like spiking detector configuration (role of the reference's ``models/tiny_yolo.py:10-89``)."""

from __future__ import annotations

from .generator import ListGen
from .modules.layer_gen import LI, LIF, Conv, Dense, Norm, Pass, Residual, Return, Tanh
from .soda import SODa


class TinyYolo(SODa):
    """C2f backbone (stride 4), three-level neck (strides 8/16/32), LI + Tanh head."""

    def backbone_cfgs(self) -> ListGen:
        return [*self._conv(64, 3, 2), *self._c2f(64, 2), *self._conv(128, 3, 2), *self._c2f(128, 3)]

    def neck_cfgs(self) -> ListGen:
        cfg = []
        for depth in (4, 3, 2):
            cfg += [*self._conv(256, 3, 2), *self._c2f(256, depth), Return()]
        return cfg

    def head_cfgs(self, box_out: int, cls_out: int) -> ListGen:
        prepare = [Conv(kernel_size=1), Norm(), LI(state_storage=self.hparams.state_storage), Tanh()]
        return [prepare, [Conv(box_out, 1)], [Conv(cls_out, 1)]]

    def _conv(self, out_channels: int = None, kernel: int = 3, stride: int = 1):
        return (
            Conv(out_channels, stride=stride, kernel_size=kernel),
            Norm(),
            LIF(state_storage=self.hparams.state_storage),
        )

    def _bottleneck(self, shortcut: bool = True):
        net = (*self._conv(),)
        return Residual([[*net], [Pass()]]) if shortcut else net

    def _rec_block(self, n: int, shortcut: bool):
        if n == 0:
            return []
        return (Dense([[self._bottleneck(shortcut), *self._rec_block(n - 1, shortcut)], [Pass()]]),)

    def _c2f(self, out_channels: int, n: int, shortcut: bool = True):
        half = int(out_channels / 2)
        return (
            Conv(out_channels, 1),
            Dense([[Conv(half, 1), *self._rec_block(n, shortcut)], [Conv(half, 1)]]),
            Conv(out_channels, 1),
        )
