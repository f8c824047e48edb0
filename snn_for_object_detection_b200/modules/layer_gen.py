"""Layer generators: tiny config objects whose ``get(in_channels)`` returns ``(module, out_channels)``.

Same names, constructor arguments and error behaviour as the reference's
``models/modules/layer_gen.py`` (:73-347) so that any ListGen written for it builds unchanged; the
modules produced are the CUDA-backed ones of this package.
"""

from __future__ import annotations

from typing import Optional, Tuple

from torch import nn

from .cells import LICell, LIFCell, SLICell, SNNCell, SynapseCell
from .conv_lstm import ConvLSTM
from .common import (SodaBatchNorm2d, SodaConv2d, SodaPool2d, SodaReLU, SodaSiLU, SodaTanh, SodaUpsample, StateStorage,
                     Storage, SumPool2d)

layers_list = (
    "Residual", "Dense", "LayerGen", "Pass", "Conv", "Norm", "LIF", "LI", "ReLU", "SiLU", "Tanh",
    "LSTM", "Pool", "Up", "Return", "Synapse", "SLI",
)

__all__ = list(layers_list)


class Residual(list):
    """List marker: the branches of this block are summed (models/modules/layer_gen.py:35-51)."""


class Dense(list):
    """List marker: the branches of this block are concatenated on channels (layer_gen.py:54-70)."""


class LayerGen:
    """Base class of the layer generators."""

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        raise NotImplementedError


class Pass(LayerGen):
    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        return nn.Identity(), in_channels


class Conv(LayerGen):
    """2-d convolution, ``bias=False``, ``padding = kernel_size // 2`` (layer_gen.py:106-136)."""

    def __init__(self, out_channels: int = None, kernel_size: int = 3, stride: int = 1):
        self.out_channels = out_channels
        self.kernel_size = kernel_size
        self.stride = stride

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        out = in_channels if self.out_channels is None else self.out_channels
        return SodaConv2d(
            in_channels, out, kernel_size=self.kernel_size, padding=int(self.kernel_size / 2),
            stride=self.stride, bias=False,
        ), out


class Pool(LayerGen):
    """``A`` average, ``M`` max, ``S`` sum pooling; default stride = kernel (layer_gen.py:139-173)."""

    def __init__(self, type: str, kernel_size: int = 2, stride: Optional[int] = None):
        self.kernel_size = kernel_size
        self.stride = stride if (stride is not None) else kernel_size
        if type not in ("A", "M", "S"):
            raise ValueError(f'[ERROR]: Non-existent pool type "{type}"!')
        self.type = type

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        if self.type == "S":
            return SumPool2d(self.kernel_size, self.stride), in_channels
        return SodaPool2d(self.kernel_size, self.stride, self.type), in_channels


class Up(LayerGen):
    def __init__(self, scale: int = 2, mode: str = "nearest"):
        self.scale = scale
        self.mode = mode

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        return SodaUpsample(scale_factor=self.scale, mode=self.mode), in_channels


class Norm(LayerGen):
    """BatchNorm2d, learnable bias off by default (layer_gen.py:197-214)."""

    def __init__(self, bias: bool = False):
        self.bias = bias

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        norm_layer = SodaBatchNorm2d(in_channels)
        if not self.bias:
            norm_layer.bias = None
        return norm_layer, in_channels


class LIF(LayerGen):
    def __init__(self, state_storage: bool = False):
        self.state_storage = state_storage

    def get(self, in_channels: int) -> Tuple[SNNCell, int]:
        module = LIFCell() if not self.state_storage else StateStorage(LIFCell())
        return module, in_channels


class LI(LayerGen):
    def __init__(self, state_storage: bool = False):
        self.state_storage = state_storage

    def get(self, in_channels: int) -> Tuple[SNNCell, int]:
        module = LICell() if not self.state_storage else StateStorage(LICell())
        return module, in_channels


class ReLU(LayerGen):
    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        return SodaReLU(), in_channels


class SiLU(LayerGen):
    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        return SodaSiLU(), in_channels


class Tanh(LayerGen):
    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        return SodaTanh(), in_channels


class Return(LayerGen):
    """Feature-pyramid tap; records its channel count as a side effect of ``get`` (layer_gen.py:305-318)."""

    out_channels: int

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        self.out_channels = in_channels
        return Storage(), in_channels


class LSTM(LayerGen):
    """ConvLSTM layer (reference ``models/modules/layer_gen.py:287-302``)."""

    def __init__(self, hidden_size: Optional[int] = None):
        self.hidden_size = hidden_size

    def get(self, in_channels: int) -> Tuple[nn.Module, int]:
        h_size = in_channels if self.hidden_size is None else self.hidden_size
        return ConvLSTM(in_channels, h_size), h_size


class Synapse(LayerGen):
    """Synaptic transmission layer (reference ``models/modules/layer_gen.py:321-328``) on ``soda_synapse_fwd``."""

    def get(self, in_channels: int) -> Tuple[SNNCell, int]:
        return SynapseCell(), in_channels


class SLI(LayerGen):
    """Saturable leaky integrator layer (reference ``models/modules/layer_gen.py:331-347``) on ``soda_sli_fwd``."""

    def __init__(self, state_storage: bool = False):
        self.state_storage = state_storage

    def get(self, in_channels: int) -> Tuple[SNNCell, int]:
        module = SLICell() if not self.state_storage else StateStorage(SLICell())
        return module, in_channels
