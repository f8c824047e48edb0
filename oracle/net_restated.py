"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

CPU restatement of the reference's network executor, in the reference's own TIME-MAJOR order:

* ``block_forward``   follows ``BlockGen.forward``       models/generator.py:169-198
  (merges: ``out[0]`` :163-164, ``torch.stack(out).sum(0)`` :145-146, ``torch.cat(out, 1)`` :157-158)
* ``neck_forward``    follows ``NeckGen.forward``        models/generator.py:343-351
* ``head_forward``    follows ``Head.forward`` / ``HeadGen.forward``  :415-457, :530-538
* ``soda_forward``    follows ``SODa.forward`` / ``_forward_impl``    models/soda.py:138-144, :235-244

It walks a module tree built by the product package only as a CONTAINER of structure and weights
(its ``state_dict`` keys equal the reference's); every number is produced here by stock torch CPU
ops (``nn.Conv2d.forward`` / ``nn.BatchNorm2d.forward`` / ``F.avg_pool2d`` — what the reference's
layers are) and by ``oracle.norse_restated`` for the neuron cells.  Pinned against the reference's
own Python by ``tests/golden/`` (``oracle/make_golden.py``).
"""

from __future__ import annotations

from typing import Any, List, Optional

import torch
import torch.nn.functional as F
from torch import nn

from . import norse_restated as nr

_LIF = nr.LIFCell()
_LI = nr.LICell()

# "fp32": nn.Conv2d.forward as the reference runs it.  "fp64": the convolution is accumulated in float64 and
# rounded once to fp32 ("the oracle should run convs in fp64 to adjudicate") - the distance
# between the two runs is the NOISE FLOOR of the comparison: what any two correct fp32 implementations with
# different summation order may differ by on this net.
CONV_DTYPE = "fp32"


def _conv(layer: nn.Conv2d, Y: torch.Tensor) -> torch.Tensor:
    if CONV_DTYPE == "fp32":
        return nn.Conv2d.forward(layer, Y)
    b = None if layer.bias is None else layer.bias.double()
    return F.conv2d(Y.double(), layer.weight.double(), b, layer.stride, layer.padding, layer.dilation,
                    layer.groups).float()


def _kind(m: nn.Module) -> str:
    return type(m).__name__


def _cell_step(layer, x, state, taps: Optional[dict]):
    """One step of a neuron cell (or StateStorage-wrapped cell), reference semantics."""
    name = _kind(layer)
    wrapped = name == "StateStorage"
    cell = layer.module if wrapped else layer
    ref = _LIF if _kind(cell) == "LIFCell" else _LI
    out, new_state = ref(x, state)
    if taps is not None:
        taps.setdefault(id(layer), []).append((out, new_state))
    return out, new_state


def block_forward(block, X, state=None, taps: Optional[dict] = None):
    out, out_state = [], []
    state = [None] * len(block.net) if state is None else state
    for branch, flags, branch_state in zip(block.net, block.branch_state, state):
        branch_state = [None] * len(branch) if branch_state is None else list(branch_state)
        Y = X
        for idx, (layer, is_state) in enumerate(zip(branch, flags)):
            name = _kind(layer)
            if taps is not None and name != "BlockGen":
                taps.setdefault(("in", id(layer)), []).append(Y)
            if name == "BlockGen":
                Y, branch_state[idx] = block_forward(layer, Y, branch_state[idx], taps)
            elif name in ("LIFCell", "LICell", "StateStorage"):
                Y, branch_state[idx] = _cell_step(layer, Y, branch_state[idx], taps)
            elif isinstance(layer, nn.Conv2d):
                Y = _conv(layer, Y)
            elif isinstance(layer, nn.BatchNorm2d):
                Y = nn.BatchNorm2d.forward(layer, Y)
            elif name == "SumPool2d":
                Y = F.avg_pool2d(Y, layer.kernel_size, layer.stride, 0) * layer.kernel_size * layer.kernel_size
            elif name == "SodaPool2d":
                fn = F.avg_pool2d if layer.mode == "A" else F.max_pool2d
                Y = fn(Y, layer.kernel_size, layer.stride)
            elif name == "Storage":
                layer.storage = Y
            elif is_state:
                raise NotImplementedError(f"oracle: stateful layer {name}")
            else:
                Y = layer(Y)  # Identity / Tanh / ReLU / SiLU / Upsample: stock torch modules
            if taps is not None and name not in ("BlockGen",):
                taps.setdefault(("act", id(layer)), []).append(Y)
        out.append(Y)
        out_state.append(branch_state)
    if block.merge == "residual":
        return torch.stack(out).sum(dim=0), out_state
    if block.merge == "dense":
        return torch.cat(out, dim=1), out_state
    return out[0], out_state


def neck_forward(neck, X, state, taps=None):
    _, state = block_forward(neck.net, X, state, taps)
    outs = [m.get_storage() for m in neck.net.modules() if _kind(m) == "Storage"]
    return outs, state


def anchors_restated(sizes, ratios, fmap):
    """utils/anchors.py:46-85, same fp32 op order."""
    in_height, in_width = fmap.shape[-2:]
    boxes_per_pixel = len(sizes) * len(ratios)
    steps_h, steps_w = 1.0 / in_height, 1.0 / in_width
    center_h = (torch.arange(in_height) + 0.5) * steps_h
    center_w = (torch.arange(in_width) + 0.5) * steps_w
    shift_y, shift_x = torch.meshgrid(center_h, center_w, indexing="ij")
    shift_y, shift_x = shift_y.reshape(-1), shift_x.reshape(-1)
    w = torch.cat([sizes * r for r in ratios]) * in_height / in_width
    h = torch.cat([sizes / r for r in ratios]) * in_width / in_height
    manip = torch.stack((-w, -h, w, h)).T.repeat(in_height * in_width, 1) / 2
    grid = torch.stack([shift_x, shift_y, shift_x, shift_y], dim=1).repeat_interleave(boxes_per_pixel, dim=0)
    return grid + manip


def head_forward(head, X: List[torch.Tensor], state, taps=None):
    state = [None] * len(X) if state is None else list(state)
    anchors, cls_preds, bbox_preds = [], [], []
    flat = lambda p: torch.flatten(torch.permute(p, (0, 2, 3, 1)), start_dim=1)  # noqa: E731
    for idx, fmap in enumerate(X):
        gen = getattr(head, f"anchor_gen_{idx}")
        anchors.append(anchors_restated(gen.sizes.detach().cpu(), gen.ratios.detach().cpu(), fmap))
        m = getattr(head, f"model_{idx}")
        st = [None] * 3 if state[idx] is None else list(state[idx])
        Y, st[0] = block_forward(m.base_net, fmap, st[0], taps)
        boxes, st[1] = block_forward(m.box_net, Y, st[1], taps)
        classes, st[2] = block_forward(m.cls_net, Y, st[2], taps)
        state[idx] = st
        bbox_preds.append(boxes)
        cls_preds.append(classes)
    anchors = torch.cat(anchors)
    cls = torch.cat([flat(p) for p in cls_preds], dim=1)
    cls = cls.reshape(cls.shape[0], -1, head.num_classes + 1)
    bbox = torch.cat([flat(p) for p in bbox_preds], dim=1)
    bbox = bbox.reshape(bbox.shape[0], -1, 4)
    return anchors, cls, bbox, state


def soda_step(model, X, state, taps=None):
    state = [None] * 3 if state is None else list(state)
    base_out, state[0] = block_forward(model.base_net.net, X, state[0], taps)
    neck_out, state[1] = neck_forward(model.neck_net, base_out, state[1], taps)
    anchors, cls, bbox, state[2] = head_forward(model.head_net, neck_out, state[2], taps)
    return (anchors, cls, bbox), state


@torch.no_grad()
def soda_forward(model, X, taps=None):
    """``for ts in X: preds, state = _forward_impl(ts, state)``; returns last preds and state."""
    state, preds = None, None
    for ts in X:
        preds, state = soda_step(model, ts, state, taps)
    return preds, state


@torch.no_grad()
def seq_block_forward(block, X, taps=None):
    """Runs a bare BlockGen over ``X[T,B,C,H,W]`` step by step; returns stacked outputs + state."""
    state, ys = None, []
    for ts in X:
        y, state = block_forward(block, ts, state, taps)
        ys.append(y)
    return torch.stack(ys), state


@torch.no_grad()
def calibrate(model: nn.Module, X: torch.Tensor, gain: float = 5.0, block=None) -> None:
    """Gives a randomly initialised net a non-trivial firing rate (with the
    reference's raw init every LIF layer is silent).  One time-major pass with every BatchNorm in
    train mode and ``momentum=None`` (cumulative average) sets the running statistics to the data
    statistics; gamma is then set to ``gain`` so LIF inputs are ~N(0, gain^2)."""
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    was_training = model.training
    for m in bns:
        m.reset_running_stats()
        m.momentum = None
        m.weight.fill_(gain)
        m.train()
    if block is not None:
        seq_block_forward(block, X)
    else:
        soda_forward(model, X)
    for m in bns:
        m.momentum = 0.1
    model.train(was_training)
    if not was_training:
        model.eval()
