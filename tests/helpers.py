"""Shared builders for the tests (configs of BASELINE.json, golden loading)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def vgg_cfg():
    """vgg_block(8)/Pool/vgg_block(32)/Pool/vgg_block(64) (reference README.md:57-64, config 1)."""
    from snn_for_object_detection_b200.modules import LIF, Conv, Norm, Pool

    def vgg_block(out_channels, kernel=3):
        return Conv(out_channels, kernel), Norm(), LIF()

    return [*vgg_block(8), Pool("S"), *vgg_block(32), Pool("S"), *vgg_block(64)]


def vgg_block_from_golden():
    from snn_for_object_detection_b200 import BlockGen

    g = golden("vgg_net.npz")
    block = BlockGen(2, vgg_cfg())
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd::")}
    block.load_state_dict(sd, strict=True)
    block.eval()
    return block, g


def tiny_yolo_from_golden():
    """Rebuilds the golden TinyYolo from its seed (weights are not stored) and checks the digest."""
    from oracle import net_restated as onr
    from oracle.make_golden import weights_digest
    from snn_for_object_detection_b200 import TinyYolo

    g = golden("tiny_yolo.npz")
    torch.manual_seed(int(g["seed"]))
    model = TinyYolo(num_classes=2)
    T, B, H, W = int(g["T"]), int(g["B"]), int(g["H"]), int(g["W"])
    X = (torch.rand(T, B, 2, H, W) < float(g["density"])).float()
    model.eval()
    onr.calibrate(model, X, gain=float(g["gain"]))
    model.eval()
    ok = abs(weights_digest(model.state_dict()) - float(g["digest"])) <= 1e-9 * abs(float(g["digest"]))
    return model, X, g, ok


def calibrated_tiny_yolo(seed=0, gain=5.0, calib_hw=(64, 64), T=6, B=2, density=0.07):
    from oracle import net_restated as onr
    from snn_for_object_detection_b200 import TinyYolo

    torch.manual_seed(seed)
    model = TinyYolo(num_classes=2)
    X = (torch.rand(T, B, 2, *calib_hw) < density).float()
    model.eval()
    onr.calibrate(model, X, gain=gain)
    model.eval()
    return model
