"""Debug aid: which ops of a TinyYolo forward see / produce tensors tagged as spike counts (ops.mark_exact)."""
import torch
from snn_for_object_detection_b200 import ops, TinyYolo

names = ["conv2d", "conv_lif", "lif_scan", "lif_scan_out", "lif_pool_scan", "li_scan", "pool2d", "add_n", "channel_affine"]
for n in names:
    f = getattr(ops, n)
    def mk(n, f):
        def w(*a, **k):
            r = f(*a, **k)
            o = r[0] if isinstance(r, tuple) else r
            xin = a[0][0] if isinstance(a[0], (list, tuple)) else a[0]
            print(f"{n:14s} in {tuple(xin.shape)} b={ops.exact_bound(xin)} -> out {tuple(o.shape)} b={ops.exact_bound(o)} "
                  f"res={None if k.get('residual') is None else ops.exact_bound(k['residual'])} out_given={k.get('out') is not None}")
            return r
        return w
    setattr(ops, n, mk(n, f))

torch.manual_seed(0)
m = TinyYolo(num_classes=2).cuda().eval()
X = (torch.rand(2, 1, 2, 64, 64, device="cuda") < 0.2).float()
with torch.no_grad():
    m(X)
