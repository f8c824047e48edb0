"""Debug: layer-major vs per-step protocol, first cell whose final state differs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from helpers import tiny_yolo_from_golden

def flat(state, acc=None, path=""):
    acc = [] if acc is None else acc
    if state is None:
        return acc
    if isinstance(state, list):
        for k, s in enumerate(state):
            flat(s, acc, path + f".{k}")
    elif hasattr(state, "v"):
        acc.append((path, state))
    return acc

model, X, g, _ = tiny_yolo_from_golden()
model = model.cuda()
Xd = X[:, :1].cuda()
with torch.no_grad():
    (anc, cls_seq, bbox_seq), s_seq = model.forward_seq(Xd)
    st, preds = None, None
    for ts in Xd:
        preds, st = model._forward_impl(ts, st)
a, b = flat(s_seq), flat(st)
print(len(a), len(b))
for (pa, sa), (pb, sb) in zip(a, b):
    dv = float((sa.v - sb.v).abs().max()); di = float((sa.i - sb.i).abs().max())
    print(pa, tuple(sa.v.shape), f"dv {dv:.3e} di {di:.3e}")
print("cls", float((preds[1] - cls_seq).abs().max()), "bbox", float((preds[2] - bbox_seq).abs().max()))
