"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

numpy restatement of the reference's event decode + binning, and the seeded synthetic ``.dat``
generator shared by tests and bench.

* ``decode_events``   follows ``prophesee_toolbox/src/io/dat_events_tools.py:100-108``
  (``x = w & 16383``, ``y = (w & 268419072) >> 14``, ``p = (w & 268435456) >> 28``)
* ``split_dat``       follows ``parse_header`` ``dat_events_tools.py:126-181``
* ``bin_st``          follows ``STPropheseeDataset.parse_data`` ``utils/datasets.py:402-433``
* ``bin_mt``          follows ``MTPropheseeDataset.parse_data`` ``utils/datasets.py:311-336``

Pinned bit-for-bit against the reference's own ``parse_data`` by ``tests/golden/binning_*.npz``.
"""

from __future__ import annotations

import numpy as np


def make_records(rng: np.random.Generator, n: int, t_lo: int, t_hi: int, H: int, W: int,
                 x_overflow: int = 0) -> np.ndarray:
    """``n`` Event2D records ``uint32 [n,2]`` (t, w), t sorted in [t_lo, t_hi); ``x_overflow`` > 0
    lets x exceed W-1 by up to that many pixels (the 1Mpx quirk the ST path clips, datasets.py:426)."""
    t = np.sort(rng.integers(t_lo, t_hi, size=n, dtype=np.int64)).astype(np.uint32)
    x = rng.integers(0, W + x_overflow, size=n, dtype=np.uint32)
    y = rng.integers(0, H, size=n, dtype=np.uint32)
    p = rng.integers(0, 2, size=n, dtype=np.uint32)
    w = x | (y << 14) | (p << 28)
    return np.stack([t, w], axis=1).astype("<u4")


def dat_bytes(records: np.ndarray, H: int, W: int) -> bytes:
    """Serialises records the way Prophesee ``*_td.dat`` files are laid out."""
    header = (
        b"% Data file containing CD events.\n% Version 2\n"
        + f"% Date 2026-01-01 00:00:00\n% Height {H}\n% Width {W}\n".encode()
    )
    return header + bytes([0x00, 0x08]) + np.ascontiguousarray(records, dtype="<u4").tobytes()


def split_dat(buf: bytes):
    """-> (records uint32 [n,2], ev_type, ev_size, (H, W))"""
    pos, H, W = 0, None, None
    while buf[pos : pos + 1] == b"%":
        end = buf.index(b"\n", pos)
        words = buf[pos:end].decode("latin-1").split()
        if len(words) > 2 and words[1] == "Height":
            H = int(words[2])
        if len(words) > 2 and words[1] == "Width":
            W = int(words[2])
        pos = end + 1
    ev_type, ev_size = buf[pos], buf[pos + 1]
    pos += 2
    n = (len(buf) - pos) // ev_size
    rec = np.frombuffer(buf, dtype="<u4", count=2 * n, offset=pos).reshape(n, 2)
    return rec, ev_type, ev_size, (H, W)


def decode_events(records: np.ndarray):
    t = records[:, 0].astype(np.int64)
    w = records[:, 1].astype(np.uint32)
    x = (w & np.uint32(16383)).astype(np.int64)
    y = ((w & np.uint32(268419072)) >> np.uint32(14)).astype(np.int64)
    p = ((w & np.uint32(268435456)) >> np.uint32(28)).astype(np.int64)
    return t, x, y, p


def bin_st(records: np.ndarray, first_event_time_us: int, step_us: int, T: int, H: int, W: int) -> np.ndarray:
    """ST path: drop t < t0, ts = (t - t0)//step, clip x to W-1, features[ts,p,y,x] = 1."""
    t, x, y, p = decode_events(records)
    keep = t >= first_event_time_us
    t, x, y, p = t[keep], x[keep], y[keep], p[keep]
    ts = (t - first_event_time_us) // step_us
    x = np.clip(x, 0, W - 1)
    f = np.zeros((T, 2, H, W), dtype=np.float32)
    f[ts, p, y, x] = 1
    return f


def bin_mt(records: np.ndarray, start_step: int, step_us: int, T: int, H: int, W: int) -> np.ndarray:
    """MT path: ts = t//step - start_step, no clipping."""
    t, x, y, p = decode_events(records)
    ts = t // step_us - start_step
    f = np.zeros((T, 2, H, W), dtype=np.float32)
    f[ts, p, y, x] = 1
    return f


def bin_batch(samples, t0_us, step_us: int, T: int, H: int, W: int, clip_x: bool = True) -> np.ndarray:
    """Batched oracle with the CUDA entry point's contract: drops t < t0 and bins >= T; returns [T,B,2,H,W]."""
    out = np.zeros((T, len(samples), 2, H, W), dtype=np.float32)
    for b, (rec, t0) in enumerate(zip(samples, t0_us)):
        t, x, y, p = decode_events(rec)
        ts = (t - t0) // step_us
        keep = (t >= t0) & (ts < T)
        if clip_x:
            x = np.clip(x, 0, W - 1)
        keep &= (x < W) & (y < H)
        out[ts[keep], b, p[keep], y[keep], x[keep]] = 1
    return out
