"""Prophesee ``*_td.dat`` streams -> occupancy frames on the GPU.

Host side: header parse and time slicing of the raw 8-byte records (file I/O; reference
``prophesee_toolbox/src/io/dat_events_tools.py:126-181`` and ``psee_loader.py:123-165``).
Device side: Event2D unpack + time binning + ``features[ts, p, y, x] = 1`` in one CUDA kernel
(``soda_bin_events``; reference ``dat_events_tools.py:100-108`` + ``utils/datasets.py:311-344``
(MT) / ``:378-435`` (ST)).  Records are NOT decoded on the host: the raw ``{u32 t, u32 w}`` pairs
are copied to the device as they lie in the file (8 B/event instead of a 583 KB fp32 frame per step).
"""

from __future__ import annotations

from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import ops

__all__ = ["parse_dat_header", "DatReader", "pack_samples", "events_to_frames", "GEOMETRY"]

GEOMETRY = {"gen1": (240, 304), "1mpx": (720, 1280)}  # (H, W), utils/datasets.py:209-217


def parse_dat_header(buf: Union[bytes, memoryview]) -> Tuple[int, int, int, Tuple[Optional[int], Optional[int]]]:
    """Returns ``(payload_offset, ev_type, ev_size, (height, width))`` of a ``.dat`` byte buffer.

    Header = ASCII lines starting with ``%``; then two bytes (event type, event size)."""
    pos, height, width = 0, None, None
    n = len(buf)
    while pos < n and buf[pos : pos + 1] == b"%":
        end = bytes(buf[pos : min(n, pos + 4096)]).find(b"\n")
        if end < 0:
            raise ValueError("unterminated .dat header line")
        words = bytes(buf[pos : pos + end]).decode("latin-1").split()
        if len(words) > 2:
            if words[1] == "Height":
                height = int(words[2])
            elif words[1] == "Width":
                width = int(words[2])
        pos += end + 1
    if pos == 0:
        return 0, 0, 8, (height, width)  # headerless file: reference defaults (dat_events_tools.py:170-172)
    if pos + 2 > n:
        raise ValueError("truncated .dat file: no event type / size bytes")
    ev_type, ev_size = int(buf[pos]), int(buf[pos + 1])
    return pos + 2, ev_type, ev_size, (height, width)


class DatReader:
    """Time-sliced reader of one ``.dat`` recording; yields RAW records (``uint32 [n, 2]`` = t, w).

    ``load_delta_t`` follows the reference loader's contract: returns the events with
    ``current_time <= t < current_time + delta_t`` and advances ``current_time`` by ``delta_t``."""

    def __init__(self, source: Union[str, bytes, np.ndarray]):
        if isinstance(source, str):
            raw = np.memmap(source, dtype=np.uint8, mode="r")
        elif isinstance(source, (bytes, bytearray, memoryview)):
            raw = np.frombuffer(source, dtype=np.uint8)
        else:
            raw = np.asarray(source, dtype=np.uint8)
        off, self.ev_type, self.ev_size, (self.height, self.width) = parse_dat_header(memoryview(raw))
        if self.ev_size != 8:
            raise ValueError(f"unsupported event size {self.ev_size} (Event2D records are 8 bytes)")
        n = (raw.size - off) // 8
        self.records = np.frombuffer(raw, dtype="<u4", count=2 * n, offset=off).reshape(n, 2)
        self.reset()

    def reset(self) -> None:
        self.current_time = 0
        self._pos = 0
        self.done = self.records.shape[0] == 0

    def event_count(self) -> int:
        return int(self.records.shape[0])

    def load_delta_t(self, delta_t: int) -> np.ndarray:
        if delta_t < 1:
            raise ValueError("load_delta_t(): delta_t must be at least 1 micro-second")
        if self.done:
            return self.records[0:0]
        final_time = self.current_time + int(delta_t)
        t = self.records[:, 0]
        end = self._pos + int(np.searchsorted(t[self._pos :], final_time, side="left"))
        out = self.records[self._pos : end]
        self._pos = end
        self.current_time = final_time
        if end >= self.records.shape[0]:
            self.done = True
        return out


def pack_samples(samples: Sequence[np.ndarray], pinned: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """Concatenates per-sample raw record arrays into one (pinned) host buffer + int64 offsets."""
    counts = [int(s.shape[0]) for s in samples]
    offs = np.zeros(len(samples) + 1, dtype=np.int64)
    np.cumsum(counts, out=offs[1:])
    buf = torch.empty((max(int(offs[-1]), 1), 2), dtype=torch.int32)
    if pinned and torch.cuda.is_available():
        buf = buf.pin_memory()
    view = buf.numpy().view(np.uint32)
    for s, a, b in zip(samples, offs[:-1], offs[1:]):
        view[a:b] = s
    return buf[: max(int(offs[-1]), 1)], torch.from_numpy(offs)


def events_to_frames(
    events: torch.Tensor,
    sample_offsets: torch.Tensor,
    t0_us: Union[torch.Tensor, Sequence[int]],
    step_us: int,
    T: int,
    H: int,
    W: int,
    clip_x: bool = True,
    sorted_by_time: bool = True,
    device: Union[str, torch.device] = "cuda",
    out: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Raw records (host or device) -> ``frames[T, B, 2, H, W]`` fp32 {0,1} on the device.

    ST dataset semantics: ``t0_us = first_event_time_us``, ``clip_x=True``;
    MT dataset semantics: ``t0_us = start_step * step_us``, ``clip_x=False``."""
    dev = torch.device(device)
    if not isinstance(t0_us, torch.Tensor):
        t0_us = torch.tensor(list(t0_us), dtype=torch.int64)
    ev = events.to(dev, non_blocking=True)
    offs = sample_offsets.to(dev, non_blocking=True)
    t0 = t0_us.to(dev, non_blocking=True)
    return ops.bin_events(ev, offs, t0, step_us, T, H, W, clip_x=clip_x, sorted_by_time=sorted_by_time, out=out)
