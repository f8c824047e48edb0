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


class STSampleStream:
    """Feature half of the reference's single-target dataset (``utils/datasets.py:346-435``) over several recordings,
    ending on the device: per recording the next labelled window is located exactly as ``STPropheseeDataset.parse_data``
    does (``:378-420``), the raw 8-byte records of that window are sliced out of the memory-mapped file (no decode on the
    host), staged in one of two pinned buffers and binned on the GPU (``soda_bin_events``) into ``X[T, B, 2, H, W]``.

    Label parsing stays with the caller (out of scope): ``label_steps[k]`` is the sorted list of time-step indices of
    recording ``k`` that carry at least one box passing the reference's size filter (``:392-399``).

    ``next_window(k)`` mirrors ``parse_data``'s control flow and returns ``(window | None, alive)``:
    ``(None, False)`` recording exhausted / no label left, ``(None, True)`` window skipped because it holds fewer than
    ``events_threshold`` events per step (``:417-418``), else ``((records, first_event_time_us, label_step), True)``."""

    def __init__(self, sources: Sequence[Union[str, bytes, np.ndarray]], label_steps: Sequence[Sequence[int]],
                 num_steps: int, time_shift: int, time_step_ms: int, height: int, width: int,
                 events_threshold: int = 4000, device: Union[str, torch.device] = "cuda"):
        if len(sources) != len(label_steps):
            raise ValueError("one label-step list per recording")
        self.readers = [DatReader(s) for s in sources]
        self.label_steps = [np.asarray(sorted(int(v) for v in ls), dtype=np.int64) for ls in label_steps]
        self.num_steps, self.time_shift = int(num_steps), int(time_shift)
        self.step_us = int(time_step_ms) * 1000
        self.height, self.width = int(height), int(width)
        self.events_threshold = int(events_threshold)
        self.device = torch.device(device)
        self.alive = [True] * len(self.readers)
        self._stage = [None, None]   # two pinned staging buffers, used alternately
        self._stage_done = [None, None]   # CUDA event recorded after the H2D copy that last read each buffer
        self._flip = 0
        self._next_k = 0             # round-robin cursor over the recordings (kept across next_batch calls)

    def next_window(self, k: int):
        rd, steps = self.readers[k], self.label_steps[k]
        if rd.done:
            return None, False
        start_time_us = rd.current_time
        start_step = start_time_us // self.step_us
        j = int(np.searchsorted(steps, start_step + self.num_steps, side="left"))
        if j >= steps.shape[0]:
            return None, False
        label_step = int(steps[j])
        first_label_time_us = label_step * self.step_us
        first_event_time_us = first_label_time_us - self.step_us * (self.num_steps - self.time_shift)
        rec = rd.load_delta_t(first_label_time_us + self.step_us * self.time_shift - start_time_us)
        lo = int(np.searchsorted(rec[:, 0], first_event_time_us, side="left")) if first_event_time_us > 0 else 0
        rec = rec[lo:]
        if (rec.shape[0] // self.num_steps) < self.events_threshold:
            return None, True
        return (rec, first_event_time_us, label_step), True

    def _pinned(self, n_rec: int) -> torch.Tensor:
        """The next staging buffer.  Its previous H2D copy is asynchronous (pinned memory, ``non_blocking``): the host
        may run several batches ahead of a GPU-bound consumer, so the buffer is only rewritten after the event
        recorded behind that copy has completed."""
        k = self._flip
        ev = self._stage_done[k]
        if ev is not None:
            ev.synchronize()
            self._stage_done[k] = None
        buf = self._stage[k]
        if buf is None or buf.shape[0] < n_rec:
            buf = torch.empty((max(n_rec, 1) * 5 // 4 + 1, 2), dtype=torch.int32)
            if torch.cuda.is_available():
                buf = buf.pin_memory()
            self._stage[k] = buf
        self._last = k
        self._flip ^= 1
        return buf

    def _staged(self) -> None:
        """Call right after the copy out of the buffer handed out last has been enqueued."""
        if torch.cuda.is_available() and self.device.type == "cuda":
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self._stage_done[self._last] = ev

    def next_batch(self, batch_size: int):
        """Up to ``batch_size`` accepted windows, visiting the live recordings round-robin (the reference's
        ``samples_generator`` without its shuffle).  Returns ``(X[T, B, 2, H, W] on the device, label_steps, recording
        indices)`` or ``None`` when every recording is exhausted."""
        picked = []
        while len(picked) < batch_size and any(self.alive):
            progressed = False
            for k in range(len(self.readers)):
                if not self.alive[k] or len(picked) >= batch_size:
                    continue
                win, self.alive[k] = self.next_window(k)
                progressed = True
                if win is not None:
                    picked.append((k,) + win)
            if not progressed:
                break
        if not picked:
            return None
        counts = [int(p[1].shape[0]) for p in picked]
        offs = np.zeros(len(picked) + 1, dtype=np.int64)
        np.cumsum(counts, out=offs[1:])
        buf = self._pinned(int(offs[-1]))
        view = buf.numpy().view(np.uint32)
        for p, a, b in zip(picked, offs[:-1], offs[1:]):
            view[a:b] = p[1]
        frames = events_to_frames(buf[: max(int(offs[-1]), 1)], torch.from_numpy(offs), [p[2] for p in picked],
                                  self.step_us, self.num_steps, self.height, self.width, clip_x=True,
                                  device=self.device)
        self._staged()
        return frames, [p[3] for p in picked], [p[0] for p in picked]


class MTSampleStream(STSampleStream):
    """Feature half of the reference's multi-target dataset (``utils/datasets.py:311-336``): consecutive windows of
    ``num_steps`` time steps per recording (a finished recording restarts, ``:321-322``), binned against
    ``start_step * step_us`` without clipping x.  ``next_window(k)`` returns ``((records, t0_us, start_step), True)``."""

    def __init__(self, sources, num_steps: int, time_step_ms: int, height: int, width: int,
                 device: Union[str, torch.device] = "cuda"):
        super().__init__(sources, [[] for _ in sources], num_steps, 0, time_step_ms, height, width, 0, device)

    def next_window(self, k: int):
        rd = self.readers[k]
        if rd.done:
            rd.reset()
        start_step = rd.current_time // self.step_us
        rec = rd.load_delta_t(self.num_steps * self.step_us)
        return (rec, start_step * self.step_us, int(start_step)), True

    def next_batch(self, batch_size: int):
        """One window from each recording in turn until ``batch_size`` samples: ``(X[T, B, 2, H, W], start_steps,
        recording indices)``."""
        picked = []
        while len(picked) < batch_size and self.readers:
            k = self._next_k % len(self.readers)     # the cursor survives the call: every recording gets its turn
            win, _ = self.next_window(k)             # (datasets.py:305-309 visits every loaded file)
            picked.append((k,) + win)
            self._next_k += 1
        counts = [int(p[1].shape[0]) for p in picked]
        offs = np.zeros(len(picked) + 1, dtype=np.int64)
        np.cumsum(counts, out=offs[1:])
        buf = self._pinned(int(offs[-1]))
        view = buf.numpy().view(np.uint32)
        for p, a, b in zip(picked, offs[:-1], offs[1:]):
            view[a:b] = p[1]
        frames = events_to_frames(buf[: max(int(offs[-1]), 1)], torch.from_numpy(offs), [p[2] for p in picked],
                                  self.step_us, self.num_steps, self.height, self.width, clip_x=False,
                                  device=self.device)
        self._staged()
        return frames, [p[3] for p in picked], [p[0] for p in picked]
