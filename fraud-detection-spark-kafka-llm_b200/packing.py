"""Pack Python strings into the (uint8 bytes, int64 offsets) layout the kernels read."""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def pack_strings(texts: Sequence) -> Tuple[np.ndarray, np.ndarray]:
    """list of str/bytes -> (uint8[total + pad], int64[n + 1]).  The byte buffer is padded with 32 zero
    bytes so that 16-byte-aligned window loads past the last dialogue stay inside the allocation."""
    enc = [t if isinstance(t, (bytes, bytearray)) else ("" if t is None else str(t)).encode("utf-8") for t in texts]
    offs = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum([len(e) for e in enc], out=offs[1:])
    buf = np.zeros(int(offs[-1]) + 32, dtype=np.uint8)
    if enc:
        joined = b"".join(enc)
        buf[:len(joined)] = np.frombuffer(joined, dtype=np.uint8)
    return buf, offs


class PackedText:
    """Packed dialogues on host (numpy) or device (torch).  ``offs`` int64[n+1]; ``data`` uint8, 16-byte
    aligned base, readable up to the next multiple of 16 past ``offs[-1]``."""

    def __init__(self, data, offs):
        self.data = data
        self.offs = offs

    @property
    def n_rows(self) -> int:
        return int(self.offs.shape[0]) - 1

    @property
    def is_device(self) -> bool:
        return hasattr(self.data, "is_cuda") and bool(self.data.is_cuda)

    @classmethod
    def from_strings(cls, texts: Sequence) -> "PackedText":
        return cls(*pack_strings(texts))

    def to_device(self, device):
        import torch

        if self.is_device:
            return self
        d = torch.from_numpy(self.data) if isinstance(self.data, np.ndarray) else self.data
        o = torch.from_numpy(self.offs) if isinstance(self.offs, np.ndarray) else self.offs
        return PackedText(d.to(device, non_blocking=True), o.to(device, non_blocking=True))

    def shard(self, rank: int, world: int) -> "PackedText":
        """Contiguous block of dialogues for ``rank`` of ``world``, balanced by BYTES rather than rows (dialogue
        lengths are skewed).  Returns a view: the byte buffer starts at the 16-byte boundary at
        or below the block's first dialogue and the offsets are rebased to it, which is all the kernels need."""
        lo, hi = shard_bounds(self.offs, world)[rank]
        first = int(self.offs[lo])
        base = first & ~15
        return PackedText(self.data[base:], self.offs[lo:hi + 1] - base)


def shard_bounds(offs, world: int):
    """[(row_lo, row_hi)] * world: consecutive row blocks whose byte counts are as equal as row boundaries allow
    (block k ends at the first row boundary at or past k/world of the bytes); every row lands in exactly one block."""
    if world < 1:
        raise ValueError("world must be >= 1")
    n = int(offs.shape[0]) - 1
    start, total = int(offs[0]), int(offs[n]) - int(offs[0])
    if hasattr(offs, "is_cuda"):
        import torch

        targets = torch.tensor([start + (total * k + world - 1) // world for k in range(1, world)], dtype=offs.dtype, device=offs.device)
        cuts = torch.searchsorted(offs.contiguous(), targets).tolist() if world > 1 else []
    else:
        targets = np.array([start + (total * k + world - 1) // world for k in range(1, world)], dtype=np.int64)
        cuts = np.searchsorted(np.asarray(offs), targets).tolist() if world > 1 else []
    edges = [0] + [min(max(int(c), 0), n) for c in cuts] + [n]
    for i in range(1, len(edges)):
        edges[i] = max(edges[i], edges[i - 1])
    return [(edges[k], edges[k + 1]) for k in range(world)]
