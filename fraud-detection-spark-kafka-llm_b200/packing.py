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
