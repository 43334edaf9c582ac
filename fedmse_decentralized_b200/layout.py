"""Flat parameter layout of one client's autoencoder.

Every client's model lives as one row of a ``[clients, P_PAD]`` fp32 slab.  The
row is the concatenation of the eight ``state_dict`` tensors in ``state_dict``
order, so packing / unpacking is a pure reshape and the on-disk checkpoint keys
stay exactly those of the reference (``/root/reference/src/Model/
Shrink_Autoencoder.py:40-42``, ``:95-97``; shapes listed in SURVEY.md §2.7).

The row is padded to a multiple of 32 floats (128 B) so that a row is a whole
number of cache lines and every row start is 16-B aligned for bulk copies and
vector loads from peer GPUs.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Tuple

import torch

STATE_KEYS: Tuple[str, ...] = (
    "encoder.encoder_network.0.weight",
    "encoder.encoder_network.0.bias",
    "encoder.encoder_network.2.weight",
    "encoder.encoder_network.2.bias",
    "decoder.decoder_network.0.weight",
    "decoder.decoder_network.0.bias",
    "decoder.decoder_network.2.weight",
    "decoder.decoder_network.2.bias",
)

#: row pitch (floats) of every device-resident data shard: 115 features padded to
#: 128 so rows are 512 B (16-B aligned for TMA / bulk copies, K padded for MMA).
D_PAD = 128


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass(frozen=True)
class ParamLayout:
    """Offsets of the 8 tensors inside a packed parameter row."""

    dim: int = 115
    hidden: int = 27
    latent: int = 7

    @property
    def shapes(self) -> List[Tuple[int, ...]]:
        D, H, Z = self.dim, self.hidden, self.latent
        return [(H, D), (H,), (Z, H), (Z,), (H, Z), (H,), (D, H), (D,)]

    @property
    def sizes(self) -> List[int]:
        out = []
        for s in self.shapes:
            n = 1
            for d in s:
                n *= d
            out.append(n)
        return out

    @property
    def offsets(self) -> List[int]:
        off, acc = [], 0
        for n in self.sizes:
            off.append(acc)
            acc += n
        return off

    @property
    def n_params(self) -> int:
        return sum(self.sizes)

    @property
    def padded(self) -> int:
        return _round_up(self.n_params, 32)

    @property
    def nbytes(self) -> int:
        return 4 * self.n_params

    # ------------------------------------------------------------------ pack
    def pack(self, state: Dict[str, torch.Tensor], out: torch.Tensor | None = None) -> torch.Tensor:
        """state_dict -> flat fp32 row of length ``padded`` (padding zeroed)."""
        if out is None:
            out = torch.zeros(self.padded, dtype=torch.float32)
        flat = out.view(-1)
        for key, off, n, shape in zip(STATE_KEYS, self.offsets, self.sizes, self.shapes):
            t = state[key]
            if tuple(t.shape) != tuple(shape):
                raise ValueError(f"{key}: expected shape {shape}, got {tuple(t.shape)}")
            flat[off:off + n].copy_(t.detach().reshape(-1).to(torch.float32))
        return out

    def unpack(self, row: torch.Tensor, clone: bool = True) -> Dict[str, torch.Tensor]:
        """flat row -> ordered state_dict with the reference's keys and shapes."""
        from collections import OrderedDict

        row = row.detach().reshape(-1)
        sd = OrderedDict()
        for key, off, n, shape in zip(STATE_KEYS, self.offsets, self.sizes, self.shapes):
            t = row[off:off + n].reshape(shape)
            sd[key] = t.clone().cpu() if clone else t
        return sd

    def views(self, row: torch.Tensor) -> List[torch.Tensor]:
        """Eight tensor views (no copy) into a packed row, state_dict order."""
        row = row.reshape(-1)
        return [row[o:o + n].reshape(s) for o, n, s in zip(self.offsets, self.sizes, self.shapes)]

    def segment_ids(self) -> torch.Tensor:
        """int64 [padded] tensor mapping each slot to its tensor index (8 = pad)."""
        ids = torch.full((self.padded,), 8, dtype=torch.int64)
        for i, (o, n) in enumerate(zip(self.offsets, self.sizes)):
            ids[o:o + n] = i
        return ids


DEFAULT_LAYOUT = ParamLayout()
assert DEFAULT_LAYOUT.n_params == 6764 and DEFAULT_LAYOUT.padded == 6784
