"""Flat, chunk-aligned layout of a list of tensors.

Every tensor starts at a multiple of CHUNK (2048) elements inside one flat vector, so that a 2048-element
chunk never straddles two tensors: per-tensor quantities (LAMB/LARS norms, weight decay, lr scale) become a
per-chunk table lookup inside the fused all-reduce kernel (csrc/fused_allreduce.cu), and every codec block
geometry of csrc/codec.cuh divides the vector. Padding elements are zero and stay zero under every update.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

CHUNK = 2048


class FlatLayout:
    def __init__(self, shapes: Sequence[torch.Size], names: Optional[Sequence[str]] = None):
        self.shapes = [torch.Size(s) for s in shapes]
        self.names = list(names) if names is not None else [f"t{i}" for i in range(len(self.shapes))]
        assert len(self.names) == len(self.shapes)
        self.numels = [int(torch.Size(s).numel()) for s in self.shapes]
        self.offsets: List[int] = []
        off = 0
        for n in self.numels:
            self.offsets.append(off)
            off += ((max(n, 1) + CHUNK - 1) // CHUNK) * CHUNK
        self.total = off
        self.n_chunks = off // CHUNK
        self.n_tensors = len(self.shapes)
        tensor_cbeg = [o // CHUNK for o in self.offsets] + [self.n_chunks]
        self.tensor_cbeg = torch.tensor(tensor_cbeg, dtype=torch.int64)
        chunk_tid = torch.empty(self.n_chunks, dtype=torch.int32)
        for t in range(self.n_tensors):
            chunk_tid[tensor_cbeg[t]:tensor_cbeg[t + 1]] = t
        self.chunk_tid = chunk_tid

    @classmethod
    def from_tensors(cls, tensors: Sequence[torch.Tensor], names=None) -> "FlatLayout":
        return cls([t.shape for t in tensors], names)

    def views(self, flat: torch.Tensor) -> List[torch.Tensor]:
        assert flat.numel() >= self.total
        return [flat[o:o + n].view(s) for o, n, s in zip(self.offsets, self.numels, self.shapes)]

    def pack(self, tensors: Sequence[torch.Tensor], out: Optional[torch.Tensor] = None,
             dtype=torch.float32, device=None) -> torch.Tensor:
        if out is None:
            out = torch.zeros(self.total, dtype=dtype, device=device if device is not None else tensors[0].device)
        for v, t in zip(self.views(out), tensors):
            v.copy_(t)
        return out

    def unpack(self, flat: torch.Tensor) -> List[torch.Tensor]:
        return [v.clone() for v in self.views(flat)]

    def shard_ranges(self, shares: Sequence[float]) -> List[tuple]:
        """Split the chunk range [0, n_chunks) proportionally to `shares` (0 = owns nothing)."""
        total = float(sum(shares))
        assert total > 0, "at least one member must own a shard"
        ranges, acc, start = [], 0.0, 0
        for i, s in enumerate(shares):
            acc += s
            end = self.n_chunks if i == len(shares) - 1 else int(round(self.n_chunks * acc / total))
            if s == 0:
                end = start
            ranges.append((start, end))
            start = end
        # the last owner absorbs rounding leftovers
        if start != self.n_chunks:
            for i in range(len(ranges) - 1, -1, -1):
                if shares[i] > 0:
                    b, _ = ranges[i]
                    ranges[i] = (b, self.n_chunks)
                    for j in range(i + 1, len(ranges)):
                        ranges[j] = (self.n_chunks, self.n_chunks)
                    break
        return ranges
