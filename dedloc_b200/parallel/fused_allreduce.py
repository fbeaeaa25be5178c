"""Python side of the fused butterfly all-reduce + optimizer kernel (csrc/fused_allreduce.cu).

`PeerBuffers` is the symmetric allocation of one peer with a fixed region layout derived from the flat
layout, so that a peer's IPC handle is all another peer needs to address any region. `GroupSpec` describes one
averaging round (members, weights, shard ownership) as agreed by matchmaking (collab/matchmaking.py), and
`fused_allreduce()` launches the kernel for this peer. Replaces hivemind's `averager.step()` + `opt.step()`
(reference call sites albert/run_trainer.py:251-264).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch

from .. import lib as native
from ..utils.compression import CompressionType, aux_floats
from .flat import CHUNK, FlatLayout
from .symm import SymmetricBuffer

MAX_PEERS = native.MAX_PEERS


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


@dataclass
class RegionLayout:
    """Byte offsets of the regions inside a peer's symmetric buffer (identical on every peer)."""
    n: int
    n_tensors: int
    flags: int = 0
    table: int = 0
    send: int = 0
    aux: int = 0
    p32: int = 0
    p16: int = 0
    m: int = 0
    v: int = 0
    total: int = 0

    @classmethod
    def build(cls, n: int, n_tensors: int, with_state: bool = True) -> "RegionLayout":
        r = cls(n=n, n_tensors=n_tensors)
        off = 0
        r.flags = off
        off = _align(off + 4 * MAX_PEERS * 4)
        r.table = off
        off = _align(off + MAX_PEERS * (n_tensors + 1) * 2 * 4)
        r.send = off
        off = _align(off + 4 * n)
        max_aux = max(aux_floats(c, n) for c in CompressionType)
        r.aux = off
        off = _align(off + 4 * max(max_aux, 1))
        r.p32 = off
        off = _align(off + 4 * n)
        r.p16 = off
        off = _align(off + 2 * n)
        if with_state:
            r.m = off
            off = _align(off + 4 * n)
            r.v = off
            off = _align(off + 4 * n)
        r.total = off
        return r


class PeerBuffers:
    """All device memory one peer contributes to averaging."""

    def __init__(self, layout: FlatLayout, device: Optional[int] = None, with_state: bool = True):
        self.layout = layout
        self.device = torch.cuda.current_device() if device is None else device
        self.regions = RegionLayout.build(layout.total, layout.n_tensors, with_state)
        self.symm = SymmetricBuffer(self.regions.total, self.device)
        n, r = layout.total, self.regions
        self.flags = self.symm.view(r.flags, 4 * MAX_PEERS, torch.int32)
        self.table = self.symm.view(r.table, MAX_PEERS * (layout.n_tensors + 1) * 2, torch.float32)
        self.send = self.symm.view(r.send, 4 * n, torch.uint8)
        self.aux = self.symm.view(r.aux, max(max(aux_floats(c, n) for c in CompressionType), 1), torch.float32)
        self.p32 = self.symm.view(r.p32, n, torch.float32)
        self.p16 = self.symm.view(r.p16, n, torch.bfloat16)
        self.m = self.symm.view(r.m, n, torch.float32) if with_state else None
        self.v = self.symm.view(r.v, n, torch.float32) if with_state else None
        dev = f"cuda:{self.device}"
        # local (never addressed by peers)
        self.acc = torch.zeros(n, dtype=torch.float32, device=dev)       # gradient accumulators (H3 step 2)
        self.gavg = torch.zeros(n, dtype=torch.float32, device=dev)      # averaged gradient of my shard
        self.chunk_stats = torch.zeros(layout.n_chunks * 2, dtype=torch.float32, device=dev)
        self.status = torch.zeros(16, dtype=torch.int32, device=dev)
        self.norm_scratch = torch.zeros(2048, dtype=torch.float32, device=dev)
        self.normsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.chunk_tid = layout.chunk_tid.to(dev)
        self.tensor_cbeg = layout.tensor_cbeg.to(dev)
        self.t_wd = torch.zeros(layout.n_tensors, dtype=torch.float32, device=dev)
        self.t_lr_scale = torch.ones(layout.n_tensors, dtype=torch.float32, device=dev)
        self.t_flags = torch.zeros(layout.n_tensors, dtype=torch.int32, device=dev)
        self._status_host = torch.zeros(16, dtype=torch.int32).pin_memory()

    @property
    def handle(self) -> bytes:
        return self.symm.handle

    def peer_base(self, handle: bytes) -> int:
        return self.symm.open_peer(handle)

    def read_status(self) -> Dict[str, int]:
        self._status_host.copy_(self.status, non_blocking=False)
        code = int(self._status_host[0])
        return {"code": code, "name": native.AR_STATUS.get(code, f"unknown_{code}"),
                "nonfinite_chunks": int(self._status_host[2])}


@dataclass
class GroupSpec:
    """One averaging round as seen by this peer."""
    bases: List[int]                    # device pointer (in this process) of every member's symmetric buffer
    me: int
    weights: List[float]                # raw sample-count weights (aux peers: 0); normalised here
    shards: List[tuple]                 # (chunk_begin, chunk_end) per member (client-mode members: empty)
    slots: List[int]                    # flag slot per member (its index on the box)
    tag: int = 4
    timeout_s: float = 30.0


@dataclass
class OptHyper:
    mode: int = native.AR_MEAN
    lr: float = 0.0
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-6
    bias_correction: float = 1.0
    clamp_value: float = 10000.0
    momentum: float = 0.9
    trust: float = 0.001
    replicate_state: bool = True


def fused_allreduce(buf: PeerBuffers, group: GroupSpec, codec: CompressionType, hyper: OptHyper,
                    stream: Optional[torch.cuda.Stream] = None, phases: int = 3, xgpu_barriers: bool = True,
                    write_p16: bool = True, mean_out32: bool = True, max_blocks: int = 0) -> None:
    """Launch K1/K2/K3 for this peer on `stream` (current stream by default). Asynchronous."""
    P = len(group.bases)
    assert 1 <= P <= MAX_PEERS
    r = buf.regions
    p = native.ARParams()
    p.P, p.me, p.mode, p.codec = P, group.me, hyper.mode, int(codec)
    p.phases, p.xgpu_barriers = phases, int(xgpu_barriers)
    p.tag = (group.tag << 2) & 0xFFFFFFFF
    p.T = buf.layout.n_tensors
    p.timeout_ns = int(group.timeout_s * 1e9)
    p.replicate_state = int(hyper.replicate_state)
    p.max_blocks = max_blocks
    p.lr, p.beta1, p.beta2, p.eps = hyper.lr, hyper.beta1, hyper.beta2, hyper.eps
    p.bias_correction, p.clamp_value = hyper.bias_correction, hyper.clamp_value
    p.momentum, p.trust = hyper.momentum, hyper.trust
    wsum = float(sum(group.weights))
    for q in range(P):
        base = group.bases[q]
        p.w[q] = (group.weights[q] / wsum) if wsum > 0 else 0.0
        p.slot[q] = group.slots[q]
        p.cbeg[q], p.cend[q] = group.shards[q]
        p.send[q] = base + r.send
        p.aux[q] = base + r.aux
        is_mean = hyper.mode == native.AR_MEAN
        p.out32[q] = (base + r.p32) if (mean_out32 or not is_mean) else None
        p.out16[q] = (base + r.p16) if write_p16 else None
        p.m[q] = (base + r.m) if r.m else None
        p.v[q] = (base + r.v) if r.v else None
        p.table[q] = base + r.table
        p.flags[q] = base + r.flags
    p.gavg = buf.gavg.data_ptr()
    p.chunk_stats = buf.chunk_stats.data_ptr()
    p.chunk_tid = buf.chunk_tid.data_ptr()
    p.tensor_cbeg = buf.tensor_cbeg.data_ptr()
    p.t_wd = buf.t_wd.data_ptr()
    p.t_lr_scale = buf.t_lr_scale.data_ptr()
    p.t_flags = buf.t_flags.data_ptr()
    p.status = buf.status.data_ptr()
    st = stream if stream is not None else torch.cuda.current_stream(buf.device)
    with torch.cuda.device(buf.device):
        native.check(native.lib().dedloc_fused_allreduce(C.byref(p), st.cuda_stream), "fused_allreduce")
