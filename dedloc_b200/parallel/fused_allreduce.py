"""Python side of the fused butterfly all-reduce + optimizer kernel (csrc/fused_allreduce.cu).

`PeerBuffers` is the symmetric allocation of one peer with a fixed region layout derived from the flat
layout, so that a peer's IPC handle is all another peer needs to address any region. `GroupSpec` describes one
averaging round (members, weights, shard ownership) as agreed by matchmaking (collab/matchmaking.py), and
`fused_allreduce()` launches the kernel for this peer. Replaces hivemind's `averager.step()` + `opt.step()`
(reference call sites albert/run_trainer.py:251-264).
"""
from __future__ import annotations

import ctypes as C
import os
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
    state_end: int = 0        # end of the regions a state download copies (p32 .. v)
    # regions of the push-based kernel (csrc/fused_allreduce_v2.cu)
    inflag: int = 0
    outflag: int = 0
    abort: int = 0
    table2: int = 0
    delta: int = 0
    aux_in: int = 0
    aux_slot_floats: int = 0
    stage_in: int = 0
    stage_bytes: int = 0
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
        r.state_end = off
        r.inflag = off
        off = _align(off + 4 * native.AR2_MAX_STAGES * MAX_PEERS)
        r.outflag = off
        off = _align(off + 4 * native.AR2_MAX_STAGES * MAX_PEERS)
        r.abort = off
        off = _align(off + 256)
        r.table2 = off
        off = _align(off + 8 * MAX_PEERS * max(n_tensors, 1) * 2)
        r.delta = off
        off = _align(off + 2 * n)
        r.aux_slot_floats = max(max_aux, 1)
        r.aux_in = off
        off = _align(off + 4 * MAX_PEERS * r.aux_slot_floats)
        # one staging slot per sender: P slots of ceil(C / K) chunks. Sized for fp32 payloads with up to twice as many
        # senders as owners (client-mode members send but own nothing), plus one chunk of rounding per slot
        r.stage_bytes = 8 * n + 4 * CHUNK * MAX_PEERS
        r.stage_in = off
        off = _align(off + r.stage_bytes)
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
        # per-chunk partial norms: [chunk][2] for the pull kernel, [chunk][8 warps][2] for the push kernel
        # partial norms: [chunk][2] for the pull kernel; [tensor][REDUCE CTA]{a, b, tag, -} for the push kernel
        self.chunk_stats = torch.zeros(max(layout.n_chunks * 2, layout.n_tensors * 4 * 640), dtype=torch.float32,
                                       device=dev)
        self._v2_cache: Dict[tuple, object] = {}
        self.status = torch.zeros(16, dtype=torch.int32, device=dev)
        self.norm_scratch = torch.zeros(2048, dtype=torch.float32, device=dev)
        self.normsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.chunk_tid = layout.chunk_tid.to(dev)
        self.tensor_cbeg = layout.tensor_cbeg.to(dev)
        self.t_wd = torch.zeros(layout.n_tensors, dtype=torch.float32, device=dev)
        self.t_lr_scale = torch.ones(layout.n_tensors, dtype=torch.float32, device=dev)
        self.t_flags = torch.zeros(layout.n_tensors, dtype=torch.int32, device=dev)
        self._status_host = torch.zeros(16, dtype=torch.int32).pin_memory()
        self.counters = torch.zeros(3 * native.AR2_MAX_STAGES, dtype=torch.int32, device=dev)
        self.last_tag = 0          # tag of the last push-kernel launch (status[1] holds it once state was modified)
        self.params_dev = torch.zeros(MAX_PEERS * C.sizeof(native.AR2Params), dtype=torch.uint8, device=dev)

    @property
    def handle(self) -> bytes:
        return self.symm.handle

    def peer_base(self, handle: bytes) -> int:
        return self.symm.open_peer(handle)

    def close_peer(self, handle: bytes) -> None:
        """Unmap a peer that left (see symm.MappingLeases); cached launch parameters may hold its address."""
        self.symm.close_peer(handle)
        self._v2_cache.clear()

    def read_status(self) -> Dict[str, int]:
        self._status_host.copy_(self.status, non_blocking=False)
        code = int(self._status_host[0])
        if code == 1 and self.last_tag and (int(self._status_host[1]) & 0xFFFFFFFF) == self.last_tag:
            code = 3      # the push-based kernel gave up after this member had begun to modify its state
        if code != 0 or int(self._status_host[2]) != 0:
            # the push-based kernel is launched without a memset: leave status and stage counters clean for the next
            # round (a round that was abandoned half-way leaves counters behind)
            self.status[0:4].zero_()
            self.counters.zero_()
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


# ----------------------------------------------------------------------------------------------------------------
# Push-based, stage-pipelined kernel (csrc/fused_allreduce_v2.cu)
# ----------------------------------------------------------------------------------------------------------------
_FLY_CODECS = (CompressionType.NONE, CompressionType.FLOAT16, CompressionType.BFLOAT16)
_MAX_CTAS: Dict[tuple, int] = {}


def _max_ctas(mode: int, codec: CompressionType) -> int:
    key = (mode, int(codec), torch.cuda.current_device())
    if key not in _MAX_CTAS:
        _MAX_CTAS[key] = int(native.lib().dedloc_ar2_max_ctas(mode, int(codec)))
    return _MAX_CTAS[key]


def _role_split(total: int, mode: int, owner: bool, K: int, weights=None) -> tuple:
    """CTAs per role (PUSH, REDUCE, UPDATE, APPLY), proportional to the bytes each role moves per element."""
    import os

    if weights is None:
        env = os.environ.get("DEDLOC_AR2_SPLIT")
        if env:
            weights = [float(x) for x in env.split(",")]
    if mode == native.AR_MEAN:
        w = list(weights) if weights else [1.0, 1.0, 0.0, 0.0]
        w[2] = 0.0
        n_apply = 1                      # one CTA waits for the completion flags
        if not owner:
            w[1] = 0.0
        rest = total - n_apply
        tot = w[0] + w[1]
        nR = max(1, int(round(rest * w[1] / tot))) if w[1] > 0 else 0
        nP = rest - nR
        return max(nP, 1), nR, 0, n_apply
    w = list(weights) if weights else [6.0, 2.0 + 16.0 / K, 2.0 + 24.0 / K, 12.0]
    if not owner:
        w[1] = w[2] = 0.0
    tot = sum(w)
    nR = max(1, int(round(total * w[1] / tot))) if w[1] > 0 else 0
    nU = max(1, int(round(total * w[2] / tot))) if w[2] > 0 else 0
    nP = max(1, int(round(total * w[0] / tot)))
    nA = total - nP - nR - nU
    assert nA >= 1, (total, nP, nR, nU)
    return nP, nR, nU, nA


def stage_geometry(n_chunks: int, ctas: int, bytes_per_elem: int = 2, sequential: bool = True) -> tuple:
    """(NS, Ls): stages of Ls chunks (a multiple of 32, so 8-bit codebook blocks never straddle stages).

    Every stage hand-over costs a system-scope fence by the warp that completes the stage, and that fence waits for
    whatever the chip still has in flight: 7-25 us while other stages are streaming, a few us on a quiet chip
    (profiles/allreduce_trace_*.txt). With the phases run one after the other there is nothing a second stage could
    overlap with, so ONE stage is fastest at every size measured (1 MB - 1 GB, 2 and 8 GPUs). With concurrent roles a
    stage carries at least DEDLOC_AR2_STAGE_MB (default 16) MB of payload. DEDLOC_AR2_STAGES forces a count."""
    stage_mb = float(os.environ.get("DEDLOC_AR2_STAGE_MB", "16"))
    by_bytes = int(stage_mb * 2 ** 20 / (CHUNK * bytes_per_elem))
    target = max(32, (max(2 * ctas, by_bytes) + 31) // 32 * 32)
    ns = min(native.AR2_MAX_STAGES, max(1, (n_chunks + target - 1) // target))
    if sequential and not os.environ.get("DEDLOC_AR2_STAGE_MB"):
        ns = 1
    if os.environ.get("DEDLOC_AR2_STAGES"):
        ns = max(1, min(native.AR2_MAX_STAGES, int(os.environ["DEDLOC_AR2_STAGES"])))
    ls = ((n_chunks + ns - 1) // ns + 31) // 32 * 32
    ns = (n_chunks + ls - 1) // ls
    return ns, ls


def v2_supported(layout: FlatLayout, regions: RegionLayout, group: "GroupSpec", codec: CompressionType) -> bool:
    """The push-based kernel needs >= 2 members, at least one owner and P staging slots that fit the region."""
    P = len(group.bases)
    K = sum(1 for b, e in group.shards if e > b)
    if P < 2 or K < 1 or not regions.stage_in:
        return False
    from ..utils.compression import payload_bytes_per_elem

    slot = _align(((layout.n_chunks + K - 1) // K) * CHUNK * payload_bytes_per_elem(codec))
    return P * slot <= regions.stage_bytes


def _fill_v2(p, buf: PeerBuffers, group: GroupSpec, codec: CompressionType, hyper: OptHyper, ctas: int,
             src_f32: bool, scale: float, write_p16: bool, mean_out32: bool, split=None) -> None:
    from ..utils.compression import payload_bytes_per_elem

    P = len(group.bases)
    r, layout = buf.regions, buf.layout
    owners = [q for q in range(P) if group.shards[q][1] > group.shards[q][0]]
    K = len(owners)
    p.P, p.me, p.mode, p.codec, p.T, p.K = P, group.me, hyper.mode, int(codec), layout.n_tensors, K
    p.tag = (group.tag & 0x7FFFFFFF) or 1
    p.src_f32 = int(src_f32 and codec in _FLY_CODECS)
    p.timeout_ns = int(group.timeout_s * 1e9)
    sequential = os.environ.get("DEDLOC_AR2_SCHEDULE", "sequential") != "concurrent"
    p.sequential = int(sequential)
    if sequential:
        # every CTA runs the four roles one after the other: each phase has the whole chip
        is_owner = group.me in owners
        p.nP, p.nA = ctas, ctas
        p.nR = ctas if is_owner else 0
        p.nU = ctas if (is_owner and hyper.mode != native.AR_MEAN) else 0
    else:
        p.nP, p.nR, p.nU, p.nA = _role_split(ctas, hyper.mode, group.me in owners, K, split)
    p.NS, p.Ls = stage_geometry(layout.n_chunks, ctas, payload_bytes_per_elem(codec), sequential)
    buf.last_tag = int(p.tag)
    p.C = layout.n_chunks
    p.replicate_state = int(hyper.replicate_state)
    p.slot_bytes = _align(((layout.n_chunks + K - 1) // K) * CHUNK * payload_bytes_per_elem(codec))
    p.aux_slot_floats = r.aux_slot_floats
    p.scale = scale
    p.lr, p.beta1, p.beta2, p.eps = hyper.lr, hyper.beta1, hyper.beta2, hyper.eps
    p.bias_correction, p.clamp_value = hyper.bias_correction, hyper.clamp_value
    p.momentum, p.trust = hyper.momentum, hyper.trust
    wsum = float(sum(group.weights))
    for q in range(MAX_PEERS):
        p.kidx[q] = -1
        p.owner_of_k[q] = 0
    for k, q in enumerate(owners):
        p.owner_of_k[k] = q
        p.kidx[q] = k
    is_mean = hyper.mode == native.AR_MEAN
    for q in range(P):
        base = group.bases[q]
        p.w[q] = (group.weights[q] / wsum) if wsum > 0 else 0.0
        p.stage_in[q] = base + r.stage_in
        p.aux_in[q] = base + r.aux_in
        p.delta[q] = base + r.delta
        p.inflag[q] = base + r.inflag
        p.outflag[q] = base + r.outflag
        p.table[q] = base + r.table2
        p.abort_word[q] = base + r.abort
        p.out32[q] = (base + r.p32) if (is_mean and mean_out32) else None
        p.out16[q] = (base + r.p16) if (is_mean and write_p16) else None
        p.m_all[q] = (base + r.m) if r.m else None
        p.v_all[q] = (base + r.v) if r.v else None
    own = buf.symm.ptr
    p.acc = buf.acc.data_ptr()
    p.send = own + r.send
    p.aux = own + r.aux
    p.p32 = own + r.p32
    p.p16 = (own + r.p16) if write_p16 else None
    p.m = (own + r.m) if r.m else None
    p.v = (own + r.v) if r.v else None
    p.gavg = buf.gavg.data_ptr()
    p.chunk_stats = buf.chunk_stats.data_ptr()
    p.chunk_tid = buf.chunk_tid.data_ptr()
    p.tensor_cbeg = buf.tensor_cbeg.data_ptr()
    p.t_wd = buf.t_wd.data_ptr()
    p.t_lr_scale = buf.t_lr_scale.data_ptr()
    p.t_flags = buf.t_flags.data_ptr()
    p.status = buf.status.data_ptr()
    p.counters = buf.counters.data_ptr()
    p.trace = None
    p.debug_flags = 0


def fused_allreduce_v2(buf: PeerBuffers, group: GroupSpec, codec: CompressionType, hyper: OptHyper,
                       stream: Optional[torch.cuda.Stream] = None, src_f32: bool = True, scale: float = 1.0,
                       write_p16: bool = True, mean_out32: bool = True, max_ctas: int = 0, split=None,
                       trace: Optional[torch.Tensor] = None, debug_flags: int = 0) -> None:
    """Launch the push-based kernel for this peer on `stream` (current stream by default). Asynchronous.

    `group.shards` only says WHO owns a share (non-empty range = owner): ownership is interleaved, chunk c belongs to
    the (c mod K)-th owner. With `src_f32` and an on-the-fly codec (NONE / FLOAT16 / BFLOAT16) the gradient is encoded
    from `buf.acc * scale` inside the kernel; otherwise `buf.send` / `buf.aux` must hold the encoded payload."""
    assert v2_supported(buf.layout, buf.regions, group, codec), "group does not fit the push-based kernel"
    with torch.cuda.device(buf.device):
        ctas = _max_ctas(hyper.mode, codec)
        if max_ctas > 0:
            ctas = min(ctas, max_ctas)
        # filling ~200 ctypes fields costs ~50 us of Python per call: the structure of a group that meets again is kept
        key = (tuple(group.bases), group.me, tuple(group.weights), tuple(e > b for b, e in group.shards), int(codec),
               hyper.mode, ctas, bool(src_f32), bool(write_p16), bool(mean_out32), None if split is None else tuple(split),
               os.environ.get("DEDLOC_AR2_STAGES"), os.environ.get("DEDLOC_AR2_STAGE_MB"), os.environ.get("DEDLOC_AR2_SPLIT"),
               os.environ.get("DEDLOC_AR2_SCHEDULE"))
        p = buf._v2_cache.get(key)
        if p is None:
            if len(buf._v2_cache) > 64:
                buf._v2_cache.clear()
            p = native.AR2Params()
            _fill_v2(p, buf, group, codec, hyper, ctas, src_f32, scale, write_p16, mean_out32, split)
            buf._v2_cache[key] = p
        else:
            p.tag = (group.tag & 0x7FFFFFFF) or 1
            buf.last_tag = int(p.tag)
            p.timeout_ns = int(group.timeout_s * 1e9)
            p.replicate_state = int(hyper.replicate_state)
            p.scale = scale
            p.lr, p.beta1, p.beta2, p.eps = hyper.lr, hyper.beta1, hyper.beta2, hyper.eps
            p.bias_correction, p.clamp_value = hyper.bias_correction, hyper.clamp_value
            p.momentum, p.trust = hyper.momentum, hyper.trust
            p.trace = None
        if trace is not None:       # int64 [4 roles, AR2_MAX_STAGES, 4] on this device: per-stage %globaltimer stamps
            assert trace.numel() >= 4 * native.AR2_MAX_STAGES * 4 + 1024 * 10 and trace.dtype == torch.int64
            p.trace = trace.data_ptr()
        p.debug_flags = debug_flags
        st = stream if stream is not None else torch.cuda.current_stream(buf.device)
        native.check(native.lib().dedloc_fused_allreduce_v2(C.byref(p), 1, buf.params_dev.data_ptr(), st.cuda_stream),
                     "fused_allreduce_v2")


def fused_allreduce_v2_emulated(bufs: Sequence[PeerBuffers], weights: Sequence[float], shards: Sequence[tuple],
                                codec: CompressionType, hyper: OptHyper, tag: int, src_f32: bool = True,
                                scale: float = 1.0, write_p16: bool = True, mean_out32: bool = True,
                                timeout_s: float = 20.0, ctas_per_rank: int = 0, split=None,
                                debug_flags: int = 0) -> None:
    """All members of a group as ONE cooperative launch over buffer sets on one GPU (tests, single-GPU soak runs):
    co-dependent ranks must not be separate launches on one device. The device code is the production code."""
    P = len(bufs)
    dev = bufs[0].device
    with torch.cuda.device(dev):
        cap = _max_ctas(hyper.mode, codec)
        ctas = cap // P if ctas_per_rank <= 0 else min(ctas_per_rank, cap // P)
        assert ctas >= 4, "not enough co-resident CTAs for this many emulated ranks"
        arr = (native.AR2Params * P)()
        bases = [b.symm.ptr for b in bufs]
        for rnk in range(P):
            g = GroupSpec(bases=bases, me=rnk, weights=list(weights), shards=list(shards), slots=list(range(P)),
                          tag=tag, timeout_s=timeout_s)
            assert v2_supported(bufs[rnk].layout, bufs[rnk].regions, g, codec)
            _fill_v2(arr[rnk], bufs[rnk], g, codec, hyper, ctas, src_f32, scale, write_p16, mean_out32, split)
            arr[rnk].debug_flags = debug_flags
        st = torch.cuda.current_stream(dev)
        native.check(native.lib().dedloc_fused_allreduce_v2(arr, P, bufs[0].params_dev.data_ptr(), st.cuda_stream),
                     "fused_allreduce_v2 (emulated ranks)")


def fused_average_parameters(buf: PeerBuffers, group: GroupSpec, stream: Optional[torch.cuda.Stream] = None) -> None:
    """Weighted mean of the members' fp32 master parameters, in place, with the bf16 mirror refreshed (MEAN mode of the
    push-based kernel, lossless fp32 payload): merges the work of groups that stepped side by side (the
    reference averages parameters in every round). In place is safe: an owner stores the average of a chunk into a
    member's parameters only after that member's PUSH of the stage has been flagged complete."""
    codec = CompressionType.NONE
    assert v2_supported(buf.layout, buf.regions, group, codec)
    with torch.cuda.device(buf.device):
        p = native.AR2Params()
        _fill_v2(p, buf, group, codec, OptHyper(mode=native.AR_MEAN), _max_ctas(native.AR_MEAN, codec), True, 1.0,
                 True, True)
        p.acc = buf.symm.ptr + buf.regions.p32          # the "gradient" of this reduction is the parameter vector
        st = stream if stream is not None else torch.cuda.current_stream(buf.device)
        native.check(native.lib().dedloc_fused_allreduce_v2(C.byref(p), 1, buf.params_dev.data_ptr(), st.cuda_stream),
                     "fused_average_parameters")
