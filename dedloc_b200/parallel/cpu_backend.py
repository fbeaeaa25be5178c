"""CPU / socket backend of the averaging round: load-balanced butterfly all-reduce with wire compression
(BASELINE.json config 1 "2 local peers on CPU").

Every member splits the flat vector into per-owner parts (share proportional to the declared bandwidth, 0 for
client-mode members), sends part j -- compressed -- to member j, the owner accumulates sum(w_i x_i) / sum(w_i)
in fp32 and answers each sender with the compressed DELTA (average - what that sender sent, after decode); the
sender adds the delta to its own decoded... local value. One hop each way: 2 (P-1)/P N bytes per peer. Any
failure aborts the round for the group (AllreduceException -> the optimizer steps on local gradients).
This is the path for CPU-only peers (auxiliary peers, tests); GPU peers use parallel/fused_allreduce.py.
"""
from __future__ import annotations

import asyncio
import logging
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..collab.matchmaking import AllreduceException, GroupInfo
from ..dht.protocol import RPCClient, RPCServer
from ..utils import compression as comp
from ..utils.compression import CHUNK, CompressionType

logger = logging.getLogger(__name__)


def load_balance_peers(n_chunks: int, throughputs: Sequence[Optional[float]], min_chunks: int = 0) -> List[int]:
    """How many chunks each member reduces (H6). The reference solves a linear program (scipy linprog) that minimises
    the time of the slowest member; this is the same optimum in closed form. In a butterfly round member i sends
    its vector minus its own part, then its averaged part to the other P - 1 members (and receives the mirror image),
    so its link carries N + (P - 2) x_i chunks and it finishes after (N + (P - 2) x_i) / b_i. Minimising the maximum
    over members with sum x_i = N, x_i >= 0:
        t* = max( max_i N / b_i ,  N (P - 2 + A) / sum_active b_i ),     x_i proportional to (t* b_i - N) / (P - 2)
    (A = members that may own a part). The first term is the floor set by the slowest link -- that member then owns
    nothing; when the second term binds everybody finishes at the same moment. P = 2 is indifferent to the split
    (P - 2 = 0) and gets the proportional one. `None` throughput = client mode = 0 chunks and unknown link speed.
    On the uniform NVSwitch fabric this is an equal split among non-client members."""
    tp = np.array([0.0 if t is None else max(float(t), 0.0) for t in throughputs], dtype=np.float64)
    if tp.sum() <= 0:
        raise ValueError("at least one member must be able to reduce (non-client with positive throughput)")
    P, N = len(tp), float(n_chunks)

    def solve(tp: np.ndarray) -> np.ndarray:
        active = tp > 0
        if P <= 2 or active.sum() == 1 or N <= 0:
            return tp / tp.sum() * N
        t_floor = float((N / tp[active]).max())
        t_even = N * (P - 2 + int(active.sum())) / float(tp[active].sum())
        t = max(t_floor, t_even)
        cap = np.where(active, np.maximum(t * tp - N, 0.0) / (P - 2), 0.0)
        if cap.sum() <= 0:            # every link is exactly at its floor: nothing to choose
            return tp / tp.sum() * N
        return cap * (N / cap.sum())

    ideal = solve(tp)
    if min_chunks > 0:  # members whose share would be tiny are not worth a round trip
        small = (ideal < min_chunks) & (tp > 0)
        if small.any() and (~small & (tp > 0)).any():
            tp = np.where(small, 0.0, tp)
            ideal = solve(tp)
    counts = np.floor(ideal + 1e-9).astype(np.int64)
    remainder = int(n_chunks - counts.sum())
    order = np.argsort(-(ideal - counts), kind="stable")
    for i in order[:max(remainder, 0)]:
        counts[i] += 1
    return counts.tolist()


def parts_from_counts(counts: Sequence[int]) -> List[Tuple[int, int]]:
    out, start = [], 0
    for c in counts:
        out.append((start, start + c))
        start += c
    return out


class _RoundState:
    def __init__(self, group: GroupInfo, weights: List[float], my_part: torch.Tensor, my_weight: float):
        self.group = group
        self.weights = weights
        self.acc = my_part.clone() * my_weight
        self.total_weight = my_weight
        self.received: Dict[int, torch.Tensor] = {}
        self.remaining = set(range(group.group_size)) - {group.my_index}
        self.done = asyncio.get_event_loop().create_future()
        self.codec: Optional[CompressionType] = None       # codec of this round's deltas (None = the averager's)

    def add(self, sender: int, decoded: torch.Tensor, weight: float) -> None:
        if sender not in self.remaining:
            raise AllreduceException(f"unexpected part from member {sender}")
        if decoded.shape != self.acc.shape:      # checked before anything is recorded: the sender may try again
            raise AllreduceException(f"member {sender} sent {decoded.numel()} elements for a part of {self.acc.numel()}")
        self.remaining.discard(sender)
        self.received[sender] = decoded
        self.acc += decoded * weight
        self.total_weight += weight
        if not self.remaining and not self.done.done():
            self.done.set_result(self.acc / max(self.total_weight, 1e-30))


class UplinkShaper:
    """Emulated uplink of one peer (the reference shapes its AWS peers with `tc`: 200 / 100 / 50 Mbit/s trainers, 500
    Mbit/s aux peers -- albert/AWS_runner.ipynb cells 1, 9, 12). Everything this peer sends in an averaging round goes
    through `send(nbytes)`, which returns when a link of `mbit_per_s` would have finished transmitting it; concurrent
    transmissions share the link (they queue). None / 0 = unshaped."""

    def __init__(self, mbit_per_s: Optional[float]):
        self.bytes_per_s = float(mbit_per_s) * 1e6 / 8 if mbit_per_s else 0.0
        self._free_at = 0.0
        self.sent_bytes = 0

    async def send(self, nbytes: int) -> None:
        self.sent_bytes += nbytes
        if self.bytes_per_s <= 0:
            return
        loop = asyncio.get_event_loop()
        start = max(loop.time(), self._free_at)
        self._free_at = start + nbytes / self.bytes_per_s
        await asyncio.sleep(max(0.0, self._free_at - loop.time()))


class CPUAverager:
    """Serves `allreduce_part` on the peer's RPC server and runs rounds on the DHT loop."""

    def __init__(self, server: Optional[RPCServer], client: RPCClient, compression: CompressionType,
                 timeout: float = 30.0, uplink_mbps: Optional[float] = None):
        self.client, self.compression, self.timeout = client, CompressionType(compression), timeout
        self.uplink = UplinkShaper(uplink_mbps)
        self.rounds: Dict[bytes, _RoundState] = {}
        self.early: Dict[bytes, List[tuple]] = {}
        if server is not None:
            server.handlers["allreduce_part"] = self.rpc_allreduce_part

    async def rpc_allreduce_part(self, peer_host: str, payload) -> bytes:
        # a group may run several reductions in one round (parameters, then gradients): the phase is part of the key
        gid, sender = bytes(payload["group_id"]) + str(payload.get("phase") or "").encode(), payload["sender"]
        enc = comp.from_bytes(payload["data"])
        decoded = comp.decode(enc)
        deadline = asyncio.get_event_loop().time() + self.timeout
        while gid not in self.rounds:  # the owner may enter the round slightly later than the sender
            if asyncio.get_event_loop().time() > deadline:
                raise AllreduceException("owner never joined the round")
            await asyncio.sleep(0.002)
        state = self.rounds[gid]
        state.add(sender, decoded, state.weights[sender])
        average = await asyncio.wait_for(asyncio.shield(state.done), self.timeout)
        delta = average - decoded
        reply = comp.to_bytes(comp.encode(delta, state.codec if state.codec is not None else self._delta_codec()))
        await self.uplink.send(len(reply))
        return reply

    def _delta_codec(self) -> CompressionType:
        return self.compression

    async def allreduce(self, group: GroupInfo, flat: torch.Tensor, weights: List[float],
                        throughputs: List[Optional[float]], phase: str = "",
                        compression: Optional[CompressionType] = None) -> torch.Tensor:
        """flat: padded fp32 vector (multiple of CHUNK). Returns the weighted average (same shape). `phase` tells
        apart several reductions of one group; `compression` overrides the wire codec (parameters travel lossless)."""
        if phase == "params" and compression is None:
            compression = CompressionType.NONE
        codec = self.compression if compression is None else CompressionType(compression)
        assert flat.dtype == torch.float32 and flat.numel() % CHUNK == 0
        P, me = group.group_size, group.my_index
        counts = load_balance_peers(flat.numel() // CHUNK, throughputs)
        parts = [(b * CHUNK, e * CHUNK) for b, e in parts_from_counts(counts)]
        gid = group.group_id + phase.encode()
        mb, me_ = parts[me]
        state = _RoundState(group, weights, flat[mb:me_], weights[me])
        state.codec = codec
        if not state.remaining and not state.done.done():
            state.done.set_result(state.acc / max(state.total_weight, 1e-30))
        self.rounds[gid] = state
        out = flat.clone()
        try:
            async def send_part(j: int):
                b, e = parts[j]
                if e == b:
                    return
                enc = comp.encode(flat[b:e], codec)
                sent = comp.decode(enc)  # what the owner will see from us
                data = comp.to_bytes(enc)
                await self.uplink.send(len(data))
                raw = await self.client.call(group.endpoints[j], "allreduce_part",
                                             {"group_id": group.group_id, "phase": phase, "sender": me, "data": data},
                                             timeout=self.timeout)
                delta = comp.decode(comp.from_bytes(raw))
                out[b:e] = sent + delta

            sends = [send_part(j) for j in range(P) if j != me]
            results = await asyncio.gather(*sends, return_exceptions=True)
            for r in results:
                if isinstance(r, Exception):
                    raise AllreduceException(f"averaging round failed: {r!r}")
            if me_ > mb:
                out[mb:me_] = await asyncio.wait_for(asyncio.shield(state.done), self.timeout)
            return out
        except asyncio.TimeoutError as e:
            raise AllreduceException("averaging round timed out") from e
        finally:
            # keep the round alive briefly so that late senders still get their delta
            asyncio.get_event_loop().call_later(self.timeout, self.rounds.pop, gid, None)
