"""Group formation for one averaging round.

Peers that want to average declare themselves under the DHT key `{prefix}_averaging.0b{group_bits}` with an
expiration `now + averaging_expiration`. The declared peer with the EARLIEST expiration is the leader; everyone
else sends it `join_group` (a long-poll RPC). The leader begins the round when `target_group_size` (or the
expected number of peers) have joined or its time is up, and answers every pending join with the same
`GroupInfo`: a random group id, the ORDERED list of member endpoints and every member's gathered blob
(weight, mode, peer-memory handle, ...). Groups are disbanded after every round; `group_bits` may be re-derived
from the group id so that large collaborations reshuffle (contributor_notebook.ipynb cell 3).
"""
from __future__ import annotations

import asyncio
import logging
import os
import random
from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Tuple

from ..dht import DHT, get_dht_time
from ..dht.protocol import RPCClient, RPCServer

logger = logging.getLogger(__name__)


class MatchmakingException(Exception):
    pass


class AllreduceException(Exception):
    """The averaging round failed; the caller proceeds with local gradients."""


@dataclass(frozen=True)
class GroupInfo:
    group_id: bytes
    endpoints: Tuple[str, ...]       # ordered: index = rank inside the group
    gathered: Tuple[Any, ...]        # one blob per member, same order
    my_index: int

    @property
    def group_size(self) -> int:
        return len(self.endpoints)


class Matchmaking:
    """Lives on the DHT's asyncio loop. `look_for_group()` is a coroutine; use `DHT.run_coroutine` to call it."""

    def __init__(self, dht: DHT, server: RPCServer, client: RPCClient, *, endpoint: str, prefix: str,
                 target_group_size: int = 256, min_group_size: int = 2, averaging_expiration: float = 5.0,
                 request_timeout: float = 3.0, client_mode: bool = False, initial_group_bits: str = ""):
        self.dht, self.client = dht, client
        self.endpoint, self.prefix = endpoint, prefix
        self.target_group_size, self.min_group_size = target_group_size, min_group_size
        self.averaging_expiration, self.request_timeout = averaging_expiration, request_timeout
        self.client_mode = client_mode
        self.group_bits = initial_group_bits
        self.lock = asyncio.Lock()
        # leader-side state of the current round
        self.looking = False
        self.declared_expiration = float("inf")
        self.followers: Dict[str, Any] = {}                      # endpoint -> gathered blob
        self.assembled: Optional[asyncio.Future] = None          # resolves to GroupInfo fields or None
        self.group_full = asyncio.Event()
        self.current_leader: Optional[str] = None
        self.my_blob: Any = None
        self.expected_size: Optional[int] = None
        self.expected_size_fn: Optional[Callable[[], Optional[int]]] = None
        self.give_up_fn: Optional[Callable[[], bool]] = None
        if server is not None:
            server.handlers["join_group"] = self.rpc_join_group

    @property
    def group_key(self) -> str:
        return f"{self.prefix}_averaging.0b{self.group_bits}"

    # ---------------------------------------------------------------- group bits (the "shuffle around")
    def num_group_bits(self, num_peers: Optional[int]) -> int:
        """How many key bits split `num_peers` into buckets of about target_group_size."""
        if not num_peers or num_peers <= self.target_group_size:
            return 0
        import math

        return max(0, math.ceil(math.log2(num_peers / self.target_group_size)))

    def assign_group_bits(self, live_keys, my_key: bytes, round_salt: int) -> Optional[str]:
        """Balanced buckets that are guaranteed to change every round: put the live peers in a fixed order (by the
        hash of their record key -- every peer that sees the same records computes the same order), keep the first one
        in place and rotate the others by `round * stride` (the circle method of round-robin tournaments), then cut the
        row into 2^bits consecutive blocks. Blocks differ by at most one peer; two consecutive rounds never have the
        same blocks, so the work of different groups is merged (parameter averaging) from one round to the next, and
        over n - 1 rounds everybody has been grouped with everybody. Against hash(group id) -> bucket (hivemind's
        scheme: random occupancy -- with 8 peers and groups of 4 a bucket of 5 leaves one peer without a group every
        other round -- and random repetition of the same groups). Returns None when this peer cannot place itself (its
        own record is not among the live ones yet): the caller falls back to `rederive_group_bits`."""
        import hashlib

        keys = list(live_keys)
        nbits = self.num_group_bits(len(keys))
        if nbits == 0:
            self.group_bits = ""
            return self.group_bits
        if my_key not in keys:
            return None
        order = sorted(keys, key=lambda k: hashlib.sha256(k).digest())
        rest = order[1:]
        stride = max(1, len(keys) // (1 << nbits) // 2)        # half a block moves per round
        if stride % len(rest) == 0:
            stride = 1
        shift = (int(round_salt) * stride) % len(rest)
        row = [order[0]] + rest[shift:] + rest[:shift]
        bucket = row.index(my_key) * (1 << nbits) // len(row)
        self.group_bits = format(bucket, f"0{nbits}b")
        return self.group_bits

    def rederive_group_bits(self, group: Optional["GroupInfo"], num_peers: Optional[int]) -> str:
        """After a round, every member moves to a bucket chosen by hash(group id, its index in the group): members of
        one group scatter over the buckets, so the next round mixes different peers (hivemind re-derives the bits from
        the group id for the same reason; contributor_notebook.ipynb cell 3 describes the effect). A peer that found
        nobody draws its bits at random."""
        import hashlib

        nbits = self.num_group_bits(num_peers)
        if nbits == 0:
            self.group_bits = ""
        elif group is None:
            self.group_bits = format(random.getrandbits(nbits), f"0{nbits}b")
        else:
            h = hashlib.sha256(group.group_id + group.my_index.to_bytes(4, "little")).digest()
            self.group_bits = format(int.from_bytes(h[:4], "little") % (1 << nbits), f"0{nbits}b")
        return self.group_bits

    # ---------------------------------------------------------------- leader side
    async def rpc_join_group(self, peer_host: str, payload) -> dict:
        """Long-poll: returns when the round begins (or the request is rejected / the group disbanded)."""
        follower, expiration, blob = payload["endpoint"], payload["expiration"], payload["blob"]
        if (not self.looking or self.client_mode or self.current_leader is not None
                or payload.get("group_key") != self.group_key):
            return {"code": "REJECTED", "suggested_leader": self.current_leader}
        if (expiration, follower) < (self.declared_expiration, self.endpoint) and not payload.get("client_mode"):
            return {"code": "REJECTED", "suggested_leader": None}   # the requester should lead, not us
        if len(self.followers) + 1 >= self.target_group_size or self.assembled is None or self.assembled.done():
            return {"code": "REJECTED", "suggested_leader": None}
        self.followers[follower] = blob
        if len(self.followers) + 1 >= self._wanted_size():
            self.group_full.set()
        assembled = self.assembled
        try:
            result = await asyncio.wait_for(asyncio.shield(assembled), self.averaging_expiration * 2 + 5)
        except asyncio.TimeoutError:
            result = None
        if result is None or follower not in result["endpoints"]:
            return {"code": "GROUP_DISBANDED"}
        return {"code": "BEGIN_ALLREDUCE", **result}

    def _wanted_size(self) -> int:
        """The group is complete at this many members: target_group_size, or fewer when the caller knows how many
        peers can still show up. `expected_size_fn` is re-read while the leader waits, so a peer that died after the
        round began to form stops being waited for as soon as the caller notices."""
        want = self.target_group_size
        expected = self.expected_size
        if self.expected_size_fn is not None:
            try:
                latest = self.expected_size_fn()
                expected = latest if expected is None else (expected if latest is None else min(expected, latest))
            except Exception:
                pass
        if expected is not None:
            want = min(want, max(expected, self.min_group_size))
        return want

    # ---------------------------------------------------------------- main entry
    async def look_for_group(self, blob: Any, *, timeout: Optional[float] = None,
                             expected_size: Optional[int] = None,
                             expected_size_fn: Optional[Callable[[], Optional[int]]] = None,
                             give_up_fn: Optional[Callable[[], bool]] = None) -> Optional[GroupInfo]:
        """Returns GroupInfo, or None when nobody else showed up before the deadline -- or when `give_up_fn` says
        that waiting is pointless (the caller saw the collaboration finish this step without us) and nobody has joined
        us yet."""
        async with self.lock:
            loop = asyncio.get_event_loop()
            now = get_dht_time()
            self.my_blob, self.expected_size, self.expected_size_fn = blob, expected_size, expected_size_fn
            self.give_up_fn = give_up_fn
            self.declared_expiration = now + self.averaging_expiration
            deadline = now + (timeout if timeout is not None else self.averaging_expiration * 2)
            self.followers = {}
            self.assembled = loop.create_future()
            self.group_full.clear()
            self.current_leader = None
            self.looking = True
            try:
                if not self.client_mode:
                    await self.dht.node.store(self.group_key, self.declared_expiration, self.declared_expiration,
                                              subkey=self.endpoint)
                found = await self._search(deadline)
                if found is None and self.group_bits and get_dht_time() < deadline:
                    # nobody else landed in this bucket: try once more under the common key
                    self.group_bits = ""
                    self.declared_expiration = get_dht_time() + self.averaging_expiration
                    self.followers = {}
                    self.assembled = loop.create_future()     # the first attempt resolved its future (to None)
                    self.group_full.clear()
                    self.current_leader = None
                    if not self.client_mode:
                        await self.dht.node.store(self.group_key, self.declared_expiration, self.declared_expiration,
                                                  subkey=self.endpoint)
                    found = await self._search(deadline)
                return found
            finally:
                self.looking = False
                if self.assembled is not None and not self.assembled.done():
                    self.assembled.set_result(None)
                if not self.client_mode:
                    # withdraw the declaration (a record that expires immediately after)
                    asyncio.ensure_future(self.dht.node.store(
                        self.group_key, None, max(self.declared_expiration, get_dht_time()) + 1e-3,
                        subkey=self.endpoint))

    async def _candidates(self) -> List[Tuple[float, str]]:
        found = await self.dht.node.get(self.group_key, latest=True)
        out = []
        now = get_dht_time()
        if found is not None and isinstance(found.value, dict):
            for ep, entry in found.value.items():
                if ep == self.endpoint or entry.value is None:
                    continue
                exp = float(entry.value)
                # client-mode peers never lead, so any declared peer is a candidate leader for them
                if exp > now and (self.client_mode or (exp, ep) < (self.declared_expiration, self.endpoint)):
                    out.append((exp, ep))
        return sorted(out)

    async def _search(self, deadline: float) -> Optional[GroupInfo]:
        rejected: set = set()
        while True:
            now = get_dht_time()
            if self.followers:
                return await self._lead()
            # (not in the first second: with several groups per round, another group may simply have been faster)
            if (self.give_up_fn is not None and now > self.declared_expiration - self.averaging_expiration + 1.0
                    and self.give_up_fn()):
                return None
            if now >= min(deadline, self.declared_expiration) and not self.client_mode:
                return await self._lead()
            if now >= deadline:
                return None
            if self.client_mode and not (await self._candidates()):
                await asyncio.sleep(0.05)   # nothing declared yet: poll until the deadline
                continue
            candidates = [c for c in await self._candidates() if c[1] not in rejected]
            if candidates and not self.followers:
                leader = candidates[0][1]
                self.current_leader = leader
                try:
                    reply = await self.client.call(
                        leader, "join_group",
                        {"endpoint": self.endpoint, "expiration": self.declared_expiration, "blob": self.my_blob,
                         "group_key": self.group_key, "client_mode": self.client_mode},
                        timeout=max(1.0, deadline - now) + self.averaging_expiration + 5)
                except Exception as e:
                    logger.debug("join_group @ %s failed: %r", leader, e)
                    reply = {"code": "REJECTED"}
                finally:
                    self.current_leader = None
                if reply["code"] == "BEGIN_ALLREDUCE":
                    endpoints = tuple(reply["endpoints"])
                    return GroupInfo(bytes(reply["group_id"]), endpoints, tuple(reply["gathered"]),
                                     endpoints.index(self.endpoint))
                rejected.add(leader)
                continue
            # nobody earlier than us: wait as a (potential) leader for followers or for our expiration
            wait = max(0.0, min(self.declared_expiration, deadline) - get_dht_time())
            try:
                await asyncio.wait_for(self.group_full.wait(), timeout=min(wait, 0.05 if not self.followers else wait))
            except asyncio.TimeoutError:
                pass
            if self.group_full.is_set():
                return await self._lead()

    async def _lead(self) -> Optional[GroupInfo]:
        """Wait until the group is full or our declaration expires, then begin."""
        while not self.group_full.is_set():
            wait = self.declared_expiration - get_dht_time()
            if wait <= 0 or len(self.followers) + 1 >= self._wanted_size():
                break
            try:
                await asyncio.wait_for(self.group_full.wait(), timeout=min(wait, 0.2))
            except asyncio.TimeoutError:
                pass
        if len(self.followers) + 1 < self.min_group_size:
            self.assembled.set_result(None)
            return None
        members = [(self.endpoint, self.my_blob)] + list(self.followers.items())
        random.shuffle(members)
        endpoints = [ep for ep, _ in members]
        result = {"group_id": os.urandom(16), "endpoints": endpoints, "gathered": [b for _, b in members]}
        self.assembled.set_result(result)
        return GroupInfo(result["group_id"], tuple(endpoints), tuple(result["gathered"]),
                         endpoints.index(self.endpoint))
