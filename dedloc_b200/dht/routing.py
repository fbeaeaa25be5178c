"""Kademlia identifiers and routing table (160-bit sha1 ids, k-buckets with k = 20)."""
from __future__ import annotations

import hashlib
import os
import time
from typing import Dict, Iterable, List, Optional, Tuple

import msgpack

Endpoint = str  # "host:port"
ID_BITS = 160


def get_dht_time() -> float:
    """Wall-clock used for every expiration in the DHT (reference: hivemind.get_dht_time,
    /root/reference/albert/run_trainer.py:164)."""
    return time.time()


class DHTID(int):
    MIN, MAX = 0, 2 ** ID_BITS

    def __new__(cls, value: int):
        assert cls.MIN <= value < cls.MAX, "DHTID out of range"
        return super().__new__(cls, value)

    @classmethod
    def generate(cls, source=None) -> "DHTID":
        """sha1 of the msgpack-serialised source; random when source is None."""
        raw = os.urandom(32) if source is None else msgpack.packb(source, use_bin_type=True)
        return cls(int.from_bytes(hashlib.sha1(raw).digest(), "big"))

    def xor_distance(self, other: int) -> int:
        return int(self) ^ int(other)

    def to_bytes(self) -> bytes:  # type: ignore[override]
        return int(self).to_bytes(ID_BITS // 8, "big")

    @classmethod
    def from_bytes(cls, raw: bytes) -> "DHTID":  # type: ignore[override]
        return cls(int.from_bytes(raw, "big"))

    def __repr__(self):
        return f"DHTID({int(self):#042x})"


class KBucket:
    def __init__(self, lower: int, upper: int, size: int):
        self.lower, self.upper, self.size = lower, upper, size
        self.nodes: Dict[DHTID, Endpoint] = {}        # insertion-ordered: oldest first
        self.replacements: Dict[DHTID, Endpoint] = {}
        self.last_updated = get_dht_time()

    def has_in_range(self, node_id: int) -> bool:
        return self.lower <= node_id < self.upper

    def add_or_update(self, node_id: DHTID, endpoint: Endpoint) -> bool:
        """Returns False when the bucket is full (caller may split)."""
        self.last_updated = get_dht_time()
        if node_id in self.nodes:
            del self.nodes[node_id]
            self.nodes[node_id] = endpoint
        elif len(self.nodes) < self.size:
            self.nodes[node_id] = endpoint
        else:
            self.replacements.pop(node_id, None)
            self.replacements[node_id] = endpoint
            return False
        return True

    def remove(self, node_id: DHTID) -> None:
        self.nodes.pop(node_id, None)
        self.replacements.pop(node_id, None)
        if self.replacements and len(self.nodes) < self.size:
            rid, rep = self.replacements.popitem()
            self.nodes[rid] = rep

    def split(self) -> Tuple["KBucket", "KBucket"]:
        mid = (self.lower + self.upper) // 2
        left, right = KBucket(self.lower, mid, self.size), KBucket(mid, self.upper, self.size)
        for nid, ep in list(self.nodes.items()) + list(self.replacements.items()):
            (left if nid < mid else right).add_or_update(nid, ep)
        return left, right


class RoutingTable:
    def __init__(self, node_id: DHTID, bucket_size: int = 20, depth_modulo: int = 5):
        self.node_id, self.bucket_size, self.depth_modulo = node_id, bucket_size, depth_modulo
        self.buckets: List[KBucket] = [KBucket(DHTID.MIN, DHTID.MAX, bucket_size)]
        self.endpoint_to_id: Dict[Endpoint, DHTID] = {}

    def _bucket_index(self, node_id: int) -> int:
        lo, hi = 0, len(self.buckets)
        while hi - lo > 1:
            mid = (lo + hi) // 2
            if self.buckets[mid].lower > node_id:
                hi = mid
            else:
                lo = mid
        return lo

    def add_or_update_node(self, node_id: DHTID, endpoint: Endpoint) -> None:
        if node_id == self.node_id:
            return
        idx = self._bucket_index(node_id)
        bucket = self.buckets[idx]
        if bucket.add_or_update(node_id, endpoint):
            self.endpoint_to_id[endpoint] = node_id
            return
        depth = ID_BITS - (bucket.upper - bucket.lower).bit_length() + 1
        if bucket.has_in_range(self.node_id) or depth % self.depth_modulo != 0:
            left, right = bucket.split()
            self.buckets[idx:idx + 1] = [left, right]
            self.add_or_update_node(node_id, endpoint)

    def remove(self, node_id: DHTID) -> None:
        bucket = self.buckets[self._bucket_index(node_id)]
        ep = bucket.nodes.get(node_id)
        bucket.remove(node_id)
        if ep is not None:
            self.endpoint_to_id.pop(ep, None)

    def get_nearest_neighbors(self, query: int, k: int, exclude: Optional[DHTID] = None) -> List[Tuple[DHTID, Endpoint]]:
        items: List[Tuple[int, DHTID, Endpoint]] = []
        for bucket in self.buckets:  # tables are small (one box / a few hundred peers): linear scan is fine
            for nid, ep in bucket.nodes.items():
                if nid != exclude:
                    items.append((nid ^ query, nid, ep))
        items.sort(key=lambda t: t[0])
        return [(nid, ep) for _, nid, ep in items[:k]]

    def __len__(self):
        return sum(len(b.nodes) for b in self.buckets)

    def __iter__(self) -> Iterable[Tuple[DHTID, Endpoint]]:
        for b in self.buckets:
            yield from b.nodes.items()
