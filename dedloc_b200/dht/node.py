"""Asynchronous Kademlia node: iterative find/store with a parallel beam, expirations,
dictionary values merged across replicas, `latest` lookups, client-only mode, record validators.
"""
from __future__ import annotations

import asyncio
import dataclasses
import logging
import time
from typing import Any, Dict, List, Optional, Sequence, Set, Tuple

import msgpack

from .protocol import RPCClient, RPCServer, split_endpoint
from .routing import DHTID, Endpoint, RoutingTable, get_dht_time
from .storage import DHTLocalStorage, DictionaryDHTValue, ValueWithExpiration
from .validation import IS_REGULAR_VALUE, CompositeValidator, DHTRecord, RecordValidatorBase

logger = logging.getLogger(__name__)


def _pack(obj: Any) -> bytes:
    return msgpack.packb(obj, use_bin_type=True)


def _unpack(raw: bytes) -> Any:
    return msgpack.unpackb(raw, raw=False, strict_map_key=False)


class DHTNode:
    def __init__(self, node_id: Optional[DHTID] = None, *, bucket_size: int = 20, num_replicas: int = 5,
                 parallel_rpc: int = 3, wait_timeout: float = 3.0, listen: bool = True,
                 record_validators: Sequence[RecordValidatorBase] = (), cache_locally: bool = True):
        self.node_id = node_id if node_id is not None else DHTID.generate()
        self.routing = RoutingTable(self.node_id, bucket_size)
        self.storage = DHTLocalStorage()
        self.cache = DHTLocalStorage(maxsize=10000)
        self.num_replicas, self.parallel_rpc, self.wait_timeout = num_replicas, parallel_rpc, wait_timeout
        self.listen = listen
        self.cache_locally = cache_locally
        self.validator = CompositeValidator(record_validators)
        self.client = RPCClient()
        self.server: Optional[RPCServer] = None
        self.port: Optional[int] = None
        self.endpoint_host = "127.0.0.1"
        # endpoint -> (do not call before, current back-off). A node that failed to answer (dead, or frozen: a stopped
        # process still accepts connections, so every call to it costs the full wait_timeout) is left alone for a
        # while; the back-off doubles with every further failure and is forgotten on the first answer.
        self._avoid: Dict[Endpoint, Tuple[float, float]] = {}
        self.backoff_first, self.backoff_max = 5.0, 60.0

    # ------------------------------------------------------------------ lifecycle
    @classmethod
    async def create(cls, *, initial_peers: Sequence[Endpoint] = (), listen: bool = True,
                     listen_on: str = "127.0.0.1:*", endpoint: Optional[str] = None, **kwargs) -> "DHTNode":
        self = cls(listen=listen, **kwargs)
        if listen:
            host, _, port = listen_on.rpartition(":")
            bind_host = host.strip("[]")
            if bind_host in ("", "*", "::"):
                bind_host = "0.0.0.0"
            self.server = RPCServer({"ping": self.rpc_ping, "store": self.rpc_store, "find": self.rpc_find})
            self.port = await self.server.start(bind_host, 0 if port in ("*", "") else int(port))
        if endpoint:
            self.endpoint_host = endpoint.rpartition(":")[0].strip("[]") or "127.0.0.1"
        if initial_peers:
            results = await asyncio.gather(*(self.call_ping(ep) for ep in initial_peers), return_exceptions=True)
            if not any(isinstance(r, DHTID) for r in results):
                logger.warning("DHTNode bootstrap failed: none of the initial peers responded: %s", initial_peers)
            # populate the routing table by looking up our own id
            await self.find_nearest_nodes([self.node_id])
        return self

    async def shutdown(self) -> None:
        if self.server is not None:
            await self.server.shutdown()
        await self.client.shutdown()

    # ------------------------------------------------------------------ sender info
    def _sender(self) -> Optional[list]:
        return [self.node_id.to_bytes(), self.port] if self.listen and self.port else None

    def _register_sender(self, peer_host: str, sender) -> None:
        if sender:
            node_id, port = DHTID.from_bytes(sender[0]), sender[1]
            self.routing.add_or_update_node(node_id, f"{peer_host}:{port}")
            self._note_success(f"{peer_host}:{port}")      # it just called us: whatever kept it silent is over

    # ------------------------------------------------------------------ RPC handlers (server side)
    async def rpc_ping(self, peer_host: str, payload) -> bytes:
        self._register_sender(peer_host, payload.get("sender"))
        return self.node_id.to_bytes()

    async def rpc_store(self, peer_host: str, payload) -> List[bool]:
        self._register_sender(peer_host, payload.get("sender"))
        oks = []
        for key, subkey, value, exp in zip(payload["keys"], payload["subkeys"], payload["values"],
                                           payload["expirations"]):
            oks.append(self._store_locally(key, subkey, value, exp))
        return oks

    def _store_locally(self, key: bytes, subkey: bytes, value: bytes, exp: float, cache: bool = False) -> bool:
        record = DHTRecord(key, subkey, value, exp)
        if not self.validator.validate(record):
            logger.debug("rejected record for key %s (validation failed)", key.hex())
            return False
        target = self.cache if cache else self.storage
        return target.store(key, value, exp, subkey=None if subkey == IS_REGULAR_VALUE else subkey)

    async def rpc_find(self, peer_host: str, payload) -> List[list]:
        self._register_sender(peer_host, payload.get("sender"))
        out = []
        for key in payload["keys"]:
            key_id = DHTID.from_bytes(key)
            found = self._lookup_local(key)
            nearest = [[nid.to_bytes(), ep] for nid, ep in
                       self.routing.get_nearest_neighbors(key_id, self.routing.bucket_size)]
            out.append([found, nearest])
        return out

    def _lookup_local(self, key: bytes):
        """-> None | ["v", value, exp] | ["d", [[subkey, value, exp], ...]]"""
        best = None
        for store in (self.storage, self.cache):
            item = store.get(key)
            if item is None:
                continue
            if isinstance(item.value, DictionaryDHTValue):
                enc = ["d", [[sk, v.value, v.expiration_time] for sk, v in item.value.items()]]
            else:
                enc = ["v", item.value, item.expiration_time]
            if best is None or item.expiration_time > best[1]:
                best = (enc, item.expiration_time)
        return best[0] if best else None

    # ------------------------------------------------------------------ RPC calls (client side)
    def _avoided(self, endpoint: Endpoint) -> bool:
        entry = self._avoid.get(endpoint)
        return entry is not None and time.monotonic() < entry[0]

    def _note_failure(self, endpoint: Endpoint) -> None:
        backoff = min(self._avoid[endpoint][1] * 2, self.backoff_max) if endpoint in self._avoid else self.backoff_first
        self._avoid[endpoint] = (time.monotonic() + backoff, backoff)
        if len(self._avoid) > 4096:     # bounded: forget the entries whose back-off ran out long ago
            now = time.monotonic()
            self._avoid = {ep: e for ep, e in self._avoid.items() if now < e[0] + self.backoff_max}

    def _note_success(self, endpoint: Endpoint) -> None:
        self._avoid.pop(endpoint, None)

    async def call_ping(self, endpoint: Endpoint) -> Optional[DHTID]:
        try:
            raw = await self.client.call(endpoint, "ping", {"sender": self._sender()}, self.wait_timeout)
            node_id = DHTID.from_bytes(raw)
            self.routing.add_or_update_node(node_id, endpoint)
            return node_id
        except Exception as e:
            logger.debug("ping %s failed: %r", endpoint, e)
            return None

    async def _call_find(self, endpoint: Endpoint, keys: List[bytes]):
        if self._avoided(endpoint):
            return None
        try:
            reply = await self.client.call(endpoint, "find", {"sender": self._sender(), "keys": keys},
                                           self.wait_timeout)
            self._note_success(endpoint)
            return reply
        except Exception as e:
            logger.debug("find @ %s failed: %r", endpoint, e)
            self._note_failure(endpoint)
            nid = self.routing.endpoint_to_id.get(endpoint)
            if nid is not None:
                self.routing.remove(nid)
            return None

    # ------------------------------------------------------------------ traversal
    async def _traverse(self, key: bytes, *, collect_values: bool, stop_on_first_value: bool):
        """Beam search towards `key`. Returns (nearest [(id, endpoint)], found payloads)."""
        key_id = DHTID.from_bytes(key)
        candidates: Dict[DHTID, Endpoint] = dict(self.routing.get_nearest_neighbors(key_id, self.routing.bucket_size))
        queried: Set[DHTID] = set()
        alive: Dict[DHTID, Endpoint] = {}
        found = []
        while True:
            pending = sorted((nid for nid in candidates if nid not in queried), key=lambda n: n ^ key_id)
            # stop when the k nearest alive nodes have all been queried
            nearest_alive = sorted(alive, key=lambda n: n ^ key_id)[: self.num_replicas]
            if not pending or (len(nearest_alive) >= self.num_replicas
                               and (pending[0] ^ key_id) > (nearest_alive[-1] ^ key_id)):
                break
            batch = pending[: self.parallel_rpc]
            queried.update(batch)
            replies = await asyncio.gather(*(self._call_find(candidates[nid], [key]) for nid in batch))
            for nid, reply in zip(batch, replies):
                if reply is None:
                    continue
                alive[nid] = candidates[nid]
                payload, nearest = reply[0]
                for raw_id, ep in nearest:
                    other = DHTID.from_bytes(raw_id)
                    if other != self.node_id and other not in candidates and not self._avoided(ep):
                        candidates[other] = ep
                        self.routing.add_or_update_node(other, ep)
                if payload is not None and collect_values:
                    found.append(payload)
            if found and stop_on_first_value:
                break
        nearest_nodes = sorted(alive.items(), key=lambda kv: kv[0] ^ key_id)
        return nearest_nodes, found

    async def find_nearest_nodes(self, key_ids: Sequence[DHTID]):
        return [(await self._traverse(k.to_bytes(), collect_values=False, stop_on_first_value=False))[0]
                for k in key_ids]

    # ------------------------------------------------------------------ public API
    async def store(self, key: Any, value: Any, expiration_time: float, subkey: Any = None) -> bool:
        key_bytes = DHTID.generate(source=key).to_bytes()
        subkey_bytes = IS_REGULAR_VALUE if subkey is None else _pack(subkey)
        record = DHTRecord(key_bytes, subkey_bytes, _pack(value), expiration_time)
        signed = self.validator.sign_value(record)
        nearest, _ = await self._traverse(key_bytes, collect_values=False, stop_on_first_value=False)
        key_id = DHTID.from_bytes(key_bytes)
        targets = [ep for _, ep in nearest[: self.num_replicas]]
        ok_any = False
        # store locally when we are among the nearest (or the network is smaller than num_replicas)
        store_self = self.listen and (len(nearest) < self.num_replicas or
                                      (self.node_id ^ key_id) < (nearest[self.num_replicas - 1][0] ^ key_id))
        if store_self or not targets:
            ok_any |= self._store_locally(key_bytes, subkey_bytes, signed, expiration_time)
        elif self.cache_locally:
            self._store_locally(key_bytes, subkey_bytes, signed, expiration_time, cache=True)
        payload = {"sender": self._sender(), "keys": [key_bytes], "subkeys": [subkey_bytes], "values": [signed],
                   "expirations": [expiration_time]}

        async def _one(ep):
            if self._avoided(ep):
                return False
            try:
                ok = (await self.client.call(ep, "store", payload, self.wait_timeout))[0]
                self._note_success(ep)
                return ok
            except Exception as e:
                logger.debug("store @ %s failed: %r", ep, e)
                self._note_failure(ep)
                return False

        results = await asyncio.gather(*(_one(ep) for ep in targets))
        return ok_any or any(results)

    async def get(self, key: Any, latest: bool = False) -> Optional[ValueWithExpiration]:
        key_bytes = DHTID.generate(source=key).to_bytes()
        payloads = []
        local = self._lookup_local(key_bytes)
        if local is not None:
            payloads.append(local)
        if latest or local is None:
            _, found = await self._traverse(key_bytes, collect_values=True, stop_on_first_value=not latest)
            payloads.extend(found)
        return self._merge(key_bytes, payloads)

    def _merge(self, key_bytes: bytes, payloads: List[list]) -> Optional[ValueWithExpiration]:
        now = get_dht_time()
        best_plain: Optional[Tuple[bytes, float]] = None
        merged: Dict[bytes, Tuple[bytes, float]] = {}
        for p in payloads:
            if p[0] == "v":
                _, value, exp = p
                if exp > now and (best_plain is None or exp > best_plain[1]):
                    if self.validator.validate(DHTRecord(key_bytes, IS_REGULAR_VALUE, value, exp)):
                        best_plain = (value, exp)
            else:
                for sk, value, exp in p[1]:
                    if exp > now and (sk not in merged or exp > merged[sk][1]):
                        if self.validator.validate(DHTRecord(key_bytes, sk, value, exp)):
                            merged[sk] = (value, exp)
        dict_exp = max((e for _, e in merged.values()), default=-float("inf"))
        if merged and (best_plain is None or dict_exp >= best_plain[1]):
            out = {}
            for sk, (value, exp) in merged.items():
                stripped = self.validator.strip_value(DHTRecord(key_bytes, sk, value, exp))
                out[_unpack(sk)] = ValueWithExpiration(_unpack(stripped), exp)
                if self.cache_locally:
                    self.cache.store(key_bytes, value, exp, subkey=sk)
            return ValueWithExpiration(out, dict_exp)
        if best_plain is not None:
            value, exp = best_plain
            stripped = self.validator.strip_value(DHTRecord(key_bytes, IS_REGULAR_VALUE, value, exp))
            return ValueWithExpiration(_unpack(stripped), exp)
        return None
