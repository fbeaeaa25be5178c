"""`P2P` -- peer-to-peer streams with relaying for peers that cannot accept inbound connections.

The reference's p2p/ tutorials (/root/reference/p2p/NAT-traversal.md, circuit-relay.md) drive hivemind's wrapper
around the Go libp2p daemon: `P2P.create(bootstrap, bootstrap_peers, dht_mode, force_reachability, auto_relay,
relay_enabled)`, `add_stream_handler(name, fn)`, `call_peer_handler(peer_id, name, bytes)`, `_client.identify()`.
libp2p (QUIC, AutoNAT hole punching) has no place on one NVSwitch box, but the API and the relay semantics are kept
so that the same scripts run: every node has a key-derived peer id; public nodes listen on a socket; private nodes
(`force_reachability="private"`) keep a long-poll open to a relay (any public node started with
`relay_enabled=True`) which forwards calls to them -- the circuit-relay pattern, over the repo's msgpack RPC.
"""
from __future__ import annotations

import asyncio
import hashlib
import itertools
import threading
from typing import Awaitable, Callable, Dict, List, Optional, Tuple

from ..dht.protocol import RPCClient, RPCServer
from ..dht.validation import RSAPrivateKey

Handler = Callable[[bytes], Awaitable[bytes]]


class _Identity:
    def __init__(self, peer_id: str, addrs: List[str]):
        self.peer_id, self.addrs = peer_id, addrs

    def __iter__(self):
        return iter((self.peer_id, self.addrs))


class _ClientView:
    """`p2p._client.identify()` of the tutorials."""

    def __init__(self, owner: "P2P"):
        self._owner = owner

    async def identify(self) -> Tuple[str, List[str]]:
        return self._owner.id, self._owner.maddrs


class P2P:
    def __init__(self):
        self._loop = asyncio.get_event_loop()
        self._key = RSAPrivateKey()
        self.id = "Qm" + hashlib.sha256(self._key.get_public_key().to_bytes()).hexdigest()[:44]
        self._handlers: Dict[str, Handler] = {}
        self._server: Optional[RPCServer] = None
        self._rpc = RPCClient()
        self.maddrs: List[str] = []
        self._peers: Dict[str, dict] = {}               # peer id -> {"addr": host:port | None, "relay": peer id}
        self._relay_queues: Dict[str, asyncio.Queue] = {}    # relay side: pending calls per private peer
        self._relay_waiters: Dict[int, asyncio.Future] = {}
        self._ids = itertools.count()
        self._poll_task: Optional[asyncio.Task] = None
        self._client = _ClientView(self)
        self.relay_enabled = False
        self.private = False

    # ------------------------------------------------------------------ construction
    @classmethod
    async def create(cls, bootstrap: bool = False, bootstrap_peers: Optional[List[str]] = None, dht_mode: str = "dht",
                     force_reachability: Optional[str] = None, auto_relay: bool = False, relay_enabled: bool = False,
                     host: str = "127.0.0.1", port: int = 0, **_ignored) -> "P2P":
        self = cls()
        self.relay_enabled = relay_enabled
        self.private = force_reachability == "private"
        if not self.private:
            self._server = RPCServer({"p2p.call": self._rpc_call, "p2p.hello": self._rpc_hello,
                                      "p2p.poll": self._rpc_poll, "p2p.respond": self._rpc_respond,
                                      "p2p.lookup": self._rpc_lookup})
            p = await self._server.start(host, port)
            self.maddrs = [f"/ip4/{host}/tcp/{p}/p2p/{self.id}"]
            self._addr = f"{host}:{p}"
        else:
            self._addr = None
        for maddr in (bootstrap_peers or []) if bootstrap else []:
            addr, pid = _parse_maddr(maddr)
            self._peers[pid] = {"addr": addr, "relay": None}
            info = await self._rpc.call(addr, "p2p.hello", {"id": self.id, "addr": self._addr, "private": self.private})
            if self.private and (auto_relay or True) and info.get("relay"):
                self._relay = (addr, pid)
                self.maddrs = [f"{maddr}/p2p-circuit/p2p/{self.id}"]
                self._poll_task = asyncio.ensure_future(self._poll_relay(addr))
        return self

    async def shutdown(self) -> None:
        if self._poll_task is not None:
            self._poll_task.cancel()
        if self._server is not None:
            await self._server.shutdown()
        await self._rpc.shutdown()

    # ------------------------------------------------------------------ public API
    async def add_stream_handler(self, name: str, handler: Handler) -> None:
        self._handlers[name] = handler

    async def call_peer_handler(self, peer_id: str, name: str, data: bytes, timeout: float = 10.0) -> bytes:
        info = self._peers.get(peer_id) or await self._lookup(peer_id)
        if info is None:
            raise KeyError(f"unknown peer {peer_id}")
        if info.get("addr"):
            return await self._rpc.call(info["addr"], "p2p.call", {"name": name, "data": data}, timeout)
        relay_addr = self._peers[info["relay"]]["addr"] if info.get("relay") in self._peers else info.get("relay_addr")
        return await self._rpc.call(relay_addr, "p2p.call", {"name": name, "data": data, "target": peer_id}, timeout)

    # ------------------------------------------------------------------ service side
    async def _dispatch(self, name: str, data: bytes) -> bytes:
        if name not in self._handlers:
            raise KeyError(f"no stream handler {name!r}")
        return await self._handlers[name](data)

    async def _rpc_call(self, peer_host, payload) -> bytes:
        target = payload.get("target")
        if target is None or target == self.id:
            return await self._dispatch(payload["name"], payload["data"])
        if not self.relay_enabled or target not in self._relay_queues:
            raise KeyError(f"cannot relay to {target}")
        req_id = next(self._ids)
        fut = asyncio.get_event_loop().create_future()
        self._relay_waiters[req_id] = fut
        await self._relay_queues[target].put({"req": req_id, "name": payload["name"], "data": payload["data"]})
        try:
            return await asyncio.wait_for(fut, 30.0)
        finally:
            self._relay_waiters.pop(req_id, None)

    async def _rpc_hello(self, peer_host, payload) -> dict:
        self._peers[payload["id"]] = {"addr": payload.get("addr"), "relay": self.id if payload.get("private") else None,
                                      "relay_addr": self._addr}
        if payload.get("private") and self.relay_enabled:
            self._relay_queues.setdefault(payload["id"], asyncio.Queue())
        return {"id": self.id, "relay": self.relay_enabled}

    async def _rpc_poll(self, peer_host, payload) -> Optional[dict]:
        q = self._relay_queues.setdefault(payload["id"], asyncio.Queue())
        try:
            return await asyncio.wait_for(q.get(), payload.get("wait", 5.0))
        except asyncio.TimeoutError:
            return None

    async def _rpc_respond(self, peer_host, payload) -> bool:
        fut = self._relay_waiters.get(payload["req"])
        if fut is not None and not fut.done():
            if payload.get("error"):
                fut.set_exception(RuntimeError(payload["error"]))
            else:
                fut.set_result(payload["data"])
        return True

    async def _rpc_lookup(self, peer_host, payload) -> Optional[dict]:
        return self._peers.get(payload["id"])

    async def _lookup(self, peer_id: str) -> Optional[dict]:
        for pid, info in list(self._peers.items()):
            if info.get("addr"):
                found = await self._rpc.call(info["addr"], "p2p.lookup", {"id": peer_id})
                if found:
                    self._peers[peer_id] = found
                    return found
        return None

    async def _poll_relay(self, relay_addr: str) -> None:
        """Private node: outbound long-poll to the relay; inbound calls arrive as poll results."""
        while True:
            try:
                item = await self._rpc.call(relay_addr, "p2p.poll", {"id": self.id, "wait": 5.0}, timeout=15.0)
                if item is None:
                    continue
                try:
                    data = await self._dispatch(item["name"], item["data"])
                    await self._rpc.call(relay_addr, "p2p.respond", {"req": item["req"], "data": data})
                except Exception as e:
                    await self._rpc.call(relay_addr, "p2p.respond", {"req": item["req"], "error": repr(e)})
            except asyncio.CancelledError:
                return
            except Exception:
                await asyncio.sleep(0.5)


def _parse_maddr(maddr: str) -> Tuple[str, str]:
    parts = maddr.strip("/").split("/")
    host, port, pid = parts[1], parts[3], parts[5]
    return f"{host}:{port}", pid
