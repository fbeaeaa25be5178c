"""`DHT` -- the synchronous façade over DHTNode, API-compatible with the `hivemind.DHT` call sites of the
reference:

    DHT(start=True, initial_peers=[...], listen=not client_mode, listen_on="[::]:*", endpoint="1.2.3.4:*",
        record_validators=[...])                    albert/run_trainer.py:236-243
    dht.store(key, subkey=..., value=..., expiration_time=..., return_future=True)    run_trainer.py:160-166
    dht.get(key, latest=True) -> None | obj.value {subkey: obj.value}                 run_first_peer.py:177-180
    dht.port                                                                          run_first_peer.py:167

hivemind forks a process for the DHT; here the node lives on a daemon thread with its own asyncio loop
(the control plane carries < 4 KiB per message, and a thread avoids forking a process that holds a CUDA context).
"""
from __future__ import annotations

import asyncio
import concurrent.futures
import threading
from typing import Any, Awaitable, Callable, Optional, Sequence

from .node import DHTNode
from .routing import DHTID, Endpoint, get_dht_time
from .storage import ValueWithExpiration
from .validation import (BytesWithPublicKey, CompositeValidator, DHTRecord, RecordValidatorBase,  # noqa: F401
                         RSAPrivateKey, RSAPublicKey, RSASignatureValidator, SchemaValidator)

__all__ = ["DHT", "DHTNode", "DHTID", "get_dht_time", "ValueWithExpiration", "RecordValidatorBase",
           "SchemaValidator", "RSASignatureValidator", "BytesWithPublicKey"]


class DHT:
    def __init__(self, start: bool = False, initial_peers: Sequence[Endpoint] = (), listen: bool = True,
                 listen_on: str = "127.0.0.1:*", endpoint: Optional[str] = None,
                 record_validators: Sequence[RecordValidatorBase] = (), daemon: bool = True, **node_kwargs):
        self.initial_peers = [p for p in initial_peers if p]
        self.listen, self.listen_on, self.endpoint = listen, listen_on, endpoint
        self.record_validators = list(record_validators)
        self.node_kwargs = node_kwargs
        self._loop: Optional[asyncio.AbstractEventLoop] = None
        self._node: Optional[DHTNode] = None
        self._thread: Optional[threading.Thread] = None
        self._ready = threading.Event()
        self._error: Optional[BaseException] = None
        self._daemon = daemon
        if start:
            self.run_in_background(await_ready=True)

    # ---------------------------------------------------------------- lifecycle
    def run_in_background(self, await_ready: bool = True, timeout: float = 30.0) -> None:
        if self._thread is not None:
            return
        self._thread = threading.Thread(target=self._run, name="dedloc-dht", daemon=self._daemon)
        self._thread.start()
        if await_ready:
            if not self._ready.wait(timeout):
                raise TimeoutError("DHT did not start in time")
            if self._error is not None:
                raise self._error

    def _run(self) -> None:
        loop = asyncio.new_event_loop()
        asyncio.set_event_loop(loop)
        self._loop = loop
        try:
            self._node = loop.run_until_complete(DHTNode.create(
                initial_peers=self.initial_peers, listen=self.listen, listen_on=self.listen_on,
                endpoint=self.endpoint, record_validators=self.record_validators, **self.node_kwargs))
        except BaseException as e:  # surfaced to the constructor
            self._error = e
            self._ready.set()
            return
        self._ready.set()
        try:
            loop.run_forever()
        finally:
            try:
                loop.run_until_complete(self._node.shutdown())
            except Exception:
                pass
            # connection handlers / pumps still parked on a socket read: cancel and reap them before closing the loop
            # (otherwise asyncio reports "Task was destroyed but it is pending" for each of them at interpreter exit)
            try:
                pending = [t for t in asyncio.all_tasks(loop) if not t.done()]
                for t in pending:
                    t.cancel()
                if pending:
                    loop.run_until_complete(asyncio.gather(*pending, return_exceptions=True))
            except Exception:
                pass
            loop.close()

    def shutdown(self) -> None:
        if self._loop is not None and self._loop.is_running():
            self._loop.call_soon_threadsafe(self._loop.stop)
        if self._thread is not None:
            self._thread.join(timeout=5)
        self._thread = None

    def is_alive(self) -> bool:
        return self._thread is not None and self._thread.is_alive()

    def __del__(self):
        try:
            self.shutdown()
        except Exception:
            pass

    # ---------------------------------------------------------------- properties
    @property
    def port(self) -> Optional[int]:
        return self._node.port if self._node is not None else None

    @property
    def node_id(self) -> DHTID:
        return self._node.node_id

    @property
    def node(self) -> DHTNode:
        assert self._node is not None, "DHT is not running"
        return self._node

    def get_visible_address(self) -> str:
        host = "127.0.0.1"
        if self.endpoint:
            host = self.endpoint.rpartition(":")[0].strip("[]") or host
        return f"{host}:{self.port}"

    # ---------------------------------------------------------------- API
    def _submit(self, coro: Awaitable, return_future: bool):
        assert self._loop is not None, "DHT is not running (pass start=True)"
        fut: concurrent.futures.Future = asyncio.run_coroutine_threadsafe(coro, self._loop)
        return fut if return_future else fut.result()

    def store(self, key: Any, value: Any, expiration_time: float, subkey: Any = None,
              return_future: bool = False):
        """Store (key[, subkey]) -> value until expiration_time (absolute `get_dht_time()` seconds)."""
        return self._submit(self._node.store(key, value, expiration_time, subkey=subkey), return_future)

    def get(self, key: Any, latest: bool = False, return_future: bool = False):
        """-> None or ValueWithExpiration(value, expiration_time); dictionary keys yield
        {subkey: ValueWithExpiration}. latest=True keeps searching for the freshest replica."""
        return self._submit(self._node.get(key, latest=latest), return_future)

    def run_coroutine(self, coro_fn: Callable[["DHT", DHTNode], Awaitable], return_future: bool = False):
        """Run `coro_fn(dht, node)` on the DHT loop (used by matchmaking / state transfer)."""
        return self._submit(coro_fn(self, self._node), return_future)

    def add_validators(self, validators: Sequence[RecordValidatorBase]) -> None:
        self._node.validator.extend(validators)

    @property
    def loop(self) -> asyncio.AbstractEventLoop:
        return self._loop
