"""Transport + RPC layer of the DHT: length-prefixed msgpack frames over asyncio TCP streams.

Replaces hivemind's gRPC `rpc_ping / rpc_store / rpc_find`. Control-plane traffic is < 4 KiB per
message, so a small pure-Python transport is the right tool; tensors never travel here (they move over NVLink
peer memory, parallel/symm.py). The same framing is reused by matchmaking and CPU state transfer.
"""
from __future__ import annotations

import asyncio
import itertools
import struct
from typing import Any, Awaitable, Callable, Dict, Optional, Tuple

import msgpack

_LEN = struct.Struct("!I")
MAX_FRAME = 1 << 30
# DEDLOC_EMULATED_RTT_MS: add that much latency to every outgoing RPC (DHT, matchmaking, socket rounds) -- the
# counterpart of the link shaping in parallel/cpu_backend.py for studying the protocol under internet round-trip times
_EMULATED_RTT_S = float(__import__("os").environ.get("DEDLOC_EMULATED_RTT_MS", "0") or 0) / 1e3


async def read_frame(reader: asyncio.StreamReader) -> Any:
    header = await reader.readexactly(_LEN.size)
    (n,) = _LEN.unpack(header)
    if n > MAX_FRAME:
        raise ValueError(f"frame too large: {n}")
    return msgpack.unpackb(await reader.readexactly(n), raw=False, strict_map_key=False)


def write_frame(writer: asyncio.StreamWriter, obj: Any) -> None:
    raw = msgpack.packb(obj, use_bin_type=True)
    writer.write(_LEN.pack(len(raw)) + raw)


def split_endpoint(endpoint: str) -> Tuple[str, int]:
    host, _, port = endpoint.rpartition(":")
    host = host.strip("[]")
    if host in ("", "*", "::", "0.0.0.0"):
        host = "127.0.0.1"
    return host, int(port)


class RPCServer:
    """Serves `handlers[method](peer_host, payload) -> result` on every accepted connection."""

    def __init__(self, handlers: Dict[str, Callable[[str, Any], Awaitable[Any]]]):
        self.handlers = handlers
        self.server: Optional[asyncio.AbstractServer] = None
        self.port: Optional[int] = None
        self._conns: set = set()

    async def start(self, host: str, port: int) -> int:
        self.server = await asyncio.start_server(self._serve, host=host, port=port)
        self.port = self.server.sockets[0].getsockname()[1]
        return self.port

    async def _serve(self, reader: asyncio.StreamReader, writer: asyncio.StreamWriter) -> None:
        peer = writer.get_extra_info("peername")
        peer_host = peer[0] if peer else "127.0.0.1"
        self._conns.add(writer)
        try:
            while True:
                try:
                    req_id, method, payload = await read_frame(reader)
                except (asyncio.IncompleteReadError, ConnectionError):
                    break
                except Exception:
                    # not a frame of ours (garbage bytes, oversized length, wrong arity): drop THIS connection, the
                    # server and every other connection keep running
                    break
                asyncio.ensure_future(self._handle(writer, peer_host, req_id, method, payload))
        finally:
            self._conns.discard(writer)
            writer.close()

    async def _handle(self, writer, peer_host, req_id, method, payload) -> None:
        try:
            handler = self.handlers[method]
            result = await handler(peer_host, payload)
            reply = [req_id, True, result]
        except Exception as e:  # the caller gets the error text; the server keeps running
            reply = [req_id, False, f"{type(e).__name__}: {e}"]
        try:
            write_frame(writer, reply)
            await writer.drain()
        except (ConnectionError, RuntimeError):
            pass

    async def shutdown(self) -> None:
        if self.server is not None:
            self.server.close()
            try:
                await asyncio.wait_for(self.server.wait_closed(), 1.0)
            except Exception:
                pass
        for w in list(self._conns):
            w.close()


class RPCError(RuntimeError):
    pass


class _Connection:
    def __init__(self, reader, writer):
        self.reader, self.writer = reader, writer
        self.pending: Dict[int, asyncio.Future] = {}
        self.ids = itertools.count()
        self.task = asyncio.ensure_future(self._pump())
        self.closed = False

    async def _pump(self) -> None:
        try:
            while True:
                req_id, ok, result = await read_frame(self.reader)
                fut = self.pending.pop(req_id, None)
                if fut is not None and not fut.done():
                    if ok:
                        fut.set_result(result)
                    else:
                        fut.set_exception(RPCError(result))
        except Exception as e:
            self.closed = True
            for fut in self.pending.values():
                if not fut.done():
                    fut.set_exception(ConnectionError(f"connection lost: {e!r}"))
            self.pending.clear()
            try:
                self.writer.close()
            except Exception:
                pass

    async def call(self, method: str, payload: Any, timeout: float) -> Any:
        req_id = next(self.ids)
        fut = asyncio.get_event_loop().create_future()
        self.pending[req_id] = fut
        write_frame(self.writer, [req_id, method, payload])
        await self.writer.drain()
        try:
            return await asyncio.wait_for(fut, timeout)
        finally:
            self.pending.pop(req_id, None)


class RPCClient:
    """Connection-pooling client: one persistent stream per endpoint, many concurrent calls per stream."""

    def __init__(self, connect_timeout: float = 3.0):
        self.connect_timeout = connect_timeout
        self._conns: Dict[str, _Connection] = {}
        self._locks: Dict[str, asyncio.Lock] = {}

    async def call(self, endpoint: str, method: str, payload: Any, timeout: float = 5.0) -> Any:
        lock = self._locks.setdefault(endpoint, asyncio.Lock())
        async with lock:
            conn = self._conns.get(endpoint)
            if conn is None or conn.closed:
                host, port = split_endpoint(endpoint)
                reader, writer = await asyncio.wait_for(asyncio.open_connection(host, port), self.connect_timeout)
                conn = self._conns[endpoint] = _Connection(reader, writer)
        if _EMULATED_RTT_S > 0:      # experiments on one machine: every call pays a network round trip
            await asyncio.sleep(_EMULATED_RTT_S)
        return await conn.call(method, payload, timeout)

    async def shutdown(self) -> None:
        for conn in self._conns.values():
            conn.task.cancel()
            try:
                conn.writer.close()
            except Exception:
                pass
        self._conns.clear()
