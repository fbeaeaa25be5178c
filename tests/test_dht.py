"""DHT behaviour: store/get with expirations, dictionary sub-keys, latest lookups across a small
swarm, client-only nodes, signature + schema validation (albert/metrics_utils.py semantics)."""
import time

import pytest

from dedloc_b200.dht import DHT, DHTID, RSASignatureValidator, SchemaValidator, get_dht_time
from dedloc_b200.dht.routing import RoutingTable
from dedloc_b200.dht.storage import DHTLocalStorage


def _swarm(n, **kw):
    first = DHT(start=True, **kw)
    peers = [first]
    for _ in range(n - 1):
        peers.append(DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"], **kw))
    return peers


def test_routing_table_nearest():
    me = DHTID.generate("me")
    table = RoutingTable(me, bucket_size=4)
    ids = [DHTID.generate(i) for i in range(200)]
    for i, nid in enumerate(ids):
        table.add_or_update_node(nid, f"127.0.0.1:{1000 + i}")
    q = DHTID.generate("query")
    got = [nid for nid, _ in table.get_nearest_neighbors(q, 5)]
    stored = [nid for nid, _ in table]
    assert got == sorted(stored, key=lambda n: n ^ q)[:5]
    assert len(table) <= 200


def test_local_storage_expiration_and_subkeys():
    st = DHTLocalStorage()
    now = get_dht_time()
    assert st.store(b"k", b"v1", now + 10)
    assert not st.store(b"k", b"v0", now + 5)          # older expiration loses
    assert st.store(b"k", b"v2", now + 20)
    assert st.get(b"k").value == b"v2"
    assert not st.store(b"gone", b"x", now - 1)
    assert st.store(b"d", b"a", now + 0.2, subkey=b"s1")
    assert st.store(b"d", b"b", now + 30, subkey=b"s2")
    assert set(dict(st.get(b"d").value.items())) == {b"s1", b"s2"}
    time.sleep(0.3)
    assert set(dict(st.get(b"d").value.items())) == {b"s2"}


def test_store_get_across_swarm():
    peers = _swarm(5)
    try:
        now = get_dht_time()
        assert peers[1].store("key1", {"a": 1, "b": [1, 2, 3]}, now + 30)
        for p in peers:
            got = p.get("key1", latest=True)
            assert got is not None and got.value == {"a": 1, "b": [1, 2, 3]}
        assert peers[0].get("missing") is None
        # dictionary value: every peer writes its own sub-key, everyone sees all of them
        for i, p in enumerate(peers):
            assert p.store("progress", {"step": i}, now + 30, subkey=f"peer{i}")
        got = peers[3].get("progress", latest=True)
        assert {k: v.value["step"] for k, v in got.value.items()} == {f"peer{i}": i for i in range(5)}
        # newer expiration wins for the same sub-key
        peers[2].store("progress", {"step": 100}, now + 60, subkey="peer2")
        assert peers[4].get("progress", latest=True).value["peer2"].value == {"step": 100}
        # futures
        fut = peers[0].store("fut", 7, now + 30, return_future=True)
        assert fut.result(timeout=5)
        assert peers[4].get("fut", return_future=True).result(timeout=5).value == 7
    finally:
        for p in peers:
            p.shutdown()


def test_expiration_over_network():
    peers = _swarm(3)
    try:
        peers[0].store("short", 1, get_dht_time() + 0.5)
        assert peers[2].get("short", latest=True).value == 1
        time.sleep(0.7)
        assert peers[2].get("short", latest=True) is None
    finally:
        for p in peers:
            p.shutdown()


def test_client_mode_node_is_never_contacted():
    first = DHT(start=True)
    client = DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"], listen=False)
    other = DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"])
    try:
        assert client.port is None
        assert client.store("from_client", "hello", get_dht_time() + 30)
        assert other.get("from_client", latest=True).value == "hello"
        assert client.get("from_client", latest=True).value == "hello"
        for p in (first, other):
            assert client.node_id not in [nid for nid, _ in p.node.routing]
    finally:
        for p in (first, client, other):
            p.shutdown()


def test_signature_and_schema_validation():
    from typing import Dict

    from pydantic import BaseModel, StrictFloat, conint

    from dedloc_b200.dht import BytesWithPublicKey
    from dedloc_b200.dht.validation import RSAPrivateKey

    class Metrics(BaseModel):
        step: conint(ge=0, strict=True)
        loss: StrictFloat

    class Schema(BaseModel):
        metrics: Dict[BytesWithPublicKey, Metrics]

    def validators(key=None):
        sig = RSASignatureValidator(key)
        return [SchemaValidator(Schema, prefix="exp"), sig], sig.local_public_key

    va, pk_a = validators(RSAPrivateKey())
    vb, pk_b = validators(RSAPrivateKey())
    a = DHT(start=True, record_validators=va)
    b = DHT(start=True, initial_peers=[f"127.0.0.1:{a.port}"], record_validators=vb)
    try:
        exp = get_dht_time() + 30
        assert a.store("exp_metrics", {"step": 3, "loss": 1.5}, exp, subkey=pk_a)
        assert b.store("exp_metrics", {"step": 4, "loss": 2.5}, exp, subkey=pk_b)
        got = b.get("exp_metrics", latest=True).value
        assert got[pk_a].value == {"step": 3, "loss": 1.5} and got[pk_b].value == {"step": 4, "loss": 2.5}
        # b cannot overwrite a's entry (signature made with the wrong key)
        assert not b.store("exp_metrics", {"step": 99, "loss": 0.0}, exp + 10, subkey=pk_a)
        assert a.get("exp_metrics", latest=True).value[pk_a].value["step"] == 3
        # schema violations are rejected: wrong type, negative step, subkey without owner tag
        assert not a.store("exp_metrics", {"step": "3", "loss": 1.5}, exp + 10, subkey=pk_a)
        assert not a.store("exp_metrics", {"step": -1, "loss": 1.5}, exp + 10, subkey=pk_a)
        assert not a.store("exp_metrics", {"step": 1, "loss": 1.5}, exp + 10, subkey=b"nobody")
        # keys outside the schema are free-form
        assert a.store("exp_other", [1, 2], exp)
    finally:
        a.shutdown()
        b.shutdown()


def test_frozen_node_is_left_alone_after_its_first_timeout():
    """A stopped process (SIGSTOP, swapped out, wedged) still accepts TCP connections; every call to it costs the whole
    wait_timeout. After the first one the node is avoided for a back-off period, so progress reads stay fast."""
    import socket

    frozen = socket.socket()
    frozen.bind(("127.0.0.1", 0))
    frozen.listen(8)                       # accepts (kernel backlog), never answers
    first = DHT(start=True, wait_timeout=0.6)
    second = DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"], wait_timeout=0.6)
    try:
        ghost_id = DHTID.generate()
        endpoint = f"127.0.0.1:{frozen.getsockname()[1]}"
        first.run_coroutine(lambda dht, node: _add(node, ghost_id, endpoint))
        assert second.store("k", 1, get_dht_time() + 30)
        t0 = time.perf_counter()
        assert first.get("k", latest=True).value == 1
        slow = time.perf_counter() - t0
        t0 = time.perf_counter()
        for _ in range(5):
            assert first.get("k", latest=True).value == 1
        fast = (time.perf_counter() - t0) / 5
        assert slow >= 0.5, "the first read runs into the frozen node"
        assert fast < 0.2, f"later reads must not wait for it again ({fast:.2f} s)"
        assert first.node._avoided(endpoint)
    finally:
        frozen.close()
        second.shutdown()
        first.shutdown()


async def _add(node, node_id, endpoint):
    node.routing.add_or_update_node(node_id, endpoint)


def test_rpc_server_survives_garbage_from_the_network():
    """Anything may connect to an open collaboration's ports: malformed frames cost the sender its connection and
    nothing else; malformed payloads of well-formed calls come back as errors."""
    import asyncio
    import random
    import socket
    import struct

    import msgpack

    node = DHT(start=True)
    other = DHT(start=True, initial_peers=[f"127.0.0.1:{node.port}"])
    rng = random.Random(0)

    def frame(obj):
        raw = msgpack.packb(obj, use_bin_type=True)
        return struct.pack("!I", len(raw)) + raw

    garbage = [
        b"GET / HTTP/1.1\r\n\r\n",                                   # somebody's browser
        struct.pack("!I", 1 << 31) + b"x" * 64,                        # absurd length
        struct.pack("!I", 16) + bytes(rng.randrange(256) for _ in range(16)),   # not msgpack
        frame([1, 2]),                                                # wrong arity
        frame({"a": 1}),                                              # not a list
        frame([7, ["find"], None]),                                   # unhashable method
        frame([8, "find", {"keys": "nope"}]),                         # bad payload of a real method
        frame([9, "store", {"keys": [b"k"], "subkeys": [b"s"], "values": [1.5], "expirations": ["soon"]}]),
        frame([10, "no_such_method", None]),
        frame([11, "ping", {"sender": [b"short", "port"]}]),
        bytes(rng.randrange(256) for _ in range(257)),
    ]
    try:
        for blob in garbage:
            with socket.create_connection(("127.0.0.1", node.port), timeout=5) as s:
                s.sendall(blob)
                s.settimeout(0.3)
                try:
                    s.recv(4096)           # an error reply, or nothing, or a close: all fine
                except (socket.timeout, ConnectionError):
                    pass
        # the node still works for well-behaved peers
        assert other.store("after_garbage", 42, get_dht_time() + 30)
        assert node.get("after_garbage", latest=True).value == 42
        assert other.get("after_garbage", latest=True).value == 42
    finally:
        other.shutdown()
        node.shutdown()
