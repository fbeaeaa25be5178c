"""Property tests (hypothesis) of the DHT's pure data structures: the routing table against a brute-force nearest-
neighbour search, and the local storage against a dictionary model of "newer expiration wins" (SURVEY H1)."""
import time

from hypothesis import given, settings, strategies as st

from dedloc_b200.dht.routing import DHTID, RoutingTable
from dedloc_b200.dht.storage import DHTLocalStorage

import os

ids = st.integers(min_value=0, max_value=DHTID.MAX - 1)
# deterministic in CI (same examples every run); HYPOTHESIS_CAMPAIGN=N runs N random examples per test instead
_N = int(os.environ.get("HYPOTHESIS_CAMPAIGN", "0"))


@settings(max_examples=_N or 60, deadline=None, derandomize=not _N)
@given(own=ids, nodes=st.lists(ids, min_size=1, max_size=120, unique=True), query=ids, k=st.integers(1, 12))
def test_routing_table_returns_the_k_nearest_of_what_it_kept(own, nodes, query, k):
    table = RoutingTable(DHTID(own), bucket_size=8)
    for n in nodes:
        if n != own:
            table.add_or_update_node(DHTID(n), f"127.0.0.1:{n % 60000 + 1}")
    kept = {int(nid): ep for nid, ep in table}
    assert own not in kept and set(kept) <= set(nodes)
    # nodes close to us are never dropped while the table has room for them (buckets around our own id split)
    got = table.get_nearest_neighbors(DHTID(query), k)
    want = sorted(kept, key=lambda n: n ^ query)[:k]
    assert [int(nid) for nid, _ in got] == want
    assert all(kept[int(nid)] == ep for nid, ep in got)
    # removal really removes
    if kept:
        victim = DHTID(next(iter(kept)))
        table.remove(victim)
        assert int(victim) not in {int(nid) for nid, _ in table}


@settings(max_examples=_N or 80, deadline=None, derandomize=not _N)
@given(ops=st.lists(st.tuples(st.sampled_from([b"k1", b"k2"]), st.sampled_from([None, b"a", b"b"]),
                              st.integers(0, 50), st.one_of(st.sampled_from([10.0, 20.0, 30.0]),       # ties on purpose
                                                            st.floats(min_value=10.0, max_value=1000.0))),
                    min_size=1, max_size=40))
def test_local_storage_keeps_the_record_with_the_latest_expiration(ops):
    now = time.time() + 100.0      # one base for every expiration of the example: equal ttl = equal expiration
    storage = DHTLocalStorage()
    model = {}          # (key, subkey) -> (value, expiration); a plain value and sub-keys exclude each other per key
    for key, subkey, value, ttl in ops:
        exp = now + ttl
        accepted = storage.store(key, value, exp, subkey=subkey)
        same_kind = {sk for (k, sk) in model if k == key}
        other_kind = [sk for sk in same_kind if (sk is None) != (subkey is None)]
        prev = model.get((key, subkey))
        if other_kind:
            # switching a key between "plain" and "dictionary" needs a strictly newer expiration than everything there
            newest = max(model[(key, sk)][1] for sk in same_kind)
            if exp > newest:
                for sk in list(same_kind):
                    del model[(key, sk)]
                model[(key, subkey)] = (value, exp)
                assert accepted
            else:
                assert not accepted
        elif prev is None or exp > prev[1]:
            model[(key, subkey)] = (value, exp)
            assert accepted
        else:
            assert not accepted
    for key in (b"k1", b"k2"):
        got = storage.get(key)
        mine = {sk: v for (k, sk), v in model.items() if k == key}
        if not mine:
            assert got is None
        elif None in mine:
            assert got.value == mine[None][0] and abs(got.expiration_time - mine[None][1]) < 1e-6
        else:
            assert {sk: (v.value, v.expiration_time) for sk, v in got.value.items()} == \
                   {sk: (val, exp) for sk, (val, exp) in mine.items()}
