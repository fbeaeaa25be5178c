"""BASELINE.json config 1: ALBERT-tiny collaborative MLM, local peers on CPU, synthetic tokens, small target batch.

Checks the plumbing end to end (DHT -> progress tracker -> matchmaking -> socket butterfly -> LAMB):
  * after every collaborative step all peers hold identical parameters;
  * the trajectory equals single-process large-batch training on the union of the contributed micro-batches;
  * a late joiner downloads the state with load_state_from_peers and joins the next rounds;
  * an auxiliary peer (weight 0) and a client-mode peer (owns no shard) take part without changing the result;
  * a peer that dies does not stall the others.
"""
import threading
import time

import pytest
import torch

from dedloc_b200.collab.optimizer import CollaborativeOptimizer
from dedloc_b200.dht import DHT
from dedloc_b200.models.albert import AlbertConfig, AlbertForPreTraining
from dedloc_b200.optim import Lamb
from dedloc_b200.utils.compression import CompressionType

CFG = dict(vocab_size=128, embedding_size=16, hidden_size=32, num_hidden_layers=2, num_attention_heads=2,
           intermediate_size=64, max_position_embeddings=16, classifier_dropout_prob=0.0)
S, BS = 16, 2


def make_model(seed=0):
    torch.manual_seed(seed)
    return AlbertForPreTraining(AlbertConfig(**CFG))


def make_batch(peer: int, k: int):
    g = torch.Generator().manual_seed(1000 * peer + k)
    ids = torch.randint(5, CFG["vocab_size"], (BS, S), generator=g)
    labels = torch.full((BS, S), -100)
    m = torch.rand(BS, S, generator=g) < 0.3
    m[:, 1] = True
    labels[m] = ids[m]
    sop = torch.randint(0, 2, (BS,), generator=g)
    return dict(input_ids=ids, labels=labels, sentence_order_label=sop)


def make_opt(model, lr=1e-2):
    no_decay = ["bias", "LayerNorm.weight"]
    named = [(n, p) for n, p in model.named_parameters()]
    groups = [{"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": 0.01},
              {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0}]
    return Lamb(groups, lr=lr, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, clamp_value=10000.0, debias=True)


def make_peer(first_dht, prefix, target, seed=0, **kw):
    dht = DHT(start=True, initial_peers=[f"127.0.0.1:{first_dht.port}"] if first_dht else [])
    model = make_model(seed)
    opt = make_opt(model)
    defaults = dict(dht=dht, prefix=prefix, target_batch_size=target, batch_size_per_step=BS,
                    compression_type=CompressionType.NONE, averaging_expiration=2.0, averaging_timeout=10.0,
                    min_refresh_period=0.05, default_refresh_period=0.1, max_refresh_period=0.5,
                    target_group_size=16, metadata_expiration=10.0)
    defaults.update(kw)
    copt = CollaborativeOptimizer(opt, **defaults)
    return dht, model, copt


def run_trainer(peer_id, model, copt, n_rounds, log, stop=None, max_batches=400):
    k = 0
    while copt.local_step < n_rounds and k < max_batches and not (stop is not None and stop.is_set()):
        batch = make_batch(peer_id, k)
        round_idx = copt.local_step
        model.zero_grad()
        model(**batch).loss.backward()
        log.append((round_idx, peer_id, k))
        copt.step()
        k += 1
        time.sleep(0.01)


def oracle(log, n_rounds, seed=0):
    model = make_model(seed)
    opt = make_opt(model)
    for r in range(n_rounds):
        items = [(p, k) for (rr, p, k) in log if rr == r]
        assert items, f"no contributions for round {r}"
        model.zero_grad()
        for p, k in items:
            (model(**make_batch(p, k)).loss / len(items)).backward()
        opt.step()
    return model


def assert_same_params(a, b, tol=2e-5):
    for (n, x), (_, y) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(x, y, atol=tol, rtol=1e-4), n


def test_two_peers_match_large_batch_training():
    first, m0, c0 = make_peer(None, "cfg1", target=16)
    d1, m1, c1 = make_peer(first, "cfg1", target=16)
    log, n_rounds = [], 3
    threads = [threading.Thread(target=run_trainer, args=(i, m, c, n_rounds, log))
               for i, (m, c) in enumerate([(m0, c0), (m1, c1)])]
    try:
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert c0.local_step == n_rounds and c1.local_step == n_rounds
        assert c0.stats["solo_rounds"] == 0 and c1.stats["solo_rounds"] == 0, (c0.stats, c1.stats)
        assert_same_params(m0, m1, tol=1e-6)
        assert_same_params(m0, oracle(log, n_rounds))
    finally:
        for c, d in ((c0, first), (c1, d1)):
            c.shutdown()
            d.shutdown()


def test_late_joiner_aux_and_client_mode():
    first, m0, c0 = make_peer(None, "cfg1b", target=12)
    d1, m1, c1 = make_peer(first, "cfg1b", target=12, client_mode=True)
    daux, maux, caux = make_peer(first, "cfg1b", target=12, auxiliary=True, allow_state_sharing=False)
    log, stop = [], threading.Event()

    def aux_loop():
        while not stop.is_set():
            caux.step_aux()
            time.sleep(0.05)

    ta = threading.Thread(target=aux_loop, daemon=True)
    threads = [threading.Thread(target=run_trainer, args=(i, m, c, 2, log)) for i, (m, c) in enumerate([(m0, c0), (m1, c1)])]
    try:
        ta.start()
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert c0.local_step == 2 and c1.local_step == 2
        assert_same_params(m0, m1, tol=1e-6)
        assert_same_params(m0, oracle(log, 2))
        # a newcomer with a DIFFERENT init adopts the collaboration's state
        d2, m2, c2 = make_peer(first, "cfg1b", target=12, seed=123)
        assert c2.load_state_from_peers()
        assert c2.local_step == 2
        assert_same_params(m2, m0, tol=1e-7)
        st2, st0 = c2.opt.state_dict()["state"], c0.opt.state_dict()["state"]
        assert torch.allclose(st2[0]["exp_avg"], st0[0]["exp_avg"])
        c2.shutdown()
        d2.shutdown()
    finally:
        stop.set()
        for c, d in ((c0, first), (c1, d1), (caux, daux)):
            c.shutdown()
            d.shutdown()


def test_dead_peer_does_not_stall_the_other():
    first, m0, c0 = make_peer(None, "cfg1c", target=8, metadata_expiration=1.0, averaging_expiration=0.5)
    d1, m1, c1 = make_peer(first, "cfg1c", target=8, metadata_expiration=1.0, averaging_expiration=0.5)
    log = []
    try:
        run_trainer(1, m1, c1, 0, log, max_batches=0)
        # peer 1 reports once, then "dies" (its record expires after metadata_expiration)
        model_out = m1(**make_batch(1, 0)).loss
        model_out.backward()
        c1.step()
        c1.shutdown()
        d1.shutdown()
        t = threading.Thread(target=run_trainer, args=(0, m0, c0, 2, log))
        t.start()
        t.join(timeout=60)
        assert c0.local_step == 2, "the surviving peer must keep stepping on its own gradients"
    finally:
        c0.shutdown()
        first.shutdown()


@pytest.mark.parametrize("average", [True, False])
def test_diverged_peer_is_healed_by_lineage(average):
    """A peer that stepped alone after a failed round holds different parameters at the same step number. Its lineage
    hash differs from the group's, so the next joint round first makes everybody equal again: by averaging the
    parameters of all members (default; the reference averages parameters in every round, and no step is thrown away),
    or -- average_diverged_parameters=False -- by copying the state of a majority member."""
    kw = dict(average_diverged_parameters=average)
    first, m0, c0 = make_peer(None, f"cfg1heal{int(average)}", target=16, **kw)
    d1, m1, c1 = make_peer(first, f"cfg1heal{int(average)}", target=16, **kw)
    d2, m2, c2 = make_peer(first, f"cfg1heal{int(average)}", target=16, **kw)
    peers = [(m0, c0), (m1, c1), (m2, c2)]
    try:
        log = []
        threads = [threading.Thread(target=run_trainer, args=(i, m, c, 1, log)) for i, (m, c) in enumerate(peers)]
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert all(c.local_step == 1 for _, c in peers)
        assert c0.lineage == c1.lineage == c2.lineage != "0" * 16
        # peer 2 "stepped alone": same step number, different parameters, different lineage
        with torch.no_grad():
            for p in m2.parameters():
                p.add_(0.01)
        c2.lineage = c2._mix(c2.lineage, "S", c2.local_step, c2.endpoint)
        threads = [threading.Thread(target=run_trainer, args=(i, m, c, 2, log)) for i, (m, c) in enumerate(peers)]
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert all(c.local_step == 2 for _, c in peers)
        if average:
            assert all(c.stats["lineage_resyncs"] == 0 and c.stats.get("parameter_averagings", 0) == 1 for _, c in peers)
        else:
            assert c2.stats["lineage_resyncs"] == 1 and c0.stats["lineage_resyncs"] == 0 and c1.stats["lineage_resyncs"] == 0
        assert c0.lineage == c1.lineage == c2.lineage
        assert_same_params(m0, m1, tol=1e-6)
        assert_same_params(m0, m2, tol=1e-6)
    finally:
        for c, d in ((c0, first), (c1, d1), (c2, d2)):
            c.shutdown()
            d.shutdown()


def test_four_peers_in_groups_of_two_merge_instead_of_discarding():
    """target_group_size < number of peers (swav/README.md:26 runs groups of 4; contributor_notebook.ipynb cell 3
    describes the reshuffle): two groups step side by side in a round, so their parameters differ afterwards. The next
    round's groups average PARAMETERS first (hivemind averages them in every round) -- nobody's step is thrown away by
    a state copy -- and the collaboration stays close to single-process large-batch training."""
    first, m0, c0 = make_peer(None, "cfg1g", target=32, target_group_size=2, averaging_expiration=1.0)
    others = [make_peer(first, "cfg1g", target=32, target_group_size=2, averaging_expiration=1.0) for _ in range(3)]
    peers = [(first, m0, c0)] + others
    log, n_rounds = [], 4
    threads = [threading.Thread(target=run_trainer, args=(i, m, c, n_rounds, log))
               for i, (_, m, c) in enumerate(peers)]
    try:
        [t.start() for t in threads]
        [t.join(timeout=240) for t in threads]
        for _, _, c in peers:
            assert c.local_step == n_rounds, [x[2].local_step for x in peers]
        # nothing is discarded by a state copy -- except by a peer that the scheduler starved until it was more than
        # step_tolerance behind (it took part in fewer rounds); on a loaded CI box that happens to at most one of them
        copied = [c for _, _, c in peers if c.stats["lineage_resyncs"] or c.stats["state_downloads"]]
        assert len(copied) <= 1 and all(c.stats["rounds"] < n_rounds for c in copied), [c.stats for _, _, c in peers]
        sizes = [c.stats["rounds"] - c.stats["solo_rounds"] for _, _, c in peers]
        assert sorted(sizes)[1] >= n_rounds - 1, [c.stats for _, _, c in peers]    # (almost) every round was a group round
        assert sum(c.stats.get("parameter_averagings", 0) for _, _, c in peers) >= 2, [c.stats for _, _, c in peers]
        # peers that were merged in the last round agree; all are close to each other and to large-batch training
        ref = oracle(log, n_rounds)
        with torch.no_grad():
            step = sum(float((p - q).abs().mean()) for p, q in zip(ref.parameters(), make_model(0).parameters()))
            for _, m, _ in peers:
                dist = sum(float((p - q).abs().mean()) for p, q in zip(m.parameters(), ref.parameters()))
                assert dist < 0.8 * step, (dist, step)      # closer to the large-batch trajectory than the trajectory is long
    finally:
        for d, _, c in peers:
            c.shutdown()
            d.shutdown()


def test_silent_peer_stops_counting_as_live_long_before_its_record_expires():
    """Matchmaking waits for `live_peers`, not for every record in the DHT: a killed peer's record stays for
    metadata_expiration (30 s in production), and every round in that window would wait averaging_expiration for it."""
    from dedloc_b200.collab.progress import ProgressTracker

    first = DHT(start=True)
    second = DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"])
    a = ProgressTracker(first, "live", b"a", 64, endpoint=None, metadata_expiration=30.0, liveness_window=0.4, start=False)
    b = ProgressTracker(second, "live", b"b", 64, endpoint=None, metadata_expiration=30.0, liveness_window=0.4, start=False)
    try:
        a.report(0, 8, update_ema_with=8, wait=True)
        b.report(0, 8, update_ema_with=8, wait=True)
        state = a.fetch_collaboration_state()
        assert (state.num_peers, state.live_peers) == (2, 2)
        time.sleep(0.6)          # b goes silent; a keeps reporting
        a.report(0, 16, update_ema_with=8, wait=True)
        state = a.fetch_collaboration_state()
        assert (state.num_peers, state.live_peers) == (2, 1)
        assert state.samples_accumulated == 24, "what the silent peer accumulated still counts; it is not extrapolated"
        # a peer whose micro-steps are long declares it: 4 x its report period outlasts the window
        b.report_period = 1.0
        b.last_reported_time = 0.0
        b.report(0, 16, update_ema_with=8, wait=True)
        time.sleep(0.6)
        assert a.fetch_collaboration_state().live_peers == 2
    finally:
        second.shutdown()
        first.shutdown()


def test_emulated_uplinks_pace_the_round_and_shift_load_to_the_fast_peer():
    """A7 link shaping (albert/AWS_runner.ipynb shapes peers to 200 / 100 / 50 Mbit/s with tc): `bandwidth` decides
    the share of the vector a peer reduces, `emulated_uplink_mbps` paces what it sends."""
    import asyncio

    from dedloc_b200.collab.matchmaking import GroupInfo
    from dedloc_b200.dht.protocol import RPCClient, RPCServer
    from dedloc_b200.parallel.cpu_backend import CPUAverager, load_balance_peers
    from dedloc_b200.utils.compression import CHUNK

    n = 64 * CHUNK                                   # 512 KiB of fp32
    rates = [8.0, 2.0]                               # Mbit/s: 1 MB/s and 0.25 MB/s
    counts = load_balance_peers(n // CHUNK, rates)
    assert counts[0] > counts[1] and sum(counts) == n // CHUNK

    async def main():
        servers = [RPCServer({}), RPCServer({})]
        ports = [await s.start("127.0.0.1", 0) for s in servers]
        clients = [RPCClient(), RPCClient()]
        avg = [CPUAverager(servers[i], clients[i], CompressionType.NONE, timeout=30.0, uplink_mbps=rates[i])
               for i in range(2)]
        endpoints = tuple(f"127.0.0.1:{p}" for p in ports)
        xs = [torch.full((n,), float(i + 1)) for i in range(2)]
        t0 = time.perf_counter()
        outs = await asyncio.gather(*(avg[i].allreduce(GroupInfo(b"g" * 16, endpoints, ({}, {}), i), xs[i],
                                                       [1.0, 1.0], rates) for i in range(2)))
        dt = time.perf_counter() - t0
        for c in clients:
            await c.shutdown()
        for s in servers:
            await s.shutdown()
        return outs, dt, [a.uplink.sent_bytes for a in avg]

    outs, dt, sent = asyncio.run(main())
    for o in outs:
        assert torch.allclose(o, torch.full((n,), 1.5))
    # each peer sends the other's part and returns the delta of its own: both directions carry about the same bytes
    part_bytes = [c * CHUNK * 4 for c in counts]
    assert abs(sent[0] - (part_bytes[1] + part_bytes[0])) < 0.05 * n * 4 and abs(sent[1] - sent[0]) < 0.05 * n * 4
    slowest = max(sent[i] / (rates[i] * 1e6 / 8) for i in range(2))
    assert 0.5 * slowest <= dt <= slowest + 2.0, (dt, slowest)


def test_authorizer_protects_group_formation_and_state_download():
    """With `authorizer=` every optimizer RPC carries a signed access token: peers of one authority average together
    and share state; a peer holding a token from somewhere else gets neither a group nor anybody's parameters."""
    from dedloc_b200.collab.auth import LocalAuthority, LocalTokenAuthorizer
    from dedloc_b200.dht.validation import RSAPrivateKey

    authority = LocalAuthority(allowlist={"alice", "bob"})
    auth = lambda name, a=authority: LocalTokenAuthorizer(a, name, RSAPrivateKey())   # noqa: E731
    first, m0, c0 = make_peer(None, "cfg1auth", target=8, authorizer=auth("alice"))
    d1, m1, c1 = make_peer(first, "cfg1auth", target=8, authorizer=auth("bob"))
    d2 = m2 = c2 = None
    log = []
    try:
        threads = [threading.Thread(target=run_trainer, args=(i, m, c, 2, log)) for i, (m, c) in enumerate(((m0, c0), (m1, c1)))]
        [t.start() for t in threads]
        [t.join(timeout=120) for t in threads]
        assert c0.local_step == c1.local_step == 2
        assert c0.stats["solo_rounds"] == 0 and c0.stats["failed_rounds"] == 0, c0.stats
        assert_same_params(m0, m1)
        # an outsider: right DHT, wrong authority
        d2, m2, c2 = make_peer(first, "cfg1auth", target=8, seed=5,
                               authorizer=auth("mallory", LocalAuthority(allowlist={"mallory"})))
        before = [p.detach().clone() for p in m2.parameters()]
        assert c2.load_state_from_peers() is False
        assert all(torch.equal(a, b) for a, b in zip(before, m2.parameters())), "nothing was downloaded"
        assert c2.stats["state_downloads"] == 0
        # ... and one without any authorizer at all
        d3, m3, c3 = make_peer(first, "cfg1auth", target=8, seed=6)
        try:
            assert c3.load_state_from_peers() is False
        finally:
            c3.shutdown()
            d3.shutdown()
        # a member with a valid token does get the state
        d4, m4, c4 = make_peer(first, "cfg1auth", target=8, seed=7, authorizer=auth("alice"))
        try:
            assert c4.load_state_from_peers() is True and c4.local_step == 2
            assert_same_params(m4, m0)
        finally:
            c4.shutdown()
            d4.shutdown()
    finally:
        for c, d in ((c0, first), (c1, d1), (c2, d2)):
            if c is not None:
                c.shutdown()
                d.shutdown()


def test_dead_peers_record_neither_delays_rounds_nor_splits_the_fleet():
    """Regression for the elasticity benchmark: a killed peer's progress record outlives it by
    metadata_expiration. While it is there (a) the leader must not wait averaging_expiration for it in every round and
    (b) it must not push the peer count over target_group_size (which would split the fleet into matchmaking buckets).
    A replacement peer joins meanwhile, as the fleet launcher does."""
    kw = dict(target=32, target_group_size=4, averaging_expiration=3.0, metadata_expiration=30.0, liveness_window=0.6)
    first, m0, c0 = make_peer(None, "cfg1live", **kw)
    peers = [(first, m0, c0)] + [make_peer(first, "cfg1live", **kw) for _ in range(3)]
    stop_victim = threading.Event()
    log, n_rounds = [], 9
    threads = [threading.Thread(target=run_trainer, args=(i, m, c, n_rounds, log, stop_victim if i == 3 else None))
               for i, (_, m, c) in enumerate(peers)]
    newcomer = None
    try:
        [t.start() for t in threads]
        while peers[3][2].local_step < 2:
            time.sleep(0.01)
        stop_victim.set()                      # peer 3 stops stepping and reporting; nothing is withdrawn from the DHT
        threads[3].join(timeout=30)
        newcomer = make_peer(first, "cfg1live", seed=3, **kw)
        assert newcomer[2].load_state_from_peers()
        t_new = threading.Thread(target=run_trainer, args=(4, newcomer[1], newcomer[2], n_rounds, log))
        t_new.start()
        [t.join(timeout=240) for t in threads[:3] + [t_new]]
        survivors = [c for _, _, c in peers[:3]]
        for c in survivors + [newcomer[2]]:
            assert c.local_step == n_rounds, [x.local_step for x in survivors]
        # the dead peer's record is still in the DHT (30 s) but, once it has gone stale (liveness window, stretched by
        # the victim's own reporting period on a loaded machine), it no longer pushes the fleet over target_group_size
        deadline = time.time() + 30
        while time.time() < deadline:
            state = survivors[0].tracker.fetch_collaboration_state()
            if state.live_peers <= 4:
                break
            time.sleep(0.2)
        assert state.num_peers == 5 and state.live_peers <= 4, state
        assert survivors[0]._matchmaking.num_group_bits(state.live_peers) == 0
        for c in survivors:
            rounds_after_death = n_rounds - 2
            # before the fix: >= averaging_expiration (3 s) in each of them
            assert c.stats["matchmaking_s"] < 0.35 * 3.0 * rounds_after_death, c.stats
            assert c.stats["solo_rounds"] <= 1 and c.stats["failed_rounds"] <= 1, c.stats
        # a starved peer may have missed a group (its step is merged by parameter averaging later): close, not equal
        assert_same_params(peers[0][1], peers[1][1], tol=5e-3)
    finally:
        for d, _, c in peers + ([newcomer] if newcomer else []):
            c.shutdown()
            d.shutdown()


def test_last_survivor_does_not_wait_for_ghost_records():
    """The only live peer of a fleet whose other records are still in the DHT (metadata_expiration 30 s) gives up
    matchmaking after about a second per round instead of averaging_expiration."""
    kw = dict(target=8, averaging_expiration=4.0, metadata_expiration=30.0, liveness_window=0.5)
    first, m0, c0 = make_peer(None, "cfg1solo", **kw)
    d1, m1, c1 = make_peer(first, "cfg1solo", **kw)
    log = []
    try:
        model_out = m1(**make_batch(1, 0)).loss
        model_out.backward()
        c1.step()                               # peer 1 reports once ...
        c1.tracker.shutdown_flag.set()          # ... and goes silent without withdrawing anything
        time.sleep(0.7)
        t0 = time.perf_counter()
        run_trainer(0, m0, c0, 3, log)
        dt = time.perf_counter() - t0
        assert c0.local_step == 3 and c0.stats["solo_rounds"] == 3
        assert c0.stats["matchmaking_s"] < 3 * 2.0 < 3 * 4.0, c0.stats       # ~1 s per round, not averaging_expiration
        assert dt < 12.0
    finally:
        c1.shutdown()
        d1.shutdown()
        c0.shutdown()
        first.shutdown()


def test_liveness_survives_a_skewed_clock():
    """A peer whose wall clock is 20 s behind writes 'old' timestamps; it still counts as live while its record keeps
    changing (observed with the reader's own clock), and stops counting once it stops changing."""
    from unittest import mock

    from dedloc_b200.collab import progress
    from dedloc_b200.collab.progress import ProgressTracker

    first = DHT(start=True)
    second = DHT(start=True, initial_peers=[f"127.0.0.1:{first.port}"])
    a = ProgressTracker(first, "skew", b"a", 64, endpoint=None, liveness_window=0.4, start=False)
    b = ProgressTracker(second, "skew", b"b", 64, endpoint=None, liveness_window=0.4, start=False)
    real = progress.get_dht_time
    try:
        a.report(0, 8, update_ema_with=8, wait=True)

        def report_b(samples):
            with mock.patch.object(progress, "get_dht_time", lambda: real() - 20.0):
                state = progress.TrainingState(endpoint=None, step=0, samples_accumulated=samples, samples_per_second=1.0,
                                               time=real() - 20.0, client_mode=False)
            second.store(b.training_progress_key, state.model_dump(), real() + 30.0, subkey=b"b")

        report_b(8)
        assert a.fetch_collaboration_state().live_peers == 1, "first look at a stale timestamp: not live yet"
        report_b(16)
        assert a.fetch_collaboration_state().live_peers == 2, "the record changed under our eyes"
        time.sleep(0.6)
        assert a.fetch_collaboration_state().live_peers == 1, "... and then stopped changing"
    finally:
        second.shutdown()
        first.shutdown()


def test_malformed_member_blob_is_a_failed_round_not_a_crash():
    from dedloc_b200.collab.matchmaking import AllreduceException

    good = {"weight": 8.0, "mode": 0, "throughput": 1.0, "step": 3, "handle": None, "total": None, "device": None,
            "lineage": "ab12", "lineage_depth": 3, "host": "h", "n_avg_tensors": 2}
    CollaborativeOptimizer._check_blobs([good, dict(good, mode=1, throughput=None), dict(good, mode=2, weight=0.0),
                                         dict(good, handle=b"\\x01" * 64, total=4096, device=0)])
    for bad in (dict(good, weight="many"), dict(good, weight=float("nan")), dict(good, weight=-1.0), dict(good, mode=7),
                dict(good, step=-1), dict(good, step="3"), dict(good, handle="not bytes"), dict(good, total=0),
                dict(good, lineage=["x"]), dict(good, throughput="fast"), "not a dict", None):
        with pytest.raises(AllreduceException):
            CollaborativeOptimizer._check_blobs([good, bad])


def test_eight_peers_in_balanced_groups_of_four():
    """More peers than target_group_size (swav/README.md:26 runs groups of 4): buckets come from the live records
    (Matchmaking.assign_group_bits), so every round has two full groups, the salt reshuffles them, parameter averaging
    merges their work and nobody's step is replaced by a state copy."""
    from dedloc_b200.collab.matchmaking import Matchmaking

    # the assignment itself: blocks differ by at most one peer, every peer computes the same split, rounds differ
    keys = [bytes([i]) * 4 for i in range(11)]
    mm = Matchmaking.__new__(Matchmaking)
    mm.target_group_size, mm.group_bits = 4, ""
    def split(keys, step):
        buckets = {}
        for k in keys:
            buckets.setdefault(mm.assign_group_bits(keys, k, step), []).append(k)
        return buckets

    splits = []
    for step in range(12):
        buckets = split(keys, step)
        sizes = sorted(len(v) for v in buckets.values())
        assert len(buckets) == 4 and sizes[-1] - sizes[0] <= 1, sizes           # 11 peers -> 3, 3, 3, 2
        splits.append(frozenset(frozenset(v) for v in buckets.values()))
    assert all(a != b for a, b in zip(splits, splits[1:])), "two consecutive rounds must never have the same groups"
    # four peers in pairs: the three possible pairings come round in turn, so everybody meets everybody in 3 rounds
    four = keys[:4]
    mm.target_group_size = 2
    pairings = [frozenset(frozenset(v) for v in split(four, step).values()) for step in range(3)]
    assert len(set(pairings)) == 3 and all(len(b) == 2 for p in pairings for b in p)
    mm.target_group_size = 4
    assert mm.assign_group_bits(keys[:4], keys[0], 0) == "" and mm.assign_group_bits(keys, b"stranger", 0) is None

    kw = dict(target=64, target_group_size=4, averaging_expiration=3.0, metadata_expiration=30.0)
    first, m0, c0 = make_peer(None, "cfg1e8", **kw)
    peers = [(first, m0, c0)] + [make_peer(first, "cfg1e8", **kw) for _ in range(7)]
    n_rounds, log = 6, []
    sizes = {i: [] for i in range(8)}
    for i, (_, _, c) in enumerate(peers):
        def wrap(orig=c._finish_global_step, c=c, i=i):
            sizes[i].append((getattr(c, "_last_round", None) or {}).get("group", 0))
            return orig()
        c._finish_global_step = wrap
    threads = [threading.Thread(target=run_trainer, args=(i, m, c, n_rounds, log)) for i, (_, m, c) in enumerate(peers)]
    try:
        [t.start() for t in threads]
        [t.join(timeout=300) for t in threads]
        assert all(c.local_step == n_rounds for _, _, c in peers), [c.local_step for _, _, c in peers]
        flat = [s for v in sizes.values() for s in v]
        assert sum(s == 4 for s in flat) >= 0.85 * len(flat), sizes          # a starved CI box may cost a round or two
        copies = sum(c.stats["lineage_resyncs"] + c.stats["state_downloads"] for _, _, c in peers)
        assert copies <= 1, [c.stats for _, _, c in peers]
        assert sum(c.stats.get("parameter_averagings", 0) for _, _, c in peers) >= 8
    finally:
        for d, _, c in peers:
            c.shutdown()
            d.shutdown()
