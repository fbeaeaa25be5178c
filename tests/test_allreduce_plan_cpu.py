"""Host-side planning of the push-based all-reduce kernel: stage geometry and the CTA role split."""


def test_stage_geometry_and_role_split():
    from dedloc_b200 import lib
    from dedloc_b200.parallel.fused_allreduce import _role_split, stage_geometry

    for n_chunks in (1, 31, 32, 33, 128, 8716, 131072, 10 ** 6):
        for ctas in (8, 74, 148, 296):
            ns, ls = stage_geometry(n_chunks, ctas, 2, sequential=False)
            assert stage_geometry(n_chunks, ctas, 2, sequential=True)[0] == 1
            assert 1 <= ns <= lib.AR2_MAX_STAGES and ls % 32 == 0 and ns * ls >= n_chunks > (ns - 1) * ls
    for total in (8, 37, 74, 148, 296):
        for mode in (0, 1, 2):
            for owner in (True, False):
                for K in (1, 2, 8, 16):
                    nP, nR, nU, nA = _role_split(total, mode, owner, K)
                    assert nP + nR + nU + nA == total and nP >= 1 and nA >= 1
                    assert (nR >= 1) == owner and (nU >= 1) == (owner and mode != 0)


def test_mapping_leases_expire_unused_peers_only():
    """parallel/symm.py MappingLeases: mappings of peers that left are closed after two record lifetimes, except those
    that still own optimizer moments."""
    from dedloc_b200.parallel.symm import MappingLeases

    leases = MappingLeases(ttl=60.0)
    a, b, c = b"a" * 64, b"b" * 64, b"c" * 64
    leases.touch([a, b, c, None, b""], now=0.0)
    leases.touch([a], now=50.0)
    assert leases.expired(now=61.0, keep=[c]) == [b]       # a was used recently, c still owns moment chunks
    assert leases.expired(now=61.0) == [c]                 # ... until it no longer does
    assert leases.expired(now=100.0) == []
    assert leases.expired(now=111.0) == [a]
    leases.touch([b], now=120.0)                           # a peer that comes back is tracked again
    assert leases.expired(now=200.0) == [b]


def test_load_balancing_matches_the_linear_program():
    """H6: the closed form of parallel/cpu_backend.load_balance_peers against scipy's linprog on the reference's
    formulation (minimise the finishing time of the slowest member of a butterfly round)."""
    import numpy as np
    from scipy.optimize import linprog

    from dedloc_b200.parallel.cpu_backend import load_balance_peers

    rng = np.random.default_rng(0)
    for trial in range(200):
        P = int(rng.integers(2, 9))
        b = rng.choice([10.0, 50.0, 100.0, 200.0, 500.0, 1000.0], size=P) * rng.uniform(0.8, 1.25, size=P)
        clients = rng.random(P) < 0.2
        if clients.all():
            clients[0] = False
        N = int(rng.integers(200, 5000))
        throughputs = [None if c else float(x) for c, x in zip(clients, b)]
        counts = load_balance_peers(N, throughputs)
        assert sum(counts) == N and all(c >= 0 for c in counts)
        assert all(c == 0 for c, cl in zip(counts, clients) if cl)
        act = [i for i in range(P) if not clients[i]]

        def finish(x):          # client members' link speed is unknown: they do not enter the objective
            return max((N + (P - 2) * x[i]) / b[i] for i in act)

        # LP: variables x_act..., t
        A = len(act)
        c = np.zeros(A + 1); c[-1] = 1.0
        A_ub = np.zeros((A, A + 1)); b_ub = np.zeros(A)
        for r, i in enumerate(act):
            A_ub[r, r] = (P - 2) / b[i]; A_ub[r, -1] = -1.0; b_ub[r] = -N / b[i]
        A_eq = np.zeros((1, A + 1)); A_eq[0, :A] = 1.0
        res = linprog(c, A_ub=A_ub, b_ub=b_ub, A_eq=A_eq, b_eq=[N], bounds=[(0, None)] * (A + 1), method="highs")
        assert res.status == 0
        ours = finish(counts)
        assert ours <= res.x[-1] * (1 + 2e-3) + (P - 2) / b[act].min(), (trial, ours, res.x[-1], counts)   # one chunk of rounding
    # the slowest link sets the floor and then owns nothing; equal links split equally; two members: proportional
    assert load_balance_peers(1000, [1000.0, 1000.0, 1000.0, 10.0])[3] == 0
    assert load_balance_peers(999, [5.0, 5.0, 5.0]) == [333, 333, 333]
    assert load_balance_peers(300, [100.0, 200.0]) == [100, 200]
    assert load_balance_peers(64, [8.0, 2.0])[0] > load_balance_peers(64, [8.0, 2.0])[1]
