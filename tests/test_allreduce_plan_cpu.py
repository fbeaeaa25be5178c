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
