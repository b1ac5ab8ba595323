"""GPU parity: cluster_bp_list / bpc2bp (through the C ABI) vs the oracle on randomized and adversarial inputs."""
import random

import pytest

pytestmark = pytest.mark.gpu


def rand_bp_list(rng, n, base, spread_modes, chrs=('chr1', 'chr2'), multi_orient=True):
    out = []
    for i in range(n):
        o1 = rng.choice('+-') if multi_orient else '-'
        o2 = rng.choice('+-') if (multi_orient and rng.random() < 0.3) else '+'
        c1 = rng.choice(chrs)
        f = rng.choice(spread_modes)
        p1 = base + int(f(rng))
        p2 = base // 2 + int(rng.choice(spread_modes)(rng))
        out.append([c1, p1, o1, 'chr1', p2, o2, ('r%d' % i, 0, 1), rng.choice([0, 0, -50, 30, 400]), rng.choice([0, 1]),
                    rng.choice([20, 60]), rng.choice([20, 60])])
    return out


MODES = [lambda r: r.gauss(0, 30), lambda r: r.gauss(0, 1500), lambda r: r.randint(0, 5), lambda r: r.gauss(0, 4000)]


def test_cluster_bp_list_matches_oracle(built_lib):
    from coral_b200 import gpu_ops
    from oracle import coral_oracle as ora
    rng = random.Random(11)
    for t in range(60):
        n = rng.randint(1, 120)
        bl = rand_bp_list(rng, n, rng.choice([1000, 10 ** 5, 2 * 10 ** 8]), MODES)
        for mcs in (3, 3.5, 1000):
            got = gpu_ops.cluster_bp_list(bl, mcs, 2000)
            want = ora.cluster_bp_list(bl, mcs, 2000)
            assert got == want, (t, mcs)
            assert all(a is b for ca, cb in zip(got, want) for a, b in zip(ca, cb))  # same record objects


def test_cluster_chain_is_order_dependent(built_lib):
    """SURVEY.md App. A.5 counter-example: the chain case where input order changes the clusters."""
    from coral_b200 import gpu_ops
    from oracle import coral_oracle as ora

    def mk(ps):
        return [['chr1', p, '+', 'chr1', 500, '-', ('r%d' % i, 0, 1), 0, 0, 60, 60] for i, p in enumerate(ps)]
    a, b = mk([100000, 101500, 103000, 100100]), mk([100000, 103000, 101500, 100100])
    ga, gb = gpu_ops.cluster_bp_list(a, 3, 2000), gpu_ops.cluster_bp_list(b, 3, 2000)
    assert len(ga) == 1 and len(gb) == 2
    assert ga == ora.cluster_bp_list(a, 3, 2000) and gb == ora.cluster_bp_list(b, 3, 2000)


def test_bpc2bp_matches_oracle(built_lib):
    from coral_b200 import gpu_ops
    from oracle import coral_oracle as ora
    rng = random.Random(5)
    n_calls = 0
    for t in range(40):
        n = rng.randint(1, 400)   # > 150 members at chr1 coordinates: sum of squares exceeds 64 bits
        bl = rand_bp_list(rng, n, rng.choice([1000, 10 ** 5, 2 * 10 ** 8, 248_000_000]), MODES[:3], chrs=('chr1',),
                          multi_orient=(t % 2 == 0))
        for c in ora.cluster_bp_list(bl, 3, 2000):
            while len(c) >= 1:
                got = gpu_ops.bpc2bp(c, 100)
                want = ora.bpc2bp(c, 100)
                assert got == want, (t, len(c))
                assert [type(x) for x in got[2]] == [type(x) for x in want[2]]
                n_calls += 1
                if not want[3] or len(want[3]) == len(c):
                    break
                c = want[3]
    assert n_calls > 100


def test_bpc2bp_mode_median_rules(built_lib):
    """Unique mode, tie -> median with ceil ('+') / floor ('-') (SURVEY.md §8c verified values)."""
    from coral_b200 import gpu_ops
    from oracle import coral_oracle as ora

    def mk(ps, o):
        return [['chr1', p, o, 'chr1', p + 7, o, ('r%d' % i, 0, 1), 0, 0, 60, 60] for i, p in enumerate(ps)]
    for o, want in (('+', 1002), ('-', 1001)):
        c = mk([1000, 1000, 1003, 1003], o)
        got = gpu_ops.bpc2bp(c, 100)
        assert got[0][1] == want
        assert got == ora.bpc2bp(c, 100)
    c = mk([1000, 1001, 1001, 1005, 1009], '+')
    assert gpu_ops.bpc2bp(c, 100)[0][1] == 1001
    c = mk([10, 20, 30], '-')
    assert gpu_ops.bpc2bp(c, 100) == ora.bpc2bp(c, 100)
