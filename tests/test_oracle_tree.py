"""CPU: DistributeOctTree (src/ORBextractor.cc:555-779) — oracle/cpp's std::list version against the independent
Python restatement (oracle/py/tree_restatement.py), plus the invariants the reference guarantees."""
import numpy as np
import pytest

from oracle.py import tree_restatement as tr


def _keys(rng, n, w, h, clustered=False):
    if clustered:
        cx, cy = rng.integers(0, w, 12), rng.integers(0, h, 12)
        i = rng.integers(0, 12, n)
        x = np.clip(cx[i] + rng.normal(0, 9, n), 0, w - 1).astype(np.int32)
        y = np.clip(cy[i] + rng.normal(0, 9, n), 0, h - 1).astype(np.int32)
    else:
        x, y = rng.integers(0, w, n), rng.integers(0, h, n)
    xy = np.unique(np.stack([x, y], 1), axis=0)
    xy = xy[np.lexsort((xy[:, 0], xy[:, 1]))]          # raster order like FAST output
    return np.concatenate([xy, rng.integers(7, 120, (len(xy), 1))], 1).astype(np.int32)


@pytest.mark.parametrize("case", [(0, 3000, 720, 448, 217, False), (1, 800, 720, 448, 217, True),
                                  (2, 150, 178, 102, 60, False), (3, 5000, 1209, 344, 434, True),
                                  (4, 40, 720, 448, 217, False), (5, 2000, 300, 300, 1086, False),
                                  (6, 1, 720, 448, 10, False), (7, 3, 100, 400 // 3, 5, False)])
def test_tree_two_restatements_agree(oracle_mod, case):
    seed, n, w, h, N, cl = case
    rng = np.random.default_rng(seed)
    keys = _keys(rng, n, w, h, cl)
    a = oracle_mod.distribute(keys, 16, 16 + w, 16, 16 + h, N)
    b = tr.distribute(keys, 16, 16 + w, 16, 16 + h, N, oracle_mod.sort_nodes)
    assert np.array_equal(a, b)
    # invariants: subset of the input, no duplicates, at most one point per final node, >= N only by the 4-way overshoot
    s = {tuple(k) for k in keys.tolist()}
    assert all(tuple(k) in s for k in a.tolist())
    assert len({tuple(k) for k in a.tolist()}) == len(a)
    assert len(a) <= max(N + 3, 4 * round(w / h) + 3) or len(a) <= len(keys)
    if len(keys) >= N * 4:
        assert len(a) >= N


def test_sort_ties_follow_libstdcxx(oracle_mod):
    """compareNodes orders by (size, UL.x) only: equal pairs are left wherever std::sort's introsort moves them."""
    rng = np.random.default_rng(0)
    for n in (5, 16, 17, 40, 300):
        size = rng.integers(2, 6, n)
        ulx = rng.integers(0, 4, n) * 45
        perm = oracle_mod.sort_nodes(size, ulx)
        assert sorted(perm.tolist()) == list(range(n))
        k = size[perm] * 1000 + ulx[perm]
        assert np.all(np.diff(k) >= 0)
