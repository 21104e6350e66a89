"""GPU: the two forms of the device quadtree (mam3slam_b200/csrc/orbx_kernels_tree.cuh) against the oracle's std::list
DistributeOctTree (src/ORBextractor.cc:555-779).  The count-pyramid form decides on counts alone and keeps one best key per
finest cell; the partitioning form moves the keys pass by pass.  Same kept keys, same order, bit for bit - and when the keys
are so clustered (or so few) that a node deeper than the pyramid would have to be split, the pyramid form must hand over."""
import numpy as np
import pytest

from tests.test_oracle_tree import _keys

pytestmark = pytest.mark.gpu

CASES = [
    # seed, n, w, h, N, clustered
    (0, 10000, 720, 448, 217, False),      # EuRoC level 0, dense: the case the pyramid is for
    (1, 3000, 720, 448, 217, False),
    (2, 800, 720, 448, 217, True),         # clustered
    (3, 150, 178, 102, 60, False),         # top level
    (4, 9000, 1209, 344, 434, False),      # KITTI: 4 roots (1209 / 344 = 3.5 -> 4), 2000 features
    (5, 5000, 1209, 344, 434, True),
    (6, 40, 720, 448, 217, False),         # fewer keys than the quota: every node ends with one key
    (7, 2000, 300, 300, 1086, False),      # one root, large quota
    (8, 1, 720, 448, 10, False),
    (9, 3, 100, 400 // 3, 5, False),
    (10, 20000, 720, 448, 217, False),     # every pixel region crowded
    (11, 600, 600, 150, 100, True),        # 4 roots, clustered
    (12, 36, 720, 448, 217, "block"),      # a 6 x 6 block of adjacent pixels: the list stops growing at once (size == prevSize)
    (14, 0, 720, 448, 1000, "pairs"),      # pairs 1, 2, 4 ... 128 px apart: some node splits usefully at EVERY depth down to 9,
                                           # so the list keeps growing below the pyramid's depth 5 -> hand-over to the partitioning form
    (13, 3000, 720, 448, 217, "block"),    # the same block inside a dense level: the block's node stays unsplit, the pyramid suffices
]


@pytest.mark.parametrize("case", CASES, ids=[f"s{c[0]}-n{c[1]}-{c[2]}x{c[3]}-N{c[4]}{'-' + str(c[5]) if c[5] else ''}" for c in CASES])
def test_pyramid_and_partition_forms_equal_the_oracle(oracle_mod, case):
    from mam3slam_b200 import ORBextractor
    seed, n, w, h, N, cl = case
    rng = np.random.default_rng(seed)
    keys = _keys(rng, n, w, h, cl is True)
    if cl == "pairs":
        pts = []
        for i in range(8):
            for rep in range(12):
                x, y = int(rng.integers(5, 580)), int(rng.integers(5, 440))
                pts += [(x, y), (x + (1 << i), y)]
        xy = np.unique(np.array(pts, np.int32), axis=0)
        xy = xy[np.lexsort((xy[:, 0], xy[:, 1]))]
        keys = np.concatenate([xy, rng.integers(7, 120, (len(xy), 1))], 1).astype(np.int32)
    if cl == "block":
        blk = np.array([[100 + i, 200 + j, 30 + 3 * i + j] for j in range(6) for i in range(6)], np.int32)
        keys = blk if n == 36 else np.concatenate([keys[(keys[:, 0] < 100) | (keys[:, 0] > 105) | (keys[:, 1] < 200) | (keys[:, 1] > 205)], blk])
        keys = keys[np.lexsort((keys[:, 0], keys[:, 1]))]
    # equal responses inside a node: the FIRST maximum in list order must win in both forms
    keys[:, 2] = np.where(rng.random(len(keys)) < 0.5, 50, keys[:, 2])
    ref = oracle_mod.distribute(keys, 16, 16 + w, 16, 16 + h, N)
    ex = ORBextractor(max(N, 100), 1.2, 1, 20, 7)
    part = ex.debug_distribute(keys, 16, 16 + w, 16, 16 + h, N)
    pyr, used = ex.debug_distribute(keys, 16, 16 + w, 16, 16 + h, N, pyramid=True)
    ex.close()
    assert np.array_equal(part, ref)
    assert np.array_equal(pyr, ref), f"pyramid form differs (used_pyramid={used})"
    if seed in (0, 1, 4, 10):
        assert used                      # dense levels stay in the pyramid
    if seed == 14:
        assert not used                  # adjacent pixels separate only at depth 8-9
