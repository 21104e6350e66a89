"""GPU: orbm_knn_device_static, the allocation-free k-NN entry point (include/orbm.h), against orbm_knn_device.

The entry point has no counterpart in the reference; it was added for callers that
cannot call cudaMalloc.  Its scratch is a fixed-size static device array shared by all calls, which keeps the previous call's
contents - hence two calls in a row with DIFFERENT inputs, each checked against the existing entry point (and the oracle).
Integer work: the bar is bit-exact, like every other k-NN test here.  Collected last on purpose (round end runs pytest -x).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(m, D, q_dev, nq, k, static):
    import torch
    dd = torch.full((nq, k), -7, dtype=torch.int16, device="cuda")
    di = torch.full((nq, k), -7, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    (m.knn_device_static if static else m.knn_device)(D, q_dev.data_ptr(), nq, k, dd.data_ptr(), di.data_ptr(), st)
    torch.cuda.synchronize()
    return dd.cpu().numpy().view(np.uint16), di.cpu().numpy()


def test_static_scratch_knn_matches_allocating_entry_point(oracle_mod):
    import torch
    from mam3slam_b200 import ORBmatcher
    from mam3slam_b200.matcher import KnnDatabase
    from oracle import matcher as om

    m = ORBmatcher(0.6, True)
    rng = np.random.default_rng(2024)
    # two calls in a row on the shared static array: a larger search first, then a smaller one with other rows, queries and k,
    # so that stale entries of the first call would show if anything read them
    cases = [(30000, 700, 2, 1000), (4099, 129, 5, 0)]
    for ndb, nq, k, row_offset in cases:
        rows = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        q = rows[rng.integers(0, ndb, nq)].copy()
        q[:, 3] ^= 0x11                                        # planted neighbours at distance 2
        q[nq // 2:] = rng.integers(0, 256, (nq - nq // 2, 32), dtype=np.uint8)
        rows[7] = rows[3]                                      # duplicate rows: ties go to the lowest row id
        rows_dev = torch.from_numpy(rows).cuda()
        q_dev = torch.from_numpy(q).cuda()
        D = KnnDatabase(m, None, row_offset=row_offset, device_ptr=rows_dev.data_ptr(), n=ndb)
        d_ref, i_ref = _run(m, D, q_dev, nq, k, static=False)
        allocs = m._L.orbm_debug_alloc_count()
        d_new, i_new = _run(m, D, q_dev, nq, k, static=True)
        assert m._L.orbm_debug_alloc_count() == allocs        # the library made no cudaMalloc / cudaMallocHost / stream creation
        assert np.array_equal(d_new, d_ref) and np.array_equal(i_new, i_ref)
        od, oi = om.knn(q, rows, k)                            # and both against the CPU oracle
        assert np.array_equal(d_new, od) and np.array_equal(i_new, oi + row_offset)
        assert (d_new[: nq // 2, 0] <= 2).all()
        # once more on the same inputs: the array now holds this very call's entries
        d_again, i_again = _run(m, D, q_dev, nq, k, static=True)
        assert np.array_equal(d_again, d_ref) and np.array_equal(i_again, i_ref)
        D.close()
    m.close()


def test_static_scratch_knn_rejects_oversized_requests():
    import torch
    from mam3slam_b200 import ORBmatcher
    from mam3slam_b200._lib import NativeError
    from mam3slam_b200.matcher import KnnDatabase

    m = ORBmatcher(0.6, True)
    ndb, nq, k = 200000, 40000, 16                             # chunks x nq x k far beyond ORBM_KNN_STATIC_MAX_ENTRIES (2^20)
    rows_dev = torch.zeros((ndb, 32), dtype=torch.uint8, device="cuda")
    q_dev = torch.zeros((nq, 32), dtype=torch.uint8, device="cuda")
    dd = torch.empty((nq, k), dtype=torch.int16, device="cuda")
    di = torch.empty((nq, k), dtype=torch.int32, device="cuda")
    D = KnnDatabase(m, None, device_ptr=rows_dev.data_ptr(), n=ndb)
    with pytest.raises(NativeError, match="ORBM_KNN_STATIC_MAX_ENTRIES"):
        m.knn_device_static(D, q_dev.data_ptr(), nq, k, dd.data_ptr(), di.data_ptr())
    m.knn_device(D, q_dev.data_ptr(), nq, k, dd.data_ptr(), di.data_ptr())       # the allocating entry point still takes it
    torch.cuda.synchronize()
    D.close()
    m.close()
