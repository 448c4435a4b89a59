"""gforb_allpairs_top2_static (include/gforb.h): the allocation-free variant of the all-pairs top-2 search -- scratch in a
file-scope device array, no handle, no cudaMalloc.  Called through the C ABI twice in a row (a larger problem first, so the
second call finds the array full of the first call's keys) and compared bit for bit with gforb_allpairs_top2.
Named test_zz_* so that pytest collects it last; it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GFORB_ERR_CAPACITY, GFORB_ERR_ARG = -3, -2
FLAGS_DEV = 1 | 2           # GFORB_IN_DEVICE | GFORB_OUT_DEVICE


def _static_call(L, qd, dd, idx_base, out, flags=FLAGS_DEV):
    return L.gforb_allpairs_top2_static(0, C.c_void_p(qd.data_ptr()), int(qd.shape[0]), C.c_void_p(dd.data_ptr()), int(dd.shape[0]),
                                        int(idx_base), C.c_void_p(out.data_ptr()), flags, C.c_void_p(0))


def test_static_allpairs_matches_allocating_entry_point(monkeypatch):
    import torch
    from gf_orb_slam_b200 import ORBmatcher, _lib
    monkeypatch.delenv("GFORB_ALLPAIRS", raising=False)
    monkeypatch.delenv("GFORB_ALLPAIRS_SPLITS", raising=False)
    L = _lib.lib()
    torch.cuda.set_device(0)
    m = ORBmatcher(max_batch=1)
    cases = [(2000, 70013, 1000), (300, 5000, 7)]                     # (n_q, n_db, idx_base): two calls in a row, the second smaller
    data = []
    for nq, ndb, base in cases:
        rng = np.random.default_rng(100 + nq)
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        db[100] = q[5]; db[ndb - 1] = q[5]                              # an exact hit twice: the tie goes to the lower row
        qd, dd = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
        exp = m.AllPairsTop2(qd, dd, idx_base=base).clone()           # the existing (allocating) entry point
        out = torch.full((nq, 2), -1, dtype=torch.int64, device="cuda")
        data.append((qd, dd, base, exp, out))
    torch.cuda.synchronize()
    for i, (qd, dd, base, exp, out) in enumerate(data):
        free0 = torch.cuda.mem_get_info()[0]
        rc = _static_call(L, qd, dd, base, out)
        assert rc == 0, rc
        torch.cuda.synchronize()
        assert torch.equal(out, exp), i                                # integer keys: bit-exact, like the other all-pairs tests
        assert int(out[5, 0]) == base + 100
        if i > 0:                                                      # after the one-time set-up of the first call: nothing is allocated
            assert torch.cuda.mem_get_info()[0] == free0
    # the same problem again, asynchronously on the default stream, then joined by the caller
    qd, dd, base, exp, out = data[0]
    out.fill_(-1)
    assert _static_call(L, qd, dd, base, out, FLAGS_DEV | 4) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, exp)


def test_static_allpairs_rejects_sizes_beyond_the_scratch():
    import torch
    from gf_orb_slam_b200 import _lib
    L = _lib.lib()
    torch.cuda.set_device(0)
    qd = torch.zeros((4097, 32), dtype=torch.uint8, device="cuda")
    dd = torch.zeros((1000, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros((4097, 2), dtype=torch.int64, device="cuda")
    assert _static_call(L, qd, dd, 0, out) == GFORB_ERR_CAPACITY       # more than 4096 queries
    assert _static_call(L, qd[:16], dd, 0, out, flags=0) == GFORB_ERR_ARG   # host pointers are not accepted
    assert _static_call(L, qd[:16], dd, 0, out) == 0
