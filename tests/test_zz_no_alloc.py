"""orbb200_match_bruteforce_top2_device_noalloc (include/orbb200.h): the brute-force top-2 search with its per-chunk scratch in a static
device array -- no cudaMalloc on the call path.  No counterpart in the reference.
Checked through the C-ABI against the existing entry point orbb200_match_bruteforce_top2_device (bit-exact: integer work) on two calls
in a row that share the array, plus the size limit.  Collected last on purpose."""
import numpy as np
import pytest

import orbb200
import synth

pytestmark = pytest.mark.gpu


def _search(fn, tq, tdb, out, stream):
    orbb200.check(fn(tq.data_ptr(), tq.shape[0], tdb.data_ptr(), tdb.shape[0], out.data_ptr(), 0, stream))


def test_noalloc_bruteforce_equals_existing_entry_point_twice_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = orbb200.lib()
    # second call smaller than the first (fewer queries, fewer database chunks): it runs over the first call's leftovers in the array
    cases = [(700, 50000, 11), (300, 9000, 12)]
    data = []
    for nq, ndb, seed in cases:
        q, db, _ = synth.descriptor_db(nq, ndb, seed=seed)
        tq, tdb = torch.from_numpy(q).to(dev), torch.from_numpy(db).to(dev)
        want = torch.empty((nq, 4), dtype=torch.int32, device=dev)
        got = torch.full((nq, 4), -7, dtype=torch.int32, device=dev)
        data.append((tq, tdb, want, got))
    st = torch.cuda.Stream(device=dev)
    for tq, tdb, want, _ in data:  # existing entry point (cudaMalloc'ed scratch)
        _search(L.orbb200_match_bruteforce_top2_device, tq, tdb, want, st.cuda_stream)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]
    for tq, tdb, _, got in data:   # the two calls in a row, same stream (stream-ordered: no overlap on the shared array)
        _search(L.orbb200_match_bruteforce_top2_device_noalloc, tq, tdb, got, st.cuda_stream)
    torch.cuda.synchronize()
    for _, _, want, got in data:
        assert torch.equal(want, got), (want[:4].cpu(), got[:4].cpu())
    # once more on the default stream (stream == NULL), first case again after the smaller one
    tq, tdb, want, got = data[0]
    got.fill_(-7)
    torch.cuda.synchronize()
    _search(L.orbb200_match_bruteforce_top2_device_noalloc, tq, tdb, got, None)
    torch.cuda.synchronize()
    assert torch.equal(want, got)
    free1 = torch.cuda.mem_get_info(dev)[0]
    assert free1 >= free0 - (2 << 20), "device memory shrank across the no-alloc calls: %d -> %d" % (free0, free1)
    # sanity: the results are real matches (planted neighbours), not an untouched buffer
    g = got.cpu().numpy()
    assert (g[:, 1] >= 0).all() and (g[:, 0] <= g[:, 2]).all() and (g[:, 0] < 256).all()


def test_noalloc_bruteforce_rejects_sizes_beyond_the_static_scratch():
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    L = orbb200.lib()
    nq = 524288 + 1  # one query more than the array holds partial records
    tq = torch.zeros((nq, 32), dtype=torch.uint8, device=dev)
    tdb = torch.zeros((128, 32), dtype=torch.uint8, device=dev)
    out = torch.full((nq, 4), -7, dtype=torch.int32, device=dev)
    rc = L.orbb200_match_bruteforce_top2_device_noalloc(tq.data_ptr(), nq, tdb.data_ptr(), 128, out.data_ptr(), 0, None)
    assert rc == -3  # ORBB200_ERR_CAPACITY
    msg = L.orbb200_last_error().decode()
    assert "static" in msg and "524288" in msg, msg
    torch.cuda.synchronize()
    assert int(out[0, 0]) == -7 and int(out[-1, 3]) == -7  # nothing was launched
    # at the limit it runs
    nq = 524288
    rc = L.orbb200_match_bruteforce_top2_device_noalloc(tq.data_ptr(), nq, tdb.data_ptr(), 128, out.data_ptr(), 0, None)
    orbb200.check(rc)
    torch.cuda.synchronize()
    o = out[:nq].cpu().numpy()
    assert (o[:, 0] == 0).all() and (o[:, 1] == 0).all() and (o[:, 2] == 0).all() and (o[:, 3] == 1).all()
