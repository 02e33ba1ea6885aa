"""GPU test of orbfe_hamming_top2_static, the entry point for callers that cannot allocate device memory
(ORB-SLAM3's per-frame Tracking thread, which must not stall the GPU work of the LocalMapping/LoopClosing threads
with cudaMalloc/cudaFree, or a host whose pool already owns all device memory).

The new entry point is called through the C ABI twice in a row — the second time with fewer queries, another database
size and the other TH variant, so anything left in the library's built-in scratch array by the first call would show —
and each result is compared with the existing allocating entry point (orbfe_hamming_top2) and, because both share their
kernels, with the CPU oracle as well.  Integer results: the tolerance is zero, as everywhere else in the suite.  Free
device memory is read before and after the second call: it must not move.  The file name sorts last on purpose, so a run that stops at the
first failure has already run every other test; it also passes when run on its own in a fresh process."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _search_both_ways(L, lib, torch, db, q, th, ratio, strict, index_base):
    from orb_slam3_detailed_comments_kor_b200._lib import MATCH_DTYPE, check, ptr
    nq, ndb = q.shape[0], db.shape[0]
    want = np.zeros(nq, MATCH_DTYPE)
    check(L.orbfe_hamming_top2(0, ptr(q), nq, ptr(db), ndb, th, ratio, int(strict), ptr(want)))
    want["idx"] += np.where(want["idx"] >= 0, index_base, 0).astype(np.int32)
    # the caller owns inputs and outputs; the entry point itself must not allocate
    d_q, d_db = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    d_out = torch.full((nq, 4), -7, dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    check(L.orbfe_hamming_top2_static(0, C.c_void_p(d_q.data_ptr()), nq, C.c_void_p(d_db.data_ptr()), ndb, index_base, th, ratio,
                                      int(strict), C.c_void_p(d_out.data_ptr()), C.c_void_p(stream.cuda_stream)))
    stream.synchronize()
    got = d_out.cpu().numpy().view(MATCH_DTYPE).reshape(nq)
    from oracle import orb_oracle as O
    ref = O.hamming_top2(q, db, th, ratio, strict)
    ref["idx"] += np.where(ref["idx"] >= 0, index_base, 0).astype(np.int32)
    for fld in ("best1", "best2", "idx", "accepted"):
        assert np.array_equal(got[fld], want[fld]), fld
        assert np.array_equal(got[fld], ref[fld]), "oracle: " + fld
    return got


def test_static_scratch_entry_point_matches_allocating_one_twice_in_a_row():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import _lib, synth
    L = _lib.lib()
    db1, q1 = synth.descriptors(41, 200000, 4096)          # the largest query count the built-in scratch takes
    first = _search_both_ways(L, _lib, torch, db1, q1, 50, 0.9, False, 0)
    assert first["accepted"].sum() > 100
    db2, q2 = synth.descriptors(42, 3001, 777)             # smaller in both dimensions, `<` variant, shard offset
    second = _search_both_ways(L, _lib, torch, db2, q2, 50, 0.9, True, 1000000)
    assert second["idx"].min() >= 1000000
    # a further call with everything already on the device leaves the amount of free device memory where it was
    d_q, d_db = torch.from_numpy(q2).cuda(), torch.from_numpy(db2).cuda()
    d_out = torch.zeros((q2.shape[0], 4), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    _lib.check(L.orbfe_hamming_top2_static(0, C.c_void_p(d_q.data_ptr()), q2.shape[0], C.c_void_p(d_db.data_ptr()), db2.shape[0], 0, 50, 0.9, 0,
                                           C.c_void_p(d_out.data_ptr()), None))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before
    # an empty database: no candidate for any query (256 / 256 / -1 / rejected), as the oracle says
    _lib.check(L.orbfe_hamming_top2_static(0, C.c_void_p(d_q.data_ptr()), q2.shape[0], None, 0, 0, 50, 0.9, 0, C.c_void_p(d_out.data_ptr()), None))
    torch.cuda.synchronize()
    empty = d_out.cpu().numpy().view(_lib.MATCH_DTYPE).reshape(-1)
    assert (empty["best1"] == 256).all() and (empty["best2"] == 256).all() and (empty["idx"] == -1).all() and not empty["accepted"].any()
    # pointers that are not 16-byte aligned are refused
    rc = L.orbfe_hamming_top2_static(0, C.c_void_p(d_q.data_ptr() + 4), 8, C.c_void_p(d_db.data_ptr()), 100, 0, 50, 0.9, 0, C.c_void_p(d_out.data_ptr()), None)
    assert rc == -2 and b"aligned" in L.orbfe_last_error()
    # beyond the stated limit the call is refused with a clear message, not truncated
    big = torch.zeros((4097, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros((4097, 4), dtype=torch.int32, device="cuda")
    rc = L.orbfe_hamming_top2_static(0, C.c_void_p(big.data_ptr()), 4097, C.c_void_p(big.data_ptr()), 4097, 0, 50, 0.9, 0,
                                     C.c_void_p(out.data_ptr()), None)
    assert rc == -4 and b"4096" in L.orbfe_last_error()
