"""Projection-gated search (ORBmatcher::SearchByProjection for Tracking::SearchLocalPoints with the frame
grid inside).  CPU: the C++ oracle against the independent Python restatement's outputs (tests/golden/projection.npz).
GPU: the kernel against the oracle and the goldens; every field must be identical."""
import os

import numpy as np
import pytest

from oracle import orb_oracle as O
from orb_slam3_detailed_comments_kor_b200.synth import projection_cases as PC

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "projection.npz")
FIELDS = ["best_idx", "best_dist", "best_level", "second_dist", "second_level", "accepted", "status"]


def _as_matrix(m):
    return np.stack([m[k] for k in FIELDS], 1)


def _golden_case(g, name):
    c = {k.split("/", 1)[1]: g[k] for k in g.files if k.startswith(name + "/")}
    c["kps"] = c["kps"].astype(PC.KP_DTYPE)
    c["queries"] = c["queries"].astype(PC.PROJ_QUERY_DTYPE)
    return c


def _run(fn, c, **kw):
    return fn(c["kps"], c["desc"], c["queries"], c["qdesc"], c["grid"], n_left=int(c["n_left"][0]), u_right=c.get("u_right"),
              partner=c.get("partner"), taken=c["taken"], th_high=int(c["th_high"][0]), nnratio=float(c["nnratio"][0]),
              check_orientation=bool(c["check_orientation"][0]), inv_level_sigma2=c.get("inv_level_sigma2"), **kw)


def test_oracle_matches_python_restatement():
    g = np.load(GOLD)
    seen_status = set()
    for name, kw in PC.CASES.items():
        c = _golden_case(g, name)
        fresh = PC.make_case(**kw)
        for k in ("kps", "desc", "queries", "qdesc", "taken"):
            assert np.array_equal(fresh[k], c[k]), "case generator drifted: %s/%s" % (name, k)
        out, taken, nm = _run(O.search_by_projection, c)
        assert np.array_equal(_as_matrix(out), c["expect"]), name
        assert np.array_equal(taken, c["taken_after"]) and nm == int(c["nmatches"][0]), name
        assert (out["status"] == 1).sum() > 100
        seen_status.update(out["status"].tolist())
    assert seen_status >= {0, 1, 2, 3, 4, 6, 7, 8}, seen_status      # every outcome occurs somewhere in the goldens


def test_oracle_known_answers():
    """hand-checked micro cases: order of the window scan, occupancy, the level rule of the ratio test"""
    kps = np.zeros(3, PC.KP_DTYPE)
    kps["x"], kps["y"], kps["octave"] = [100, 101, 130], [100, 101, 100], [1, 1, 2]
    desc = np.zeros((3, 32), np.uint8)
    desc[1, 0] = 0x0F          # distance 4 from a zero query; rows 0 and 2 are at distance 0
    grid = np.array([0, 0, 64 / 752, 48 / 480], np.float32)
    q = np.zeros(3, PC.PROJ_QUERY_DTYPE)
    q["x"], q["y"], q["radius"], q["min_level"], q["max_level"], q["flags"] = [100.5, 100.5, 130], [100.5, 100.5, 100], 4, 0, 2, 2
    qd = np.zeros((3, 32), np.uint8)
    out, taken, nm = O.search_by_projection(kps, desc, q, qd, grid, nnratio=0.8)
    # query 0: best row 0 (0), second row 1 (4), same level, 0 > 0.8*4 is false -> accepted, row 0 occupied
    assert tuple(out[0])[:6] == (0, 0, 1, 4, 1, 1)
    # query 1: row 0 is gone: best row 1 (4), no second (256, level -1) -> levels differ -> accepted
    assert tuple(out[1])[:6] == (1, 4, 1, 256, -1, 1)
    assert tuple(out[2])[:6] == (2, 0, 2, 256, -1, 1) and nm == 3 and taken.tolist() == [1, 1, 1]
    # same level and too close: rejected
    desc[0, 0] = 0x07
    out, taken, nm = O.search_by_projection(kps, desc, q[:1], qd[:1], grid, nnratio=0.7)
    assert tuple(out[0]) == (0, 3, 1, 4, 1, 0, 3) and nm == 0 and taken.sum() == 0


@pytest.mark.gpu
def test_gpu_search_matches_oracle_and_goldens():
    from orb_slam3_detailed_comments_kor_b200 import projection as P
    g = np.load(GOLD)
    for name in PC.CASES:
        c = _golden_case(g, name)
        out, taken, nm = _run(P.search_by_projection, c)
        bad = np.nonzero((_as_matrix(out) != c["expect"]).any(1))[0]
        assert bad.size == 0, (name, bad[:5], _as_matrix(out)[bad[:3]], c["expect"][bad[:3]])
        assert np.array_equal(taken, c["taken_after"]) and nm == int(c["nmatches"][0]), name
    # larger and harder: the per-frame capacity, wide windows and crowds (long chains of map points after the same keypoints)
    for kw in (dict(seed=21, n=4096, nq=6000, th=1.0), dict(seed=22, n=3000, nq=5000, th=5.0, crowd=0.6, stereo_gate=True),
               dict(seed=23, n=2600, nq=4000, th=3.0, crowd=0.8, fisheye=True), dict(seed=24, n=40, nq=900, th=15.0, crowd=1.0),
               dict(seed=25, n=1, nq=5), dict(seed=26, n=1500, nq=1, th=2.0),
               dict(seed=29, n=2400, nq=3000, th=7.0, stereo_gate=True, crowd=0.3, last_frame="still"),
               dict(seed=30, n=3000, nq=3500, th=15.0, fisheye=True, crowd=0.5, last_frame="forward"),
               dict(seed=31, n=60, nq=800, th=15.0, crowd=1.0, last_frame="backward"),
               dict(seed=32, n=3500, nq=5000, th=3.0, stereo_gate=True, crowd=0.3, fuse=True)):
        c = PC.make_case(**kw)
        want, wtaken, wnm = _run(O.search_by_projection, c)
        out, taken, nm = _run(P.search_by_projection, c)
        bad = np.nonzero((_as_matrix(out) != _as_matrix(want)).any(1))[0]
        assert bad.size == 0, (kw, bad[:5], out[bad[:3]], want[bad[:3]])
        assert np.array_equal(taken, wtaken) and nm == wnm, kw
    # the worst case for the speculative rounds: every map point has the same descriptor and sees every keypoint, so each one
    # depends on all the ones before it (one query resolves per round); must still be exact, and finish
    c = PC.make_case(seed=33, n=1000, nq=1500, th=1.0)
    c["qdesc"][:] = c["desc"][7]
    c["queries"]["radius"] = 2000.0
    c["queries"]["min_level"], c["queries"]["max_level"] = 0, -1
    c["queries"]["flags"] = PC.PQ_BLOCKS | PC.PQ_BEST_ONLY
    c["th_high"] = np.array([256], np.int32)               # every map point takes the best keypoint that is left
    want, wtaken, wnm = _run(O.search_by_projection, c)
    out, taken, nm = _run(P.search_by_projection, c)
    assert np.array_equal(_as_matrix(out), _as_matrix(want)) and np.array_equal(taken, wtaken) and nm == wnm and nm > 800
    # no queries, and a frame beyond the capacity
    c = PC.make_case(seed=27, n=50, nq=4)
    out, taken, nm = P.search_by_projection(c["kps"], c["desc"], c["queries"][:0], c["qdesc"][:0], c["grid"])
    assert out.shape == (0,) and nm == 0
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    big = PC.make_case(seed=28, n=4097, nq=4)
    with pytest.raises(OrbfeError):
        P.search_by_projection(big["kps"], big["desc"], big["queries"], big["qdesc"], big["grid"])


@pytest.mark.gpu
def test_gpu_search_batch_device():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import projection as P
    cases = [PC.make_case(seed=40 + i, n=900 + 150 * i, nq=1100 + 100 * i, th=1.0 + 0.5 * (i % 3), crowd=0.2 * (i % 3), stereo_gate=True,
                          last_frame="still")
             for i in range(7)]
    cases[3]["kps"] = np.concatenate([cases[3]["kps"]] * 3)[:2000]     # frame 3 breaks the max_keypoints promise below
    cases[3]["desc"] = np.concatenate([cases[3]["desc"]] * 3)[:2000]
    cases[3]["taken"] = np.concatenate([cases[3]["taken"]] * 3)[:2000]
    cases[3]["u_right"] = np.concatenate([cases[3]["u_right"]] * 3)[:2000]
    cat = lambda k: np.concatenate([c[k] for c in cases])  # noqa: E731
    koff = np.cumsum([0] + [len(c["kps"]) for c in cases]).astype(np.int32)
    qoff = np.cumsum([0] + [len(c["queries"]) for c in cases]).astype(np.int32)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()  # noqa: E731
        d_kps, d_desc, d_q, d_qd = dev(cat("kps")), dev(cat("desc")), dev(cat("queries")), dev(cat("qdesc"))
        d_ur, d_tk, d_grid = dev(cat("u_right")), dev(cat("taken")), dev(np.stack([c["grid"] for c in cases]))
        d_koff, d_qoff = dev(koff), dev(qoff)
        d_out = torch.zeros(int(qoff[-1]) * 28, dtype=torch.uint8, device="cuda")
        d_tko = torch.zeros(int(koff[-1]), dtype=torch.uint8, device="cuda")
        P.search_by_projection_device(len(cases), 1999, d_kps.data_ptr(), d_desc.data_ptr(), d_koff.data_ptr(), d_grid.data_ptr(), d_q.data_ptr(),
                                      d_qd.data_ptr(), d_qoff.data_ptr(), d_out.data_ptr(), d_u_right=d_ur.data_ptr(), d_taken_in=d_tk.data_ptr(),
                                      d_taken_out=d_tko.data_ptr(), check_orientation=True, stream=st.cuda_stream)
        out = d_out.cpu().numpy().view(P.PROJ_MATCH_DTYPE)
        tko = d_tko.cpu().numpy()
    st.synchronize()
    for i, c in enumerate(cases):
        got = out[qoff[i]:qoff[i + 1]]
        if i == 3:
            assert (got["status"] == P.PS_CAPACITY).all() and (got["accepted"] == 0).all()
            continue
        want, wtaken, _ = _run(O.search_by_projection, c)
        assert np.array_equal(_as_matrix(got), _as_matrix(want)), "frame %d" % i
        assert np.array_equal(tko[koff[i]:koff[i + 1]], wtaken), "frame %d occupancy" % i


@pytest.mark.gpu
def test_gpu_host_entry_points_from_several_threads_and_release():
    """The one-frame host entry points stage their arguments through a block that belongs to the calling thread.  Four threads call at the
    same time with sizes that change from call to call (the block grows and is reused), give the block back in between
    (orbfe_thread_release) and call again; every result must equal the oracle's.  A large batch of frames (more than two per SM: the
    512-thread kernel without the shared-memory cache at 4096 keypoints) and a small one go through the device entry point too."""
    from concurrent.futures import ThreadPoolExecutor
    from orb_slam3_detailed_comments_kor_b200 import projection as P
    from orb_slam3_detailed_comments_kor_b200 import _lib
    cases = [PC.make_case(seed=60 + i, n=n, nq=nq, th=th, crowd=crowd, stereo_gate=True, last_frame=lf)
             for i, (n, nq, th, crowd, lf) in enumerate([(300, 200, 1.0, 0.0, None), (2500, 3000, 3.0, 0.3, None), (40, 900, 15.0, 1.0, "still"),
                                                         (4096, 5000, 1.0, 0.0, None), (1200, 1200, 7.0, 0.2, "forward"), (1, 3, 1.0, 0.0, None)])]
    wants = [_run(O.search_by_projection, c) for c in cases]

    def worker(t):
        bad = []
        for rep in range(3):
            for k in range(len(cases)):
                i = (k + t) % len(cases)
                out, taken, nm = _run(P.search_by_projection, cases[i])
                if not (np.array_equal(_as_matrix(out), _as_matrix(wants[i][0])) and np.array_equal(taken, wants[i][1]) and nm == wants[i][2]):
                    bad.append((t, rep, i))
            _lib.lib().orbfe_thread_release()
        return bad
    with ThreadPoolExecutor(4) as tp:
        res = list(tp.map(worker, range(4)))
    assert not any(res), res
    _lib.lib().orbfe_thread_release()     # a thread that holds nothing: no-op
    out, taken, nm = _run(P.search_by_projection, cases[0])
    assert np.array_equal(_as_matrix(out), _as_matrix(wants[0][0])) and nm == wants[0][2]
    # more than two frames per SM: the 512-thread kernel, without the shared-memory cache (4096 keypoints promised) and with it (1300)
    import torch
    for sel, maxkp in (([0, 1, 2, 3, 4, 5], 4096), ([0, 2, 4, 5], 1300)):
        order = [sel[i % len(sel)] for i in range(330)]
        cat = lambda k: np.concatenate([cases[i][k] for i in order])  # noqa: E731
        koff = np.cumsum([0] + [len(cases[i]["kps"]) for i in order]).astype(np.int32)
        qoff = np.cumsum([0] + [len(cases[i]["queries"]) for i in order]).astype(np.int32)
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()  # noqa: E731
        d_kps, d_desc, d_q, d_qd = dev(cat("kps")), dev(cat("desc")), dev(cat("queries")), dev(cat("qdesc"))
        d_ur, d_tk, d_grid = dev(cat("u_right")), dev(cat("taken")), dev(np.stack([cases[i]["grid"] for i in order]))
        d_koff, d_qoff = dev(koff), dev(qoff)
        d_out = torch.zeros(int(qoff[-1]) * 28, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        # check_orientation differs between the cases (the frame-to-frame ones have it): one launch per setting, compare the frames it belongs to
        for ori in (False, True):
            P.search_by_projection_device(len(order), maxkp, d_kps.data_ptr(), d_desc.data_ptr(), d_koff.data_ptr(), d_grid.data_ptr(), d_q.data_ptr(),
                                          d_qd.data_ptr(), d_qoff.data_ptr(), d_out.data_ptr(), d_u_right=d_ur.data_ptr(), d_taken_in=d_tk.data_ptr(),
                                          check_orientation=ori)
            torch.cuda.synchronize()
            out = d_out.cpu().numpy().view(P.PROJ_MATCH_DTYPE)
            for f, i in enumerate(order):
                if bool(cases[i]["check_orientation"][0]) == ori:
                    assert np.array_equal(_as_matrix(out[qoff[f]:qoff[f + 1]]), _as_matrix(wants[i][0])), (maxkp, f, i)
