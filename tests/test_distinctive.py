"""MapPoint::ComputeDistinctiveDescriptors for batches of map points (/root/reference/src/MapPoint.cc:349-423).
CPU: the C++ oracle against a numpy statement of the same rule and hand-made cases; GPU: the kernel against the oracle."""
import numpy as np
import pytest

from oracle import orb_oracle as O

POPC = np.array([bin(i).count("1") for i in range(256)], np.int32)


def make_points(seed, npoints, max_obs):
    """Map points as clusters: a base descriptor observed n times with 0..40 flipped bits, sometimes exact repeats (ties)."""
    rng = np.random.RandomState(seed)
    rows, off = [], [0]
    for p in range(npoints):
        n = int(rng.randint(0, max_obs + 1)) if p % 7 else int(rng.randint(0, 3))
        base = rng.randint(0, 256, 32).astype(np.uint8)
        for i in range(n):
            d = base.copy()
            if rows and i and rng.uniform() < 0.2:
                d = rows[-1].copy()
            else:
                for b in rng.choice(256, size=int(rng.randint(0, 41)), replace=False):
                    d[b >> 3] ^= np.uint8(1 << (b & 7))
            rows.append(d)
        off.append(len(rows))
    desc = np.stack(rows) if rows else np.zeros((0, 32), np.uint8)
    return desc, np.array(off, np.int32)


def numpy_rule(desc, off):
    idx, med = [], []
    for p in range(len(off) - 1):
        d = desc[off[p]:off[p + 1]]
        n = len(d)
        if n == 0:
            idx.append(-1); med.append(np.iinfo(np.int32).max)
            continue
        dist = POPC[d[:, None, :] ^ d[None, :, :]].sum(2)
        m = np.sort(dist, 1)[:, int(0.5 * (n - 1))]
        idx.append(int(np.argmin(m))); med.append(int(m.min()))       # argmin returns the first minimum, like the strict `<`
    return np.array(idx, np.int32), np.array(med, np.int32)


def test_oracle_matches_numpy_rule_and_known_answers():
    desc, off = make_points(3, 300, 40)
    idx, med = O.distinctive_descriptors(desc, off)
    widx, wmed = numpy_rule(desc, off)
    assert np.array_equal(idx, widx) and np.array_equal(med, wmed)
    assert (np.diff(off) == 0).any() and (np.diff(off) == 1).any() and (np.diff(off) == 2).any()
    # three rows: a, a with 2 bits flipped, a with 6 bits flipped (all inside the first byte pair) -> distances
    # row0: 0,2,6  row1: 2,0,4  row2: 6,4,0; medians (index int(0.5*2) = 1): 2, 2, 4 -> first of the tie = row 0
    a = np.zeros((3, 32), np.uint8)
    a[1, 0] = 0b11
    a[2, 0] = 0b111111
    idx, med = O.distinctive_descriptors(a, np.array([0, 3], np.int32))
    assert idx[0] == 0 and med[0] == 2
    # two rows: median index int(0.5*1) = 0 -> the self distance 0 for both -> row 0
    idx, med = O.distinctive_descriptors(a[1:], np.array([0, 2], np.int32))
    assert idx[0] == 0 and med[0] == 0


@pytest.mark.gpu
def test_gpu_matches_oracle():
    import torch
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    from orb_slam3_detailed_comments_kor_b200 import _lib
    for seed, npts, mx in ((5, 1, 1), (6, 2000, 12), (7, 300, 150), (8, 3, 700)):
        desc, off = make_points(seed, npts, mx)
        if desc.shape[0] == 0:
            desc = np.zeros((1, 32), np.uint8)
        want = O.distinctive_descriptors(desc, off)
        got = M.distinctive_descriptors(desc, off)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (seed, npts, mx)
    # device entry point on a stream
    desc, off = make_points(9, 5000, 20)
    want = O.distinctive_descriptors(desc, off)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        d_desc, d_off = torch.from_numpy(desc).cuda(), torch.from_numpy(off).cuda()
        d_idx = torch.zeros(len(off) - 1, dtype=torch.int32, device="cuda")
        _lib.check(_lib.lib().orbfe_distinctive_descriptors_device(0, d_desc.data_ptr(), d_off.data_ptr(), len(off) - 1, d_idx.data_ptr(), None, st.cuda_stream))
        got = d_idx.cpu().numpy()
    st.synchronize()
    assert np.array_equal(got, want[0])
