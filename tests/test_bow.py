"""Bag-of-words assignment: Frame::ComputeBoW -> DBoW2 TemplatedVocabulary::transform.  The reference's
vocabulary file is not in the repository, so the trees are synthetic.  CPU: the C++ oracle against a plain-Python restatement
written from the DBoW2 headers; GPU: the kernels against the oracle — word ids, node ids and feature vectors identical, BowVector
values bit-identical doubles."""
import numpy as np
import pytest

from oracle import orb_oracle as O

POPC = np.array([bin(i).count("1") for i in range(256)], np.int32)


def make_vocabulary(seed, k, L, prune=0.1, stop=0.02):
    """Rows of a vocabulary text file in DBoW2's creation order (HKmeansStep: a node's children get consecutive ids, then each
    child is expanded in turn).  prune: chance that a node above the last level stays a leaf or gets fewer than k children."""
    rng = np.random.RandomState(seed)
    parent, leaf, desc, weight = [], [], [], []

    def expand(pid, pdesc, level):
        nch = k if rng.uniform() > prune else int(rng.randint(1, k + 1))
        ids = []
        for _ in range(nch):
            d = pdesc.copy()
            for b in rng.choice(256, size=int(256 >> (level + 1)) // 2 + 2, replace=False):
                d[b >> 3] ^= np.uint8(1 << (b & 7))
            parent.append(pid); desc.append(d); leaf.append(0); weight.append(0.0)
            ids.append(len(parent))                       # node id = row + 1
        for nid in ids:
            if level < L and rng.uniform() > prune * 0.5:
                expand(nid, desc[nid - 1], level + 1)
            else:
                leaf[nid - 1] = 1
                weight[nid - 1] = 0.0 if rng.uniform() < stop else float(rng.uniform(0.1, 9.0))
        # a duplicate child descriptor now and then: ties in the descent must go to the first child
        if nch > 1 and rng.uniform() < 0.2:
            desc[ids[-1] - 1] = desc[ids[0] - 1].copy()

    expand(0, rng.randint(0, 256, 32).astype(np.uint8), 1)
    return (np.array(parent, np.int32), np.array(leaf, np.uint8), np.stack(desc).astype(np.uint8), np.array(weight, np.float64))


def make_features(seed, voc_desc, n):
    rng = np.random.RandomState(seed)
    base = voc_desc[rng.randint(0, len(voc_desc), n)].copy()
    for i in range(n):
        for b in rng.choice(256, size=int(rng.randint(0, 30)), replace=False):
            base[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return base


def python_transform(k, L, scoring, weighting, rows, feats, levelsup):
    """TemplatedVocabulary::transform(features, v, fv, levelsup) in plain Python (dicts stand for the std::maps)."""
    parent, leaf, desc, weight = rows
    children = {}
    word_of, nwords = {}, 0
    for r in range(len(parent)):
        children.setdefault(int(parent[r]), []).append(r + 1)
        if leaf[r]:
            word_of[r + 1] = nwords
            nwords += 1
    v, fv = {}, {}
    wids, nids = [], []
    for i, f in enumerate(feats):
        node, level, nid = 0, 0, 0
        while True:
            level += 1
            ch = children[node]
            d = [int(POPC[desc[c - 1] ^ f].sum()) for c in ch]
            node = ch[int(np.argmin(d))]                  # first minimum
            if level == L - levelsup:
                nid = node
            if node not in children:
                break
        wid, w = word_of[node], float(weight[node - 1])
        wids.append(wid); nids.append(nid)
        if w > 0:
            if weighting in (0, 1):
                v[wid] = v[wid] + w if wid in v else w
            elif wid not in v:
                v[wid] = w
            fv.setdefault(nid, []).append(i)
    must = scoring != 5
    if weighting in (0, 1) and v and not must:
        nd = float(len(v))
        for key in v:
            v[key] /= nd
    if must:
        norm = 0.0
        if scoring != 1:
            for key in sorted(v):
                norm += abs(v[key])
        else:
            for key in sorted(v):
                norm += v[key] * v[key]
            norm = float(np.sqrt(np.float64(norm)))
        if norm > 0:
            for key in v:
                v[key] /= norm
    keys = sorted(v)
    return {"word_ids": np.array(wids, np.int32), "node_ids": np.array(nids, np.int32), "bow_words": np.array(keys, np.int32),
            "bow_values": np.array([v[key] for key in keys], np.float64), "feat_vec": {key: np.array(fv[key], np.int32) for key in sorted(fv)}}


def same(a, b):
    ok = all(np.array_equal(a[key], b[key]) for key in ("word_ids", "node_ids", "bow_words"))
    ok = ok and np.array_equal(a["bow_values"].view(np.uint64), b["bow_values"].view(np.uint64))      # bit-identical doubles
    ok = ok and list(a["feat_vec"].keys()) == list(b["feat_vec"].keys())
    return ok and all(np.array_equal(a["feat_vec"][key], b["feat_vec"][key]) for key in a["feat_vec"])


CONFIGS = [(4, 3, 0, 0, 2, 11), (10, 3, 0, 0, 1, 12), (5, 4, 1, 1, 4, 13), (3, 5, 5, 0, 2, 14), (6, 3, 2, 2, 3, 15), (4, 4, 5, 3, 0, 16)]
# k, L, scoring, weighting, levelsup, seed


def test_oracle_matches_python_restatement():
    for k, L, scoring, weighting, levelsup, seed in CONFIGS:
        rows = make_vocabulary(seed, k, L)
        voc = O.OracleVocabulary(k, L, scoring, weighting, *rows)
        assert voc.words() == int(rows[1].sum())
        feats = make_features(seed + 100, rows[2], 300)
        got = voc.transform(feats, levelsup)
        want = python_transform(k, L, scoring, weighting, rows, feats, levelsup)
        assert same(got, want), (k, L, scoring, weighting)
        assert len(got["bow_words"]) > 20 and len(got["feat_vec"]) >= 1
        if scoring == 0:
            assert abs(got["bow_values"].sum() - 1.0) < 1e-12
    # no features: both vectors empty
    got = voc.transform(np.zeros((0, 32), np.uint8), 2)
    assert len(got["bow_words"]) == 0 and len(got["feat_vec"]) == 0


@pytest.mark.gpu
def test_gpu_transform_matches_oracle():
    from orb_slam3_detailed_comments_kor_b200.bow import ORBVocabulary
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    for k, L, scoring, weighting, levelsup, seed in CONFIGS + [(10, 4, 0, 0, 2, 17)]:
        rows = make_vocabulary(seed, k, L)
        orc = O.OracleVocabulary(k, L, scoring, weighting, *rows)
        voc = ORBVocabulary(k, L, scoring, weighting, *rows, max_features=4096)
        assert voc.words() == orc.words()
        for n in (0, 1, 33, 1200, 4096):
            feats = make_features(seed + n, rows[2], n)
            assert same(voc.transform(feats, levelsup), orc.transform(feats, levelsup)), (k, L, scoring, weighting, n)
        with pytest.raises(OrbfeError):
            voc.transform(np.zeros((4097, 32), np.uint8), levelsup)
        voc.close()
    bad = list(make_vocabulary(11, 4, 3))
    bad[1] = 1 - bad[1]                                   # leaf flags that contradict the tree
    with pytest.raises(OrbfeError):
        ORBVocabulary(4, 3, 0, 0, *bad)


@pytest.mark.gpu
def test_gpu_transform_chained_after_extraction():
    """Frame::Frame then Frame::ComputeBoW: the extractor's device buffers go straight into the vocabulary."""
    import ctypes as C
    import torch
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth, _lib
    from orb_slam3_detailed_comments_kor_b200.bow import ORBVocabulary
    W, H, params, nb = 752, 480, (1200, 1.2, 8, 20, 7), 5
    imgs = np.stack([synth.frame(60 + i, W, H) for i in range(nb)])
    imgs[2] = 128                                        # a frame without keypoints inside the batch
    ex = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cnt, mono, kps, desc = ex.extract_batch_host(imgs, (0, 0))
    rows = make_vocabulary(31, 10, 4)
    voc = ORBVocabulary(10, 4, 0, 0, *rows, max_features=ex.capacity, max_batch=nb)
    orc = O.OracleVocabulary(10, 4, 0, 0, *rows)
    d_kps, d_desc, d_cnt, d_mono = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    _lib.check(_lib.lib().orbfe_batch_device_views(ex.handle, C.byref(d_kps), C.byref(d_desc), C.byref(d_cnt), C.byref(d_mono)))
    st = torch.cuda.Stream()
    voc.transform_batch_device(d_desc.value, ex.capacity, d_cnt.value, nb, 2, stream=st.cuda_stream)
    st.synchronize()
    for f in range(nb):
        assert same(voc.result(f, int(cnt[f])), orc.transform(desc[f, :cnt[f]], 2)), "frame %d" % f
    assert cnt[2] == 0 and len(voc.result(2, 0)["bow_words"]) == 0
    voc.close(); ex.close()


@pytest.mark.gpu
def test_gpu_transform_batch_reports_frames_beyond_max_features():
    """DBoW2 uses every feature of a frame; a device batch whose frame has more descriptors than the object holds must say so
    (ORBFE_E_CAPACITY for that frame), not return vectors of a truncated feature set.  The other frames stay correct."""
    import torch
    from orb_slam3_detailed_comments_kor_b200.bow import ORBVocabulary
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    rows = make_vocabulary(7, 10, 4)
    voc = ORBVocabulary(10, 4, 0, 0, *rows, max_features=64, max_batch=2)
    orc = O.OracleVocabulary(10, 4, 0, 0, *rows)
    feats = make_features(3, rows[2], 200).reshape(2, 100, 32)
    d_desc, d_cnt = torch.from_numpy(feats).cuda(), torch.tensor([40, 100], dtype=torch.int32, device="cuda")
    voc.transform_batch_device(d_desc.data_ptr(), 100, d_cnt.data_ptr(), 2, 2)
    torch.cuda.synchronize()
    assert same(voc.result(0, 40), orc.transform(feats[0, :40], 2))
    with pytest.raises(OrbfeError, match="max_features"):
        voc.result(1, 64)
    voc.close()


# ------------------------------------------------------------------------------------------------------ SearchByBoW
def make_pair(seed, rows, n_kf, n_f, turn=25.0):
    """A keyframe and a frame looking at the same scene: most frame descriptors are keyframe descriptors with a few bits flipped."""
    from oracle.orb_oracle import KP_DTYPE
    rng = np.random.RandomState(seed)
    d_kf = make_features(seed, rows[2], n_kf)
    src = rng.randint(0, max(n_kf, 1), n_f)
    d_f = d_kf[src].copy() if n_kf else np.zeros((n_f, 32), np.uint8)
    for i in range(n_f):
        nb = int(rng.randint(0, 24)) if rng.uniform() < 0.8 else 120
        for b in rng.choice(256, size=nb, replace=False):
            d_f[i, b >> 3] ^= np.uint8(1 << (b & 7))
    k_kf, k_f = np.zeros(n_kf, KP_DTYPE), np.zeros(n_f, KP_DTYPE)
    k_kf["angle"] = rng.uniform(0, 360, n_kf)
    k_f["angle"] = (k_kf["angle"][src] - turn + np.where(rng.uniform(size=n_f) < 0.85, rng.uniform(-5, 5, n_f), rng.uniform(0, 360, n_f))) % 360 if n_kf else 0
    has_mp = (rng.uniform(size=n_kf) < 0.8).astype(np.uint8)
    return d_kf, k_kf, has_mp, d_f, k_f


def python_search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th_low, ratio, check_ori, kf_kf=False, has_mp_f=None, n_left_f=-1):
    """ORBmatcher::SearchByBoW(pKF, F, ...) (F.Nleft == -1) in plain Python, written from src/ORBmatcher.cc:269-470."""
    f32 = np.float32
    match = np.full(len(d_f), -1, np.int32)
    hist = [[] for _ in range(30)]
    nm = 0
    for node in sorted(set(fv_kf) & set(fv_f)):          # the lockstep walk visits exactly the common nodes, in key order
        for ikf in fv_kf[node]:
            if not has_mp[ikf]:
                continue
            b1, b2, bi = 256, 256, -1
            r1, r2, ri = 256, 256, -1                      # the right camera's rows of a fisheye rig compete separately
            for i_f in fv_f[node]:
                if match[i_f] >= 0 or (kf_kf and has_mp_f is not None and not has_mp_f[i_f]):
                    continue
                d = int(POPC[d_kf[ikf] ^ d_f[i_f]].sum())
                if n_left_f < 0 or kf_kf or i_f < n_left_f:
                    if d < b1:
                        b2, b1, bi = b1, d, i_f
                    elif d < b2:
                        b2 = d
                else:
                    if d < r1:
                        r2, r1, ri = r1, d, i_f
                    elif d < r2:
                        r2 = d
            if (b1 < th_low if kf_kf else b1 <= th_low) and f32(b1) < f32(ratio) * f32(b2):
                match[bi] = ikf
                nm += 1
                if check_ori:
                    rot = f32(k_kf["angle"][ikf]) - f32(k_f["angle"][bi])
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    b = int(np.floor(abs(float(f32(rot * f32(f32(1) / f32(30))))) + 0.5))
                    hist[0 if b == 30 else b].append(bi)
            if not kf_kf and n_left_f >= 0 and b1 <= th_low and r1 <= th_low:      # ORBmatcher.cc:388-416 (`|| true`)
                match[ri] = ikf
                nm += 1
                if check_ori:
                    rot = f32(k_kf["angle"][ikf]) - f32(k_f["angle"][ri])
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    b = int(np.floor(abs(float(f32(rot * f32(f32(1) / f32(30))))) + 0.5))
                    hist[0 if b == 30 else b].append(ri)
    if check_ori:
        ind, mx = [-1, -1, -1], [0, 0, 0]
        for i, h in enumerate(hist):
            sz = len(h)
            if sz > mx[0]:
                mx = [sz, mx[0], mx[1]]; ind = [i, ind[0], ind[1]]
            elif sz > mx[1]:
                mx = [mx[0], sz, mx[1]]; ind = [ind[0], i, ind[1]]
            elif sz > mx[2]:
                mx[2] = sz; ind[2] = i
        if f32(mx[1]) < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif f32(mx[2]) < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i not in ind:
                for bi in hist[i]:
                    match[bi] = -1
                    nm -= 1
    return match, nm


SB_CONFIGS = [(4, 3, 1, 21, 400, 450), (10, 3, 2, 22, 800, 700), (3, 4, 3, 23, 600, 600), (10, 4, 2, 24, 1200, 1300)]   # k, L, levelsup, seed, nKF, nF


def test_oracle_search_by_bow_matches_python():
    for k, L, levelsup, seed, n_kf, n_f in SB_CONFIGS:
        rows = make_vocabulary(seed, k, L)
        voc = O.OracleVocabulary(k, L, 0, 0, *rows)
        d_kf, k_kf, has_mp, d_f, k_f = make_pair(seed, rows, n_kf, n_f)
        fv_kf, fv_f = voc.transform(d_kf, levelsup)["feat_vec"], voc.transform(d_f, levelsup)["feat_vec"]
        for ori in (True, False):
            match, kf_out, nm = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori)
            want, wnm = python_search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori)
            assert np.array_equal(match, want) and nm == wnm, (k, L, levelsup, ori)
            assert nm > 50 and (not ori or (kf_out["accepted"] == 2).sum() > 0)
        # a fisheye rig: the second half of the frame's rows belongs to the right camera
        for ori in (True, False):
            match, kf_out, nm = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori, n_left_f=n_f // 2)
            want, wnm = python_search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori, n_left_f=n_f // 2)
            assert np.array_equal(match, want) and nm == wnm and (match[n_f // 2:] >= 0).sum() > 10, (k, L, "fisheye", ori)
        # the keyframe-keyframe overload: candidates need a map point as well, threshold strict (49 here, so that ties at the limit occur)
        mp_f = (np.random.RandomState(seed).uniform(size=n_f) < 0.7).astype(np.uint8)
        match, kf_out, nm = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 20, 0.7, True, kf_kf=True, has_map_point_f=mp_f)
        want, wnm = python_search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 20, 0.7, True, kf_kf=True, has_mp_f=mp_f)
        assert np.array_equal(match, want) and nm == wnm and nm > 20, (k, L, "kf-kf")
        assert mp_f[match >= 0].all()


@pytest.mark.gpu
def test_gpu_search_by_bow_matches_oracle():
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    for k, L, levelsup, seed, n_kf, n_f in SB_CONFIGS + [(2, 2, 1, 25, 900, 1000), (10, 4, 2, 26, 4096, 4096), (10, 3, 1, 27, 0, 50), (10, 3, 1, 28, 40, 0)]:
        rows = make_vocabulary(seed, k, L)
        voc = O.OracleVocabulary(k, L, 0, 0, *rows)
        d_kf, k_kf, has_mp, d_f, k_f = make_pair(seed, rows, n_kf, n_f)
        fv_kf, fv_f = voc.transform(d_kf, levelsup)["feat_vec"], voc.transform(d_f, levelsup)["feat_vec"]
        for ori in (True, False):
            want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori)
            got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori)
            assert np.array_equal(got[0], want[0]) and got[2] == want[2], (k, L, levelsup, n_kf, n_f, ori)
            assert np.array_equal(got[1], want[1]), (k, L, levelsup, n_kf, n_f, ori)
        if n_f > 1:
            for ori in (True, False):
                want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori, n_left_f=n_f // 2)
                got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 50, 0.7, ori, n_left_f=n_f // 2)
                assert np.array_equal(got[0], want[0]) and got[2] == want[2] and np.array_equal(got[1], want[1]), (k, L, n_kf, n_f, "fisheye", ori)
        mp_f = (np.random.RandomState(seed).uniform(size=n_f) < 0.7).astype(np.uint8)
        want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 20, 0.7, True, kf_kf=True, has_map_point_f=mp_f)
        got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, 20, 0.7, True, kf_kf=True, has_map_point_f=mp_f)
        assert np.array_equal(got[0], want[0]) and got[2] == want[2] and np.array_equal(got[1], want[1]), (k, L, levelsup, n_kf, n_f, "kf-kf")


@pytest.mark.gpu
def test_gpu_search_by_bow_under_heavy_competition():
    """Descriptors from a handful of clusters a few bits apart, two vocabulary nodes: every keyframe feature could take dozens of frame
    features and most of them compete for the same ones, so the out-of-order resolution of the kernel (a query is final as soon as no
    earlier unresolved one can take a feature it cares about; a feature counts as matched only for the queries behind its taker) has to
    reproduce the serial loop exactly — including the best / second-best records of the queries that end up without a match."""
    from oracle.orb_oracle import KP_DTYPE
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    seen = []
    for seed, n_kf, n_f, clusters, noise, ratio, th in [(1, 600, 700, 3, 8, 0.9, 50), (2, 1500, 1500, 6, 14, 0.7, 50), (3, 900, 400, 2, 5, 0.95, 30), (4, 2000, 2500, 12, 20, 0.8, 50)]:
        rng = np.random.RandomState(seed)
        rows = make_vocabulary(seed + 40, 2, 1)
        voc = O.OracleVocabulary(2, 1, 0, 0, *rows)
        centres = rng.randint(0, 256, (clusters, 32)).astype(np.uint8)

        def feats(n):
            d = centres[rng.randint(0, clusters, n)].copy()
            for i in range(n):
                for b in rng.randint(0, 256, int(rng.randint(0, noise + 1))):
                    d[i, b >> 3] ^= np.uint8(1 << (b & 7))
            k = np.zeros(n, KP_DTYPE)
            k["x"], k["y"], k["angle"] = rng.uniform(0, 700, n), rng.uniform(0, 400, n), rng.uniform(0, 360, n)
            return d, k
        (d_kf, k_kf), (d_f, k_f) = feats(n_kf), feats(n_f)
        has_mp = (rng.uniform(size=n_kf) < 0.8).astype(np.uint8)
        fv_kf, fv_f = voc.transform(d_kf, 0)["feat_vec"], voc.transform(d_f, 0)["feat_vec"]
        for ori in (False, True):
            want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, ori)
            got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, ori)
            assert np.array_equal(got[0], want[0]) and got[2] == want[2] and np.array_equal(got[1], want[1]), (seed, ori)
            want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, ori, n_left_f=n_f // 2)
            got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, ori, n_left_f=n_f // 2)
            assert np.array_equal(got[0], want[0]) and got[2] == want[2] and np.array_equal(got[1], want[1]), (seed, ori, "rig")
        mp_f = (rng.uniform(size=n_f) < 0.7).astype(np.uint8)
        want = O.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, True, kf_kf=True, has_map_point_f=mp_f)
        got = B.search_by_bow(d_kf, k_kf, fv_kf, has_mp, d_f, k_f, fv_f, th, ratio, True, kf_kf=True, has_map_point_f=mp_f)
        assert np.array_equal(got[0], want[0]) and got[2] == want[2] and np.array_equal(got[1], want[1]), (seed, "kf-kf")
        seen.append(int(want[2]))
    print("matches per configuration (kf-kf):", seen)
    assert sum(seen) > 20


@pytest.mark.gpu
def test_gpu_extract_bow_search_chain_on_device():
    """Tracking's relocalisation path end to end on the device: extract a batch, ComputeBoW, SearchByBoW between frame pairs."""
    import ctypes as C
    import torch
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, synth, _lib
    from orb_slam3_detailed_comments_kor_b200.bow import ORBVocabulary
    W, H, params, nb = 752, 480, (1200, 1.2, 8, 20, 7), 6
    imgs = np.stack([synth.stereo_pair(80 + i // 2, W, H)[i % 2] for i in range(nb)])     # (L, R) of three scenes: 0-1, 2-3, 4-5 overlap
    ex = ORBextractor(*params, width=W, height=H, max_batch=nb)
    cnt, mono, kps, desc = ex.extract_batch_host(imgs, (0, 0))
    rows = make_vocabulary(33, 10, 4)
    voc = ORBVocabulary(10, 4, 0, 0, *rows, max_features=ex.capacity, max_batch=nb)
    orc = O.OracleVocabulary(10, 4, 0, 0, *rows)
    views = [C.c_void_p() for _ in range(4)]
    _lib.check(_lib.lib().orbfe_batch_device_views(ex.handle, *[C.byref(v) for v in views]))
    d_kps, d_desc, d_cnt = views[0].value, views[1].value, views[2].value
    st = torch.cuda.Stream()
    voc.transform_batch_device(d_desc, ex.capacity, d_cnt, nb, 2, stream=st.cuda_stream)
    bv = [C.c_void_p() for _ in range(9)]
    _lib.check(_lib.lib().orbfe_bow_device_views(voc.handle, *[C.byref(v) for v in bv]))
    pairs_kf, pairs_f = np.array([0, 2, 4, 1, 0], np.int32), np.array([1, 3, 5, 0, 4], np.int32)
    with torch.cuda.stream(st):
        d_pk, d_pf = torch.from_numpy(pairs_kf).cuda(), torch.from_numpy(pairs_f).cuda()
        d_match = torch.zeros((len(pairs_kf), ex.capacity), dtype=torch.int32, device="cuda")
        d_nm = torch.zeros(len(pairs_kf), dtype=torch.int32, device="cuda")
        _lib.check(_lib.lib().orbfe_search_by_bow_device(0, len(pairs_kf), ex.capacity, d_pk.data_ptr(), d_pf.data_ptr(), d_desc, d_kps, ex.capacity,
                                                         bv[5].value, bv[6].value, bv[7].value, bv[8].value, None, None, 50, C.c_float(0.7), 1, 0,
                                                         d_match.data_ptr(), None, d_nm.data_ptr(), st.cuda_stream))
        match, nm = d_match.cpu().numpy(), d_nm.cpu().numpy()
    st.synchronize()
    for p, (a, b) in enumerate(zip(pairs_kf, pairs_f)):
        fa, fb = orc.transform(desc[a, :cnt[a]], 2)["feat_vec"], orc.transform(desc[b, :cnt[b]], 2)["feat_vec"]
        want, _, wnm = O.search_by_bow(desc[a, :cnt[a]], kps[a, :cnt[a]], fa, None, desc[b, :cnt[b]], kps[b, :cnt[b]], fb, 50, 0.7, True)
        assert np.array_equal(match[p, :cnt[b]], want) and nm[p] == wnm, "pair %d" % p
        assert (match[p, cnt[b]:] == -1).all()
    assert nm[:3].sum() > 0                              # (a random vocabulary separates real descriptors poorly: few, but some, matches)
    voc.close(); ex.close()


# ------------------------------------------------------------------------------------------- SearchForTriangulation
def make_triangulation_pair(seed, rows, n):
    """Two pinhole keyframes one baseline apart: keyframe 2 sees keyframe 1's features shifted along x (the epipolar lines are
    the image rows), with some vertical noise so that the epipolar test cuts both ways."""
    from oracle.orb_oracle import KP_DTYPE
    rng = np.random.RandomState(seed)
    f32 = np.float32
    d1 = make_features(seed, rows[2], n)
    k1 = np.zeros(n, KP_DTYPE)
    k1["x"], k1["y"] = rng.uniform(20, 730, n), rng.uniform(20, 460, n)
    k1["octave"] = np.minimum(rng.geometric(0.35, n) - 1, 7)
    k1["angle"] = rng.uniform(0, 360, n)
    perm = rng.permutation(n)
    d2, k2 = d1[perm].copy(), k1[perm].copy()
    for i in range(n):
        nb = int(rng.randint(0, 30)) if rng.uniform() < 0.8 else 110
        for b in rng.choice(256, size=nb, replace=False):
            d2[i, b >> 3] ^= np.uint8(1 << (b & 7))
    k2["x"] = (k2["x"] - rng.uniform(2, 60, n)).astype(f32)
    k2["y"] = (k2["y"] + np.where(rng.uniform(size=n) < 0.7, rng.uniform(-1.5, 1.5, n), rng.uniform(-12, 12, n))).astype(f32)
    k2["angle"] = (k2["angle"] - 10 + np.where(rng.uniform(size=n) < 0.85, rng.uniform(-4, 4, n), rng.uniform(0, 360, n))) % 360
    u1 = np.where(rng.uniform(size=n) < 0.5, k1["x"] - 5, -1).astype(f32)
    u2 = np.where(rng.uniform(size=n) < 0.5, k2["x"] - 5, -1).astype(f32)
    mp1, mp2 = (rng.uniform(size=n) < 0.3).astype(np.uint8), (rng.uniform(size=n) < 0.3).astype(np.uint8)
    K = np.array([[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], np.float64)
    tx = np.array([[0, 0, 0], [0, 0, -0.11], [0, 0.11, 0]], np.float64)          # [t12]x for t12 = (0.11, 0, 0), R12 = I
    F = (np.linalg.inv(K).T @ tx @ np.linalg.inv(K)).astype(f32)
    ep = np.array([300.0, 240.0], f32)                                            # any point: the rule only needs to be applied identically
    scale = np.array([f32(1.2) ** i for i in range(8)], f32)
    sigma2 = (scale * scale).astype(f32)
    return d1, k1, u1, mp1, d2, k2, u2, mp2, F, ep, scale, sigma2


def python_search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, th_low, only_stereo, coarse, check_ori):
    """ORBmatcher::SearchForTriangulation + Pinhole::epipolarConstrain in plain Python (written from the reference sources)."""
    f32 = np.float32
    F = np.asarray(F, f32).reshape(3, 3)
    m12 = np.full(len(d1), -1, np.int32)
    hist = [[] for _ in range(30)]
    nm = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if mp1[i1]:
                continue
            st1 = u1[i1] >= 0
            if only_stereo and not st1:
                continue
            x1, y1 = f32(k1["x"][i1]), f32(k1["y"][i1])
            best, bi = th_low, -1
            for i2 in fv2[node]:
                if mp2[i2]:
                    continue
                st2 = u2[i2] >= 0
                if only_stereo and not st2:
                    continue
                d = int(POPC[d1[i1] ^ d2[i2]].sum())
                if d > th_low or d > best:
                    continue
                x2, y2, o2 = f32(k2["x"][i2]), f32(k2["y"][i2]), int(k2["octave"][i2])
                if not st1 and not st2:
                    ex, ey = f32(ep[0] - x2), f32(ep[1] - y2)
                    if f32(f32(ex * ex) + f32(ey * ey)) < f32(f32(100) * scale[o2]):
                        continue
                ok = coarse
                if not ok:
                    a = f32(f32(f32(x1 * F[0, 0]) + f32(y1 * F[1, 0])) + F[2, 0])
                    b = f32(f32(f32(x1 * F[0, 1]) + f32(y1 * F[1, 1])) + F[2, 1])
                    c = f32(f32(f32(x1 * F[0, 2]) + f32(y1 * F[1, 2])) + F[2, 2])
                    num = f32(f32(f32(a * x2) + f32(b * y2)) + c)
                    den = f32(f32(a * a) + f32(b * b))
                    if den != 0:
                        ok = float(f32(f32(num * num) / den)) < 3.84 * float(sigma2[o2])
                if ok:
                    best, bi = d, i2
            if bi >= 0:
                m12[i1] = bi
                nm += 1
                if check_ori:
                    rot = f32(k1["angle"][i1]) - f32(k2["angle"][bi])
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    b = int(np.floor(abs(float(f32(rot * f32(f32(1) / f32(30))))) + 0.5))
                    hist[0 if b == 30 else b].append(i1)
    if check_ori:
        ind, mx = [-1, -1, -1], [0, 0, 0]
        for i, h in enumerate(hist):
            sz = len(h)
            if sz > mx[0]:
                mx = [sz, mx[0], mx[1]]; ind = [i, ind[0], ind[1]]
            elif sz > mx[1]:
                mx = [mx[0], sz, mx[1]]; ind = [ind[0], i, ind[1]]
            elif sz > mx[2]:
                mx[2] = sz; ind[2] = i
        if f32(mx[1]) < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif f32(mx[2]) < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i not in ind:
                for i1 in hist[i]:
                    m12[i1] = -1
                    nm -= 1
    return m12, nm


ST_CONFIGS = [(4, 3, 1, 41, 700), (10, 3, 2, 42, 1200), (3, 3, 2, 43, 500)]   # k, L, levelsup, seed, n


def _triangulation_inputs(k, L, levelsup, seed, n):
    rows = make_vocabulary(seed, k, L)
    voc = O.OracleVocabulary(k, L, 0, 0, *rows)
    t = make_triangulation_pair(seed, rows, n)
    fv1, fv2 = voc.transform(t[0], levelsup)["feat_vec"], voc.transform(t[4], levelsup)["feat_vec"]
    return t, fv1, fv2


def test_oracle_search_for_triangulation_matches_python():
    for cfg in ST_CONFIGS:
        (d1, k1, u1, mp1, d2, k2, u2, mp2, F, ep, scale, sigma2), fv1, fv2 = _triangulation_inputs(*cfg)
        seen = 0
        for only_stereo, coarse, ori in ((False, False, True), (True, False, True), (False, True, False), (False, False, False)):
            got, nm = O.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, only_stereo, coarse, ori)
            want, wnm = python_search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, only_stereo, coarse, ori)
            assert np.array_equal(got, want) and nm == wnm, (cfg, only_stereo, coarse, ori)
            seen += nm
        assert seen > 100
        # the epipolar test must cut: coarse accepts more than the strict run
        strict = O.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, False, False, False)[1]
        loose = O.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, False, True, False)[1]
        assert loose > strict > 0


@pytest.mark.gpu
def test_gpu_search_for_triangulation_matches_oracle():
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    for cfg in ST_CONFIGS + [(2, 2, 1, 44, 3000), (10, 4, 2, 45, 4096)]:
        (d1, k1, u1, mp1, d2, k2, u2, mp2, F, ep, scale, sigma2), fv1, fv2 = _triangulation_inputs(*cfg)
        for only_stereo, coarse, ori in ((False, False, True), (True, False, True), (False, True, False), (False, False, False)):
            want = O.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, only_stereo, coarse, ori)
            got = B.search_for_triangulation(d1, k1, u1, mp1, fv1, d2, k2, u2, mp2, fv2, F, ep, scale, sigma2, 50, only_stereo, coarse, ori)
            assert np.array_equal(got[0], want[0]) and got[1] == want[1], (cfg, only_stereo, coarse, ori)
    # monocular keyframes (no mvuRight at all), empty sides
    (d1, k1, u1, mp1, d2, k2, u2, mp2, F, ep, scale, sigma2), fv1, fv2 = _triangulation_inputs(4, 3, 1, 46, 300)
    want = O.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2)
    got = B.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2)
    assert np.array_equal(got[0], want[0]) and got[1] == want[1]
    got = B.search_for_triangulation(d1[:0], k1[:0], None, None, {}, d2, k2, None, mp2, fv2, F, ep, scale, sigma2)
    assert got[0].shape == (0,) and got[1] == 0


# ----------------------------------------------------------------------- SearchForTriangulation on two-camera (fisheye) rigs
def make_rig_pair(seed, rows, n):
    """Two keyframes of a fisheye rig (left + right camera rows each), the second one moved sideways: scene points are projected
    through KannalaBrandt8 into the four cameras, so the TriangulateMatches constraint accepts true correspondences and rejects
    scrambled ones.  Returns the two frames' arrays and the orbfe_triangulation_rig bytes."""
    from oracle.orb_oracle import KP_DTYPE
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    rng = np.random.RandomState(seed)
    f32 = np.float32
    camL = np.array([190.978, 190.973, 254.93, 256.90, 0.00348, 0.000715, -0.00205, 0.000203], f32)
    camR = np.array([190.442, 190.434, 252.60, 254.92, 0.00340, 0.001766, -0.00266, 0.00033], f32)
    Rlr, tlr = np.eye(3), np.array([0.101, 0.002, 0.001])                 # right camera in the left camera's frame: x_l = Rlr x_r + tlr
    T21 = (np.eye(3), np.array([-0.25, 0.0, 0.02]))                       # keyframe 2 (left camera) from keyframe 1: x_2 = R x_1 + t
    pts = np.stack([rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(1.0, 5.0, n)], 1)

    def to_cam(X, kf2, right):                                             # keyframe-1-left coordinates -> the camera's coordinates
        if kf2:
            X = X @ T21[0].T + T21[1]
        return (X - tlr) @ Rlr if right else X

    def proj(cam, X):
        return np.array([O.kb8_project(cam, p) for p in X.astype(f32)], np.float64)

    side1, side2 = rng.randint(0, 2, n), rng.randint(0, 2, n)             # which camera of each keyframe sees the point
    d1 = make_features(seed, rows[2], n)
    d2 = d1.copy()
    for i in range(n):
        nb = int(rng.randint(0, 30)) if rng.uniform() < 0.8 else 110
        for b in rng.choice(256, size=nb, replace=False):
            d2[i, b >> 3] ^= np.uint8(1 << (b & 7))
    uv1 = np.where(side1[:, None] == 0, proj(camL, to_cam(pts, False, False)), proj(camR, to_cam(pts, False, True)))
    uv2 = np.where(side2[:, None] == 0, proj(camL, to_cam(pts, True, False)), proj(camR, to_cam(pts, True, True)))
    uv2 += np.where(rng.uniform(size=(n, 1)) < 0.7, rng.uniform(-0.4, 0.4, (n, 2)), rng.uniform(-15, 15, (n, 2)))   # some gross outliers

    def frame(uv, side, d, perm):
        order = np.concatenate([perm[side[perm] == 0], perm[side[perm] == 1]])           # left rows first, then right rows
        k = np.zeros(n, KP_DTYPE)
        k["x"], k["y"] = uv[order, 0], uv[order, 1]
        k["octave"] = np.minimum(rng.geometric(0.35, n) - 1, 7)
        k["angle"] = rng.uniform(0, 360, n)
        return k, d[order], int((side == 0).sum())

    k1, d1o, nl1 = frame(uv1, side1, d1, np.arange(n))
    k2, d2o, nl2 = frame(uv2, side2, d2, rng.permutation(n))
    k2["angle"] = rng.uniform(0, 360, n)
    mp1, mp2 = (rng.uniform(size=n) < 0.2).astype(np.uint8), (rng.uniform(size=n) < 0.2).astype(np.uint8)
    # relative poses x_1side = R x_2side + t for the four side combinations (src/ORBmatcher.cc:995-1003)
    R12, t12 = T21[0].T, -T21[0].T @ T21[1]                                # keyframe-1-left from keyframe-2-left
    def side_pose(r1, r2):
        # x_1left = R12 x_2left + t12; x_left = Rlr x_right + tlr
        R, t = R12.copy(), t12.copy()
        if r2:
            R, t = R @ Rlr, R @ tlr + t                                    # from keyframe-2-right
        if r1:
            R, t = Rlr.T @ R, Rlr.T @ (t - tlr)                            # into keyframe-1-right
        return R, t
    poses = [side_pose(0, 0), side_pose(0, 1), side_pose(1, 0), side_pose(1, 1)]
    rig = B.make_triangulation_rig([camL, camR, camL, camR], [p[0] for p in poses], [p[1] for p in poses], nl1, nl2)
    return d1o, k1, mp1, d2o, k2, mp2, rig, (camL, camR, poses, nl1, nl2)


def python_search_for_triangulation_rig(d1, k1, mp1, fv1, d2, k2, mp2, fv2, geo, sigma2, th_low, coarse, check_ori):
    """The two-camera branch of ORBmatcher::SearchForTriangulation in plain Python; the constraint calls the (cv2-pinned) oracle of
    KannalaBrandt8::TriangulateMatches pair by pair."""
    camL, camR, poses, nl1, nl2 = geo
    f32 = np.float32
    m12 = np.full(len(d1), -1, np.int32)
    hist = [[] for _ in range(30)]
    nm = 0
    for node in sorted(set(fv1) & set(fv2)):
        for i1 in fv1[node]:
            if mp1[i1]:
                continue
            best, bi = th_low, -1
            for i2 in fv2[node]:
                if mp2[i2]:
                    continue
                d = int(POPC[d1[i1] ^ d2[i2]].sum())
                if d > th_low or d > best:
                    continue
                ok = coarse
                if not ok:
                    r1, r2 = int(i1 >= nl1), int(i2 >= nl2)
                    R, t = poses[2 * r1 + r2]
                    dep, _ = O.kb8_triangulate_matches(camR if r1 else camL, camR if r2 else camL, R, t, [[k1["x"][i1], k1["y"][i1]]], [[k2["x"][i2], k2["y"][i2]]],
                                                       [sigma2[k1["octave"][i1]]], [sigma2[k2["octave"][i2]]])
                    ok = dep[0] > f32(0.0001)
                if ok:
                    best, bi = d, i2
            if bi >= 0:
                m12[i1] = bi
                nm += 1
                if check_ori:
                    rot = f32(k1["angle"][i1]) - f32(k2["angle"][bi])
                    if rot < 0:
                        rot = f32(rot + f32(360))
                    b = int(np.floor(abs(float(f32(rot * f32(f32(1) / f32(30))))) + 0.5))
                    hist[0 if b == 30 else b].append(i1)
    if check_ori:
        ind, mx = [-1, -1, -1], [0, 0, 0]
        for i, h in enumerate(hist):
            sz = len(h)
            if sz > mx[0]:
                mx = [sz, mx[0], mx[1]]; ind = [i, ind[0], ind[1]]
            elif sz > mx[1]:
                mx = [mx[0], sz, mx[1]]; ind = [ind[0], i, ind[1]]
            elif sz > mx[2]:
                mx[2] = sz; ind[2] = i
        if f32(mx[1]) < f32(0.1) * f32(mx[0]):
            ind[1] = ind[2] = -1
        elif f32(mx[2]) < f32(0.1) * f32(mx[0]):
            ind[2] = -1
        for i in range(30):
            if i not in ind:
                for i1 in hist[i]:
                    m12[i1] = -1
                    nm -= 1
    return m12, nm


def _rig_inputs(k, L, levelsup, seed, n):
    rows = make_vocabulary(seed, k, L)
    voc = O.OracleVocabulary(k, L, 0, 0, *rows)
    d1, k1, mp1, d2, k2, mp2, rig, geo = make_rig_pair(seed, rows, n)
    scale = np.array([np.float32(1.2) ** i for i in range(8)], np.float32)
    return d1, k1, mp1, d2, k2, mp2, rig, geo, voc.transform(d1, levelsup)["feat_vec"], voc.transform(d2, levelsup)["feat_vec"], scale, (scale * scale).astype(np.float32)


def test_oracle_search_for_triangulation_rig_matches_python():
    F, ep = np.zeros(9, np.float32), np.zeros(2, np.float32)               # unused with a rig
    for cfg in [(4, 3, 1, 51, 500), (10, 3, 2, 52, 700)]:
        d1, k1, mp1, d2, k2, mp2, rig, geo, fv1, fv2, scale, sigma2 = _rig_inputs(*cfg)
        counts = []
        for coarse, ori in ((False, True), (False, False), (True, False)):
            got, nm = O.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, False, coarse, ori, rig=rig)
            want, wnm = python_search_for_triangulation_rig(d1, k1, mp1, fv1, d2, k2, mp2, fv2, geo, sigma2, 50, coarse, ori)
            assert np.array_equal(got, want) and nm == wnm, (cfg, coarse, ori)
            counts.append(nm)
        assert counts[2] > counts[1] > 20                                  # the triangulation test cuts, and lets true pairs through
        # bOnlyStereo with a rig: nothing is stereo, nothing matches (:1041-1043)
        assert O.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, True, False, True, rig=rig)[1] == 0


@pytest.mark.gpu
def test_gpu_search_for_triangulation_rig_matches_oracle():
    """The rig branch on the GPU (KannalaBrandt8::TriangulateMatches inside the candidate loop) against the oracle.  The device's tanf /
    atan2f differ from the host's by an ulp or two, so a pair whose reprojection error sits exactly on the 5.991 sigma^2 threshold could
    flip: the test allows at most 2 per thousand rows to differ and asserts that, in practice, none does on these inputs."""
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    F, ep = np.zeros(9, np.float32), np.zeros(2, np.float32)
    for cfg in [(4, 3, 1, 51, 500), (10, 3, 2, 52, 700), (3, 4, 2, 53, 3000)]:
        d1, k1, mp1, d2, k2, mp2, rig, geo, fv1, fv2, scale, sigma2 = _rig_inputs(*cfg)
        for coarse, ori in ((False, True), (False, False), (True, False)):
            want = O.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, False, coarse, ori, rig=rig)
            got = B.search_for_triangulation(d1, k1, None, mp1, fv1, d2, k2, None, mp2, fv2, F, ep, scale, sigma2, 50, False, coarse, ori, rig=rig)
            ndiff = int((got[0] != want[0]).sum())
            assert ndiff <= max(1, len(d1) // 500), (cfg, coarse, ori, ndiff)
            assert ndiff == 0 and got[1] == want[1], (cfg, coarse, ori, ndiff)
            assert want[1] > 20


# ----------------------------------------------------------------------- many pairs over host memory in one call
@pytest.mark.gpu
def test_gpu_frames_entry_points_equal_the_oracle_pair_by_pair():
    """orbfe_search_for_triangulation_frames / orbfe_search_by_bow_frames: the loops of CreateNewMapPoints, Relocalization and LoopClosing as one
    call.  Frames of different sizes (one empty), a pair of a frame with itself, every pair checked against the oracle on that pair alone."""
    from orb_slam3_detailed_comments_kor_b200 import bow as B
    from orb_slam3_detailed_comments_kor_b200._lib import OrbfeError
    k, L, levelsup, seed = 10, 3, 1, 71
    rows = make_vocabulary(seed, k, L)
    voc = O.OracleVocabulary(k, L, 0, 0, *rows)
    # keyframes 0..4 share a scene (triangulation pairs against keyframe 0), 5 is empty
    base = make_triangulation_pair(seed, rows, 900)
    d1, k1, u1, mp1, d2, k2, u2, mp2, F, ep, scale, sigma2 = base
    frames = [dict(desc=d1, kps=k1, u_right=u1, has_map_point=mp1, feat_vec=voc.transform(d1, levelsup)["feat_vec"])]
    Fs, eps = [], []
    rng = np.random.RandomState(seed)
    for i, n in enumerate((900, 650, 1200, 300)):
        keep = rng.permutation(len(d2))[:min(n, len(d2))]
        dd, kk, uu, mm = d2[keep].copy(), k2[keep].copy(), u2[keep].copy(), mp2[keep].copy()
        frames.append(dict(desc=dd, kps=kk, u_right=uu, has_map_point=mm, feat_vec=voc.transform(dd, levelsup)["feat_vec"]))
        Fs.append((F * np.float32(1.0 + 0.01 * i)).astype(np.float32)); eps.append(ep)
    frames.append(dict(desc=d1[:0], kps=k1[:0], u_right=u1[:0], has_map_point=mp1[:0], feat_vec={}))
    p1, p2 = [0, 0, 0, 0, 0, 2], [1, 2, 3, 4, 5, 2]
    Fs += [F, F]; eps += [ep, ep]
    for only_stereo, coarse, ori in ((False, False, True), (False, True, False)):
        got, nm = B.search_for_triangulation_frames(frames, p1, p2, np.stack(Fs), np.stack(eps), scale, sigma2, 50, only_stereo, coarse, ori)
        for p, (a, b) in enumerate(zip(p1, p2)):
            fa, fb = frames[a], frames[b]
            want = O.search_for_triangulation(fa["desc"], fa["kps"], fa["u_right"], fa["has_map_point"], fa["feat_vec"], fb["desc"], fb["kps"], fb["u_right"],
                                              fb["has_map_point"], fb["feat_vec"], Fs[p], eps[p], scale, sigma2, 50, only_stereo, coarse, ori)
            assert np.array_equal(got[p], want[0]) and nm[p] == want[1], (p, a, b, only_stereo, coarse, ori)
        assert nm[:4].min() > 20
    # SearchByBoW: frame 0 (the current frame) against candidate keyframes, both overloads, with and without kf_out
    for kf_kf in (False, True):
        pk, pf = [1, 2, 3, 4, 5, 0], [0, 0, 0, 0, 0, 0]
        got, ko, nm = B.search_by_bow_frames(frames, pk, pf, 50 if not kf_kf else 20, 0.7, True, kf_kf, want_kf_out=True)
        got2, _, nm2 = B.search_by_bow_frames(frames, pk, pf, 50 if not kf_kf else 20, 0.7, True, kf_kf)
        for p, (a, b) in enumerate(zip(pk, pf)):
            fa, fb = frames[a], frames[b]
            want = O.search_by_bow(fa["desc"], fa["kps"], fa["feat_vec"], fa["has_map_point"], fb["desc"], fb["kps"], fb["feat_vec"], 50 if not kf_kf else 20, 0.7,
                                   True, kf_kf=kf_kf, has_map_point_f=fb["has_map_point"])
            assert np.array_equal(got[p], want[0]) and nm[p] == want[2] and np.array_equal(ko[p], want[1]), (p, a, b, kf_kf)
            assert np.array_equal(got2[p], want[0]) and nm2[p] == want[2]
    assert nm.max() > 20
    with pytest.raises(OrbfeError):
        B.search_by_bow_frames(frames, [0, 9], [1, 1])
