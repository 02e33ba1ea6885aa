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
