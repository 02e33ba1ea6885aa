"""Experiment: one SearchByBoW pair on the device, five launches (the target of a single-launch ncu capture)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch, ctypes as C
from test_bow import make_vocabulary, make_pair
from oracle import orb_oracle as O
from oracle.orb_oracle import KP_DTYPE
from orb_slam3_detailed_comments_kor_b200 import bow as B, _lib
L = _lib.lib()
rows = make_vocabulary(7, 10, 3)
voc = O.OracleVocabulary(10, 3, 0, 0, *rows)
S, NF = 1536, 64
frames = []
for i in range(NF // 2):
    d_kf, k_kf, has_mp, d_f, k_f = make_pair(100 + i, rows, 1500, 1500)
    frames += [(d_kf, k_kf), (d_f, k_f)]
fvs = [voc.transform(d, 1)["feat_vec"] for d, _ in frames]
desc = np.zeros((NF, S, 32), np.uint8); kps = np.zeros((NF, S), KP_DTYPE)
nodes = np.zeros((NF, S), np.int32); starts = np.zeros((NF, S + 1), np.int32); feats = np.zeros((NF, S), np.int32); cnt = np.zeros(NF, np.int32)
for i, ((d, k), fv) in enumerate(zip(frames, fvs)):
    desc[i, :len(d)] = d; kps[i, :len(k)] = k
    n, s, f = B._fv_arrays(fv)
    nodes[i, :len(n)] = n; starts[i, :len(s)] = s; feats[i, :len(f)] = f; cnt[i] = len(n)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
NP = 1
pk = np.array([2 * (i % (NF // 2)) for i in range(NP)], np.int32); pf = pk + 1
d = dict(desc=dev(desc), kps=dev(kps), nodes=dev(nodes), starts=dev(starts), feats=dev(feats), cnt=dev(cnt), pk=dev(pk), pf=dev(pf))
out = torch.zeros((NP, S), dtype=torch.int32, device="cuda"); nm = torch.zeros(NP, dtype=torch.int32, device="cuda")
def step(a, b):
    _lib.check(L.orbfe_search_by_bow_device(0, NP, S, d[a].data_ptr(), d[b].data_ptr(), d["desc"].data_ptr(), d["kps"].data_ptr(), S, d["nodes"].data_ptr(),
                                            d["starts"].data_ptr(), d["feats"].data_ptr(), d["cnt"].data_ptr(), None, None, 50, C.c_float(0.7), 1, 0, out.data_ptr(), None,
                                            nm.data_ptr(), None))
for _ in range(5): step("pk", "pf")
torch.cuda.synchronize()
print("ok")
