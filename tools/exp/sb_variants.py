"""Experiment: k_search_by_bow on a batch of pairs with deep competition (device entry), on 1 and 8 pairs, and on one pair through the host entry."""
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
NP = 512
pk = np.array([2 * (i % (NF // 2)) for i in range(NP)], np.int32); pf = pk + 1
d = dict(desc=dev(desc), kps=dev(kps), nodes=dev(nodes), starts=dev(starts), feats=dev(feats), cnt=dev(cnt), pk=dev(pk), pf=dev(pf))
out = torch.zeros((NP, S), dtype=torch.int32, device="cuda"); nm = torch.zeros(NP, dtype=torch.int32, device="cuda")
def step(a, b):
    _lib.check(L.orbfe_search_by_bow_device(0, NP, S, d[a].data_ptr(), d[b].data_ptr(), d["desc"].data_ptr(), d["kps"].data_ptr(), S, d["nodes"].data_ptr(),
                                            d["starts"].data_ptr(), d["feats"].data_ptr(), d["cnt"].data_ptr(), None, None, 50, C.c_float(0.7), 1, 0, out.data_ptr(), None,
                                            nm.data_ptr(), None))
for name, a, b in (("moved pairs", "pk", "pf"), ("same frame", "pk", "pk")):
    for _ in range(3): step(a, b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): step(a, b)
    e1.record(); torch.cuda.synchronize()
    print("%-12s %.3f ms per %d pairs, %.0f matches per pair" % (name, e0.elapsed_time(e1) / 10, NP, nm.float().mean().item()), flush=True)
NP1 = NP
for NPX in (1, 8):
    NP = NPX
    for name, a, b in (("moved pairs", "pk", "pf"), ("same frame", "pk", "pk")):
        for _ in range(3): step(a, b)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20): step(a, b)
        e1.record(); torch.cuda.synchronize()
        print("%-12s %.3f ms per %d pairs on the device" % (name, e0.elapsed_time(e1) / 20, NP), flush=True)
NP = NP1
(d0, k0), (d1, k1) = frames[0], frames[1]
for name, fa, fb in (("moved pair", 0, 1), ("same frame", 0, 0)):
    fn = lambda: B.search_by_bow(frames[fa][0], frames[fa][1], fvs[fa], None, frames[fb][0], frames[fb][1], fvs[fb], 50, 0.7, True)
    fn(); ts = []
    for _ in range(20):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    print("one host call, %-10s %.3f ms" % (name, np.median(ts) * 1e3), flush=True)
