"""How many speculative rounds does k_search_by_projection need on the bench's frames and on crowded ones?"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from orb_slam3_detailed_comments_kor_b200 import _lib, projection as PJ
from orb_slam3_detailed_comments_kor_b200.synth import projection_cases as PC
L = _lib.lib()
MAXKP = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
for name, kw in [("bench", dict(n=1200, nq=1500, th=1.0, stereo_gate=True)), ("th3", dict(n=1200, nq=1500, th=3.0)), ("crowd", dict(n=1200, nq=1500, th=3.0, crowd=0.5)),
                 ("last_frame", dict(n=1200, nq=1200, th=7.0, last_frame="still"))]:
    cases = [PC.make_case(seed=900 + i, W=752, H=480, **kw) for i in range(8)]
    PF = 512
    order = [cases[i % 8] for i in range(PF)]
    cat = lambda k: np.concatenate([c[k] for c in order])
    koff = np.cumsum([0] + [len(c["kps"]) for c in order]).astype(np.int32); qoff = np.cumsum([0] + [len(c["queries"]) for c in order]).astype(np.int32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
    d = {k: dev(cat(k)) for k in ("kps", "desc", "queries", "qdesc", "taken")}
    d["grid"] = dev(np.stack([c["grid"] for c in order])); d["koff"], d["qoff"] = dev(koff), dev(qoff)
    ur = dev(cat("u_right")) if "u_right" in cases[0] else None
    out = torch.zeros(int(qoff[-1]) * 28, dtype=torch.uint8, device="cuda")
    def step():
        PJ.search_by_projection_device(PF, MAXKP, d["kps"].data_ptr(), d["desc"].data_ptr(), d["koff"].data_ptr(), d["grid"].data_ptr(), d["queries"].data_ptr(),
                                       d["qdesc"].data_ptr(), d["qoff"].data_ptr(), out.data_ptr(), d_u_right=ur.data_ptr() if ur is not None else None,
                                       d_taken_in=d["taken"].data_ptr(), check_orientation=bool(cases[0]["check_orientation"][0]))
    for _ in range(3): step()
    torch.cuda.synchronize()
    a, b = C.c_int(0), C.c_int(0)
    L.orbfe_debug_projection_rounds(C.byref(a), C.byref(b))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): step()
    e1.record(); torch.cuda.synchronize()
    L.orbfe_debug_projection_rounds(C.byref(a), C.byref(b))
    def step1():
        PJ.search_by_projection_device(1, MAXKP, d["kps"].data_ptr(), d["desc"].data_ptr(), d["koff"].data_ptr(), d["grid"].data_ptr(), d["queries"].data_ptr(),
                                       d["qdesc"].data_ptr(), d["qoff"].data_ptr(), out.data_ptr(), d_u_right=ur.data_ptr() if ur is not None else None,
                                       d_taken_in=d["taken"].data_ptr(), check_orientation=bool(cases[0]["check_orientation"][0]))
    for _ in range(3): step1()
    torch.cuda.synchronize()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(20): step1()
    f1.record(); torch.cuda.synchronize()
    one_kernel = f0.elapsed_time(f1) / 20
    L.orbfe_debug_projection_rounds(C.byref(a), C.byref(b))
    print("   one frame on the device: %.3f ms (20 launches back to back), rounds %d" % (one_kernel, a.value))
    c0 = cases[0]
    def one():
        PJ.search_by_projection(c0["kps"], c0["desc"], c0["queries"], c0["qdesc"], c0["grid"], u_right=c0.get("u_right"), taken=c0["taken"],
                                check_orientation=bool(c0["check_orientation"][0]))
    import time
    one(); ts = []
    for _ in range(20):
        t0 = time.perf_counter(); one(); ts.append(time.perf_counter() - t0)
    print("%-10s %.3f ms per %d frames; rounds max %d mean %.1f; one host call %.3f ms" % (name, e0.elapsed_time(e1) / 10, PF, a.value, b.value / (10 * PF), np.median(ts) * 1e3), flush=True)
