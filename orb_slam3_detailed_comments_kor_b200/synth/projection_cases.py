"""Seeded inputs for the projection-gated search, shared by tools/make_goldens.py, the tests and bench.py.
A case is a dict of arrays in the layouts of include/orbfe.h (orbfe_keypoint, orbfe_proj_query, ...)."""
import numpy as np

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
PROJ_QUERY_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("radius", "<f4"), ("xr", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"), ("flags", "<i4"),
                             ("angle", "<f4")])
PQ_RIGHT, PQ_BLOCKS, PQ_SKIP, PQ_AFTER_LEFT, PQ_BEST_ONLY, PQ_NEEDS_LEFT_WINDOW, PQ_ERROR_GATE = 1, 2, 4, 8, 16, 32, 64


def _flip(rng, d, k):
    d = d.copy()
    for b in rng.choice(256, size=k, replace=False):
        d[b >> 3] ^= np.uint8(1 << (b & 7))
    return d


def make_case(seed, n=600, nq=700, W=752, H=480, th=1.0, fisheye=False, stereo_gate=False, crowd=0.0, last_frame=None, fuse=False):
    """crowd: fraction of the map points aimed at a small set of keypoints (long chains of map points competing for the same
    keypoints); th: the search-window factor of SearchByProjection.  last_frame: None = the SearchLocalPoints overload;
    "still" / "forward" / "backward" = SearchByProjection(CurrentFrame, LastFrame, th, bMono) with the level range of that
    motion (src/ORBmatcher.cc:2253-2258), best-only acceptance and the rotation check (case["check_orientation"] = 1).
    fuse: the search of ORBmatcher::Fuse (:1643-1841): best only, TH_LOW, chi-square gate on the reprojection error, nothing blocks."""
    rng = np.random.RandomState(seed)
    f32 = np.float32
    kps = np.zeros(n, KP_DTYPE)
    kps["x"] = rng.uniform(-4, W + 4, n).astype(f32)          # a few fall outside the image bounds (PosInGrid drops them)
    kps["y"] = rng.uniform(-4, H + 4, n).astype(f32)
    # clusters: a third of the keypoints sit within a few pixels of another one
    src = rng.randint(0, n, n // 3)
    dst = rng.choice(n, n // 3, replace=False)
    kps["x"][dst] = kps["x"][src] + rng.uniform(-3, 3, n // 3).astype(f32)
    kps["y"][dst] = kps["y"][src] + rng.uniform(-3, 3, n // 3).astype(f32)
    kps["octave"] = np.minimum(rng.geometric(0.3, n) - 1, 7)
    kps["size"] = 31
    kps["class_id"] = -1
    kps["angle"] = rng.uniform(0, 360, n).astype(f32)
    desc = rng.randint(0, 256, (n, 32)).astype(np.uint8)
    desc[dst[: n // 8]] = desc[src[: n // 8]]                   # exact descriptor duplicates next to each other: distance ties
    n_left = -1
    if fisheye:
        n_left = n // 2 + 7
    case = {"kps": kps, "desc": desc, "n_left": np.array([n_left], np.int32),
            "grid": np.array([0, 0, f32(64) / f32(W), f32(48) / f32(H)], f32), "th_high": np.array([100], np.int32),
            "nnratio": np.array([0.8], f32), "taken": (rng.uniform(size=n) < 0.08).astype(np.uint8),
            "check_orientation": np.array([1 if last_frame else 0], np.int32)}
    if fuse:
        case["th_high"] = np.array([50], np.int32)
        case["taken"] = np.zeros(n, np.uint8)
        case["inv_level_sigma2"] = np.array([f32(1) / (f32(1.2) ** i * f32(1.2) ** i) for i in range(16)], f32)
    turn = f32(rng.uniform(-40, 40))                          # camera roll between the two frames (last_frame cases)
    if stereo_gate:
        ur = kps["x"] - rng.uniform(1, 40, n).astype(f32)
        ur[rng.uniform(size=n) < 0.4] = -1
        case["u_right"] = ur.astype(f32)
    if fisheye:
        partner = np.full(n, -1, np.int32)
        nl, nr = n_left, n - n_left
        for _ in range(nl // 2):                               # like ComputeStereoFishEyeMatches: later matches overwrite one side
            a, b = rng.randint(0, nl), rng.randint(0, nr)
            partner[a] = nl + b
            partner[nl + b] = a
        case["partner"] = partner
    scale = np.array([f32(1.2) ** i for i in range(8)], f32)
    hot = rng.choice(n, min(n, max(4, n // 40)), replace=False)
    q = []
    qd = []
    count = 0
    while count < nq:
        aimed = rng.uniform() < 0.75
        if aimed:
            t = int(hot[rng.randint(len(hot))]) if rng.uniform() < crowd else int(rng.randint(n))
            lvl = int(min(7, kps["octave"][t] + rng.randint(0, 2)))
            r = f32(2.5 if rng.uniform() < 0.5 else 4.0) * f32(th) * scale[lvl]
            jit = 0.25 if fuse else 0.7
            x = f32(kps["x"][t] + rng.uniform(-jit, jit) * r)
            y = f32(kps["y"][t] + rng.uniform(-jit, jit) * r)
            d = _flip(rng, desc[t], int(rng.randint(0, 75)))
        else:
            t = -1
            lvl = int(rng.randint(0, 8))
            r = f32(2.5 if rng.uniform() < 0.5 else 4.0) * f32(th) * scale[lvl]
            x, y = f32(rng.uniform(-30, W + 30)), f32(rng.uniform(-30, H + 30))
            d = rng.randint(0, 256, 32).astype(np.uint8)
        flags = PQ_BLOCKS if rng.uniform() < 0.9 else 0
        if rng.uniform() < 0.05:
            flags |= PQ_SKIP
        xr = f32(x - rng.uniform(1, 40))
        if stereo_gate and t >= 0 and case["u_right"][t] > 0:
            xr = f32(case["u_right"][t] + rng.uniform(-1.3, 1.3) * r)
        minl, maxl = lvl - 1, lvl
        if rng.uniform() < 0.03:
            minl, maxl = 0, -1                                  # bCheckLevels == false
        ang = f32(0)
        if fuse:
            flags = (flags & ~PQ_BLOCKS) | PQ_BEST_ONLY | PQ_ERROR_GATE
        if last_frame:
            flags |= PQ_BEST_ONLY
            minl, maxl = {"still": (lvl - 1, lvl + 1), "forward": (lvl, -1), "backward": (0, lvl)}[last_frame]
            # most matches turn by the camera roll (one or two histogram bins); the rest are scattered and get filtered
            ang = f32((kps["angle"][t] if t >= 0 else 0) + (turn + rng.uniform(-4, 4) if rng.uniform() < 0.8 else rng.uniform(0, 360)))
            ang = f32(ang % 360)
        if fisheye:
            right_target = t >= n_left
            # left-camera half, then (for most points) the right-camera half of the same map point
            q.append((x, y, r, xr, minl, maxl, flags | (PQ_RIGHT if right_target and rng.uniform() < 0.5 else 0), ang))
            qd.append(d)
            count += 1
            if not (q[-1][6] & PQ_RIGHT) and rng.uniform() < 0.7:
                dep = PQ_NEEDS_LEFT_WINDOW if last_frame else PQ_AFTER_LEFT
                q.append((f32(x + rng.uniform(-2, 2)), f32(y + rng.uniform(-2, 2)), r, xr, minl, maxl, flags | PQ_RIGHT | dep, ang))
                qd.append(d)
                count += 1
        else:
            q.append((x, y, r, xr, minl, maxl, flags, ang))
            qd.append(d)
            count += 1
    case["queries"] = np.array(q, PROJ_QUERY_DTYPE)
    case["qdesc"] = np.stack(qd).astype(np.uint8)
    return case


CASES = {
    # name: make_case arguments
    "mono": dict(seed=11, n=600, nq=700),
    "stereo_gate": dict(seed=12, n=800, nq=900, stereo_gate=True),
    "wide": dict(seed=13, n=500, nq=600, th=3.0, crowd=0.5),
    "fisheye": dict(seed=14, n=900, nq=800, fisheye=True, th=1.5, crowd=0.3),
    "last_still": dict(seed=15, n=800, nq=700, th=7.0, stereo_gate=True, crowd=0.2, last_frame="still"),
    "last_forward": dict(seed=16, n=700, nq=600, th=15.0, crowd=0.4, last_frame="forward"),
    "last_backward_fisheye": dict(seed=17, n=900, nq=800, th=7.0, fisheye=True, crowd=0.3, last_frame="backward"),
    "fuse": dict(seed=18, n=900, nq=1000, th=3.0, stereo_gate=True, crowd=0.3, fuse=True),
}
