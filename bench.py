#!/usr/bin/env python3
"""bench.py — ORB front-end throughput on B200 (BASELINE.json metric: ORB frames/s at 752x480 / 1200 features,
plus Hamming matches/s), one JSON line on stdout.

Workload (config.workload): BASELINE.json configs[1] — EuRoC-shaped stereo pairs 2x752x480, ORBextractor(1200, 1.2, 8,
20, 7), left + right extraction followed by Frame::ComputeStereoMatches — batched: one "step" = one pass of the hot
path over `pairs` stereo pairs per GPU (frames are independent, so N GPUs each take their own pairs: weak scaling,
no collective).  `value` = frames/s with the images already in HBM; `e2e` = the same through the host-buffer entry
points (H2D of the images and D2H of keypoints/descriptors/uRight/depth inside the timed region).
The second north-star metric, Hamming matches/s (config 5: 4096 queries x 2^20 database rows, top-2 + TH_LOW/ratio,
database sharded over the ranks with an NCCL all-gather + merge), is reported under "hamming".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--pairs P] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 752, 480
PARAMS = (1200, 1.2, 8, 20, 7)
MBF, FX = 47.90639, 435.2047             # Examples/Stereo-Inertial/EuRoC.yaml:9,26
WORKLOAD = "euroc_stereo_2x752x480_1200feat_extract+ComputeStereoMatches"
METRIC = "orb_frames_per_sec_752x480_1200feat"
LAP = (0, 0)
# BASELINE.json configs: [1] EuRoC (the headline, default), [2] TUM-VI fisheye stereo, [3] KITTI x 4096 images sharded by frame
WORKLOADS = {
    "euroc": dict(W=752, H=480, params=(1200, 1.2, 8, 20, 7), mbf=47.90639, fx=435.2047, lap=(0, 0), match="stereo", scaling="weak",
                  name="euroc_stereo_2x752x480_1200feat_extract+ComputeStereoMatches", metric="orb_frames_per_sec_752x480_1200feat"),
    # Examples/Stereo/KITTI00-02.yaml:9,21-22,28,41; 4096 images = 2048 pairs in total, split over the ranks (strong scaling)
    "kitti": dict(W=1241, H=376, params=(2000, 1.2, 8, 20, 7), mbf=386.1448, fx=718.856, lap=(0, 0), match="stereo", scaling="strong", total_pairs=2048,
                  name="kitti_stereo_1241x376_2000feat_4096_images_extract+ComputeStereoMatches", metric="orb_frames_per_sec_1241x376_2000feat"),
    # Examples/Stereo-Inertial/TUM_512.yaml:42-46 (lapping 0..511 on both cameras), knn-2 + 0.7 of ComputeStereoFishEyeMatches
    "tumvi": dict(W=512, H=512, params=(1500, 1.2, 8, 20, 7), mbf=0.0, fx=0.0, lap=(0, 511), match="fisheye", scaling="weak",
                  name="tumvi_fisheye_stereo_2x512x512_1500feat_lapping_extract+ComputeStereoFishEyeMatches", metric="orb_frames_per_sec_512x512_1500feat"),
}


def set_workload(name):
    global W, H, PARAMS, MBF, FX, WORKLOAD, METRIC, LAP
    w = WORKLOADS[name]
    W, H, PARAMS, MBF, FX, WORKLOAD, METRIC, LAP = w["W"], w["H"], w["params"], w["mbf"], w["fx"], w["name"], w["metric"], w["lap"]
    return w
STAGES = ["pyramid", "fast", "blur", "quadtree", "assemble", "orient_describe", "stereo", "stereo_median"]


def level_sizes():
    s, out = 1.0, []
    scale = np.float32(1.0)
    for l in range(PARAMS[2]):
        inv = np.float32(1.0) / scale
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
        scale = np.float32(float(scale) * float(np.float32(PARAMS[1])))
    return out


def algorithmic_bytes(nkp):
    """SURVEY.md §8d staged model, per image, split per stage (DESIGN.md 'algorithmic bytes')."""
    lv = level_sizes()
    px = sum(w * h for w, h in lv)
    padded = sum((w + 38) * (h + 38) for w, h in lv)
    src = sum(w * h for w, h in lv[:-1])
    per = {"pyramid": W * H + padded + src, "fast": px, "blur": 2 * px,
           "orient_describe": nkp * (961 + 512 + 32 + 28), "quadtree": 0, "assemble": 0}
    return per, sum(per.values())


def gen_pairs(n, seed0, pin, distinct=None):
    """n synthetic stereo pairs in (pinned) host tensors; with `distinct` < n only that many are generated and then repeated."""
    import torch
    from orb_slam3_detailed_comments_kor_b200 import synth
    tl = torch.empty((n, H, W), dtype=torch.uint8, pin_memory=pin)
    tr = torch.empty((n, H, W), dtype=torch.uint8, pin_memory=pin)
    L, R = tl.numpy(), tr.numpy()
    nd = n if distinct is None else min(n, distinct)
    with ThreadPoolExecutor(os.cpu_count() or 4) as ex:
        list(ex.map(lambda i: synth.stereo_pair(seed0 + i, W, H, L[i], R[i]), range(nd)))
    for i in range(nd, n):
        L[i] = L[i % nd]; R[i] = R[i % nd]
    return tl, tr


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled while the timed region runs (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []
        self.nvml, self.samples, self.stop_flag = None, [], False
        try:   # NVML answers in microseconds; the nvidia-smi loop needs ~100 ms to start, longer than a short timed region
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _poll(self):
        n = self.nvml
        reasons_fn = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while True:
            try:
                self.samples.append((n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM), n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM),
                                     int(reasons_fn(self.h))))
            except Exception:
                pass
            if self.stop_flag:
                break
            time.sleep(0.002)

    def start(self):
        if self.nvml:
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self):
        if self.nvml:
            self.stop_flag = True
            self.t.join(timeout=2)
            bits = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}   # nvmlClocksEventReason*
            sm = [x[0] for x in self.samples]
            reasons = sorted(nm for nm, b in bits.items() if any(x[2] & b for x in self.samples))
            return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(x[1] for x in self.samples)) if sm else None,
                    "reasons": reasons, "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


def reference_available():
    """oracle/_ref/liborb_ref.so: the reference's own ORBextractor.cc + Frame::ComputeStereoMatches compiled by oracle/Makefile
    (prebuilt where /root/reference exists; it travels to the GPU box with the repository snapshot)."""
    try:
        from oracle import ref_oracle as Rf
        return Rf.available()
    except Exception:
        return False


def cpu_oracle_rate(npairs, threads, impl="auto"):
    """The reference's CPU path on `npairs` stereo pairs with `threads` host threads, timed the way the reference instruments
    itself (around extraction and ComputeStereoMatches, src/Frame.cc:116-127,136-144).  impl "reference" = the reference's own code
    (oracle/_ref), "port" = the CPU oracle restatement (oracle/liborb_oracle.so); "auto" takes the reference when it is there."""
    from oracle import orb_oracle as O
    from orb_slam3_detailed_comments_kor_b200 import synth
    pairs = [synth.stereo_pair(900000 + i, W, H) for i in range(npairs)]
    local = threading.local()
    if impl == "auto":
        impl = "reference" if reference_available() else "port"
    if impl == "reference":
        from oracle import ref_oracle as Rf
        ref = Rf.ReferenceExtractor(*PARAMS)
        mb = float(np.float32(MBF) / np.float32(FX)) if FX > 0 else 0.0

        def work(i):   # noqa: F811
            if FX > 0:   # both extractions + Frame::ComputeStereoMatches in the reference's code
                return len(ref.stereo(pairs[i][0], pairs[i][1], MBF, mb)[0])
            mL, kL, dL = ref(pairs[i][0], LAP)
            mR, kR, dR = ref(pairs[i][1], LAP)
            if len(kL) > mL and len(kR) > mR:
                O.knn2_ratio(dL[mL:], dR[mR:], 0.7)          # BFMatcher.knnMatch is OpenCV code: the oracle's restatement stands in
            return len(kL)
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, range(min(threads, npairs))))
            t0 = time.perf_counter()
            list(ex.map(work, range(npairs)))
            dt = time.perf_counter() - t0
        return 2 * npairs / dt, dt

    def work(i):
        if not hasattr(local, "ex"):
            local.ex = (O.OracleExtractor(*PARAMS), O.OracleExtractor(*PARAMS))
        exL, exR = local.ex
        mL, kL, dL = exL(pairs[i][0], LAP)
        mR, kR, dR = exR(pairs[i][1], LAP)
        if FX > 0:
            O.stereo_match(exL, exR, kL, dL, kR, dR, MBF, 0.0, FX)
        elif len(kL) > mL and len(kR) > mR:
            O.knn2_ratio(dL[mL:], dR[mR:], 0.7)              # ComputeStereoFishEyeMatches' matching step, src/Frame.cc:1137-1144
        return len(kL)

    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(min(threads, npairs))))           # warm the per-thread extractors
        t0 = time.perf_counter()
        list(ex.map(work, range(npairs)))
        dt = time.perf_counter() - t0
    return 2 * npairs / dt, dt


def cpu_two_thread_rate(npairs):
    """The reference's own threading for one stereo frame (src/Frame.cc:119-122): left and right extraction on two threads, join, then
    ComputeStereoMatches; one pair at a time, CPU oracle restatement."""
    from oracle import orb_oracle as O
    from orb_slam3_detailed_comments_kor_b200 import synth
    pairs = [synth.stereo_pair(910000 + i, W, H) for i in range(npairs)]
    exL, exR = O.OracleExtractor(*PARAMS), O.OracleExtractor(*PARAMS)
    with ThreadPoolExecutor(2) as tp:
        def one(i):
            fl, fr = tp.submit(exL, pairs[i][0], LAP), tp.submit(exR, pairs[i][1], LAP)
            (mL, kL, dL), (mR, kR, dR) = fl.result(), fr.result()
            if FX > 0:
                O.stereo_match(exL, exR, kL, dL, kR, dR, MBF, 0.0, FX)
        one(0)
        t0 = time.perf_counter()
        for i in range(npairs):
            one(i)
        dt = time.perf_counter() - t0
    return 2 * npairs / dt, dt


def cv2_lines(nframes, threads):
    """Two more CPU figures with the real OpenCV (the cv2 wheel: SIMD FAST / GaussianBlur / resize, as a real ORB-SLAM3 binary has):
    `cv2_primitives_frames_per_s` = only the pixel primitives of one extraction (pyramid resize + border, FAST per 35 px cell with the
    20/7 retry, 7x7 blur of every level; frame-parallel over `threads`, one OpenCV thread each), called through cv2's Python binding — its
    700 detect() calls per frame each build a Python list of KeyPoint objects, so binding overhead is part of the figure;
    `cv2_frames_per_s` = tools/make_goldens.py::CvExtractor, the same primitives with the glue in Python (one thread: the glue holds
    the interpreter lock).  Neither is a C++ OpenCV build of the reference (its headers are not in this image); they are the closest
    thing to "real OpenCV arithmetic, timed" that the box can run."""
    try:
        import cv2
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import make_goldens as G
    except Exception as exc:
        return {"cv2": "unavailable: %r" % (exc,)}
    from orb_slam3_detailed_comments_kor_b200 import synth
    cv2.setNumThreads(1)
    imgs = [synth.frame(910000 + i, W, H) for i in range(nframes)]
    local = threading.local()

    def primitives(i):
        if not hasattr(local, "ex"):
            local.ex = G.CvExtractor(*PARAMS)
        ex = local.ex
        ex.compute_pyramid(imgs[i])
        ncorner = 0
        for l in range(ex.nlevels):
            im = ex.pyr[l]
            minB, maxBX, maxBY = 16, im.shape[1] - 16, im.shape[0] - 16
            width, height = np.float32(maxBX - minB), np.float32(maxBY - minB)
            nCols, nRows = int(width / np.float32(35)), int(height / np.float32(35))
            wCell, hCell = int(np.ceil(float(width / np.float32(nCols)))), int(np.ceil(float(height / np.float32(nRows))))
            for ci in range(nRows):
                iniY = minB + ci * hCell
                if iniY >= maxBY - 3:
                    continue
                maxY = min(iniY + hCell + 6, maxBY)
                for cj in range(nCols):
                    iniX = minB + cj * wCell
                    if iniX >= maxBX - 6:
                        continue
                    roi = im[iniY:maxY, iniX:min(iniX + wCell + 6, maxBX)]
                    kps = ex.fast[ex.ini_th].detect(roi)
                    if not kps:
                        kps = ex.fast[ex.min_th].detect(roi)
                    ncorner += len(kps)
            cv2.GaussianBlur(im.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        return ncorner

    with ThreadPoolExecutor(threads) as tp:
        list(tp.map(primitives, range(min(threads, nframes))))
        t0 = time.perf_counter()
        list(tp.map(primitives, range(nframes)))
        dt = time.perf_counter() - t0
    full = G.CvExtractor(*PARAMS)
    nfull = min(4, nframes)
    t0 = time.perf_counter()
    for i in range(nfull):
        full(imgs[i], LAP)
    dtf = time.perf_counter() - t0
    return {"cv2_version": cv2.__version__, "cv2_primitives_frames_per_s": nframes / dt,
            "cv2_primitives_sample": "%d frames (%.1f s wall), frame-parallel over %d threads, cv2.setNumThreads(1)" % (nframes, dt, threads),
            "cv2_frames_per_s": nfull / dtf, "cv2_sample": "%d frames (%.1f s wall), one thread, Python glue" % (nfull, dtf)}


def peaks_for_issue():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("sm_max_mhz", 1965.0))
    except (OSError, ValueError):
        return 1965.0


def cpu_hamming_rate(nq_sample, ndb, threads):
    from oracle import orb_oracle as O
    from orb_slam3_detailed_comments_kor_b200 import synth
    db, q = synth.descriptors(5, ndb, max(nq_sample, threads))
    chunks = np.array_split(np.arange(nq_sample), threads)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(lambda c: O.hamming_top2(q[c], db, 50, 0.9, False) if len(c) else None, chunks))
    dt = time.perf_counter() - t0
    return nq_sample * ndb / dt, dt


def run_reference(args, rank):
    """--impl reference: the reference's CPU path for this workload.  The reference itself cannot be compiled here
    (needs OpenCV/Eigen/Pangolin), so this is the oracle port on all host cores."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample = max(4, 4 * cores)
    kind = "reference" if reference_available() else "port"
    for _ in range(args.warmup):
        cpu_oracle_rate(min(sample, cores), cores, kind)
    times = []
    for _ in range(args.steps):
        _, dt = cpu_oracle_rate(sample, cores, kind)
        times.append(dt)
    port_rate, _ = cpu_oracle_rate(sample, cores, "port")
    extra = cv2_lines(2 * cores, cores)
    ms = 1e3 * float(np.mean(times))
    value = 2 * sample / (ms / 1e3)
    hq = 1024
    hv, hdt = cpu_hamming_rate(hq, 1 << 20, cores)
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step": sample, "threads": cores},
            "cpu_baseline": dict({"value": value, "unit": "frames/s", "cores": cores, "kind": kind,
                                  "sample": "%d stereo pairs per step, frame-parallel over %d threads; %s" % (
                                      sample, cores, "the reference's ORBextractor.cc + Frame::ComputeStereoMatches compiled by oracle/Makefile (OpenCV primitives: "
                                      "scalar restatements pinned to cv2 4.13)" if kind == "reference" else "CPU oracle restatement"),
                                  "port_frames_per_s": port_rate}, **extra),
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "hamming": {"value": hv, "unit": "descriptor_pairs/s", "sample": "%d queries x 2^20 rows (%.1f s wall)" % (hq, hdt),
                        "cores": cores}}
    print(json.dumps(line), flush=True)


def run_workload(args, wl, rank, world, local_rank, barrier, max_over_ranks):
    """BASELINE.json configs 3 (TUM-VI) and 4 (KITTI x 4096 images): the same measurement as the headline config — device-resident
    throughput with CUDA events, per-stage times, roofline of the longest stage, end-to-end from pinned host memory, CPU sample —
    on the workload's own shape, features and matching step.  A step = every pair of this rank once, in sub-batches of at most 256
    pairs (the extractor objects' batch size); kitti splits 2048 pairs over the ranks (strong scaling), tumvi gives every rank its
    own 256 pairs (weak)."""
    import ctypes as C
    import torch
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, _lib
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    L = _lib.lib()
    P = wl["total_pairs"] // world if wl["scaling"] == "strong" else args.pairs
    SUB = min(P, 256)
    assert P % SUB == 0
    fisheye = wl["match"] == "fisheye"
    hL, hR = gen_pairs(P, rank * 1000003, pin=True, distinct=64)
    dL, dR = hL.cuda(non_blocking=True), hR.cuda(non_blocking=True)
    exL = ORBextractor(*PARAMS, width=W, height=H, device=local_rank, max_batch=SUB)
    exR = ORBextractor(*PARAMS, width=W, height=H, device=local_rank, max_batch=SUB)
    cap = exL.capacity
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    st = tstream.cuda_stream
    d_match = torch.empty((SUB, cap), dtype=torch.int32, device="cuda") if fisheye else None
    fb = W * H
    if fisheye:   # Examples/Stereo-Inertial/TUM_512.yaml:9-35: the two KannalaBrandt8 cameras and Tlr
        cam1 = [190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504,
                0.0034823894022493434, 0.0007150348452162257, -0.0020532361418706202, 0.00020293673591811182]
        cam2 = [190.44236969414825, 190.4344384721956, 252.59949716835982, 254.91723064636983,
                0.0034003170790442797, 0.001766278153469831, -0.00266312569781606, 0.0003299517423931039]
        Rlr = [0.999999445773493, 0.000791687752817, 0.000694034010224, -0.000823363992158, 0.998899461915674, 0.046895490788700,
               -0.000656143613644, -0.046896036240590, 0.998899560146304]
        tlr = [0.101063427414194, 0.001946204678584, 0.001015350132563]
        d_rl = torch.empty((SUB, cap), dtype=torch.int32, device="cuda")
        d_depth, d_p3d = torch.empty((SUB, cap), device="cuda"), torch.empty((SUB, cap, 3), device="cuda")

    def match(stream):
        if fisheye:   # Frame::ComputeStereoFishEyeMatches: knn-2 + 0.7 ratio, then TriangulateMatches and the depth test
            M.fisheye_match_batch_device(exL, exR, SUB, 0.7, d_match.data_ptr(), None, stream)
            M.fisheye_triangulate_batch_device(exL, exR, SUB, cam1, cam2, Rlr, tlr, d_match.data_ptr(), d_rl.data_ptr(), d_depth.data_ptr(), d_p3d.data_ptr(), stream)
        else:
            M.compute_stereo_matches(exL, exR, SUB, MBF, 0.0, FX, stream)

    def step_serial():   # one stream: what the per-stage event pairs see
        for s0 in range(0, P, SUB):
            exL.extract_batch_device(dL.data_ptr() + s0 * fb, SUB, LAP, st)
            exR.extract_batch_device(dR.data_ptr() + s0 * fb, SUB, LAP, st)
            match(st)

    def step():          # the product calls: left / right side by side on the two objects' streams
        for s0 in range(0, P, SUB):
            if fisheye:
                M.extract_pair_batch_device(exL, exR, dL.data_ptr() + s0 * fb, dR.data_ptr() + s0 * fb, SUB, LAP, st)
                match(st)
            else:
                M.stereo_frames_batch_device(exL, exR, dL.data_ptr() + s0 * fb, dR.data_ptr() + s0 * fb, SUB, MBF, 0.0, FX, LAP, st)

    for _ in range(args.warmup):
        step_serial()
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    value = 2 * P * world / (ms_step / 1e3)
    for ex in (exL, exR):
        _lib.check(L.orbfe_set_profiling(ex.handle, 1))
    for _ in range(args.steps):
        step_serial()
    torch.cuda.synchronize()
    stage_ms = np.zeros(8)
    launches = C.c_int64(0)
    for ex in (exL, exR):
        buf = (C.c_double * 8)()
        _lib.check(L.orbfe_stage_times(ex.handle, buf, C.byref(launches)))
        stage_ms += np.array(list(buf))
        _lib.check(L.orbfe_set_profiling(ex.handle, 0))
    clocks = sampler.stop()
    cnt, _ = exL.batch_counts(SUB)
    cntR, _ = exR.batch_counts(SUB)
    nkp = float(np.mean(np.concatenate([cnt, cntR])))
    per_stage_bytes, total_bytes = algorithmic_bytes(nkp)
    stage_ms_per_step = stage_ms / args.steps
    top = int(np.argmax(stage_ms_per_step[:6]))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    top_bytes = per_stage_bytes[STAGES[top]] * 2 * P
    achieved = top_bytes / (stage_ms_per_step[top] / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": STAGES[top], "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650 GB/s",
                "algorithmic_bytes_per_launch": top_bytes, "launch_ms": float(stage_ms_per_step[top]),
                "stage_ms_per_step": {s: float(v) for s, v in zip(STAGES, stage_ms_per_step)},
                "whole_step": {"algorithmic_bytes_per_frame": total_bytes, "achieved": total_bytes * 2 * P / (ms_step / 1e3) / 1e9,
                               "frac": total_bytes * 2 * P / (ms_step / 1e3) / 1e9 / peak}}

    # end to end from pinned host memory: every sub-batch uploads its images and downloads its results
    def out_set():
        side = lambda: (np.zeros(SUB, np.int32), np.zeros(SUB, np.int32),  # noqa: E731
                        torch.empty((SUB, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(_lib.KP_DTYPE).reshape(SUB, cap),
                        torch.empty((SUB, cap, 32), dtype=torch.uint8, pin_memory=True).numpy())
        return side() + side() + (torch.empty((SUB, cap), dtype=torch.float32, pin_memory=True).numpy(), torch.empty((SUB, cap), dtype=torch.float32, pin_memory=True).numpy())
    hLn, hRn = hL.numpy(), hR.numpy()
    subs = list(range(0, P, SUB))
    if fisheye:
        pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True).numpy()  # noqa: E731
        fish_set = lambda: out_set()[:8] + (pin((SUB, cap), torch.int32), pin((SUB, cap), torch.int32), pin((SUB, cap), torch.float32), pin((SUB, cap, 3), torch.float32))  # noqa: E731
        outs = [fish_set(), fish_set()]
        rig = M.make_fisheye_rig(cam1, cam2, np.array(Rlr, np.float32).reshape(3, 3), tlr)
        call = "orbfe_fisheye_frames_submit / orbfe_stereo_frames_wait, two sub-batches in flight (keypoints, descriptors, both match vectors, depths and 3-D points out)"

        def e2e_run(nsteps):
            seq = [s0 for _ in range(nsteps) for s0 in subs]
            M.fisheye_frames_submit(exL, exR, hLn[seq[0]:seq[0] + SUB], hRn[seq[0]:seq[0] + SUB], rig, 0.7, LAP, outs[0])
            for k in range(1, len(seq)):
                M.fisheye_frames_submit(exL, exR, hLn[seq[k]:seq[k] + SUB], hRn[seq[k]:seq[k] + SUB], rig, 0.7, LAP, outs[k & 1])
                M.fisheye_frames_wait(exL, exR, outs[(k - 1) & 1])
            M.fisheye_frames_wait(exL, exR, outs[(len(seq) - 1) & 1])
        d2h = 2 * P * (cap * 60 + 8) + P * cap * (4 + 4 + 4 + 12)
    else:
        outs = [out_set(), out_set()]
        call = "orbfe_stereo_frames_submit / _wait, two sub-batches in flight"

        def e2e_run(nsteps):
            seq = [s0 for _ in range(nsteps) for s0 in subs]
            M.stereo_frames_submit(exL, exR, hLn[seq[0]:seq[0] + SUB], hRn[seq[0]:seq[0] + SUB], MBF, 0.0, FX, LAP, outs[0])
            for k in range(1, len(seq)):
                M.stereo_frames_submit(exL, exR, hLn[seq[k]:seq[k] + SUB], hRn[seq[k]:seq[k] + SUB], MBF, 0.0, FX, LAP, outs[k & 1])
                M.stereo_frames_wait(exL, exR, outs[(k - 1) & 1])
            M.stereo_frames_wait(exL, exR, outs[(len(seq) - 1) & 1])
        d2h = 2 * P * (cap * 60 + 8) + 2 * P * cap * 4
    e2e_run(2)
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.steps)
    barrier()
    e2e = {"value": 2 * P * world / (e2e_ms / 1e3), "unit": "frames/s", "ms_per_step": e2e_ms, "call": call,
           "fraction_of_device_resident": ms_step / e2e_ms, "h2d_gbs_reached_in_step": 2 * P * W * H / (e2e_ms / 1e3) / 1e9,
           "h2d_bytes_per_step": 2 * P * W * H, "d2h_bytes_per_step": d2h}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        sample = max(8 * cores, 32)
        kind = "reference" if reference_available() else "port"
        rate, dt = cpu_oracle_rate(sample, cores, kind)
        cpu = dict({"value": rate, "unit": "frames/s", "cores": cores, "kind": kind,
                    "sample": "%d stereo pairs (%.1f s wall) frame-parallel over %d threads; built -O3 -march=x86-64-v3" % (sample, dt, cores)},
                   **cv2_lines(2 * cores, cores))
    if rank == 0:
        nsub = len(subs)
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
                "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": WORKLOAD, "pairs_per_gpu_per_step": P, "frames_per_step": 2 * P * world, "sub_batch_pairs": SUB,
                           "distinct_pairs": min(P, 64), "lapping": list(LAP), "matching": wl["match"],
                           "l2": "inputs_larger_than_L2 (%d MB of images per step per GPU)" % (2 * P * W * H >> 20),
                           "mean_keypoints_per_frame": nkp, "parallelism": "frames sharded by rank, no collective",
                           "schedule": "left and right extraction on the two objects' streams (the reference's two ExtractORB threads), blur beside "
                                       "quadtree + assemble, matching behind both; timed by events on the caller's stream"},
                "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": nsub * (2 * (PARAMS[2] + 8) + (2 if fisheye else 3)) * args.steps}
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=512, help="stereo pairs per GPU per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--hamming-queries", type=int, default=4096)
    ap.add_argument("--hamming-db", type=int, default=1 << 20)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="euroc", choices=sorted(WORKLOADS), help="euroc = BASELINE.json's headline config (default); kitti, tumvi = configs 4 and 3")
    args = ap.parse_args()
    wl = set_workload(args.workload)
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from orb_slam3_detailed_comments_kor_b200 import ORBextractor, DescriptorDatabase, synth
    from orb_slam3_detailed_comments_kor_b200 import matcher as M
    from orb_slam3_detailed_comments_kor_b200 import _lib
    import ctypes as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.workload != "euroc":
        run_workload(args, wl, rank, world, local_rank, barrier, max_over_ranks)
        if world > 1:
            dist.destroy_process_group()
        return

    P = args.pairs
    hL, hR = gen_pairs(P, rank * 1000003, pin=True)
    dL, dR = hL.cuda(non_blocking=True), hR.cuda(non_blocking=True)
    exL = ORBextractor(*PARAMS, width=W, height=H, device=local_rank, max_batch=P)
    exR = ORBextractor(*PARAMS, width=W, height=H, device=local_rank, max_batch=P)
    # a real (non-default) stream: every kernel of the step is enqueued on it and the CUDA events below are recorded
    # on it, so the events see exactly the timed work (a NULL stream would make the library use its own streams)
    tstream = torch.cuda.Stream()
    torch.cuda.set_stream(tstream)
    st = tstream.cuda_stream
    assert st != 0
    L = _lib.lib()

    def step_serial():   # every kernel of both images on the one stream: what the per-stage event pairs and the ncu launch list see
        exL.extract_batch_device(dL.data_ptr(), P, (0, 0), st)
        exR.extract_batch_device(dR.data_ptr(), P, (0, 0), st)
        M.compute_stereo_matches(exL, exR, P, MBF, 0.0, FX, st)

    def step():          # the product call for a batch of device-resident pairs: left / right side by side like the reference's two threads
        M.stereo_frames_batch_device(exL, exR, dL.data_ptr(), dR.data_ptr(), P, MBF, 0.0, FX, (0, 0), st)

    for _ in range(args.warmup):
        step_serial()
    torch.cuda.synchronize()
    es0, es1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    es0.record()
    for _ in range(args.steps):
        step_serial()
    es1.record()
    torch.cuda.synchronize()
    ms_serial = es0.elapsed_time(es1) / args.steps
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_step = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    value = 2 * P * world / (ms_step / 1e3)

    # per-stage device time: the same steps again with the library's CUDA-event brackets switched on
    for ex in (exL, exR):
        _lib.check(L.orbfe_set_profiling(ex.handle, 1))
    for _ in range(args.steps):
        step_serial()
    torch.cuda.synchronize()
    stage_ms = np.zeros(8)
    launches = C.c_int64(0)
    for ex in (exL, exR):
        buf = (C.c_double * 8)()
        _lib.check(L.orbfe_stage_times(ex.handle, buf, C.byref(launches)))
        stage_ms += np.array(list(buf))
        _lib.check(L.orbfe_set_profiling(ex.handle, 0))
    clocks = sampler.stop()
    cnt, _ = exL.batch_counts(P)
    cntR, _ = exR.batch_counts(P)
    nkp = float(np.mean(np.concatenate([cnt, cntR])))
    per_stage_bytes, total_bytes = algorithmic_bytes(nkp)
    stage_ms_per_step = stage_ms / args.steps            # both extractors together: 2P images per step
    top = int(np.argmax(stage_ms_per_step[:6]))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    top_bytes = per_stage_bytes[STAGES[top]] * 2 * P
    achieved = top_bytes / (stage_ms_per_step[top] / 1e3) / 1e9 if stage_ms_per_step[top] > 0 else 0.0
    traffic, issue = None, None
    try:   # DRAM bytes and executed warp instructions of the stage from the committed `ncu --set full` capture, per launch
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        traffic = tj["dram_bytes_per_frame"].get(STAGES[top])
        traffic = None if traffic is None else traffic * 2 * P
        # the issue roofline: every SM sub-partition issues at most one warp instruction per clock, so a stage that executes
        # I warp instructions cannot finish before I / (148 SM x 4 x SM clock); all extraction kernels sit against this bound, not HBM
        clk = float(clocks.get("sm_mhz") or peaks_for_issue()) * 1e6
        issue_peak = 148 * 4 * clk
        wi = tj["warp_instructions_per_frame"]
        ext = sum(wi[k] for k in ("pyramid", "fast", "blur", "quadtree", "assemble", "orient_describe"))
        issue = {"bound": "instruction issue (148 SM x 4 sub-partitions x SM clock)", "peak_warp_instructions_per_s": issue_peak,
                 "stage_warp_instructions_per_frame": wi.get(STAGES[top]), "stage_frac": wi[STAGES[top]] * 2 * P / (stage_ms_per_step[top] / 1e3) / issue_peak,
                 "whole_step_warp_instructions_per_frame": ext + wi["stereo"] / 2,
                 "whole_step_frac": (ext * 2 * P + wi["stereo"] * P) / (ms_step / 1e3) / issue_peak,
                 "frames_per_s_at_full_issue": issue_peak / (ext + wi["stereo"] / 2),
                 "source": "smsp__inst_executed.sum per kernel, profiles/r2_traffic.json (ncu --set full of tools/prof_step.py, 64 pairs)"}
    except (OSError, KeyError, ValueError, TypeError):
        pass
    stage_kernels = {"pyramid": "k_pyramid_copy0 + 7 x k_pyramid_level + k_pyramid_fringe",
                     "fast": "k_fast_stream + k_fast_nms + k_fast_cells_list", "blur": "k_blur7_tma (all levels in one launch)", "quadtree": "k_quadtree",
                     "assemble": "k_assemble", "orient_describe": "k_orient_describe"}
    roofline = {"bound": "hbm", "kernel": STAGES[top], "kernels": stage_kernels.get(STAGES[top]), "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "issue": issue,
                "limiter": "instruction issue (ncu: issue slots 72 % busy, ALU pipe 66 %, DRAM 6 % of peak; profiles/r2_kernels_full.txt), not HBM",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if peaks else "fallback 6650 GB/s",
                "algorithmic_bytes_per_launch": top_bytes,
                "launch_ms": float(stage_ms_per_step[top]),
                "stage_ms_per_step": {s: float(v) for s, v in zip(STAGES, stage_ms_per_step)},
                "whole_step": {"algorithmic_bytes_per_frame": total_bytes,
                               "achieved": total_bytes * 2 * P / (ms_step / 1e3) / 1e9,
                               "frac": total_bytes * 2 * P / (ms_step / 1e3) / 1e9 / peak}}

    # end to end through the host-buffer entry points (pinned images in, results out)
    cap = exL.capacity
    outL = (np.zeros(P, np.int32), np.zeros(P, np.int32), torch.empty((P, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(_lib.KP_DTYPE).reshape(P, cap),
            torch.empty((P, cap, 32), dtype=torch.uint8, pin_memory=True).numpy())
    outR = (np.zeros(P, np.int32), np.zeros(P, np.int32), torch.empty((P, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(_lib.KP_DTYPE).reshape(P, cap),
            torch.empty((P, cap, 32), dtype=torch.uint8, pin_memory=True).numpy())
    hu = torch.empty((P, cap), dtype=torch.float32, pin_memory=True).numpy()
    hd = torch.empty((P, cap), dtype=torch.float32, pin_memory=True).numpy()
    hLn, hRn = hL.numpy(), hR.numpy()

    pool2 = ThreadPoolExecutor(2)

    def e2e_step():
        # left and right extraction from two host threads on two extractor objects, as Frame's constructor does
        # (src/Frame.cc:119-122); each call pipelines upload / kernels / download on its own streams
        fl = pool2.submit(exL.extract_batch_host, hLn, (0, 0), outL)
        fr = pool2.submit(exR.extract_batch_host, hRn, (0, 0), outR)
        fl.result(); fr.result()
        M.compute_stereo_matches(exL, exR, P, MBF, 0.0, FX)
        _lib.check(L.orbfe_stereo_result_batch(exL.handle, P, _lib.ptr(hu), _lib.ptr(hd), cap))

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_two_calls_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.steps)
    barrier()

    # the same work through the one-call stereo entry point (uploads, both extractions, matching and downloads pipelined inside)
    e2e_out = outL + outR + (hu, hd)

    def e2e_step():   # noqa: F811
        M.stereo_frames_batch_host(exL, exR, hLn, hRn, MBF, 0.0, FX, (0, 0), e2e_out)

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_sync_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.steps)
    barrier()

    # the same call split into submit / wait with two batches in flight: the upload of step k+1 overlaps the kernels of step k and
    # the download of step k overlaps the kernels of step k+1 (every step still uploads its own images from pinned memory and
    # downloads its own results; the outputs of consecutive steps go to two alternating sets of pinned arrays)
    def pinned_out():
        side = lambda: (np.zeros(P, np.int32), np.zeros(P, np.int32),  # noqa: E731
                        torch.empty((P, cap, 28), dtype=torch.uint8, pin_memory=True).numpy().view(_lib.KP_DTYPE).reshape(P, cap),
                        torch.empty((P, cap, 32), dtype=torch.uint8, pin_memory=True).numpy())
        return side() + side() + (torch.empty((P, cap), dtype=torch.float32, pin_memory=True).numpy(), torch.empty((P, cap), dtype=torch.float32, pin_memory=True).numpy())
    outs = [e2e_out, pinned_out()]

    def e2e_pipelined(nsteps):
        M.stereo_frames_submit(exL, exR, hLn, hRn, MBF, 0.0, FX, (0, 0), outs[0])
        for k in range(1, nsteps):
            M.stereo_frames_submit(exL, exR, hLn, hRn, MBF, 0.0, FX, (0, 0), outs[k & 1])
            M.stereo_frames_wait(exL, exR, outs[(k - 1) & 1])
        M.stereo_frames_wait(exL, exR, outs[(nsteps - 1) & 1])

    e2e_pipelined(3)
    barrier()
    t0 = time.perf_counter()
    e2e_pipelined(args.steps)
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3 / args.steps)
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][3][0], outs[1][3][0]), "the two output sets of the pipelined run disagree"
    barrier()
    # the host link's own speed for the same upload (pinned -> device), to read the e2e figure against
    t0 = time.perf_counter()
    for _ in range(3):
        dL.copy_(hL, non_blocking=True); dR.copy_(hR, non_blocking=True)
    torch.cuda.synchronize()
    h2d_gbs = 3 * 2 * P * W * H / (time.perf_counter() - t0) / 1e9
    e2e = {"value": 2 * P * world / (e2e_ms / 1e3), "unit": "frames/s", "ms_per_step": e2e_ms,
           "call": "orbfe_stereo_frames_submit / _wait, two batches in flight (pinned host images in, keypoints / descriptors / uRight / depth out, every step)",
           "one_synchronous_call_ms_per_step": e2e_sync_ms,
           "two_extract_calls_plus_stereo_ms_per_step": e2e_two_calls_ms,
           "fraction_of_device_resident": ms_step / e2e_ms,
           "h2d_gbs_reached_in_step": 2 * P * W * H / (e2e_ms / 1e3) / 1e9,
           "pinned_h2d_gbs_measured": h2d_gbs, "h2d_only_ms_per_step": 2 * P * W * H / h2d_gbs / 1e6,
           "h2d_bytes_per_step": 2 * P * W * H,
           "d2h_bytes_per_step": 2 * P * (cap * 60 + 8) + 2 * P * cap * 4}

    # ---- single-frame latency through the reference-facing call (config 1: one 752x480 frame, host image in, results out)
    ex1 = ORBextractor(*PARAMS, width=W, height=H, device=local_rank, max_batch=1)
    one = hL.numpy()[0]
    for _ in range(5):
        ex1(one)
    t0 = time.perf_counter()
    for _ in range(20):
        ex1(one)
    single_ms = (time.perf_counter() - t0) * 1e3 / 20
    ex1.close()
    # the same frame through the C++ drop-in as shipped (include/orbfe_shim/ORBextractor.h::operator(), compiled from
    # tests/shim_smoke.cpp by build()): with the eight mvImagePyramid host mirrors, and built with -DORBFE_SHIM_LAZY_PYRAMID
    shim_ms = {}
    if rank == 0:
        import re
        import subprocess
        pkg = os.path.join(ROOT, "orb_slam3_detailed_comments_kor_b200")
        for key, exe in (("pyramid_mirrored", "shim_probe.bin"), ("pyramid_lazy", "shim_probe_lazy.bin")):
            try:
                out = subprocess.run([os.path.join(pkg, exe), "--latency", "200"], capture_output=True, text=True, timeout=120,
                                     env=dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", str(local_rank)))).stdout
                shim_ms[key] = float(re.search(r"shim_latency_us=([\d.]+)", out).group(1)) / 1e3
            except Exception as exc:   # the probe is optional evidence, not part of the timed region
                shim_ms[key] = None
                shim_ms["error"] = repr(exc)[:200]

    # ---- input conditioning: rectification of the same frames, device to device
    from orb_slam3_detailed_comments_kor_b200.conditioning import Rectifier
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    xn, yn = (xx - 367.215) / 458.654, (yy - 248.375) / 457.296            # EuRoC cam0 intrinsics / distortion (EuRoC.yaml)
    r2 = xn * xn + yn * yn
    fac = (1 - 0.28340811 * r2 + 0.07395907 * r2 * r2).astype(np.float32)
    rect = Rectifier((xn * fac * 458.654 + 367.215).astype(np.float32), (yn * fac * 457.296 + 248.375).astype(np.float32), device=local_rank)
    d_rect = torch.empty_like(dL)
    for _ in range(3):
        rect.batch_device(dL.data_ptr(), d_rect.data_ptr(), P, st)
    barrier()
    e0.record()
    for _ in range(args.steps):
        rect.batch_device(dL.data_ptr(), d_rect.data_ptr(), P, st)
    e1.record()
    barrier()
    rect_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    rectify = {"value": P * world / (rect_ms / 1e3), "unit": "frames/s", "ms_per_step": rect_ms, "frames_per_step": P * world,
               "achieved_gbs": P * W * H * 2 / (rect_ms / 1e3) / 1e9, "note": "cv::remap INTER_LINEAR, read + write of the frame counted"}

    # ---- Hamming top-2 (config 5): database sharded over the ranks, NCCL all-gather of the per-shard top-2 + merge
    from orb_slam3_detailed_comments_kor_b200 import sharding
    nq, ndb = args.hamming_queries, args.hamming_db
    lo, hi = sharding.shard_rows(ndb, world, rank)
    rows = hi - lo
    db, q = synth.descriptors(5, ndb, nq)
    shard = DescriptorDatabase(rows, nq, device=local_rank)
    shard.set(db[lo:hi], index_base=lo)
    d_q = torch.from_numpy(q).cuda()
    d_part = torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
    d_out = torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ham_out = [d_out]

    comm = sharding.create_nccl_comm(rank, world, local_rank)

    def ham_step():   # the C entry point: shard search, ncclAllGather, merge on one stream (no Python between them)
        shard.top2_sharded(comm, world, d_q.data_ptr(), nq, 50, 0.9, False, d_out.data_ptr(), st)

    def ham_step_harness():   # the same through torch.distributed (round 1's path), kept as a cross-check and a comparison
        shard.top2_device(d_q.data_ptr(), nq, 50, 0.9, False, d_part.data_ptr(), st)
        gathered = sharding.gather_top2(d_part, world)
        return sharding.merge_gathered(gathered, 50, 0.9, False, device=local_rank, stream=st)

    for _ in range(3):
        ham_step()
    harness_ms = None
    if world > 1:
        check_out = ham_step_harness()
        torch.cuda.synchronize()
        assert torch.equal(check_out, d_out), "orbfe_hamming_top2_sharded disagrees with the torch.distributed all-gather + merge"
        barrier()
        e0.record()
        for _ in range(args.steps):
            ham_step_harness()
        e1.record()
        barrier()
        harness_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    barrier()
    e0.record()
    for _ in range(args.steps):
        ham_step()
    e1.record()
    barrier()
    ham_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    pairs_s = nq * ndb / (ham_ms / 1e3)
    popc_nominal = 148 * 16 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6       # POPC/s/GPU: 16 per clock per SM (nominal)
    popc_measured = C.c_double(0.0)
    _lib.check(L.orbfe_probe_popc(local_rank, C.byref(popc_measured)))            # POPC/s/GPU measured on this GPU just now (k_probe_popc)
    popc_peak = popc_measured.value / 8.0                                         # pairs/s/GPU at 8 algorithmic POPC per pair
    hamming = {"value": pairs_s, "unit": "descriptor_pairs/s", "ms_per_step": ham_ms, "scaling": "strong",
               "queries": nq, "db_rows": ndb, "shards": world, "accepted": int(ham_out[0][:, 3].sum().item()),
               "call": "orbfe_hamming_top2_sharded (shard search + ncclAllGather + merge enqueued by one C call)",
               "torch_distributed_all_gather_path_ms": harness_ms,
               "roofline": {"bound": "int_popc", "achieved": pairs_s / world, "peak": popc_peak, "unit": "pairs/s/GPU",
                            "frac": pairs_s / world / popc_peak,
                            "peak_source": "measured in this run: orbfe_probe_popc (eight independent POPC chains per thread) / 8 POPC per pair",
                            "popc_per_s_measured": popc_measured.value, "popc_per_s_nominal": popc_nominal,
                            "frac_of_issued_popc_bound": pairs_s / world / (popc_measured.value / 6.0),
                            "note": "accounting stays at 8 algorithmic POPC per pair; the kernel issues 6 (two carry-save adders), so frac > 1.0 is possible; "
                                    "frac_of_issued_popc_bound divides by the POPC rate / 6, the bound of the instruction stream as written"}}

    # ---- projection-gated search: Tracking::SearchLocalPoints-sized frames, one CTA per frame
    from orb_slam3_detailed_comments_kor_b200 import projection as PJ
    from orb_slam3_detailed_comments_kor_b200.synth import projection_cases as PC
    pj_cases = [PC.make_case(seed=900 + i, n=1200, nq=1500, W=W, H=H, th=1.0, stereo_gate=True) for i in range(8)]
    PF = 512                                                              # frames per step per GPU (the 8 distinct ones, tiled)
    pj_order = [pj_cases[i % 8] for i in range(PF)]
    pj_cat = lambda k: np.concatenate([c[k] for c in pj_order])  # noqa: E731
    pj_koff = np.cumsum([0] + [len(c["kps"]) for c in pj_order]).astype(np.int32)
    pj_qoff = np.cumsum([0] + [len(c["queries"]) for c in pj_order]).astype(np.int32)
    to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()  # noqa: E731
    pj_d = {k: to_dev(pj_cat(k)) for k in ("kps", "desc", "queries", "qdesc", "u_right", "taken")}
    pj_d["grid"] = to_dev(np.stack([c["grid"] for c in pj_order]))
    pj_d["koff"], pj_d["qoff"] = to_dev(pj_koff), to_dev(pj_qoff)
    pj_out = torch.zeros(int(pj_qoff[-1]) * 28, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    def pj_step():
        PJ.search_by_projection_device(PF, 1200, pj_d["kps"].data_ptr(), pj_d["desc"].data_ptr(), pj_d["koff"].data_ptr(), pj_d["grid"].data_ptr(),
                                       pj_d["queries"].data_ptr(), pj_d["qdesc"].data_ptr(), pj_d["qoff"].data_ptr(), pj_out.data_ptr(),
                                       d_u_right=pj_d["u_right"].data_ptr(), d_taken_in=pj_d["taken"].data_ptr(), device=local_rank, stream=st)

    for _ in range(3):
        pj_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        pj_step()
    e1.record()
    barrier()
    pj_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    pj_acc = int(pj_out.cpu().numpy().view(PJ.PROJ_MATCH_DTYPE)["accepted"].sum())
    projection = {"value": int(pj_qoff[-1]) * world / (pj_ms / 1e3), "unit": "map_point_searches/s", "frames_per_s": PF * world / (pj_ms / 1e3),
                  "ms_per_step": pj_ms, "frames_per_step": PF * world, "keypoints_per_frame": 1200, "map_points_per_frame": int(pj_qoff[1]),
                  "accepted_per_step": pj_acc, "note": "ORBmatcher::SearchByProjection(Frame&, vpMapPoints, th=1) incl. the frame grid; device resident"}

    # the tracker's per-frame call: SearchByProjection(CurrentFrame, LastFrame, th = 7, bMono = false) with the rotation check
    lf_cases = [PC.make_case(seed=950 + i, n=1200, nq=1200, W=W, H=H, th=7.0, stereo_gate=True, last_frame="still") for i in range(8)]
    lf_order = [lf_cases[i % 8] for i in range(PF)]
    lf_cat = lambda k: np.concatenate([c[k] for c in lf_order])  # noqa: E731
    lf_koff = np.cumsum([0] + [len(c["kps"]) for c in lf_order]).astype(np.int32)
    lf_qoff = np.cumsum([0] + [len(c["queries"]) for c in lf_order]).astype(np.int32)
    lf_d = {k: to_dev(lf_cat(k)) for k in ("kps", "desc", "queries", "qdesc", "u_right", "taken")}
    lf_d["grid"], lf_d["koff"], lf_d["qoff"] = to_dev(np.stack([c["grid"] for c in lf_order])), to_dev(lf_koff), to_dev(lf_qoff)
    lf_out = torch.zeros(int(lf_qoff[-1]) * 28, dtype=torch.uint8, device="cuda")

    def lf_step():
        PJ.search_by_projection_device(PF, 1200, lf_d["kps"].data_ptr(), lf_d["desc"].data_ptr(), lf_d["koff"].data_ptr(), lf_d["grid"].data_ptr(),
                                       lf_d["queries"].data_ptr(), lf_d["qdesc"].data_ptr(), lf_d["qoff"].data_ptr(), lf_out.data_ptr(),
                                       d_u_right=lf_d["u_right"].data_ptr(), d_taken_in=lf_d["taken"].data_ptr(), check_orientation=True, device=local_rank, stream=st)
    for _ in range(3):
        lf_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        lf_step()
    e1.record()
    barrier()
    lf_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    projection["last_frame"] = {"frames_per_s": PF * world / (lf_ms / 1e3), "ms_per_step": lf_ms, "frames_per_step": PF * world, "map_points_per_frame": 1200,
                                "accepted_per_step": int(lf_out.cpu().numpy().view(PJ.PROJ_MATCH_DTYPE)["accepted"].sum()),
                                "note": "ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th=7, bMono=false) with the rotation check (TrackWithMotionModel)"}

    # ---- ORBmatcher::SearchForInitialization: pairs of frames from the initial extractor (5 x nFeatures keypoints, of
    #      which about 22 % are on level 0 and take part), window 100 px, one CTA per pair
    def init_case(seed, n=5 * PARAMS[0], clusters=150, noise=20):
        rng = np.random.RandomState(seed)
        centres = rng.randint(0, 256, (clusters, 32)).astype(np.uint8)
        def frame():
            k = np.zeros(n, _lib.KP_DTYPE)
            k["x"], k["y"] = rng.uniform(0, W, n).astype(np.float32), rng.uniform(0, H, n).astype(np.float32)
            k["octave"] = np.where(rng.rand(n) < 0.22, 0, rng.randint(1, 8, n))
            k["angle"] = rng.uniform(0, 40, n).astype(np.float32)
            flip = np.zeros((n, 256), np.uint8)
            for i, m in enumerate(rng.randint(0, noise + 1, n)):
                flip[i, rng.randint(0, 256, m)] = 1
            return k, centres[rng.randint(0, clusters, n)] ^ np.packbits(flip, axis=1, bitorder="little")
        (k1, d1), (k2, d2) = frame(), frame()
        return k1, d1, k2, d2, PJ.frame_grid(0, W, 0, H), np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    in_cases = [init_case(700 + i) for i in range(8)]
    IP = 148
    in_order = [in_cases[i % 8] for i in range(IP)]
    in_o1 = np.cumsum([0] + [len(c[0]) for c in in_order]).astype(np.int32)
    in_o2 = np.cumsum([0] + [len(c[2]) for c in in_order]).astype(np.int32)
    in_d = [to_dev(np.concatenate([c[i] for c in in_order])) for i in range(4)] + [to_dev(np.stack([c[4] for c in in_order])), to_dev(in_o1), to_dev(in_o2)]
    in_prev0 = to_dev(np.concatenate([c[5] for c in in_order]))
    in_prev = in_prev0.clone()
    in_m = torch.zeros(int(in_o1[-1]), dtype=torch.int32, device="cuda")
    in_nm, in_rounds = torch.zeros(IP, dtype=torch.int32, device="cuda"), torch.zeros(IP, dtype=torch.int32, device="cuda")

    def in_step():
        in_prev.copy_(in_prev0, non_blocking=True)                        # vbPrevMatched is updated in place: every step starts from the same one
        PJ.search_for_initialization_device(IP, in_d[0].data_ptr(), in_d[1].data_ptr(), in_d[5].data_ptr(), in_d[2].data_ptr(), in_d[3].data_ptr(), in_d[6].data_ptr(),
                                            in_d[4].data_ptr(), in_prev.data_ptr(), in_m.data_ptr(), in_nm.data_ptr(), 100, 0.9, True, 50, in_rounds.data_ptr(),
                                            device=local_rank, stream=st)
    for _ in range(3):
        in_step()
    barrier()
    e0.record()
    for _ in range(args.steps):
        in_step()
    e1.record()
    barrier()
    in_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    initialization = {"value": IP * world / (in_ms / 1e3), "unit": "frame_pairs/s", "ms_per_step": in_ms, "pairs_per_step": IP * world,
                      "keypoints_per_frame": 5 * PARAMS[0], "level0_keypoints_per_frame": int((in_cases[0][0]["octave"] == 0).sum()),
                      "matches_per_pair": float(in_nm.float().mean().item()), "rounds_max": int(in_rounds.max().item()),
                      "note": "ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, 100) incl. F2's grid; device resident"}

    # ---- bag-of-words assignment: Frame::ComputeBoW on the left frames the extractor just produced (device
    #      buffers chained), ORBvoc-shaped synthetic vocabulary (k = 10, L = 6, 1.1 M nodes, TF-IDF, L1)
    from orb_slam3_detailed_comments_kor_b200.bow import ORBVocabulary
    vrng = np.random.RandomState(77)
    v_parent, v_desc = [], []
    level_ids, level_desc, next_id = np.array([0]), vrng.randint(0, 256, (1, 32)).astype(np.uint8), 1
    for lvl in range(1, 7):
        nch = len(level_ids) * 10
        flip = vrng.randint(0, 256, (nch, 32)).astype(np.uint8) & vrng.randint(0, 256, (nch, 32)).astype(np.uint8)
        for _ in range(lvl):
            flip &= vrng.randint(0, 256, (nch, 32)).astype(np.uint8)                  # fewer flipped bits further down the tree
        d = np.repeat(level_desc, 10, axis=0) ^ flip
        v_parent.append(np.repeat(level_ids, 10)); v_desc.append(d)
        level_ids, level_desc, next_id = np.arange(next_id, next_id + nch), d, next_id + nch
    v_parent = np.concatenate(v_parent).astype(np.int32); v_desc = np.concatenate(v_desc)
    v_leaf = np.zeros(len(v_parent), np.uint8); v_leaf[-1000000:] = 1
    v_weight = np.where(v_leaf == 1, vrng.uniform(0.5, 9.0, len(v_parent)), 0.0)
    voc = ORBVocabulary(10, 6, 0, 0, v_parent, v_leaf, v_desc, v_weight, max_features=exL.capacity, max_batch=P, device=local_rank)
    vk, vd, vc, vm = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
    _lib.check(L.orbfe_batch_device_views(exL.handle, C.byref(vk), C.byref(vd), C.byref(vc), C.byref(vm)))
    for _ in range(3):
        voc.transform_batch_device(vd.value, exL.capacity, vc.value, P, 4, stream=st)
    barrier()
    e0.record()
    for _ in range(args.steps):
        voc.transform_batch_device(vd.value, exL.capacity, vc.value, P, 4, stream=st)
    e1.record()
    barrier()
    bow_ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
    bow = {"value": P * world / (bow_ms / 1e3), "unit": "frames/s", "ms_per_step": bow_ms, "frames_per_step": P * world,
           "features_per_frame": nkp, "vocabulary": "synthetic k=10 L=6 (1111110 nodes, 10^6 words), TF-IDF, L1",
           "words_in_frame0": int(len(voc.result(0, int(exL.batch_counts(1)[0][0]))["bow_words"])),
           "note": "Frame::ComputeBoW = DBoW2 transform(features, BowVec, FeatVec, 4); extractor buffers chained on the device"}

    # ---- the consumers of the feature vectors: SearchByBoW and SearchForTriangulation between consecutive frames of the
    #      batch, chained on the device behind the transform above; and MapPoint::ComputeDistinctiveDescriptors (N4) for 20000 map points
    bv = [C.c_void_p() for _ in range(9)]
    _lib.check(L.orbfe_bow_device_views(voc.handle, *[C.byref(v) for v in bv]))
    NP = P - 1
    d_pk, d_pf = torch.arange(0, NP, dtype=torch.int32, device="cuda"), torch.arange(1, P, dtype=torch.int32, device="cuda")
    d_sm, d_snm = torch.zeros((NP, exL.capacity), dtype=torch.int32, device="cuda"), torch.zeros(NP, dtype=torch.int32, device="cuda")
    d_snm2 = torch.zeros(NP, dtype=torch.int32, device="cuda")
    Kc = np.array([[458.654, 0, 367.215], [0, 457.296, 248.375], [0, 0, 1]], np.float64)
    F12 = (np.linalg.inv(Kc).T @ np.array([[0, 0, 0], [0, 0, -0.11], [0, 0.11, 0]]) @ np.linalg.inv(Kc)).astype(np.float32)
    sf16 = np.array([np.float32(1.2) ** i for i in range(16)], np.float32)
    d_f12, d_ep = to_dev(np.tile(F12.reshape(9), NP)), to_dev(np.tile(np.array([-3000.0, 240.0], np.float32), NP))
    d_sf, d_sig = to_dev(sf16), to_dev((sf16 * sf16).astype(np.float32))

    def sbb_step():
        _lib.check(L.orbfe_search_by_bow_device(local_rank, NP, exL.capacity, d_pk.data_ptr(), d_pf.data_ptr(), vd.value, vk.value, exL.capacity, bv[5].value,
                                                bv[6].value, bv[7].value, bv[8].value, None, None, 50, C.c_float(0.7), 1, 0, d_sm.data_ptr(), None,
                                                d_snm.data_ptr(), st))

    def sbb_self_step():     # every frame against itself: each feature finds its own row at distance 0, so nearly all of them take one
        _lib.check(L.orbfe_search_by_bow_device(local_rank, NP, exL.capacity, d_pk.data_ptr(), d_pk.data_ptr(), vd.value, vk.value, exL.capacity, bv[5].value,
                                                bv[6].value, bv[7].value, bv[8].value, None, None, 50, C.c_float(0.7), 1, 0, d_sm.data_ptr(), None,
                                                d_snm2.data_ptr(), st))

    def sft_step():
        _lib.check(L.orbfe_search_for_triangulation_device(local_rank, NP, exL.capacity, d_pk.data_ptr(), d_pf.data_ptr(), vd.value, vk.value, None, exL.capacity,
                                                           bv[5].value, bv[6].value, bv[7].value, bv[8].value, None, d_f12.data_ptr(), d_ep.data_ptr(),
                                                           d_sf.data_ptr(), d_sig.data_ptr(), 50, 0, 0, 1, d_sm.data_ptr(), d_snm.data_ptr(), st, None))
    drng = np.random.RandomState(5)
    dd_off = np.concatenate([[0], np.cumsum(drng.randint(2, 16, 20000))]).astype(np.int32)
    dd_desc = drng.randint(0, 256, (int(dd_off[-1]), 32)).astype(np.uint8)
    d_dd, d_doff = to_dev(dd_desc), to_dev(dd_off)
    d_didx, d_dmed = torch.zeros(20000, dtype=torch.int32, device="cuda"), torch.zeros(20000, dtype=torch.int32, device="cuda")

    def dd_step():
        _lib.check(L.orbfe_distinctive_descriptors_device(local_rank, d_dd.data_ptr(), d_doff.data_ptr(), 20000, d_didx.data_ptr(), d_dmed.data_ptr(), st))
    consumers = {}
    for name, fn, units, unit in (("search_by_bow", sbb_step, NP, "frame_pairs/s"), ("search_by_bow_same_frame", sbb_self_step, NP, "frame_pairs/s"),
                                  ("search_for_triangulation", sft_step, NP, "keyframe_pairs/s"),
                                  ("distinctive_descriptors", dd_step, 20000, "map_points/s")):
        for _ in range(3):
            fn()
        barrier()
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1) / args.steps)
        consumers[name] = {"value": units * world / (ms / 1e3), "unit": unit, "ms_per_step": ms, "per_step": units * world}
    consumers["search_by_bow"]["note"] = "ORBmatcher::SearchByBoW(pKF, F, ...) between consecutive frames of the batch, feature vectors from the transform above (device chain)"
    consumers["search_by_bow"]["matches_per_pair"] = float(d_snm.float().mean().item())
    consumers["search_by_bow_same_frame"]["note"] = "the same call with every frame paired with itself: the load of a tracked frame, where most features take a row"
    consumers["search_by_bow_same_frame"]["matches_per_pair"] = float(d_snm2.float().mean().item())
    consumers["search_for_triangulation"]["note"] = "ORBmatcher::SearchForTriangulation (pinhole, Pinhole::epipolarConstrain on a sideways F12) on the same pairs"
    consumers["distinctive_descriptors"]["note"] = "MapPoint::ComputeDistinctiveDescriptors, 2-15 observations per map point"

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        sample = max(16 * cores, 64)
        kind = "reference" if reference_available() else "port"
        rate, dt = cpu_oracle_rate(sample, cores, kind)
        port_rate, _ = cpu_oracle_rate(sample, cores, "port") if kind == "reference" else (rate, dt)
        hrate, hdt = cpu_hamming_rate(nq, ndb, cores)
        cpu = {"value": rate, "unit": "frames/s", "cores": cores, "kind": kind,
               "sample": "%d stereo pairs (%.1f s wall) frame-parallel over %d threads; %s, built -O3 -march=x86-64-v3" % (
                   sample, dt, cores, "the reference's own ORBextractor.cc + Frame::ComputeStereoMatches (oracle/_ref; OpenCV primitives = scalar "
                   "restatements pinned to cv2 4.13)" if kind == "reference" else "CPU oracle restatement"),
               "port_frames_per_s": port_rate,
               "one_thread_frames_per_s": cpu_oracle_rate(6, 1, kind)[0],
               "two_threads_per_stereo_frame_frames_per_s": cpu_two_thread_rate(8)[0],
               "hamming": {"value": hrate, "unit": "descriptor_pairs/s", "sample": "%d queries x %d rows (%.1f s wall)" % (nq, ndb, hdt)}}
        cpu.update(cv2_lines(2 * cores, cores))
        from oracle import orb_oracle as O
        reps = 64 * cores

        def pj_cpu(i):
            c = pj_cases[i % 8]
            O.search_by_projection(c["kps"], c["desc"], c["queries"], c["qdesc"], c["grid"], u_right=c["u_right"], taken=c["taken"])

        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as tp:
            list(tp.map(pj_cpu, range(reps)))
        pdt = time.perf_counter() - t0
        ovoc = O.OracleVocabulary(10, 6, 0, 0, v_parent, v_leaf, v_desc, v_weight)
        cntL0, _, kpsL0, descL0 = exL.extract_batch_host(hLn[:8], (0, 0))
        bsample = [np.ascontiguousarray(descL0[i, :cntL0[i]]) for i in range(8)]
        ksample = [np.ascontiguousarray(kpsL0[i, :cntL0[i]]) for i in range(8)]
        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as tp:
            list(tp.map(lambda i: ovoc.transform(bsample[i % 8], 4), range(8 * cores)))
        bdt = time.perf_counter() - t0
        cpu["bow"] = {"value": 8 * cores / bdt, "unit": "frames/s", "sample": "%d frames (%.2f s wall) frame-parallel over %d threads" % (8 * cores, bdt, cores)}
        fvs = [ovoc.transform(bsample[i], 4)["feat_vec"] for i in range(8)]
        scale8 = sf16[:8].copy()
        for name, call in (("search_by_bow", lambda i: O.search_by_bow(bsample[i % 7], ksample[i % 7], fvs[i % 7], None, bsample[i % 7 + 1], ksample[i % 7 + 1], fvs[i % 7 + 1], 50, 0.7, True)),
                           ("search_for_triangulation", lambda i: O.search_for_triangulation(bsample[i % 7], ksample[i % 7], None, None, fvs[i % 7], bsample[i % 7 + 1], ksample[i % 7 + 1], None,
                                                                                             None, fvs[i % 7 + 1], F12, [-3000.0, 240.0], scale8, scale8 * scale8, 50, False, False, True))):
            t0 = time.perf_counter()
            with ThreadPoolExecutor(cores) as tp:
                list(tp.map(call, range(8 * cores)))
            cdt = time.perf_counter() - t0
            cpu[name] = {"value": 8 * cores / cdt, "unit": consumers[name]["unit"], "kind": "port", "sample": "%d pairs (%.2f s wall) pair-parallel over %d threads" % (8 * cores, cdt, cores)}
        t0 = time.perf_counter()
        O.distinctive_descriptors(dd_desc[:int(dd_off[4000])], dd_off[:4001])
        cdt = time.perf_counter() - t0
        cpu["distinctive_descriptors"] = {"value": 4000 / cdt, "unit": "map_points/s", "kind": "port", "sample": "4000 map points (%.2f s wall), one thread" % cdt}
        from oracle import ref_oracle as _RO
        init_fn = _RO.search_for_initialization if reference_available() else O.search_for_initialization
        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as tp:
            list(tp.map(lambda i: init_fn(*in_cases[i % 8][:6], 100, 0.9, True), range(4 * cores)))
        idt = time.perf_counter() - t0
        cpu["initialization"] = {"value": 4 * cores / idt, "unit": "frame_pairs/s", "kind": "reference" if reference_available() else "port",
                                 "sample": "%d pairs (%.2f s wall) pair-parallel over %d threads" % (4 * cores, idt, cores)}
        t0 = time.perf_counter()
        with ThreadPoolExecutor(cores) as tp:
            list(tp.map(lambda i: O.search_by_projection(lf_cases[i % 8]["kps"], lf_cases[i % 8]["desc"], lf_cases[i % 8]["queries"], lf_cases[i % 8]["qdesc"],
                                                         lf_cases[i % 8]["grid"], u_right=lf_cases[i % 8]["u_right"], taken=lf_cases[i % 8]["taken"], check_orientation=True),
                        range(16 * cores)))
        ldt = time.perf_counter() - t0
        cpu["projection_last_frame"] = {"value": 16 * cores / ldt, "unit": "frames/s", "kind": "port", "sample": "%d frames (%.2f s wall) frame-parallel over %d threads" % (16 * cores, ldt, cores)}
        # ---- one call at a time, as the C++ shims make them for one frame or keyframe pair: wall time of the host entry point (copies
        #      included) beside the same inputs on one CPU thread
        from orb_slam3_detailed_comments_kor_b200 import bow as BW

        def med_ms(fn, n):
            fn()
            ts = []
            for _ in range(n):
                t0 = time.perf_counter()
                fn()
                ts.append(time.perf_counter() - t0)
            return float(np.median(ts) * 1e3)
        c0, l0, i0 = pj_cases[0], lf_cases[0], in_cases[0]
        cpu_ex = _RO.ReferenceExtractor(*PARAMS) if reference_available() else O.OracleExtractor(*PARAMS)
        img0 = np.ascontiguousarray(hLn[0])
        one_call = {}
        for name, gpu_fn, cpu_fn in (
                ("extract", lambda: exL(img0, (0, 0)), lambda: cpu_ex(img0, (0, 0))),
                ("stereo_frame", lambda: M.stereo_frames_batch_host(exL, exR, hLn[:1], hRn[:1], MBF, 0.0, FX, (0, 0)),
                 (lambda: cpu_ex.stereo(img0, np.ascontiguousarray(hRn[0]), MBF, float(np.float32(MBF) / np.float32(FX)))) if reference_available() else None),
                ("compute_bow", lambda: voc.transform(bsample[0], 4), lambda: ovoc.transform(bsample[0], 4)),
                ("search_by_projection", lambda: PJ.search_by_projection(c0["kps"], c0["desc"], c0["queries"], c0["qdesc"], c0["grid"], u_right=c0["u_right"], taken=c0["taken"]),
                 lambda: O.search_by_projection(c0["kps"], c0["desc"], c0["queries"], c0["qdesc"], c0["grid"], u_right=c0["u_right"], taken=c0["taken"])),
                ("search_by_projection_last_frame",
                 lambda: PJ.search_by_projection(l0["kps"], l0["desc"], l0["queries"], l0["qdesc"], l0["grid"], u_right=l0["u_right"], taken=l0["taken"], check_orientation=True),
                 lambda: O.search_by_projection(l0["kps"], l0["desc"], l0["queries"], l0["qdesc"], l0["grid"], u_right=l0["u_right"], taken=l0["taken"], check_orientation=True)),
                ("search_by_bow", lambda: BW.search_by_bow(bsample[0], ksample[0], fvs[0], None, bsample[0], ksample[0], fvs[0], 50, 0.7, True),
                 lambda: O.search_by_bow(bsample[0], ksample[0], fvs[0], None, bsample[0], ksample[0], fvs[0], 50, 0.7, True)),
                ("search_for_triangulation",
                 lambda: BW.search_for_triangulation(bsample[0], ksample[0], None, None, fvs[0], bsample[1], ksample[1], None, None, fvs[1], F12, [-3000.0, 240.0], scale8,
                                                     scale8 * scale8, 50, False, False, True),
                 lambda: O.search_for_triangulation(bsample[0], ksample[0], None, None, fvs[0], bsample[1], ksample[1], None, None, fvs[1], F12, [-3000.0, 240.0], scale8,
                                                    scale8 * scale8, 50, False, False, True)),
                ("search_for_initialization", lambda: PJ.search_for_initialization(*[a.copy() if hasattr(a, "copy") else a for a in i0[:6]], 100, 0.9, True),
                 lambda: init_fn(*[a.copy() if hasattr(a, "copy") else a for a in i0[:6]], 100, 0.9, True))):
            one_call[name] = {"gpu_host_call_ms": med_ms(gpu_fn, 20), "cpu_one_thread_ms": med_ms(cpu_fn, 3) if cpu_fn else None}
        # the loops the reference runs over independent pairs, as ONE call: CreateNewMapPoints' 20 neighbours, Relocalization's candidates
        hframes = [dict(desc=bsample[i], kps=ksample[i], feat_vec=fvs[i]) for i in range(8)]
        nb20 = [1 + (i % 7) for i in range(20)]
        F20, ep20 = np.tile(F12.reshape(1, 9), (20, 1)).astype(np.float32), np.tile(np.array([[-3000.0, 240.0]], np.float32), (20, 1))
        one_call["search_for_triangulation_20_neighbours"] = {
            "gpu_host_call_ms": med_ms(lambda: BW.search_for_triangulation_frames(hframes, [0] * 20, nb20, F20, ep20, scale8, scale8 * scale8, 50, False, False, True), 20),
            "cpu_one_thread_ms": med_ms(lambda: [O.search_for_triangulation(bsample[0], ksample[0], None, None, fvs[0], bsample[j], ksample[j], None, None, fvs[j], F12,
                                                                            [-3000.0, 240.0], scale8, scale8 * scale8, 50, False, False, True) for j in nb20], 3)}
        one_call["search_by_bow_20_candidates"] = {
            "gpu_host_call_ms": med_ms(lambda: BW.search_by_bow_frames(hframes, nb20, [0] * 20, 50, 0.7, True), 20),
            "cpu_one_thread_ms": med_ms(lambda: [O.search_by_bow(bsample[j], ksample[j], fvs[j], None, bsample[0], ksample[0], fvs[0], 50, 0.7, True) for j in nb20], 3)}
        cpu["one_call_latency"] = one_call
        cpu["one_call_latency_note"] = ("median wall time of ONE call of the host entry point a shim makes (host arrays in, host arrays out, copies included, Python "
                                        "ctypes glue on both sides) beside the CPU oracle / reference on one thread with the same inputs; extract = one 752x480 "
                                        "frame; search_by_bow = a frame against itself (every feature takes a row)")
        cpu["projection"] = {"value": reps * int(pj_qoff[1]) / pdt, "unit": "map_point_searches/s",
                             "sample": "%d frames (%.2f s wall) frame-parallel over %d threads" % (reps, pdt, cores)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": WORKLOAD, "pairs_per_gpu_per_step": P, "frames_per_step": 2 * P * world,
                           "l2": "inputs_larger_than_L2 (%d MB of images per step per GPU)" % (2 * P * W * H >> 20),
                           "mean_keypoints_per_frame": nkp, "parallelism": "frames sharded by rank, no collective",
                           "schedule": "orbfe_stereo_frames_batch_device: left and right extraction on the two objects' streams (the reference's two "
                                       "ExtractORB threads), blur beside quadtree + assemble, ComputeStereoMatches behind both; timed by events on the caller's stream",
                           "serial_ms_per_step": ms_serial},
                "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "single_frame_latency_ms": single_ms, "shim_single_frame_latency_ms": shim_ms, "rectify": rectify, "projection": projection, "initialization": initialization, "bow": bow, "consumers": consumers,
                "gpu_launches": (2 * (PARAMS[2] + 8) + 3) * args.steps,
                "hamming": hamming}
        print(json.dumps(line), flush=True)
    if world > 1:
        sharding.destroy_nccl_comm(comm)
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
