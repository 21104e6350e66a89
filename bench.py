#!/usr/bin/env python3
"""bench.py — ORB front-end throughput on B200 (BASELINE.json metric: ORB extract frames/s at 752x480 / 1000 kp,
plus Hamming matches/s).

One "step" = one pass of the hot path over one batch of synthetic EuRoC-shaped frames (config[1] of BASELINE.json:
per-frame ORB extraction + SearchByProjection matching against the previous frame).
  value  : frames/s with the input frames already resident in HBM (device pipeline only)
  e2e    : frames/s through the C-ABI call with HOST buffers (pinned H2D of the frames + D2H of keypoints/descriptors
           inside the timed region)
  roofline / cpu_baseline / clocks / gpu_launches : see DESIGN.md "Measurement".
N > 1 (torchrun): one agent stream per GPU, no data-path collective -> weak scaling.
`--impl reference` times the CPU restatement of the reference path (oracle/, all host threads) on the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, NFEAT, NLEVELS, SCALE, INI_TH, MIN_TH = 752, 480, 1000, 8, 1.2, 20, 7
ALG_BYTES_PER_FRAME = 1_933_774          # input + 2 x derived levels + 60 B per keypoint
METRIC = "ORB extract frames/s (752x480, 1000 kp) + Hamming matches/s, 1-8 B200"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def make_frames(seed, n):
    from mam3slam_b200 import synth
    return synth.stream(W, H, seed, n)


TH_TRACK = 15.0                          # TrackWithMotionModel's first window (src/Tracking.cc:2789-2812)
BOUNDS = (0.0, 0.0, float(W), float(H))
WORKLOAD = ("single-agent EuRoC-shaped 752x480 synthetic mono stream per GPU, 1000 kp, 8 levels @1.2, FAST 20/7: per frame "
            "ORBextractor::operator() + SearchByProjection(Cur, Last, th=15, mono)")
I8_MAC_PEAK = 8192.0 * 148 * 1.965e9        # tcgen05.mma kind::i8, measured issue rate (tools/tc_rate.cu)
POPC_PEAK = 4.635e12                     # measured on this pool's B200 (profiles/int_peaks_b200.txt): 16.5 popc/clk/SM


def common_config(args):
    """`config` of the JSON line: the SAME dict in both arms (the driver compares them).  What differs between the arms - the
    bounded CPU sample, the measured keypoint / match counts, the schedules - lives under other keys (`cpu_baseline.sample`,
    `workload_stats`, `schedule`)."""
    nbytes = args.sets * args.batch * W * H
    return {"workload": WORKLOAD, "frames_per_batch_per_gpu": args.batch, "batches_per_step": args.batches_per_step,
            "frames_per_step_per_gpu": args.batch * args.batches_per_step,
            "l2_policy": f"{args.sets} distinct input batches rotated ({nbytes / 1e6:.0f} MB > 126 MB L2); all intermediates "
                         "rewritten every batch"}


def cpu_reference_run(frames, iters, threads):
    """Times the CPU restatement (oracle/) of the hot path — ORBextractor::operator() on every frame followed by
    SearchByProjection(Cur, Last) against the previous frame — on `iters` frames split over `threads` host threads
    (one independent agent stream per thread, like the reference's one Tracking thread per agent)."""
    import oracle
    from oracle import matcher as om
    from tests.matcher_scenarios import to_kp
    exs = [oracle.Extractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH) for _ in range(threads)]
    sf = exs[0].tables()["scale"]
    per = [iters // threads + (1 if t < iters % threads else 0) for t in range(threads)]
    evals = [0] * threads

    def work(t):
        last = None
        for i in range(per[t]):
            _, k, d = exs[t](frames[i % len(frames)])
            k = to_kp(k)
            if last is not None:
                k1, d1 = last
                om.search_frame(k, d, BOUNDS, np.zeros(len(k), np.uint8), sf, np.ones(len(k1), np.uint8), k1["x"], k1["y"],
                                k1["octave"], k1["angle"], d1, TH_TRACK, 100, True)
            last = (k, d)

    for ex in exs[:1]:
        ex(frames[0])                                # warm the allocator / caches
    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    t0 = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    wall = time.perf_counter() - t0
    return wall, sum(per)


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores (oracle port: the
    reference itself needs OpenCV C++ / Eigen, which are not installed)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    frames = make_frames(0, 8)
    per_step = threads * 8                  # every thread runs 8 consecutive frames of its stream: all but its first are matched
    for _ in range(min(args.warmup, 1)):
        cpu_reference_run(frames, threads, threads)
    tot_t, tot_f = 0.0, 0
    for _ in range(args.steps):
        t, f = cpu_reference_run(frames, per_step, threads)
        tot_t += t
        tot_f += f
    fps = tot_f / tot_t
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": common_config(args),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "port",
                             "sample": f"{per_step} frames per step x {args.steps} steps, {threads} threads (one agent stream of 8 "
                                       "consecutive frames each per step), oracle/cpp restatement of ORBextractor + "
                                       "SearchByProjection (the extractor restatement is bit-identical to the reference's own "
                                       "ORBextractor.cc, tests/test_oracle_vs_ref.py; ORBmatcher.cc needs Eigen/Sophus/DBoW2: "
                                       "not buildable here)"},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def latency_bench(args, torch, local, frames):
    """What ONE agent's Tracking thread sees (SURVEY.md 8d 'GPU timing (a)'): one synchronous ORBextractor::operator() call per
    frame with host buffers (H2D of the image and D2H of keypoints + descriptors inside the call), then one
    SearchByProjection(Cur, Last, th = 15) call with host arrays.  Median over 100 calls after 10 warm-up calls."""
    from mam3slam_b200 import ORBextractor, ORBmatcher
    ex = ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local)
    m = ORBmatcher(0.9, True, device=local)
    sf = ex.GetScaleFactors()
    pin = torch.from_numpy(np.ascontiguousarray(frames[:8])).pin_memory().numpy()
    t_ex, t_sp = [], []
    last, F = None, None
    for i in range(110):
        t0 = time.perf_counter()
        _, k, d = ex(pin[i % len(pin)])
        t1 = time.perf_counter()
        if last is not None and len(k):
            F = m.frame(k, d, BOUNDS) if F is None else (F.update(k, d, BOUNDS) or F)
            lk, ld = last
            m.SearchByProjection_Frame(F, np.zeros(len(k), np.uint8), sf, np.ones(len(lk), np.uint8), lk["x"], lk["y"], lk["octave"],
                                       lk["angle"], ld, TH_TRACK)
        t2 = time.perf_counter()
        if i >= 10:
            t_ex.append(t1 - t0); t_sp.append(t2 - t1)
        last = (k, d)
    if F is not None:
        F.close()
    m.close(); ex.close()
    e, s_ = float(np.median(t_ex)), float(np.median(t_sp))
    return {"operator_call_ms": e * 1e3, "search_by_projection_call_ms": s_ * 1e3, "single_stream_frames_per_s": 1.0 / (e + s_),
            "note": "one frame in flight, synchronous calls, host buffers (pinned image)"}


def agents_in_flight(local, frames, counts=(1, 4, 16)):
    """How the reference feeds frames (src/Agent.cc:92-154: every agent's thread grabs ONE image, tracks it, grabs the next):
    A agents on one GPU, each a host thread with its own extractor + matcher (mam3slam_b200.agents.AgentFrontEnd), one frame in
    flight per agent, synchronous calls with host buffers.  Aggregate frames/s and the median per-frame latency an agent sees."""
    import threading
    from mam3slam_b200.agents import AgentFrontEnd
    rows = []
    nframes = 60
    for A in counts:
        agents = [AgentFrontEnd(a, local, NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH) for a in range(A)]
        lat = [[] for _ in range(A)]
        go = threading.Barrier(A + 1)

        def run(a):
            ag = agents[a]
            for i in range(6):
                ag.GrabImageMonocular(frames[(a + i) % len(frames)], TH_TRACK)
            go.wait()
            for i in range(nframes):
                t0 = time.perf_counter()
                ag.GrabImageMonocular(frames[(a + i) % len(frames)], TH_TRACK)
                lat[a].append(time.perf_counter() - t0)
        th = [threading.Thread(target=run, args=(a,)) for a in range(A)]
        for t in th:
            t.start()
        go.wait()
        t0 = time.perf_counter()
        for t in th:
            t.join()
        wall = time.perf_counter() - t0
        for ag in agents:
            ag.close()
        rows.append({"agents": A, "frames_per_s": A * nframes / wall,
                     "median_frame_latency_ms": float(np.median(np.concatenate(lat))) * 1e3})
    return {"rows": rows, "note": "A agents x one frame in flight on ONE GPU, one host thread per agent (Python threads: the calls "
                                  "release the GIL, the numpy marshalling between them does not), extraction + SearchByProjection "
                                  "per frame, host buffers"}


def agents_native(local, frames, counts=(1, 4, 8, 16), per_agent=300):
    """The same loop from NATIVE host threads (tools/agents_native.cc, built by mam3slam_b200/csrc/Makefile): one std::thread per
    agent, each with its own extractor + matcher handle, orbx_extract + orbm_frame_update + orbm_search_frame per frame over
    the C ABI with host buffers - the reference's shape (one C++ Tracking thread per agent) without the interpreter in the
    loop.  With and without the tracking search."""
    import subprocess
    import tempfile
    exe = os.path.join(ROOT, "mam3slam_b200", "agents_native")
    if not os.path.exists(exe):
        return {"unavailable": "mam3slam_b200/agents_native not built (make -C mam3slam_b200/csrc)"}
    out = {"note": "A agents x one frame in flight on ONE GPU, one native host thread per agent over the C ABI, host buffers "
                   "(pageable image in, keypoints / descriptors / matches out), %d frames per agent" % per_agent}
    with tempfile.NamedTemporaryFile(suffix=".u8") as f:
        np.stack(frames[:16]).astype(np.uint8).tofile(f.name)
        nfile = min(len(frames), 16)
        # *_pinned_frames: the frame array page-locked once (orbx_pin_host, a registered camera ring buffer): frames go up in place
        for track, pin, key in ((1, 0, "extract_and_track"), (0, 0, "extract_only"), (1, 1, "extract_and_track_pinned_frames"),
                                (0, 1, "extract_only_pinned_frames")):
            r = subprocess.run([exe, f.name, str(W), str(H), str(nfile), str(local), str(per_agent), ",".join(map(str, counts)), str(track),
                                str(pin)], capture_output=True, text=True, timeout=600)
            if r.returncode != 0:
                out[key] = {"error": r.stderr[-300:]}
                continue
            out[key] = json.loads(r.stdout.strip().splitlines()[-1])["rows"]
    return out


def search_latency_rows(local):
    """One row per ORBmatcher search member: per-call time through the product the way the C++ adapter calls it
    - an ORBmatcher CONSTRUCTED PER CALL, as the reference does (`ORBmatcher matcher(0.9,true);` inside every tracking /
    mapping / loop-closing function, e.g. src/Tracking.cc:2789), its frame slot(s) re-bound, host arrays in, host arrays out -
    beside the CPU port's (oracle/cpp) time for the same call on one host thread.  Median over 30 / 8 calls after 3 warm-ups.
    Inputs: two consecutive EuRoC-shaped frames of the synthetic stream (tests/matcher_scenarios.py)."""
    from mam3slam_b200 import ORBmatcher
    from oracle import matcher as om
    from tests import matcher_scenarios as sc
    (k1, d1), (k2, d2), bounds, tabs = sc.frame_pair(W, H, NFEAT, 0)
    rng = np.random.default_rng(3)
    n1, n2 = len(k1), len(k2)
    scale, sig2, isig2 = tabs["scale"], tabs["sigma2"], tabs["inv_sigma2"]
    jit = lambda a, amp: (a + rng.uniform(-amp, amp, len(a))).astype(np.float32)
    u, v = jit(k1["x"], 4), jit(k1["y"], 4)
    u2, v2 = jit(k2["x"], 2), jit(k2["y"], 2)
    lvl = np.clip(k1["octave"] + rng.integers(0, 2, n1), 0, NLEVELS - 1).astype(np.int32)
    lvl2 = np.clip(k2["octave"] + rng.integers(0, 2, n2), 0, NLEVELS - 1).astype(np.int32)
    ones1, ones2 = np.ones(n1, np.uint8), np.ones(n2, np.uint8)
    occ = np.zeros(n2, np.uint8)
    vc = np.ones(n1, np.float32)
    fv1, fv2 = sc.feat_vec(d1), sc.feat_vec(d2)
    fv1t, fv2t = sc.feat_vec(d1, 5), sc.feat_vec(d2, 5)
    F12 = np.array([[0, 0, 0.02], [0, 0, -1.0], [-0.02, 1.0, 0]], np.float32)
    ep = np.array([W * 0.8, H * 0.5], np.float32)
    prev = np.stack([k1["x"], k1["y"]], 1)

    def gpu(nn, ori, fn, two=False):
        def call():
            m = ORBmatcher(nn, ori, device=local)
            A = m.slot(0, k2, d2, bounds) if not two else m.slot(0, k1, d1, bounds)
            B = m.slot(1, k2, d2, bounds) if two else None
            r = fn(m, A, B)
            m.close()
            return r
        return call

    rows = [
        ("M1 SearchByProjection(Frame, local map points)", "src/ORBmatcher.cc:43-213",
         gpu(0.8, True, lambda m, A, B: m.SearchByProjection_LocalMap(A, occ, scale, ones1, u, v, lvl, vc, d1, 3.0)),
         lambda: om.search_local_map(k2, d2, bounds, occ, scale, ones1, u, v, lvl, vc, d1, 3.0, 0.8)),
        ("M2 SearchByProjection(Cur, Last)", ":1676-1887",
         gpu(0.9, True, lambda m, A, B: m.SearchByProjection_Frame(A, occ, scale, ones1, u, v, k1["octave"], k1["angle"], d1, 15.0)),
         lambda: om.search_frame(k2, d2, bounds, occ, scale, ones1, u, v, k1["octave"], k1["angle"], d1, 15.0, 100, True)),
        ("M3 SearchByProjection(Frame, KeyFrame, found)", ":1889-2010",
         gpu(0.9, True, lambda m, A, B: m.SearchByProjection_Frame(A, occ, scale, ones1, u, v, lvl, k1["angle"], d1, 10.0, 64)),
         lambda: om.search_frame(k2, d2, bounds, occ, scale, ones1, u, v, lvl, k1["angle"], d1, 10.0, 64, True)),
        ("M4 SearchByBoW(KeyFrame, Frame)", ":223-425",
         gpu(0.7, True, lambda m, A, B: m.SearchByBoW_KF_Frame(A, fv1, ones1, B, fv2), True),
         lambda: om.search_bow_kf_frame(k1, d1, fv1, ones1, k2, d2, fv2, 0.7, True)),
        ("M4 SearchByBoW(KeyFrame, KeyFrame)", ":765-905",
         gpu(0.8, True, lambda m, A, B: m.SearchByBoW_KF_KF(A, fv1, ones1, B, fv2, ones2), True),
         lambda: om.search_bow_kf_kf(k1, d1, fv1, ones1, k2, d2, fv2, ones2, 0.8, True)),
        ("M5 SearchForInitialization", ":648-763",
         gpu(0.9, True, lambda m, A, B: m.SearchForInitialization(A, B, prev, 100), True),
         lambda: om.search_initialization(k1, d1, k2, d2, bounds, prev, 100, 0.9, True)),
        ("M6 SearchForTriangulation (Pinhole)", ":907-1146",
         gpu(0.6, False, lambda m, A, B: m.SearchForTriangulation(A, fv1t, ones1, B, fv2t, ones2, ep, scale, F12, sig2 * 9, False), True),
         lambda: om.search_triangulation(k1, d1, fv1t, ones1, k2, d2, fv2t, ones2, ep, scale, F12, sig2 * 9, False, False)),
        ("M7 SearchByProjection(KeyFrame, Sim3, points)", ":427-646",
         gpu(0.9, True, lambda m, A, B: m.SearchByProjection_Sim3(A, occ, scale, ones1, u, v, lvl, d1, 8, 1.5)),
         lambda: om.search_sim3(k2, d2, bounds, occ, scale, ones1, u, v, lvl, d1, 8, 1.5)),
        ("M8 SearchBySim3", ":1457-1674",
         gpu(0.75, True, lambda m, A, B: m.SearchBySim3(A, B, scale, scale, ones1, u, v, lvl, d1, ones2, u2, v2, lvl2, d2, 7.5), True),
         lambda: om.search_by_sim3(k1, d1, bounds, k2, d2, bounds, scale, scale, ones1, u, v, lvl, d1, ones2, u2, v2, lvl2, d2, 7.5)),
        ("M9 Fuse (candidate part)", ":1148-1455",
         gpu(0.6, True, lambda m, A, B: m.Fuse_candidates(A, scale, isig2, ones1, u, v, lvl, d1, 3.0, True)),
         lambda: om.fuse_candidates(k2, d2, bounds, scale, isig2, ones1, u, v, lvl, d1, 3.0, True)),
    ]

    def med(fn, warm, reps):
        for _ in range(warm):
            fn()
        t = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); t.append(time.perf_counter() - t0)
        return float(np.median(t)) * 1e3

    out = []
    for name, ref, g, c in rows:
        g_ms, c_ms = med(g, 3, 30), med(c, 1, 8)
        out.append({"member": name, "reference": ref, "gpu_call_ms": g_ms, "cpu_port_call_ms": c_ms, "ratio": c_ms / g_ms})
    ORBmatcher.drain_pool()
    return {"note": "ORBmatcher constructed per call (pooled handle + frame slots), frame(s) re-uploaded and grid rebuilt per call, "
                    f"host arrays in / out, synchronous; {n1} / {n2} keypoints; CPU = oracle/cpp on one thread through ctypes",
            "rows": out}


def sparse_bench(args, torch, local, rank):
    """the dense synthetic texture (26 k FAST candidates per frame) loads FAST's sparse paths and the
    quadtree unrepresentatively.  The same device-resident extraction on a SPARSE-texture stream (synth.stream(sparse=True):
    smooth shading + shape edges, ~11 k candidates), strictly serial on one stream, with the exact compass early-out of the
    FAST strength pass off and on (ORBX_FAST_COMPASS; byte-identical results, tests/test_env_switches_gpu.py)."""
    from mam3slam_b200 import ORBextractor, synth
    B = args.batch
    fr = synth.stream(W, H, 500 + rank, min(B, 32), sparse=True)
    host = np.concatenate([np.roll(fr, 3 * r, axis=2) for r in range((2 * B + len(fr) - 1) // len(fr))])[:2 * B]
    dev = torch.from_numpy(host).cuda()
    ts = torch.cuda.Stream()
    out = {"workload": f"sparse-texture 752x480 stream, extraction only, {B} device-resident frames per batch, one stream"}
    for name, flag in (("compass_off", "0"), ("compass_on", "1")):
        os.environ["ORBX_FAST_COMPASS"] = flag
        ex = ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B)
        os.environ.pop("ORBX_FAST_COMPASS")
        def step(i):
            ex.extract_device(dev.data_ptr() + (i % 2) * B * W * H, W * H, B, W, H, W, (0, 1000), ts.cuda_stream)
        for i in range(4):
            step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 20
        e0.record(ts)
        for i in range(iters):
            step(i)
        e1.record(ts)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        out[name] = {"frames_per_s": B / (ms * 1e-3), "ms_per_batch": ms}
        if name == "compass_off":
            ex.set_stage_timing(True)
            step(0); torch.cuda.synchronize(); ex.synchronize()
            out["stage_ms_per_batch"] = ex.stage_times()
            ex.set_stage_timing(False)
            out["keypoints_per_frame"] = float(np.mean([len(ex.fetch(f)[1]) for f in range(4)]))
            out["fast_candidates_per_frame"] = float(np.mean([sum(len(ex.candidates(l, f)) for l in range(NLEVELS)) for f in range(2)]))
        ex.close()
    return out


def kitti_bench(args, torch, dist, rank, world, local):
    """BASELINE config 3: KITTI-shaped 1241x376 frames, 2000 kp: device-resident extraction of 32 frames per step, and
    SearchForTriangulation (LocalMapping::CreateNewMapPoints: ratio 0.6, no orientation check) on keyframe pairs (t, t+5)
    through the host-buffer C ABI.  BoW node ids are a hash of descriptor bits (the ORB vocabulary file is absent from the
    reference checkout).  Returns a dict for the JSON line."""
    from mam3slam_b200 import ORBextractor, ORBmatcher, synth
    KW, KH, KN, KB = 1241, 376, 2000, 32
    frames = synth.stream(KW, KH, 100 + rank, 8)
    host = np.concatenate([np.roll(frames, 5 * r, axis=2) for r in range(KB // len(frames))])
    dev = torch.from_numpy(host).cuda()
    ex = ORBextractor(KN, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=KB)
    ts = torch.cuda.Stream()
    def step():
        ex.extract_device(dev.data_ptr(), KW * KH, KB, KW, KH, KW, (0, 1000), ts.cuda_stream)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 10
    e0.record(ts)
    for _ in range(iters):
        step()
    e1.record(ts)
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    res = [ex.fetch(f) for f in range(8)]                   # (mono, keypoints, descriptors) of the 8 distinct frames
    nkp = float(np.mean([len(r[1]) for r in res]))
    # keyframe pairs (t, t+5): unmatched keypoints (60 %) searched under a sideways-translation epipolar geometry
    m = ORBmatcher(0.6, False, device=local)
    bounds = (0.0, 0.0, float(KW), float(KH))
    sf = ex.GetScaleFactors()
    sig2 = ex.GetScaleSigmaSquares()
    rng = np.random.default_rng(5)
    def featvec(desc):
        node = ((desc[:, 0].astype(np.int32) ^ (desc[:, 7].astype(np.int32) << 1)) & 63) * 3 + 5
        nodes = np.unique(node)
        order = np.argsort(node, kind="stable")
        off = np.concatenate([[0], np.cumsum([(node == nd).sum() for nd in nodes])]).astype(np.int32)
        return nodes.astype(np.int32), off, order.astype(np.int32)
    F12 = np.array([[0, 0, 0.02], [0, 0, -1.0], [-0.02, 1.0, 0]], np.float32)
    ep = np.array([KW * 0.8, KH * 0.5], np.float32)
    pairs = []
    for a, b in ((0, 5), (1, 6), (2, 7)):
        (_, k1, d1), (_, k2, d2) = res[a], res[b]
        pairs.append((m.frame(k1, d1, bounds), featvec(d1), (rng.random(len(k1)) < 0.6).astype(np.uint8),
                      m.frame(k2, d2, bounds), featvec(d2), (rng.random(len(k2)) < 0.6).astype(np.uint8)))
    def tri():
        tot, ev = 0, 0
        for K1, fv1, f1, K2, fv2, f2 in pairs:
            nm, _ = m.SearchForTriangulation(K1, fv1, f1, K2, fv2, f2, ep, sf, F12, sig2 * 9, False)
            tot += nm; ev += m.last_evaluations()
        return tot, ev
    tri()
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        nm, ev = tri()
    dt = (time.perf_counter() - t0) / (reps * len(pairs))
    out = {"workload": f"KITTI-shaped {KW}x{KH}, {KN} kp: extraction of {KB} device-resident frames per step; SearchForTriangulation "
                       f"on keyframe pairs (t, t+5), host buffers, synchronous calls",
           "extract_frames_per_s": KB * world / (ms * 1e-3), "ms_per_step": ms, "keypoints_per_frame": nkp,
           "triangulation_calls_per_s": world / dt, "triangulation_matches_per_call": nm / len(pairs),
           "triangulation_evaluations_per_call": ev / len(pairs)}
    for P in pairs:
        P[0].close(); P[3].close()
    m.close(); ex.close()
    return out


def knn_bench(args, torch, dist, rank, world, local):
    """BASELINE config 5: 1 M x 2000 brute-force k-NN (k = 2), database rows sharded over the ranks, per-rank top-k lists
    all-gathered (NCCL) and merged with the (dist, global row) rule.  Returns a dict for the JSON line."""
    from mam3slam_b200 import ORBmatcher
    from mam3slam_b200.matcher import KnnDatabase
    NDB, NQ, K = 1_000_000, 2000, 2
    g = torch.Generator(device="cuda").manual_seed(1234)
    db = torch.randint(0, 256, (NDB, 32), dtype=torch.uint8, device="cuda", generator=g)
    q = db[torch.randint(0, NDB, (NQ,), device="cuda", generator=g)].clone()
    q[:, 0] ^= 1
    lo, hi = NDB * rank // world, NDB * (rank + 1) // world
    shard = db[lo:hi].contiguous()
    m = ORBmatcher(0.6, True, device=local)
    D = KnnDatabase(m, None, row_offset=lo, device_ptr=shard.data_ptr(), n=hi - lo)
    # distances (u16) and row ids (i32) of a rank sit in ONE buffer so that a single all-gather moves both (the exchange is
    # latency-bound: 24 KB per rank)
    nb_i, nb_d = NQ * K * 4, NQ * K * 2
    out = torch.empty(nb_i + nb_d, dtype=torch.uint8, device="cuda")
    di = out[:nb_i].view(torch.int32).view(NQ, K)
    dd = out[nb_i:].view(torch.int16).view(NQ, K)
    st = torch.cuda.current_stream().cuda_stream
    gath = torch.empty((world, nb_i + nb_d), dtype=torch.uint8, device="cuda")
    gath_i = [gath[r, :nb_i].view(torch.int32).view(NQ, K) for r in range(world)]
    gath_d = [gath[r, nb_i:] for r in range(world)]

    md_dev = torch.empty((NQ, K), dtype=torch.int16, device="cuda")
    mi_dev = torch.empty((NQ, K), dtype=torch.int32, device="cuda")

    def step_nccl():
        m.knn_device(D, q.data_ptr(), NQ, K, dd.data_ptr(), di.data_ptr(), st)
        if world > 1:                  # the library-collective form: local top-k, one NCCL all-gather, k-way merge on every rank
            dist.all_gather_into_tensor(gath, out)
            m.knn_merge_device(gath.data_ptr() + nb_i, gath.data_ptr(), nb_i + nb_d, world, NQ, K, md_dev.data_ptr(),
                               mi_dev.data_ptr(), st)

    # the product path on N > 1 GPUs: the scan kernel stores its merged lists into every rank's exchange buffer over NVLink peer
    # stores and the merge kernel waits for the arrival counters (orbm_knn_sharded): no collective call at all
    xchg = None
    xd_dev = torch.empty((NQ, K), dtype=torch.int16, device="cuda")
    xi_dev = torch.empty((NQ, K), dtype=torch.int32, device="cuda")
    if world > 1:
        from mam3slam_b200.matcher import KnnExchange

        def gather(b):
            outl = [None] * world
            dist.all_gather_object(outl, b)
            return outl
        xchg = KnnExchange(m, rank, world, NQ, K).connect(gather)
        dist.barrier()

    def step_peer():
        m.knn_sharded(xchg, D, q.data_ptr(), NQ, K, xd_dev.data_ptr(), xi_dev.data_ptr(), st)

    def timed(step):
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        iters = 10
        e0.record()
        for _ in range(iters):
            step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms_nccl = timed(step_nccl)
    ms = timed(step_peer) if world > 1 else ms_nccl
    sharded_knn_tc = os.environ.get("ORBM_KNN_TC") == "1"       # A/B: the shard scanned by the tensor-core kernel inside orbm_knn_sharded

    # A/B beside the POPC kernel (never instead of it: north_star's design is the integer pipe): the same search with the dot
    # products on the tensor cores, tcgen05.mma kind::i8 (orbm_knn_device_tc), on this rank's shard; results must be identical
    td_dev = torch.empty((NQ, K), dtype=torch.int16, device="cuda")
    ti_dev = torch.empty((NQ, K), dtype=torch.int32, device="cuda")
    ms_tc = timed(lambda: m.knn_device_tc(D, q.data_ptr(), NQ, K, td_dev.data_ptr(), ti_dev.data_ptr(), st))
    m.knn_device(D, q.data_ptr(), NQ, K, dd.data_ptr(), di.data_ptr(), st)
    torch.cuda.synchronize()
    if not (torch.equal(td_dev, dd) and torch.equal(ti_dev, di)):
        raise SystemExit("bench.py: tensor-core k-NN differs from the POPC kernel")
    # correctness of the merged result on rank 0: the planted neighbour (distance 1) is found for every query
    if world > 1:
        md, mi = ORBmatcher.knn_merge(torch.stack(gath_d).cpu().numpy().view(np.uint16).reshape(world, NQ, K), torch.stack(gath_i).cpu().numpy())
        if not (np.array_equal(md, md_dev.cpu().numpy().view(np.uint16)) and np.array_equal(mi, mi_dev.cpu().numpy())):
            raise SystemExit("bench.py: device merge of the sharded k-NN lists differs from the host merge")
        if xchg.status() != 0 or not (np.array_equal(md, xd_dev.cpu().numpy().view(np.uint16)) and np.array_equal(mi, xi_dev.cpu().numpy())):
            raise SystemExit("bench.py: peer-exchange sharded k-NN differs from the all-gather + host merge")
    else:
        md, mi = dd.cpu().numpy().view(np.uint16), di.cpu().numpy()
    ok = bool((md[:, 0] <= 1).all())
    evals = float(NDB) * NQ
    eps = evals / (ms * 1e-3)
    out = {"workload": f"brute-force k-NN {NDB} x {NQ}, k={K}, db rows sharded over {world} GPU(s)" +
                       ("; exchange of the per-rank lists by NVLink peer stores from the scan kernel + merge kernel waiting on arrival "
                        "counters (orbm_knn_sharded), all inside the timed step" if world > 1 else ""),
           "ms": ms, "ms_nccl_allgather_variant": ms_nccl if world > 1 else None,
           "sharded_scan_kernel": ("k_knn_tc (ORBM_KNN_TC=1)" if sharded_knn_tc else "k_knn (POPC)") if world > 1 else None,
           "evaluations_per_s": eps, "planted_neighbours_found": ok,
           "tensor_core_variant": {
               "ms_local_shard": ms_tc, "evaluations_per_s_per_gpu": float(hi - lo) * NQ / (ms_tc * 1e-3),
               "identical_to_popc_kernel": True,
               "what": "orbm_knn_device_tc on this rank's rows (scan + chunk merge, no exchange): Hamming = popc(q) + popc(r) - 2<q,r>, "
                       "dot products by tcgen05.mma kind::i8 (M=128 queries, N=256 rows, K=288 = 256 bit-bytes + 32 popcount bytes) on "
                       "bits expanded to bytes in shared memory, s32 accumulators in TMEM, top-k per TMEM lane",
               "roofline_tensor": {"bound": "tensor", "unit": "TMAC/s", "achieved": float(hi - lo) * NQ * 288 / (ms_tc * 1e-3) / 1e12,
                                   "peak": I8_MAC_PEAK / 1e12, "frac": float(hi - lo) * NQ * 288 / (ms_tc * 1e-3) / I8_MAC_PEAK,
                                   "peak_source": "tools/tc_rate.cu on this pool: 128 cycles per 128x256x32 MMA = 8192 MAC/clk/SM x 148 SMs x 1.965 GHz",
                                   "note": "tile time = MMA (1152 cycles) + accumulator read-out (128 KB at the 64 B/clk TMEM -> register rate): "
                                           "the read-out, not the MMA, is the larger part (DESIGN.md 4)"}},
           "roofline_int": {"bound": "popc", "popc_issued_per_evaluation": 4, "achieved_popc_per_s": 4 * eps,
                            "peak_popc_per_s": POPC_PEAK * world, "frac": 4 * eps / (POPC_PEAK * world),
                            "popc8_equivalent_frac": 8 * eps / (POPC_PEAK * world),
                            "note": "frac = POPC instructions the kernel ISSUES (4 per 256-bit evaluation: carry-save adder tree in "
                                    "LOP3 halves the naive 8) against the measured POPC rate (16.5 /clk/SM, "
                                    "profiles/int_peaks_b200.txt); popc8_equivalent_frac charges the naive 8 and may exceed 1"}}
    if rank == 0 and not args.no_cpu:
        # CPU baseline for Hamming matches/s: the oracle's popcount brute force on ALL host threads over a
        # 50 k-row subsample of the same database and the same 2000 queries; evaluations/s do not depend on the row count, so the
        # full-size figure is the same rate (labelled extrapolated: the 1 M x 2000 search itself is not run on the CPU)
        from oracle import matcher as om
        sub, threads = 50_000, os.cpu_count() or 1
        db_h, q_h = db[:sub].cpu().numpy(), q.cpu().numpy()
        om.knn(q_h[:64], db_h, K, threads)
        t0 = time.perf_counter()
        om.knn(q_h, db_h, K, threads)
        dt = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": sub * NQ / dt, "unit": "evaluations/s", "cores": threads, "kind": "port",
                               "sample": f"{sub} x {NQ} of the same rows / queries, k={K}, oracle/cpp popcount scan "
                                         "(ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2058-2074, in a strict-< k-NN loop)",
                               "extrapolated_full_search_s": NDB * NQ / (sub * NQ / dt)}
    if xchg is not None:
        xchg.close()
    D.close(); m.close()
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from mam3slam_b200 import ORBextractor, ORBmatcher, KEYPOINT_DTYPE

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    # host placement: this rank's threads and pinned staging buffers go to the NUMA node of its GPU (mam3slam_b200/numa.py);
    # the original CPU mask comes back before the CPU baselines run.  BENCH_NUMA=0 switches it off (A/B).
    from mam3slam_b200 import numa
    cpu_mask0 = os.sched_getaffinity(0)
    placement = (numa.bind_to_device_node(local) if os.environ.get("BENCH_NUMA", "1") != "0"
                 else {"bound": False, "node": None, "cpus": 0, "reason": "BENCH_NUMA=0"})
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    B = args.batch
    nsets = args.sets                      # distinct input batches: nsets*B*361 KB > L2 (126 MB)
    frames = make_frames(rank, min(64, B))                 # agent `rank`'s stream
    reps = (B * nsets + len(frames) - 1) // len(frames)
    # page-locked host frames: the e2e leg copies from pinned memory as the contract asks
    host_t = torch.empty((B * nsets, H, W), dtype=torch.uint8).pin_memory()
    host = host_t.numpy()
    host[:] = np.concatenate([frames] * reps)[:B * nsets]
    # make the batches distinct so no step re-reads cached input: roll each copy by a few pixels
    for i in range(len(frames), len(host)):
        host[i] = np.roll(host[i], (i // len(frames)) * 3, axis=1)
    dev = torch.from_numpy(host).cuda()
    img_bytes = W * H
    # `pipes` independent (extractor, matcher, stream) triples work on alternating batches of the SAME frame stream, so that the
    # tracking step and the kernel tails of batch i overlap the extraction of batch i+1 (each handle owns its buffers, so
    # nothing is overwritten while it is still being read).  pipes = 1 is the strictly serial schedule.
    npipes = max(1, args.pipes)
    epipes = max(1, args.e2e_pipes)
    pipes = []
    for _ in range(max(npipes, epipes)):
        pex = ORBextractor(NFEAT, SCALE, NLEVELS, INI_TH, MIN_TH, device=local, max_batch=B)
        pmt = ORBmatcher(0.9, True, device=local)
        pts = torch.cuda.Stream()          # a real (non-NULL) stream: CUDA events below are recorded on it
        pex.extract_device(dev.data_ptr(), img_bytes, B, W, H, W, (0, 1000), pts.cuda_stream)   # allocates the shape-dependent buffers
        torch.cuda.synchronize()
        pdk, pdd, pdc, pstride = pex.results_batch_device()
        pipes.append(dict(ex=pex, mt=pmt, ts=pts, st=pts.cuda_stream, dk=pdk, dd=pdd, dc=pdc,
                          kpm=torch.empty((B, pstride), dtype=torch.int32, device="cuda"),
                          nmt=torch.empty(B, dtype=torch.int32, device="cuda")))
    ex, mt, tstream = pipes[0]["ex"], pipes[0]["mt"], pipes[0]["ts"]
    sf = ex.GetScaleFactors()
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    dk, dd, dc, stride = pipes[0]["dk"], pipes[0]["dd"], pipes[0]["dc"], ex.results_batch_device()[3]
    kpm, nmt = pipes[0]["kpm"], pipes[0]["nmt"]

    bps = max(1, args.batches_per_step)

    def batch_device(j):
        s = j % nsets
        P = pipes[j % npipes]
        P["ex"].extract_device(dev.data_ptr() + s * B * img_bytes, img_bytes, B, W, H, W, (0, 1000), P["st"])
        P["mt"].track_batch_device(B, P["dk"], P["dd"], P["dc"], stride, BOUNDS, sf, TH_TRACK, P["kpm"].data_ptr(), P["nmt"].data_ptr(),
                                   stream=P["st"])

    def step_device(i):                    # one step = bps batches of B frames (consecutive batches alternate between the pipes)
        for b in range(bps):
            batch_device(i * bps + b)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: device-resident input ----
    # initialisation, not part of the W warm-up steps: every (handle pair, input set) combination runs once so that the CUDA
    # graph of each argument set exists before anything is timed (the extractor captures one graph per input base address)
    import math
    for i in range(nsets * npipes // math.gcd(nsets, npipes)):
        batch_device(i)
    torch.cuda.synchronize()
    for i in range(args.warmup):
        step_device(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    def launches_now():
        return sum(P["ex"].launch_count() + P["mt"].launch_count() for P in pipes[:npipes])
    l0 = launches_now()
    e0 = torch.cuda.Event(enable_timing=True)
    e0.record(pipes[(args.warmup * bps) % npipes]["ts"])      # every stream is idle here (barrier above); this pipe takes the first timed batch
    for i in range(args.steps):
        step_device(args.warmup + i)
    ends = []
    for P in pipes[:npipes]:
        e = torch.cuda.Event(enable_timing=True)
        e.record(P["ts"])
        ends.append(e)
    barrier()
    ms = max(e0.elapsed_time(e) for e in ends)
    launches = launches_now() - l0
    nkp = float(np.mean([ex.fetch(f)[1].shape[0] for f in range(min(B, 4))]))
    track_evals = mt.track_evaluations()
    nmatch = float(nmt[1:].float().mean().item())

    # ---- per-kernel durations (CUDA events on the launching stream, same work as the timed region) ----
    ex.set_stage_timing(True)
    acc = {}
    nrep = min(args.steps, 5)
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(nrep):
        s = i % nsets
        ex.extract_device(dev.data_ptr() + s * B * img_bytes, img_bytes, B, W, H, W, (0, 1000), stream)
        t0e.record()
        mt.track_batch_device(B, dk, dd, dc, stride, BOUNDS, sf, TH_TRACK, kpm.data_ptr(), nmt.data_ptr(), stream=stream)
        t1e.record()
        torch.cuda.synchronize()
        ex.synchronize()
        for k, v in ex.stage_times().items():
            acc[k] = acc.get(k, 0.0) + v / nrep
        acc["track"] = acc.get("track", 0.0) + t0e.elapsed_time(t1e) / nrep
    acc["total"] = acc["total"] + acc["track"]
    ex.set_stage_timing(False)

    # ---- e2e: page-locked host buffers through the C-ABI calls (H2D of frames, D2H of keypoints / descriptors / matches) ----
    cap = stride
    host_batches = [host[s * B:(s + 1) * B] for s in range(nsets)]
    for P in pipes:
        P["kps_h"] = torch.empty((B, cap * KEYPOINT_DTYPE.itemsize), dtype=torch.uint8).pin_memory().numpy().view(KEYPOINT_DTYPE).reshape(B, cap)
        P["desc_h"] = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory().numpy()
        P["n_h"], P["mono_h"] = np.zeros(B, np.int32), np.zeros(B, np.int32)
        P["kpm_h"] = torch.empty((B, stride), dtype=torch.int32).pin_memory()
        P["nm_h"] = torch.empty(B, dtype=torch.int32).pin_memory()

    def batch_e2e(i):
        # the calls a user makes, with HOST buffers: frames in, keypoints + descriptors + matches out, then a stream sync
        P = pipes[i % epipes]
        P["ex"].extract_batch_into(host_batches[i % nsets], P["kps_h"], P["desc_h"], P["n_h"], P["mono_h"])
        with torch.cuda.stream(P["ts"]):
            P["mt"].track_batch_device(B, P["dk"], P["dd"], P["dc"], stride, BOUNDS, sf, TH_TRACK, P["kpm"].data_ptr(),
                                       P["nmt"].data_ptr(), stream=P["st"])
            P["kpm_h"].copy_(P["kpm"], non_blocking=True)
            P["nm_h"].copy_(P["nmt"], non_blocking=True)
        P["ts"].synchronize()

    def run_e2e(first, count):
        # one host thread per handle pair, like one Tracking thread per extractor in the reference; thread p takes steps
        # p, p + epipes, ... so that the copies of one batch overlap the kernels of the other
        if epipes == 1:
            for i in range(first, first + count):
                batch_e2e(i)
            return
        import threading

        def worker(p):
            torch.cuda.set_device(local)
            for i in range(first + ((p - first) % epipes), first + count, epipes):
                batch_e2e(i)
        th = [threading.Thread(target=worker, args=(p,)) for p in range(epipes)]
        for t in th:
            t.start()
        for t in th:
            t.join()

    run_e2e(0, max(min(args.warmup, 3), epipes))
    barrier()
    t0 = time.perf_counter()
    run_e2e(0, args.steps * bps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e ceiling: the same pinned 23.1 MB batches, host -> device copies ONLY, every rank at once ----
    # What the end-to-end figure could reach if kernels and result copies were free: PCIe plus host-memory bandwidth shared by
    # all ranks of the node.  Same streams and the same alternation over the input sets as the e2e leg.
    land = [torch.empty((B, H, W), dtype=torch.uint8, device="cuda") for _ in range(epipes)]
    host_sets = [host_t[s * B:(s + 1) * B] for s in range(nsets)]
    def h2d_only(count):
        for i in range(count):
            with torch.cuda.stream(pipes[i % epipes]["ts"]):
                land[i % epipes].copy_(host_sets[i % nsets], non_blocking=True)
        torch.cuda.synchronize()
    h2d_only(2 * epipes)
    barrier()
    t0 = time.perf_counter()
    h2d_only(args.steps * bps)
    h2d_s = time.perf_counter() - t0
    del land
    os.sched_setaffinity(0, cpu_mask0)

    tms = torch.tensor([ms, e2e_s * 1e3, h2d_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms, e2e_ms, h2d_ms = tms.tolist()
    bound = torch.tensor([1.0 if placement["bound"] else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(bound, op=dist.ReduceOp.SUM)
    total_frames = B * bps * args.steps * world
    hamming = None if args.no_knn else knn_bench(args, torch, dist, rank, world, local)
    kitti = None if args.no_kitti else kitti_bench(args, torch, dist, rank, world, local)
    latency = latency_bench(args, torch, local, frames) if rank == 0 and not args.no_latency else None
    if latency is not None:
        latency["agents_one_frame_in_flight"] = agents_in_flight(local, frames)
        latency["agents_one_frame_in_flight_native"] = agents_native(local, frames)
    if latency is not None and not args.no_cpu:
        latency["search_members"] = search_latency_rows(local)
    sparse = sparse_bench(args, torch, local, rank) if rank == 0 and not args.no_sparse else None
    if rank == 0:
        pk, pk_kind = peaks()
        fps = total_frames / (ms * 1e-3)
        stage_bytes = stage_algorithmic_bytes(nkp, track_evals / max(B - 1, 1))
        dom = max((k for k in acc if k != "total"), key=lambda k: acc[k])
        ach = stage_bytes[dom] * B / (acc[dom] * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic(dom) if B == 64 else (None, None)
        line = {
            "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": common_config(args),
            "workload_stats": {"keypoints_per_frame": nkp, "matches_per_frame": nmatch,
                               "hamming_evaluations_per_frame": track_evals / max(B - 1, 1),
                               "texture": "dense synthetic texture: ~26 k FAST candidates per frame, far denser than EuRoC imagery "
                                          "(SURVEY A8); see `sparse` for a sparse-texture stream"},
            "schedule": {"device": (f"{npipes} handle pairs on {npipes} CUDA streams take alternating batches of the one frame stream "
                                    "(tracking of batch i overlaps extraction of batch i+1)") if npipes > 1 else "serial, one stream",
                         "e2e": f"{epipes} host threads, one handle pair each, take alternating batches; every call is synchronous"},
            "e2e": {"value": total_frames / (e2e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(bps * B * W * H), "d2h_bytes_per_step": int(bps * B * (cap * 60 + 8 + stride * 4 + 4)),
                    "ceiling_frames_per_s": total_frames / (h2d_ms * 1e-3),
                    "ceiling_h2d_gbs": total_frames * W * H / (h2d_ms * 1e-3) / 1e9,
                    "ceiling_note": "the same pinned batches copied host -> device and nothing else, all ranks at once, max over ranks "
                                    "(PCIe + host memory shared by the node's ranks); value / ceiling = how much of the link the "
                                    "end-to-end path uses",
                    "host_placement": {"ranks_bound_to_gpu_numa_node": int(bound.item()), "rank0": placement}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s",
                         "frac": ach / pk["hbm_gbs"], "traffic": traffic, "traffic_source": traffic_src,
                         "peak_kind": pk_kind,
                         "alu": fast_alu_roofline(acc.get("fast"), B),
                         "stage_ms_per_batch": acc,
                         "note": "per launch = one batch of 64 frames; integer/issue-bound path: see DESIGN.md 4; pipeline_* = 1.93 MB "
                                 "algorithmic bytes per frame",
                         "pipeline_alg_gbs": ALG_BYTES_PER_FRAME * fps / 1e9 / world,
                         "pipeline_frac": ALG_BYTES_PER_FRAME * fps / 1e9 / world / pk["hbm_gbs"]},
            "hamming": {"track_evaluations_per_s": track_evals * world / (acc["track"] * 1e-3), "knn": hamming},
            "kitti": kitti,
            "latency": latency,
            "sparse": sparse,
        }
        if world == 1 and not args.no_cpu:
            cf = make_frames(0, 8)
            wall, n = cpu_reference_run(cf, args.cpu_frames, 1)
            line["cpu_baseline"] = {"value": n / wall, "unit": "frames/s", "cores": 1, "kind": "port",
                                    "sample": f"{n} frames of the same stream (extraction + SearchByProjection), 1 thread, "
                                              "oracle/cpp restatement (extractor bit-identical to the reference's own "
                                              "ORBextractor.cc compiled against a cv:: shim, tests/test_oracle_vs_ref.py)"}
            line["cpu_baseline_cv2"] = cv2_baseline(cf)
        emit(line)
    ex.close(); mt.close()
    if world > 1:
        dist.destroy_process_group()


STAGE_KERNELS = {"pyramid": ("k_pyr_level",), "fast": ("k_fast_score", "k_fast_nms", "k_fast_select", "k_fast_cells"),
                 "octree": ("k_octree",), "blur": ("k_blur7",), "finalize": ("k_finalize",), "orient_brief": ("k_orient_brief",),
                 "track": ("k_track_",)}


def ncu_traffic(stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch (one batch of 64 EuRoC frames) of a stage's kernels, summed over
    the kernels of ONE batch, read from the newest committed `ncu --set full` summary under profiles/ (tools/ncu_summary.py).
    Returns (bytes or None, source file or None): nothing is hard-coded, so a stale number cannot survive a new profile."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_summary.csv")))
    if not files:
        return None, None
    try:
        with open(files[-1]) as f:
            rows = list(csv.reader(f))
        hdr, units = rows[0], rows[1]
        ik, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot, batches = 0.0, 0
        for r in rows[2:]:
            batches += "k_octree" in r[ik]                  # one quadtree launch per batch: the capture may hold several batches
            if any(k in r[ik] for k in STAGE_KERNELS[stage]):
                tot += float(r[ir]) * scale[units[ir]] + float(r[iw]) * scale[units[iw]]
        if tot == 0.0:
            return None, None
        return tot / max(batches, 1), os.path.relpath(files[-1], ROOT)
    except Exception:
        return None, None


def cv2_baseline(frames):
    """Secondary CPU baseline with the REAL OpenCV (SIMD), to show the scalar port is not a straw man.
    (a) cv2.ORB with the reference's settings - OpenCV's own ORB: the same pixel stages (pyramid, FAST-9/16 + NMS, IC angle,
    7x7 blur, rBRIEF) but NOT the reference's per-cell FAST fallback and quadtree; (b) the reference's pixel primitives alone
    on whole levels (cv2.resize + copyMakeBorder, cv2.FAST at 20, GaussianBlur).  One thread (cv2.setNumThreads(1))."""
    try:
        import cv2
    except Exception as e:                                   # cv2 is part of the image; never fatal for the bench line
        return {"unavailable": str(e)}
    from oracle.py import cv2_pipeline as cp
    old = cv2.getNumThreads()
    cv2.setNumThreads(1)
    try:
        orb = cv2.ORB_create(nfeatures=NFEAT, scaleFactor=SCALE, nlevels=NLEVELS, edgeThreshold=19, firstLevel=0, WTA_K=2,
                             scoreType=cv2.ORB_FAST_SCORE, patchSize=31, fastThreshold=INI_TH)
        orb.detectAndCompute(frames[0], None)
        n, t0 = 0, time.perf_counter()
        while time.perf_counter() - t0 < 3.0:
            orb.detectAndCompute(frames[n % len(frames)], None)
            n += 1
        t_orb = (time.perf_counter() - t0) / n
        st = {"pyramid": 0.0, "fast_whole_level": 0.0, "blur": 0.0}
        det = cv2.FastFeatureDetector_create(INI_TH, True)
        reps = 10
        for i in range(reps):
            img = frames[i % len(frames)]
            t0 = time.perf_counter(); pyr = cp.pyramid(img, SCALE, NLEVELS); t1 = time.perf_counter()
            for _, roi in pyr:
                det.detect(np.ascontiguousarray(roi))
            t2 = time.perf_counter()
            for _, roi in pyr:
                cp.blur(roi)
            t3 = time.perf_counter()
            st["pyramid"] += (t1 - t0) / reps; st["fast_whole_level"] += (t2 - t1) / reps; st["blur"] += (t3 - t2) / reps
    finally:
        cv2.setNumThreads(old)
    return {"value": 1.0 / t_orb, "unit": "frames/s", "cores": 1, "kind": "cv2.ORB (OpenCV %s, SIMD), extraction only" % cv2.__version__,
            "ms_per_frame": t_orb * 1e3, "primitive_ms_per_frame": {k: v * 1e3 for k, v in st.items()},
            "sample": f"{n} calls of cv2.ORB.detectAndCompute on the same 752x480 stream (1000 kp, 8 levels @1.2, FAST 20, FAST_SCORE); "
                      "OpenCV's ORB has no per-cell minThFAST fallback and no quadtree, so it is a lower bound on the CPU cost of "
                      "the reference's extractor with real OpenCV primitives"}


def fast_alu_roofline(fast_ms, B):
    """FAST is bound by the min/max network, not by HBM.  Per pixel PAIR the strength network issues 48 two-input
    VIMNMX.U16x2 + 2 (final reduction) and 22 + 1 three-input VIMNMX3.U16x2 = 94 ALU-pipe slots of a two-input op (a
    three-input op costs two: tools/microbench.cu, profiles/int_peaks_b200.txt: vminu2 36264 G/s, vimin3_u16x2 18417 G/s)."""
    if not fast_ms:
        return None
    lv = [(752, 480), (627, 400), (522, 333), (435, 278), (363, 231), (302, 193), (252, 161), (210, 134)]
    tested = sum((w - 32) * (h - 32) for w, h in lv)          # FAST runs 16 px inside every level (EDGE_THRESHOLD - 3)
    slots = (50 + 2 * 23) * 1.0
    ops = tested / 2 * slots * B
    ach = ops / (fast_ms * 1e-3)
    return {"op": "VIMNMX.U16x2 slots (3-input = 2)", "per_pixel_pair": slots, "instructions_per_pixel_pair": 73,
            "achieved_ops_per_s": ach, "peak_ops_per_s": 36264.3e9, "frac": ach / 36264.3e9}


def stage_algorithmic_bytes(nkp, evals_per_frame=20000.0):
    """Algorithmic bytes per FRAME of each kernel (DESIGN.md 'Kernels'): pixels of the 8 EuRoC levels etc."""
    lv = [752 * 480, 627 * 400, 522 * 333, 435 * 278, 363 * 231, 302 * 193, 252 * 161, 210 * 134]
    allpx, derived = sum(lv), sum(lv[1:])
    ncand = 26000                               # typical FAST candidates per synthetic frame (4 B each)
    return {
        "pyramid": sum(lv[:-1]) + derived,      # read every source level once, write every derived level once
        "fast": allpx + 4 * ncand,              # read every level once, write packed candidates
        "octree": 4 * ncand + 8 * ncand + 8 * nkp,   # read candidates, write 64-bit keys once, write kept keys
        "blur": 2 * allpx,                      # read + write every level
        "finalize": 8 * nkp + 28 * nkp + 16 * nkp,
        "orient_brief": nkp * (749 + 512 + 32 + 4),  # patch reads + 512 blurred samples + descriptor
        "track": 52 * nkp + 48 * evals_per_frame + 4 * nkp,   # grid build + (32-B row + 16-B kp4) per evaluation + matches
    }


_JSON_FD = None


def emit(line):
    """The ONE JSON line goes to the process's original stdout; everything else written to fd 1 while the bench runs (e.g.
    the "NCCL version ..." banner NCCL prints on communicator creation) has been diverted to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="frames per batch (one extractor pass) per GPU")
    ap.add_argument("--batches-per-step", type=int, default=16,
                    help="batches per step: a step is 1024 frames (~12 ms on a B200) so that the driver's 20 steps time >= 0.2 s")
    ap.add_argument("--sets", type=int, default=8, help="distinct input batches (input working set > L2)")
    ap.add_argument("--cpu-frames", type=int, default=400, help="frames timed for cpu_baseline")
    ap.add_argument("--pipes", type=int, default=3, help="extractor/matcher handle pairs working on alternating batches (1 = serial)")
    ap.add_argument("--e2e-pipes", type=int, default=3, help="host threads (one handle pair each) of the end-to-end leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-kitti", action="store_true", help="skip the KITTI-shaped leg (config 3)")
    ap.add_argument("--no-knn", action="store_true", help="skip the 1M x 2000 k-NN leg (config 5)")
    ap.add_argument("--no-sparse", action="store_true", help="skip the sparse-texture leg")
    ap.add_argument("--no-latency", action="store_true", help="skip the one-frame-in-flight legs (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
