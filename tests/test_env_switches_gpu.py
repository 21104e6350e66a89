"""GPU: the A/B switches of the extractor's launch schedule only move kernels between streams / launches;
every setting must produce byte-identical keypoints and descriptors (the non-default paths are otherwise never exercised)."""
import hashlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROG = r'''
import hashlib, sys, numpy as np
sys.path.insert(0, %r)
from mam3slam_b200 import ORBextractor, synth
h = hashlib.sha256()
frames = synth.stream(752, 480, 3, 6)
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=6)
for f in frames[:2]:
    mono, k, d = ex(f)
    h.update(np.int32(mono).tobytes()); h.update(k.tobytes()); h.update(d.tobytes())
res = ex.extract_batch(list(frames))
for mono, k, d in res:
    h.update(np.int32(mono).tobytes()); h.update(k.tobytes()); h.update(d.tobytes())
small = ORBextractor(300, 1.2, 5, 20, 7)
mono, k, d = small(synth.frame(333, 257, 1))
h.update(k.tobytes()); h.update(d.tobytes())
mono, k, d = ex(synth.frame(752, 480, 9, sparse=True))          # sparse texture: cells decided by the minThFAST fallback
h.update(np.int32(mono).tobytes()); h.update(k.tobytes()); h.update(d.tobytes())
print(h.hexdigest())
''' % ROOT


def _digest(env_extra):
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-c", PROG], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.strip().splitlines()[-1]


@pytest.mark.gpu
def test_schedule_switches_do_not_change_results():
    base = _digest({})
    for env in ({"ORBX_FAST_GROUPS": "0"}, {"ORBX_BLUR_EARLY": "0"}, {"ORBX_FAST_EARLY": "1"}, {"ORBX_PRIO": "0"},
                {"ORBX_GRAPHS": "0"}, {"ORBX_OVERLAP": "0"}, {"ORBX_FAST_EARLY": "1", "ORBX_BLUR_EARLY": "0", "ORBX_GRAPHS": "0"},
                # FAST: first-generation per-cell kernel, compass early-out of the strength pass, plain-load tile staging (no TMA)
                {"ORBX_FAST_GEN": "1"}, {"ORBX_FAST_GEN": "1", "ORBX_FAST_GROUPS": "0"}, {"ORBX_FAST_COMPASS": "1"},
                {"ORBX_FAST_TMA": "0"}, {"ORBX_FAST_TMA": "0", "ORBX_FAST_COMPASS": "1", "ORBX_GRAPHS": "0"},
                # blur: register-only kernel instead of the TMA-staged one; quadtree CTAs capped per SM
                {"ORBX_BLUR_TMA": "0"}, {"ORBX_BLUR_TMA": "0", "ORBX_OVERLAP": "0"}, {"ORBX_TREE_CTAS": "1"},
                # quadtree in its partitioning form only (the default tries the count pyramid first)
                {"ORBX_TREE_PYRAMID": "0"},
                # quadtree CTA sizes (launch parameter: 512 threads for few frames, 128 for full batches by default)
                {"ORBX_TREE_THREADS_SMALL": "256"}, {"ORBX_TREE_THREADS_SMALL": "128", "ORBX_TREE_PYRAMID": "0"},
                # orientation + BRIEF with global gathers instead of TMA-staged windows
                {"ORBX_DESC_TMA": "0"}, {"ORBX_DESC_TMA": "0", "ORBX_BLUR_TMA": "0", "ORBX_FAST_GEN": "1"},
                # pyramid levels with direct global loads instead of TMA-staged source tiles; everything pre-TMA at once
                {"ORBX_PYR_TMA": "0"},
                # one launch per pyramid level even for few frames (the default chains all levels of <= ~4 frames into one launch)
                {"ORBX_PYR_CHAIN": "0"}, {"ORBX_PYR_CHAIN": "0", "ORBX_GRAPHS": "0"},
                {"ORBX_PYR_TMA": "0", "ORBX_DESC_TMA": "0", "ORBX_BLUR_TMA": "0", "ORBX_FAST_TMA": "0", "ORBX_GRAPHS": "0"}):
        assert _digest(env) == base, env


# ---- matcher: the two resolvers of the stateful window searches -----------------------------------------------------------
# Default: orbm_search_frame / orbm_search_sim3 resolve in claim rounds (orbm_kernels_rounds.cuh), the batch tracking step
# in query order on one CTA per pair.  The switches swap them; results must not move (and the defaults are checked against the
# oracle in tests/test_matcher_gpu.py).
MATCH_PROG = r'''
import hashlib, sys, numpy as np, torch
sys.path.insert(0, %r)
from mam3slam_b200 import ORBextractor, ORBmatcher, synth
from tests import matcher_scenarios as sc
h = hashlib.sha256()
M = ORBmatcher(0.9, True)
bounds = (0.0, 0.0, 752.0, 480.0)
scale = (1.2 ** np.arange(8)).astype(np.float32)
# (a) two consecutive frames, tracking windows, with and without the rotation histogram
(k1, d1), (k2, d2), _, _ = sc.frame_pair(752, 480, 1000, 0)
F = M.frame(k2, d2, bounds)
for th, ori in ((15.0, True), (30.0, False), (3.0, True)):
    M.mbCheckOrientation = ori
    nm, km = M.SearchByProjection_Frame(F, np.zeros(len(k2), np.uint8), scale, np.ones(len(k1), np.uint8), k1["x"], k1["y"],
                                        k1["octave"], k1["angle"], d1, th, 100)
    h.update(np.int32(nm).tobytes()); h.update(km.tobytes())
F.close()
# (b) contention: 900 queries over one cluster of 600 near-identical keypoints, some pre-occupied, some non-occupying
rng = np.random.default_rng(5)
n2, nq = 600, 900
k = np.zeros(n2, sc.KP)
k["x"] = 300 + rng.random(n2).astype(np.float32) * 60; k["y"] = 200 + rng.random(n2).astype(np.float32) * 60
k["octave"] = rng.integers(0, 3, n2); k["angle"] = rng.random(n2).astype(np.float32) * 360
base = rng.integers(0, 256, 32, dtype=np.uint8)
d = np.repeat(base[None], n2, 0); fl = rng.integers(0, 256, (n2, 6))
for j in range(6):
    d[np.arange(n2), fl[:, j] >> 3] ^= (1 << (fl[:, j] & 7)).astype(np.uint8)
qd = np.repeat(base[None], nq, 0); f2 = rng.integers(0, 256, (nq, 3))
for j in range(3):
    qd[np.arange(nq), f2[:, j] >> 3] ^= (1 << (f2[:, j] & 7)).astype(np.uint8)
u = (300 + rng.random(nq) * 60).astype(np.float32); v = (200 + rng.random(nq) * 60).astype(np.float32)
octave = rng.integers(0, 3, nq).astype(np.int32); angle = (rng.random(nq) * 360).astype(np.float32)
F = M.frame(k, d, bounds)
for th in (15.0, 40.0):
    for ori in (False, True):
        M.mbCheckOrientation = ori
        occ = (rng.random(n2) < 0.05).astype(np.uint8)
        for qocc in (None, (rng.random(nq) < 0.6).astype(np.uint8)):
            nm, km = M.SearchByProjection_Frame(F, occ, scale, np.ones(nq, np.uint8), u, v, octave, angle, qd, th, 100, q_occupies=qocc)
            h.update(np.int32(nm).tobytes()); h.update(km.tobytes())
F.close()
# (c) the batch tracking step over device-resident extractor results
nf = 5
frames = synth.stream(752, 480, 3, nf)
ex = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=nf)
dev = torch.from_numpy(frames).cuda()
ex.extract_device(dev.data_ptr(), 752 * 480, nf, 752, 480, 752)
ex.synchronize()
dk, dd, dc, stride = ex.results_batch_device()
kpm = torch.empty((nf, stride), dtype=torch.int32, device="cuda"); nmt = torch.empty(nf, dtype=torch.int32, device="cuda")
for ori in (True, False):
    M.mbCheckOrientation = ori
    for rep in range(2):                 # twice: the per-pair tickets of the claim-rounds launch must come back to zero
        M.track_batch_device(nf, dk, dd, dc, stride, bounds, ex.GetScaleFactors(), 15.0, kpm.data_ptr(), nmt.data_ptr())
        torch.cuda.synchronize()
        h.update(kpm.cpu().numpy().tobytes()); h.update(nmt.cpu().numpy().tobytes())
print(h.hexdigest())
''' % ROOT


def _match_digest(env_extra):
    env = dict(os.environ, **env_extra)
    r = subprocess.run([sys.executable, "-c", MATCH_PROG], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.strip().splitlines()[-1]


@pytest.mark.gpu
def test_matcher_resolver_switches_do_not_change_results():
    base = _match_digest({})
    for env in ({"ORBM_ROUNDS": "0"}, {"ORBM_TRACK_ROUNDS": "1"}, {"ORBM_ROUNDS": "0", "ORBM_TRACK_ROUNDS": "1"},
                {"ORBM_TRACK_STAGE": "0"}):
        assert _match_digest(env) == base, env
