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
