"""The native multi-agent driver over the C ABI (tools/agents_native.cc -> mam3slam_b200/agents_native, built by
mam3slam_b200/csrc/Makefile): it links against the library with nothing but include/orbx.h and include/orbm.h, which is what a
C++ host (the reference's Tracking threads) would do."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "mam3slam_b200", "agents_native")


def _built():
    if not os.path.exists(EXE):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "mam3slam_b200", "csrc")])
    return EXE


def test_driver_builds_and_explains_itself():
    r = subprocess.run([_built()], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_agents_share_one_gpu_and_track(tmp_path):
    """Three agents x one frame in flight on device 0: every agent extracts ~1000 keypoints per frame and tracks most of them
    against its previous frame; the same stream through one agent gives the same match count per frame (agents are independent)."""
    from mam3slam_b200 import synth
    frames = np.stack(synth.stream(752, 480, 0, 8)).astype(np.uint8)
    path = tmp_path / "frames.u8"
    frames.tofile(path)

    def run(counts, track=1, pin=0):
        r = subprocess.run([_built(), str(path), "752", "480", "8", "0", "20", counts, str(track), str(pin)], capture_output=True, text=True,
                           timeout=300)
        assert r.returncode == 0, r.stderr[-500:]
        return json.loads(r.stdout.strip().splitlines()[-1])["rows"]

    rows = run("1,3")
    assert [r["agents"] for r in rows] == [1, 3]
    for r in rows:
        assert r["frames_per_s"] > 100 and r["median_frame_latency_ms"] < 20
        assert r["matches_per_frame"] > 300          # consecutive frames of the synthetic stream share most keypoints
        assert r["gpu_launches"] > 0
    only = run("2", track=0)
    assert only[0]["matches_per_frame"] == 0 and only[0]["frames_per_s"] > 100
    # frames page-locked once with orbx_pin_host (uploaded in place instead of through the staging buffer): same matches
    pinned = run("1,3", pin=1)
    assert [r["matches_per_frame"] for r in pinned] == [r["matches_per_frame"] for r in rows]
