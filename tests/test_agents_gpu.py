"""Config 4: N agents in ONE process, one host thread + one handle set per agent, bound to GPUs round-robin
(reference src/MultiAgentSystem.cc:56-66, src/Tracking.cc:600-606).  The agents are independent units: results of the
concurrent run must equal each agent's own serial run, and the oracle's."""
import numpy as np
import pytest

from mam3slam_b200 import synth


def test_agent_binding_is_round_robin_host_logic():
    from mam3slam_b200.agents import MultiAgentFrontEnd
    mas = MultiAgentFrontEnd(n_gpus=4)
    assert mas.n_gpus == 4 and mas.mvpAgents == []
    # the binding rule itself (no handle is created without a GPU: there is no CPU fallback)
    assert [i % mas.n_gpus for i in range(10)] == [0, 1, 2, 3, 0, 1, 2, 3, 0, 1]
    import torch
    if not torch.cuda.is_available():
        from mam3slam_b200 import NativeError
        with pytest.raises(NativeError):
            mas.addAgent()


@pytest.mark.gpu
def test_agents_concurrent_equals_serial_and_oracle():
    import oracle
    from oracle import matcher as om
    from mam3slam_b200.agents import AgentFrontEnd, MultiAgentFrontEnd
    from tests.matcher_scenarios import to_kp
    n_agents, n_frames = 6, 4
    streams = [synth.stream(752, 480, seed, n_frames) for seed in range(n_agents)]
    mas = MultiAgentFrontEnd()
    for _ in range(n_agents):
        mas.addAgent(with_ini_extractor=True)
    assert [a.device for a in mas.mvpAgents] == [i % mas.n_gpus for i in range(n_agents)]
    for rep in range(2):                                   # twice: handles are reused across calls
        conc = mas.run(streams)
        for a in mas.mvpAgents:
            a._last = None
    mas.Shutdown()
    for i in (0, n_agents - 1):
        solo = AgentFrontEnd(i, 0)
        for t, f in enumerate(streams[i]):
            k, d, nm, km = solo.GrabImageMonocular(f)
            ck, cd, cnm, ckm = conc[i][t]
            assert np.array_equal(k, ck) and np.array_equal(d, cd) and nm == cnm and np.array_equal(km, ckm)
        solo.close()
    # agent 1 against the oracle, frame by frame
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    tabs = ex.tables()
    last = None
    for t, f in enumerate(streams[1]):
        _, k, d = ex(f)
        k = to_kp(k)
        ck, cd, cnm, ckm = conc[1][t]
        assert np.array_equal(ck["x"], k["x"]) and np.array_equal(ck["y"], k["y"]) and np.array_equal(cd, d)
        if last is not None:
            lk, ld = last
            onm, okm = om.search_frame(k, d, (0, 0, 752, 480), np.zeros(len(k), np.uint8), tabs["scale"], np.ones(len(lk), np.uint8),
                                       lk["x"], lk["y"], lk["octave"], lk["angle"], ld, 15.0, 100, True)
            assert cnm == onm and np.array_equal(ckm, okm) and onm > 100
        last = (k, d)


@pytest.mark.gpu
def test_device_resident_tracking_on_every_device_of_one_process():
    """One process driving several GPUs (agent i -> GPU i mod G): the opt-in shared-memory limits of the search kernels are set
    per device, so the device-resident tracking step must work - and give the same matches - on every device, not only on the
    one that happened to be current when the first handle was made.  Needs >= 2 GPUs (skipped on a one-GPU box)."""
    import torch
    from mam3slam_b200 import ORBextractor, ORBmatcher
    from mam3slam_b200._lib import device_count
    if device_count() < 2:
        pytest.skip("one GPU")
    B, W, H = 8, 752, 480
    frames = np.stack(synth.stream(W, H, 3, B))
    ref = None
    for dev_id in range(min(device_count(), 4)):
        with torch.cuda.device(dev_id):
            dev = torch.from_numpy(frames).cuda()
            ex = ORBextractor(1000, 1.2, 8, 20, 7, device=dev_id, max_batch=B)
            mt = ORBmatcher(0.9, True, device=dev_id)
            ts = torch.cuda.Stream()
            ex.extract_device(dev.data_ptr(), W * H, B, W, H, W, (0, 1000), ts.cuda_stream)
            dk, dd, dc, stride = ex.results_batch_device()
            kpm = torch.empty((B, stride), dtype=torch.int32, device="cuda")
            nmt = torch.empty(B, dtype=torch.int32, device="cuda")
            mt.track_batch_device(B, dk, dd, dc, stride, (0., 0., float(W), float(H)), ex.GetScaleFactors(), 15.0, kpm.data_ptr(),
                                  nmt.data_ptr(), stream=ts.cuda_stream)
            ts.synchronize()
            got = (nmt.cpu().numpy().copy(), kpm.cpu().numpy().copy())
            mt.close(); ex.close()
        if ref is None:
            ref = got
            assert got[0][1:].min() > 100
        else:
            assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]), f"device {dev_id}"
