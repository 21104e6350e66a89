"""Host-side mirror of how MAM3SLAM spreads its front end over agents (reference src/MultiAgentSystem.cc:56-66 addAgent,
src/Agent.cc:41-46, src/Tracking.cc:600-606): every agent owns a Tracking thread with two extractor objects
(mpORBextractorLeft, and mpIniORBextractor with 5x the features until the map is initialised) and creates matcher objects
on the fly.  Here an agent's extractors / matcher are CUDA handles bound to ONE GPU (agent i -> device i mod n_gpus,
SURVEY.md 8e: independent units, no data-path collective) and every agent is driven by its own host thread, the way the
reference runs one Tracking thread per agent inside one process."""
import threading

import numpy as np

from .extractor import ORBextractor
from .matcher import ORBmatcher
from ._lib import device_count


class AgentFrontEnd:
    """One agent's per-frame hot path: ORBextractor::operator() + SearchByProjection(Cur, Last, th, mono)."""

    def __init__(self, agent_id, device, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7,
                 with_ini_extractor=False):
        self.agent_id, self.device = agent_id, device
        self.mpORBextractorLeft = ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=device)
        self.mpIniORBextractor = (ORBextractor(5 * nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=device)
                                  if with_ini_extractor else None)
        self.matcher = ORBmatcher(0.9, True, device=device)            # TrackWithMotionModel (src/Tracking.cc:2789)
        self._scale = self.mpORBextractorLeft.GetScaleFactors()
        self._last = None
        self._frame = None

    def GrabImageMonocular(self, im, th=15.0, initialised=True):
        """Returns (keypoints, descriptors, nmatches, kpMatch) for this frame; kpMatch[i] = index of the LAST frame's
        keypoint matched to keypoint i, -1 = none (the reference's CurrentFrame.mvpMapPoints)."""
        ex = self.mpORBextractorLeft if (initialised or self.mpIniORBextractor is None) else self.mpIniORBextractor
        _, kps, desc = ex(im)
        h, w = im.shape
        bounds = (0.0, 0.0, float(w), float(h))
        nm, km = 0, np.full(len(kps), -1, np.int32)
        if self._last is not None and len(kps):
            lk, ld = self._last
            if self._frame is None:
                self._frame = self.matcher.frame(kps, desc, bounds)
            else:
                self._frame.update(kps, desc, bounds)
            nm, km = self.matcher.SearchByProjection_Frame(self._frame, np.zeros(len(kps), np.uint8), self._scale,
                                                           np.ones(len(lk), np.uint8), lk["x"], lk["y"], lk["octave"], lk["angle"],
                                                           ld, th)
        self._last = (kps, desc)
        return kps, desc, nm, km

    def close(self):
        if self._frame is not None:
            self._frame.close()
        self.matcher.close()
        self.mpORBextractorLeft.close()
        if self.mpIniORBextractor is not None:
            self.mpIniORBextractor.close()


class MultiAgentFrontEnd:
    """addAgent() binds agent i to GPU i mod n_gpus; run() drives every agent's frame stream on its own host thread."""

    def __init__(self, n_gpus=None):
        self.n_gpus = n_gpus if n_gpus else max(device_count(), 1)
        self.mvpAgents = []

    def addAgent(self, **extractor_settings):
        i = len(self.mvpAgents)
        self.mvpAgents.append(AgentFrontEnd(i, i % self.n_gpus, **extractor_settings))
        return self.mvpAgents[-1]

    def run(self, streams, th=15.0):
        """streams[i] = iterable of frames of agent i.  Returns per agent the list of GrabImageMonocular results."""
        assert len(streams) == len(self.mvpAgents)
        out = [None] * len(streams)
        err = []

        def work(i):
            try:
                out[i] = [self.mvpAgents[i].GrabImageMonocular(f, th) for f in streams[i]]
            except BaseException as e:          # surfaced to the caller below
                err.append(e)

        threads = [threading.Thread(target=work, args=(i,)) for i in range(len(streams))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if err:
            raise err[0]
        return out

    def Shutdown(self):
        for a in self.mvpAgents:
            a.close()
        self.mvpAgents = []
