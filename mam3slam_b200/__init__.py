"""mam3slam_b200 — B200-native ORB front end (extraction + Hamming matching) behind MAM3SLAM's ORBextractor /
ORBmatcher interfaces.  Compute lives in libmam3slam_b200.so (hand-written sm_100a CUDA behind a C ABI); this
package is the Python host-side mirror used by tests and bench.py."""
from .extractor import ORBextractor, KEYPOINT_DTYPE  # noqa: F401
from .matcher import ORBmatcher, MatchFrame, KnnDatabase, ORBVocabulary  # noqa: F401
from .agents import AgentFrontEnd, MultiAgentFrontEnd  # noqa: F401
from ._lib import NativeError, SO_PATH  # noqa: F401
