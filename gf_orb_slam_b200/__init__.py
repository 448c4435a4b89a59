"""gf_orb_slam_b200 -- B200-native (sm_100a) tracking front end of GF-ORB-SLAM behind the reference's operator API.

The compute path is libgforb.so (hand-written CUDA, C ABI in include/gforb.h); this package is the thin host-side
mirror of ORB_SLAM::ORBextractor / ORBmatcher / Observability used by the tests and bench.py.  There is no CPU
fallback: without the built library and a CUDA device the operators raise."""
from ._lib import KP_DTYPE, GforbError, LIB_PATH  # noqa: F401
from .extractor import ORBextractor, FAST_SCORE, HARRIS_SCORE  # noqa: F401
from .matcher import ORBmatcher, Frames, MapPoints, WorldPoints  # noqa: F401
from .observability import Observability  # noqa: F401
from .optimizer import Optimizer  # noqa: F401
