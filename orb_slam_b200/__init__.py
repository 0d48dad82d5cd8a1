"""orb_slam_b200 — B200-native (sm_100a) ORB feature front-end: extraction + Hamming matching.

Product surface: the C ABI in include/orbx.h / include/orbm.h (liborbx.so) and these thin host mirrors of the
reference classes.  Nothing here imports oracle/ and nothing falls back to a CPU path.
"""
from .extractor import KP_DTYPE, ORBextractor, OrbxError, synth_frames_device  # noqa: F401
from .matcher import (BowDatabase, Camera, FrameSet, FrustumView, Matcher, image_bounds, MatchFrame, ORBmatcher, OrbmError, ORBVocabulary,
                      ProjectedPoints, ShardedKnn2,
                      shard_range)  # noqa: F401,E402
