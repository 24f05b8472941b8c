"""orb_slam2_b200 — B200 (sm_100a) implementation of ORB-SLAM2's per-frame feature front end.

Only the hot path is here: ORBextractor::operator(), Frame::ComputeStereoMatches and the
ORBmatcher Hamming searches, behind the C ABI in include/orbx.h (csrc/ builds liborbx.so).
"""
from .api import (KP_DTYPE, ComputeDistinctiveDescriptors, ComputeStereoMatches, Frame, FrameGrids, ORBVocabulary, ORBextractor, ORBmatcher, OrbxError,  # noqa: F401
                  kernel_launch_count, lib, library_path)
