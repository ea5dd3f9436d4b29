"""sfm_b200 -- B200-native bundle adjustment / nonlinear refinement behind the reference's
BaProcessor / TriangulationProcessor / CamposeProcessor class APIs.

Hot path: hand-written sm_100a CUDA kernels in libsfmba.so, reached through the C ABI of
include/sfmba.h via ctypes.  No CPU fallback: every compute call raises if the library or the GPU
is missing.
"""
from ._capi import SfmbaError, load_library, LIB_PATH                     # noqa: F401
from .engine import Engine, BAProblem, SolveResult, get_engine              # noqa: F401
from .utils import (KeyPt, TriPt, convert_rotation_to_quaternion,          # noqa: F401
                    convert_quaternion_to_rotation, verify_rotation_mat, LMConfig, RansacConfig)
from .index import build_visibility_host, problem_from_tracks               # noqa: F401
from .triangulation_processor import TriangulationProcessor                 # noqa: F401
from .campose_processor import CamposeProcessor                             # noqa: F401
from .epipolar_processor import EpipolarProcessor                           # noqa: F401
from .ba_processor import BaProcessor                                       # noqa: F401

__version__ = '0.1.0'
