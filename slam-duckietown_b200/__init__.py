"""B200-native EKF-SLAM filter core (drop-in for the predict/update hot path of SLAM-Duckietown).

Host side = Python over a C-ABI shared library (``csrc/libekf_b200.so``, hand-written sm_100a CUDA).
There is no CPU fallback: anything that computes raises if the library or a CUDA device is missing.
"""
from . import capi  # noqa: F401  (ctypes loader; import never touches the GPU)
