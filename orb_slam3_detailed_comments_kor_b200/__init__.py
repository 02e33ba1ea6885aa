"""B200-native ORB front-end (extractor + Hamming matching) behind ORB-SLAM3's class surface.

Only the hot path named by BASELINE.json lives here: csrc/ (CUDA kernels + C ABI, built to liborbfe.so),
the ctypes host mirror of the reference classes (extractor.py, matcher.py), and synth/ (deterministic inputs)."""
from .extractor import ORBextractor  # noqa: F401
from .matcher import ORBmatcher, DescriptorDatabase  # noqa: F401
