"""B200-native inner Levenberg-Marquardt iteration of VIBA behind a C ABI (include/viba_b200.h)."""
from . import graph, synth  # noqa: F401
