"""acmmp_b200 - B200-native (sm_100a) implementation of ACMMP's per-view PatchMatch
MVS hot path.  The product is the CUDA library `libacmmp_b200.so` behind the C ABI
of include/acmmp_b200.h; this package is the Python host-side mirror of the
reference's operator interface (class ACMMP, ProcessProblem, RunJBU ...) used by
tests and bench.py.  There is no CPU fallback: importing `acmmp_b200.lib` fails
loudly when the CUDA extension has not been built."""
__version__ = "0.1"
