"""svmb200 -- B200-native RBF-kernel SMO trainer: Python-side handle on the C ABI (include/svmb200.h).

The product is libsvmb200.so (CUDA, sm_100a) and the C++ host layer under host/.  This package
only binds the C ABI with ctypes for the tests and bench.py, and holds the synthetic workload
generator.  There is no CPU implementation here: importing works anywhere, but creating a Solver
without the built library or without a B200 raises.
"""
from .binding import (LIB_PATH, Params, Solver, Stats, SvmB200Error, STOP_NAMES, build_library,  # noqa: F401
                      declared_symbols, load_library)
from . import binding, datagen  # noqa: F401
