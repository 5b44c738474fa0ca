"""ultragroth_b200 -- B200-native (sm_100a) Groth16 / UltraGroth proving backend.

The product is the C-ABI shared library `libultragroth_b200.so` (CUDA kernels + C++ host layer, sources in
csrc/, headers in ../include/).  This Python package is the harness-side binding used by the tests and
bench.py; it mirrors the reference's C API (src/prover.h) one-to-one.
"""
from ._lib import LIB_PATH, UltraGrothError, lib  # noqa: F401
from .device import Context  # noqa: F401
