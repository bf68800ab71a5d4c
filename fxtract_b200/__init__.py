"""fxtract_b200 -- B200-native implementation of fxtract's per-record search path.

The product is `libfxgpu.so` (hand-written sm_100a CUDA behind the C ABI in
include/fxgpu.h) and the `fxtract` host CLI (fxtract_b200/host, C++).  This
Python package is only a ctypes binding for tests and bench.py; it never computes
anything itself and raises if the CUDA library is missing (no CPU fallback).
"""
from .lib import (FxgError, Context, Config, SynthSpec, load_library, find_record_cut,  # noqa: F401
                  MODE_LITERAL, MODE_MULTI, MODE_EXACT, TARGET_SEQ, TARGET_NAME, TARGET_COMMENT, TARGET_QUAL)
