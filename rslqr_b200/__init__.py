"""rslqr_b200 -- B200-native rsLQR (nested-dissection LQR solve).

The product is the C library ``librslqr.so`` (C host API of the reference + sm_100a kernels behind
the C-ABI of ``include/rslqr_cuda.h``). This package is only the thin ctypes binding the tests and
``bench.py`` use; it contains no arithmetic and no fallback path: importing it without the built
library raises.
"""
from .lib import (  # noqa: F401
    LIB, LIB_PATH, RslqrError, CudaSolver, NdLqrSolver, load_problem_json, synth_problems,
    knot_stride, ShardSolver, shard_exchange, MODE_AUTO, MODE_STAGED, MODE_FUSED, MODE_SCAN, KEEP_FACT, PROFILE, STAGE,
)
