"""rslqr_b200 -- B200-native rsLQR (nested-dissection LQR solve).

The product is the C library ``librslqr.so`` (C host API of the reference + sm_100a kernels behind
the C-ABI of ``include/rslqr_cuda.h``). This package is only the thin ctypes binding the tests and
``bench.py`` use; it contains no arithmetic and no fallback path: the first use of any name below loads the
library and raises if it has not been built. (The load is deferred to first use so that
``rslqr_b200.ctypes_api`` -- struct layouts only -- can be imported by the CPU reference arm without mapping the
product library.)
"""
_NAMES = ("LIB", "LIB_PATH", "RslqrError", "CudaSolver", "NdLqrSolver", "load_problem_json", "synth_problems",
          "knot_stride", "ShardSolver", "shard_exchange", "MODE_AUTO", "MODE_STAGED", "MODE_FUSED", "MODE_SCAN",
          "KEEP_FACT", "PROFILE", "STAGE")


def __getattr__(name):
    if name in _NAMES or name == "lib":
        import importlib
        mod = importlib.import_module(".lib", __name__)   # raises ImportError when librslqr.so is missing
        return mod if name == "lib" else getattr(mod, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def __dir__():
    return sorted(list(globals()) + list(_NAMES))
