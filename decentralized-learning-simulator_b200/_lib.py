"""ctypes binding of the C-ABI library `csrc/libdls_b200.so` (declared in include/dls_b200.h).

There is no CPU fallback: if the shared library is missing or a compute entry point is called
without a CUDA device, an exception is raised.
"""
import ctypes
import os
from ctypes import c_int, c_int32, c_int64, c_float, c_void_p, c_char_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libdls_b200.so")

ERRORS = {-1: "DLS_E_ARG (null pointer / negative size)",
          -2: "DLS_E_ALIGN (pointer or leading dimension not 16-byte aligned)",
          -3: "DLS_E_SHAPE (unsupported shape / model / precision)",
          -4: "DLS_E_WS (workspace too small)",
          -5: "DLS_E_TOO_BIG (request exceeds a compiled-in maximum)"}

# symbol -> (restype, argtypes); must list every function include/dls_b200.h declares
SIGNATURES = {
    "dls_version": (c_int, []),
    "dls_build_info": (c_char_p, []),
    "dls_launch_count": (c_int64, []),
    "dls_tune": (c_int, [c_int32, c_int32]),
    "dls_agg_variant_fits": (c_int, [c_int32, c_int32, c_int32, c_int32]),
    "dls_agg_spmm_f32": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int64, c_int32, c_void_p]),
    "dls_chunk_mean_f32": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_int32, c_int32, c_int32, c_void_p]),
    "dls_copy_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int32, c_int64,
                              c_void_p]),
    "dls_push_rows": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int32, c_int64, c_void_p]),
    "dls_broadcast_row": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int32, c_int64, c_void_p]),
    "dls_gradient_update": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int32, c_int64, c_float,
                                    c_float, c_void_p]),
    "dls_train_workspace_bytes": (c_int64, [c_int32, c_int32, c_int32]),
    "dls_train_wave": (c_int, [c_int32, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_int32,
                               c_float, c_float, c_float, c_void_p, c_int32, c_void_p, c_int64, c_void_p]),
    "dls_grad_wave": (c_int, [c_int32, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_int32, c_void_p, c_int64,
                              c_void_p]),
    "dls_eval_workspace_bytes": (c_int64, [c_int32, c_int32, c_int32]),
    "dls_eval_wave": (c_int, [c_int32, c_void_p, c_int64, c_void_p, c_int32, c_void_p, c_void_p, c_int32,
                              c_void_p, c_void_p, c_void_p, c_int32, c_void_p, c_int64, c_void_p]),
    "dls_tc_probe": (c_int, [c_void_p, c_int32, c_void_p, c_int32, c_void_p, c_void_p, c_int32, ctypes.c_uint32,
                             c_int32, c_void_p, c_int32, c_void_p, c_void_p]),
    "dls_tc_contention": (c_int, [c_int32, c_int32, c_int32, c_void_p, c_void_p, c_void_p]),
    "dls_eval_wave_noalloc": (c_int, [c_int32, c_void_p, c_int64, c_void_p, c_int32, c_void_p, c_void_p, c_int32,
                                      c_void_p, c_void_p, c_void_p]),
}

_lib = None


class DlsError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile the CUDA library in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
    import subprocess
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if not verbose:
        cmd.append("-s")
    subprocess.check_call(cmd)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DlsError("CUDA extension %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc: int, what: str):
    if rc == 0:
        return
    if rc < 0:
        raise DlsError("%s failed: %s" % (what, ERRORS.get(rc, rc)))
    raise DlsError("%s failed: cudaError %d" % (what, rc))


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else t.data_ptr()
