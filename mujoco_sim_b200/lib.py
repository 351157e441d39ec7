"""ctypes binding of libmjsim_b200.so (include/mjsim_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is
visible, constructing a simulator raises.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

from .model.cstruct import MjsEnvState, MjsModel, MjsTask

_PKG = Path(__file__).resolve().parent
# MJS_LIB: tuning experiments load an alternative build of the same library
LIB_PATH = Path(os.environ["MJS_LIB"]).resolve() if os.environ.get("MJS_LIB") else _PKG / "libmjsim_b200.so"

PRECISION_F64 = 0
PRECISION_MIXED = 1


class MjsimError(RuntimeError):
    pass


def build(force: bool = False, verbose: bool = False):
    """Compile libmjsim_b200.so in-tree with nvcc for sm_100a (no GPU needed)."""
    args = ["make", "-j6", "-C", str(_PKG / "csrc")]
    if not verbose:
        args.append("-s")
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return LIB_PATH


_lib = None

_SIGS = {
    "mjs_create": (ctypes.c_int, [ctypes.POINTER(MjsModel), ctypes.POINTER(MjsTask), ctypes.c_int, ctypes.c_int, ctypes.c_ulonglong, ctypes.c_uint, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]),
    "mjs_destroy": (None, [ctypes.c_void_p]),
    "mjs_reset": (ctypes.c_int, [ctypes.c_void_p] * 4),
    "mjs_step": (ctypes.c_int, [ctypes.c_void_p] * 6 + [ctypes.c_int, ctypes.c_void_p]),
    "mjs_status": (ctypes.c_int, [ctypes.c_void_p] * 3),
    "mjs_step_noalloc": (ctypes.c_int, [ctypes.c_void_p] * 6 + [ctypes.c_int, ctypes.c_void_p]),
    "mjs_reseed": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_ulonglong]),
    "mjs_set_output64": (ctypes.c_int, [ctypes.c_void_p] * 4),
    "mjs_render": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "mjs_step_host": (ctypes.c_int, [ctypes.c_void_p] * 6 + [ctypes.c_int]),
    "mjs_reset_host": (ctypes.c_int, [ctypes.c_void_p] * 3),
    "mjs_set_state": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(MjsEnvState), ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "mjs_get_state": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(MjsEnvState), ctypes.c_int, ctypes.c_int]),
    "mjs_stats_device": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "mjs_stats": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.c_int]),
    "mjs_join_aux": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "mjs_debug_forward": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double), ctypes.c_int]),
    "mjs_debug_substeps": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "mjs_demo_policy": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "mjs_debug_times": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_longlong), ctypes.c_int]),
    "mjs_launch_count": (ctypes.c_longlong, [ctypes.c_void_p]),
    "mjs_num_envs": (ctypes.c_int, [ctypes.c_void_p]),
    "mjs_obs_dim": (ctypes.c_int, [ctypes.c_void_p]),
    "mjs_action_dim": (ctypes.c_int, [ctypes.c_void_p]),
    "mjs_last_error": (ctypes.c_char_p, []),
    "mjs_version": (ctypes.c_char_p, []),
}


def exported_symbols():
    return list(_SIGS)


def load():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise MjsimError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (no CPU fallback exists)")
        lib = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int):
    if rc != 0:
        raise MjsimError(load().mjs_last_error().decode())
