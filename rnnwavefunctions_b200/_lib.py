"""ctypes binding of ``librnnwf.so`` (declared in ``include/rnnwf.h``).

The shared library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no
Python / CPU fallback: if the library is missing, importing this module raises, and every compute
entry point fails loudly without a CUDA device.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RNNWF_LIB_PATH selects another build of the same library (e.g. the cycle-counter variant used by tools/tc_cycles.py)
LIB_PATH = os.environ.get("RNNWF_LIB_PATH") or os.path.join(_HERE, "librnnwf.so")

CELL_GRU1D, CELL_VANILLA2D, CELL_TENS2D, CELL_TENSGRU2D, CELL_GRU4 = range(5)
HAM_TFIM1D, HAM_TFIM2D, HAM_J1J2, HAM_HEIS2D, HAM_RYDBERG = range(5)


class Config(C.Structure):
    _fields_ = [("cell", C.c_int32), ("complex_head", C.c_int32), ("nx", C.c_int32), ("ny", C.c_int32),
                ("num_units", C.c_int32), ("num_layers", C.c_int32), ("u1", C.c_int32), ("magnetization", C.c_int32),
                ("reserved", C.c_int32 * 8)]


def _load():
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp, i64, u64, i32 = C.c_void_p, C.c_int64, C.c_uint64, C.c_int32
    sig = {
        "rnnwf_last_error": (C.c_char_p, []),
        "rnnwf_version": (C.c_int, []),
        "rnnwf_create": (C.c_int, [C.POINTER(Config), C.POINTER(vp)]),
        "rnnwf_destroy": (C.c_int, [vp]),
        "rnnwf_count_params": (i64, [C.POINTER(Config)]),
        "rnnwf_num_params": (i64, [vp]),
        "rnnwf_num_tensors": (C.c_int, [vp]),
        "rnnwf_param_info": (C.c_int, [vp, C.c_int, C.c_char_p, C.c_int, C.POINTER(i64), C.POINTER(i32), C.POINTER(i64)]),
        "rnnwf_set_weights": (C.c_int, [vp, vp, vp]),
        "rnnwf_workspace_bytes": (i64, [vp]),
        "rnnwf_saved_bytes": (i64, [vp, i64]),
        "rnnwf_sample": (C.c_int, [vp, i64, vp, u64, u64, vp, vp, vp, vp, vp, i64, vp]),
        "rnnwf_logpsi": (C.c_int, [vp, i64, vp, vp, vp, vp, vp, i64, vp]),
        "rnnwf_eloc_workspace_bytes": (i64, [vp, C.c_int, i64]),
        "rnnwf_local_energies": (C.c_int, [vp, C.c_int, C.POINTER(C.c_double), C.c_int, i64, vp, vp, vp, vp, vp, vp, vp, i64, vp]),
        "rnnwf_backward_workspace_bytes": (i64, [vp, i64]),
        "rnnwf_backward": (C.c_int, [vp, i64, vp, vp, vp, vp, vp, vp, i64, vp]),
        "rnnwf_logpsi_static": (C.c_int, [vp, i64, vp, vp, vp, vp]),
        "rnnwf_adam_step": (C.c_int, [vp, vp, vp, vp, i64, C.c_double, C.c_double, C.c_double, C.c_double, i64, C.c_double, vp]),
        "rnnwf_sign_sgd_step": (C.c_int, [vp, vp, i64, C.c_double, u64, i64, vp]),
        "rnnwf_energy_moments": (C.c_int, [i64, vp, vp, vp, vp]),
        "rnnwf_cost_seeds": (C.c_int, [i64, vp, vp, vp, C.c_int, C.c_double, C.c_int, vp, vp, vp, vp]),
        "rnnwf_symmetry_epilogue": (C.c_int, [C.c_int, i64, vp, vp, C.POINTER(C.c_double), vp, vp, vp]),
        "rnnwf_symmetry_seeds": (C.c_int, [C.c_int, i64, vp, vp, C.POINTER(C.c_double), vp, vp, vp, vp, vp]),
        "rnnwf_symmetry_create": (C.c_int, [vp, C.c_int, C.POINTER(i32), C.POINTER(C.c_int8), C.POINTER(C.c_double), C.POINTER(vp)]),
        "rnnwf_symmetry_destroy": (C.c_int, [vp]),
        "rnnwf_symmetry_order": (C.c_int, [vp]),
        "rnnwf_symmetry_apply": (C.c_int, [vp, i64, vp, vp, vp]),
        "rnnwf_eloc_sym_workspace_bytes": (i64, [vp, vp, C.c_int, i64]),
        "rnnwf_local_energies_sym": (C.c_int, [vp, vp, C.c_int, C.POINTER(C.c_double), C.c_int, i64, vp, vp, vp, vp, vp, vp, i64, vp]),
        "rnnwf_describe": (C.c_int, [vp, C.c_char_p, C.c_int]),
        "rnnwf_launch_count": (C.c_longlong, []),
        "rnnwf_ffma_probe": (C.c_int, [vp, C.c_int, C.c_int, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


class RnnwfError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != 0:
        msg = lib.rnnwf_last_error().decode(errors="replace")
        if rc == -1:
            raise ValueError(f"{what}: {msg}")
        raise RnnwfError(f"{what} failed (code {rc}): {msg}")
