"""ctypes binding of lib3dba_b200.so (include/b3d.h).  The library is the product: if it is missing or fails to load this
module raises — there is no Python/CPU fallback for any entry point."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib3dba_b200.so")
_LIB = None

NJ, NS, NPF, MAXKP, NPD = 24, 10, 76, 17, 207
F32, F64 = 0, 1
OK = 0
FIT_POSE, FIT_POSE_SHAPE = 0, 1
FWD_POSEDIRS = 1

KP_DTYPE = np.dtype([("jid", "<i4"), ("pad", "<i4"), ("u", "<f8"), ("v", "<f8")])
ITER_DTYPE = np.dtype([("cost", "<f8"), ("cost_change", "<f8"), ("gradient_max_norm", "<f8"), ("step_norm", "<f8"),
                       ("relative_decrease", "<f8"), ("radius", "<f8"), ("flags", "<i4"), ("aux", "<i4")])


class ModelDesc(C.Structure):
    _fields_ = [("num_verts", C.c_int32), ("num_joints", C.c_int32), ("num_shape", C.c_int32), ("num_faces", C.c_int32),
                ("dtype", C.c_int32), ("reserved", C.c_int32), ("v_template", C.c_void_p), ("shapedirs", C.c_void_p),
                ("J_regressor", C.c_void_p), ("weights", C.c_void_p), ("kintree_row0", C.c_void_p), ("faces", C.c_void_p),
                ("posedirs", C.c_void_p)]


class FitOpts(C.Structure):
    _fields_ = [("mode", C.c_int32), ("max_iters", C.c_int32), ("beta_pose", C.c_double), ("beta_shape", C.c_double),
                ("lambda_temp", C.c_double), ("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("huber", C.c_double), ("scale_lo", C.c_double), ("scale_hi", C.c_double)]


class SeqOpts(C.Structure):
    _fields_ = [("skip", C.c_int32), ("wsize", C.c_int32), ("overlap", C.c_int32), ("it1", C.c_int32), ("it2", C.c_int32),
                ("num_shards", C.c_int32), ("beta_pose", C.c_double), ("beta_shape", C.c_double), ("lambda_temp", C.c_double),
                ("beta_shape_lock", C.c_double)]


class IcpSolveOpts(C.Structure):
    _fields_ = [("max_iterations", C.c_int32), ("flags", C.c_int32), ("function_tolerance", C.c_double),
                ("gradient_tolerance", C.c_double), ("parameter_tolerance", C.c_double), ("beta_pose", C.c_double),
                ("beta_shape", C.c_double)]


class B3DError(RuntimeError):
    def __init__(self, code: int, where: str):
        self.code = code
        super().__init__(f"{where}: status {code}: {last_error()}")


def build(force: bool = False) -> str:
    """Compile the CUDA library in-tree (nvcc, sm_100a).  Works without a GPU (cross-compilation)."""
    src_dir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir) if f.endswith((".cu", ".cuh", ".hpp"))]
    srcs.append(os.path.join(_HERE, "..", "include", "b3d.h"))
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        res = subprocess.run(["make", "-C", src_dir, "-j", "8"], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("building lib3dba_b200.so failed:\n" + res.stdout[-4000:] + res.stderr[-4000:])
    return LIB_PATH


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        _LIB = C.CDLL(LIB_PATH)
        _LIB.b3d_last_error.restype = C.c_char_p
        _LIB.b3d_version.restype = C.c_char_p
        _LIB.b3d_launch_count.restype = C.c_uint64
    return _LIB


def last_error() -> str:
    return lib().b3d_last_error().decode()


def check(rc: int, where: str):
    if rc != OK:
        raise B3DError(rc, where)


def ptr(a):
    """void* of a numpy array, a raw integer device address, or None."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return C.c_void_p(int(a))
    return a.ctypes.data_as(C.c_void_p)
