"""ctypes binding of the C++ scene compiler (include/t4c_compiler.h, csrc/compiler/t4_scene_compiler.cpp):
URDF / SRDF / STL / PLY / octomap .bt -> t4c_scene_desc + t4c_fk_model_desc, without any Python-side geometry work."""
import ctypes as C
import os
import subprocess

import numpy as np

from . import capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libt4c_compiler.so")
SRC = os.path.join(HERE, "csrc", "compiler", "t4_scene_compiler.cpp")
_lib = None


def build(force=False):
    hdr = os.path.join(HERE, "..", "include", "t4c_compiler.h")
    stale = force or not os.path.exists(LIB_PATH) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH)
                                                         for s in (SRC, hdr))
    if stale:
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", SRC, "-o", LIB_PATH])
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        VP, I32, I64, DP = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_double)
        L.t4c_compiler_last_error.restype = C.c_char_p
        L.t4c_compile_urdf.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, I32, C.c_double, C.POINTER(VP)]
        L.t4c_compiled_add_octomap.argtypes = [VP, C.c_char_p, C.c_char_p, I32, DP]
        L.t4c_compiled_add_object.argtypes = [VP, C.c_char_p, I32, C.POINTER(capi.Shape), I32, DP, DP]
        L.t4c_compiled_scene.argtypes = [VP]
        L.t4c_compiled_scene.restype = C.POINTER(capi.SceneDesc)
        L.t4c_compiled_fk.argtypes = [VP]
        L.t4c_compiled_fk.restype = C.POINTER(capi.FkModelDesc)
        L.t4c_compiled_link_index.argtypes = [VP, C.c_char_p]
        L.t4c_compiled_joint_name.argtypes = [VP, I32]
        L.t4c_compiled_joint_name.restype = C.c_char_p
        L.t4c_compiled_limits.argtypes = [VP, DP, DP]
        L.t4c_compiled_destroy.argtypes = [VP]
        L.t4c_compiled_destroy.restype = None
        L.t4c_convex_hull_vertices.argtypes = [DP, I32, C.POINTER(I32), C.POINTER(I32)]
        L.t4c_read_octomap_bt.argtypes = [C.c_char_p, I64, DP, C.POINTER(I64)]
        _lib = L
    return _lib


class CompilerError(RuntimeError):
    pass


def _check(st):
    if st != 0:
        raise CompilerError(lib().t4c_compiler_last_error().decode())


class CompiledScene:
    """owner of a t4c_compiled; .scene_desc() / .fk_desc() are what t4c_scene_create / t4c_fk_model_create take"""

    def __init__(self, urdf, srdf=None, package_root="", convex=True, margin=0.0):
        self.L = lib()
        self.h = C.c_void_p()
        _check(self.L.t4c_compile_urdf(urdf.encode(), srdf.encode() if srdf else None, package_root.encode(),
                                       1 if convex else 0, float(margin), C.byref(self.h)))

    def __del__(self):
        try:
            self.L.t4c_compiled_destroy(self.h)
        except Exception:
            pass

    def scene_desc(self):
        return self.L.t4c_compiled_scene(self.h).contents

    def fk_desc(self):
        return self.L.t4c_compiled_fk(self.h).contents

    def link_index(self, name):
        return int(self.L.t4c_compiled_link_index(self.h, name.encode()))

    def limits(self):
        nq = self.fk_desc().n_q
        lo, hi = np.zeros(nq), np.zeros(nq)
        dp = C.POINTER(C.c_double)
        _check(self.L.t4c_compiled_limits(self.h, lo.ctypes.data_as(dp), hi.ctypes.data_as(dp)))
        return lo, hi

    def add_octomap(self, link, bt_path, sub_type=0, pose=None):
        p = None if pose is None else np.ascontiguousarray(pose, dtype=np.float64).reshape(12)
        _check(self.L.t4c_compiled_add_octomap(self.h, link.encode(), bt_path.encode(), int(sub_type),
                                               None if p is None else p.ctypes.data_as(C.POINTER(C.c_double))))

    def add_object(self, link, shapes, vertices=None, pose=None):
        """shapes: list of dicts as in tesseract_b200.scene.SceneBuilder.shapes (link ignored)"""
        arr = (capi.Shape * max(len(shapes), 1))()
        for i, s in enumerate(shapes):
            arr[i].type, arr[i].shape_id, arr[i].subshape_id = s["type"], s["shape_id"], s["subshape_id"]
            arr[i].params[:] = [float(x) for x in s["params"]]
            arr[i].vertex_offset, arr[i].vertex_count = s.get("vertex_offset", 0), s.get("vertex_count", 0)
            arr[i].pose[:] = np.asarray(s["pose"], dtype=np.float64).reshape(-1).tolist()
        v = np.ascontiguousarray(vertices if vertices is not None else np.zeros((0, 3)), dtype=np.float64)
        p = None if pose is None else np.ascontiguousarray(pose, dtype=np.float64).reshape(12)
        dp = C.POINTER(C.c_double)
        _check(self.L.t4c_compiled_add_object(self.h, link.encode(), len(shapes), arr, len(v), v.ctypes.data_as(dp),
                                              None if p is None else p.ctypes.data_as(dp)))


def convex_hull_vertices(points):
    pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
    idx = np.zeros(len(pts), dtype=np.int32)
    n = C.c_int32()
    _check(lib().t4c_convex_hull_vertices(pts.ctypes.data_as(C.POINTER(C.c_double)), len(pts),
                                          idx.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(n)))
    return idx[:n.value]


def read_octomap_bt(path):
    n = C.c_int64()
    _check(lib().t4c_read_octomap_bt(path.encode(), 0, None, C.byref(n)))
    out = np.zeros((n.value, 4))
    _check(lib().t4c_read_octomap_bt(path.encode(), n.value, out.ctypes.data_as(C.POINTER(C.c_double)), C.byref(n)))
    return out
