"""ctypes mirror of include/t4c.h (the C ABI) + loader of the in-tree CUDA library.

The product path fails loudly when the CUDA extension is missing: there is no CPU
fallback anywhere in this package (the CPU oracle lives under oracle/ and is only
ever loaded by tests/, smoke() and bench.py's cpu_baseline / --impl reference legs).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libt4c.so")

T4C_OK = 0
T4C_ERR_INVALID, T4C_ERR_CUDA, T4C_ERR_CAPACITY, T4C_ERR_UNSUPPORTED, T4C_ERR_NOMEM = -1, -2, -3, -4, -5
SPHERE, BOX, CYLINDER, CONE, CAPSULE, HULL, TRIANGLE = range(7)
JOINT_FIXED, JOINT_REVOLUTE, JOINT_PRISMATIC = 0, 1, 2
TEST_FIRST, TEST_CLOSEST, TEST_ALL = 0, 1, 2
CC_NONE, CC_TIME0, CC_TIME1, CC_BETWEEN = 0, 1, 2, 3
Q_F32_SOA, Q_F64_ROWMAJOR, Q_F32_ROWMAJOR = 0, 1, 2


class Shape(C.Structure):
    _fields_ = [("type", C.c_int32), ("link", C.c_int32), ("shape_id", C.c_int32), ("subshape_id", C.c_int32),
                ("params", C.c_double * 3), ("vertex_offset", C.c_int32), ("vertex_count", C.c_int32),
                ("pose", C.c_double * 12)]


class Link(C.Structure):
    _fields_ = [("name", C.c_char_p), ("type_id", C.c_int32), ("enabled", C.c_int32), ("active", C.c_int32),
                ("pose", C.c_double * 12)]


class SceneDesc(C.Structure):
    _fields_ = [("n_links", C.c_int32), ("links", C.POINTER(Link)), ("n_shapes", C.c_int32),
                ("shapes", C.POINTER(Shape)), ("n_vertices", C.c_int32), ("vertices", C.POINTER(C.c_double)),
                ("n_acm_pairs", C.c_int32), ("acm_pairs", C.POINTER(C.c_int32)), ("default_margin", C.c_double),
                ("n_margin_pairs", C.c_int32), ("margin_pairs", C.POINTER(C.c_int32)),
                ("margin_values", C.POINTER(C.c_double))]


class Joint(C.Structure):
    _fields_ = [("type", C.c_int32), ("parent_link", C.c_int32), ("child_link", C.c_int32), ("q_index", C.c_int32),
                ("origin", C.c_double * 12), ("axis", C.c_double * 3)]


class FkModelDesc(C.Structure):
    _fields_ = [("n_links", C.c_int32), ("n_joints", C.c_int32), ("joints", C.POINTER(Joint)), ("n_q", C.c_int32),
                ("root_link", C.c_int32)]


class Request(C.Structure):
    _fields_ = [("type", C.c_int32), ("calculate_penetration", C.c_int32), ("calculate_distance", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("states", "bvh_nodes", "candidates", "culled", "gjk_pairs",
                                          "gjk_iterations", "epa_pairs", "epa_iterations", "contacts",
                                          "colliding_states", "cull_iterations", "exact_pairs", "phase1_iterations",
                                          "phase2_iterations", "epa_gjk_iterations", "support_vertices",
                                          "support_vertices_fp64", "bvh_stack_overflows", "phase1_skipped",
                                          "support_vertices_epa", "phase1_skip_violations",
                                          "closest_pruned")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class Result(C.Structure):
    _fields_ = [("hit", C.POINTER(C.c_uint8)), ("contacts", C.c_void_p), ("capacity", C.c_int64),
                ("n_contacts", C.c_int64), ("stats", Stats), ("cc_local", C.POINTER(C.c_float))]


class DeviceIO(C.Structure):
    _fields_ = [("q", C.c_void_p), ("q_layout", C.c_int32), ("hit", C.c_void_p), ("contacts", C.c_void_p),
                ("capacity", C.c_int64), ("cc_local", C.c_void_p), ("n_contacts", C.c_int64), ("stats", Stats)]


class Gathered(C.Structure):
    _fields_ = [("hit", C.c_void_p), ("contacts", C.c_void_p), ("capacity", C.c_int64), ("n_contacts", C.c_int64),
                ("n_states", C.c_int64), ("stats", Stats)]


FK_NOALLOC_SCRATCH_BYTES = 8 << 20  # include/t4c.h T4C_FK_NOALLOC_SCRATCH_BYTES

# numpy view of t4c_contact (80 bytes)
CONTACT_DTYPE = np.dtype([("state", "<i4"), ("link", "<i4", 2), ("subshape_id", "<i4", 2), ("shape_id", "<i2", 2),
                          ("distance", "<f4"), ("point", "<f4", (2, 3)), ("normal", "<f4", 3), ("cc_time", "<f4", 2),
                          ("cc_type", "u1", 2), ("reserved", "u1", 2), ("flags", "<u4")], align=False)
assert CONTACT_DTYPE.itemsize == 80

_lib = None


def load_library(path=None):
    """Load libt4c.so (built in-tree by __graft_entry__.build()). Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("T4C_LIB") or LIB_PATH  # T4C_LIB: alternative build of the same library (experiments)
    if not os.path.exists(p):
        raise RuntimeError(f"tesseract_b200 CUDA library not built: {p} (run __graft_entry__.build()); "
                           "there is no CPU fallback")
    lib = C.CDLL(p)
    VP, I32, I64, DP = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_double)
    lib.t4c_last_error.restype = C.c_char_p
    lib.t4c_abi_version.restype = I32
    lib.t4c_build_info.restype = C.c_char_p
    if path is None and lib.t4c_build_info() != b"cuda sm_100a":
        # only the tests hand over the host simulator, and they do it explicitly (path=...): never through T4C_LIB
        raise RuntimeError(f"{p} is not the CUDA build of libt4c.so ({lib.t4c_build_info().decode()}); "
                           "there is no CPU fallback")
    lib.t4c_device_count.restype = I32
    lib.t4c_scene_create.argtypes = [C.POINTER(SceneDesc), I32, C.POINTER(VP)]
    lib.t4c_scene_clone.argtypes = [VP, C.POINTER(VP)]
    lib.t4c_scene_destroy.argtypes = [VP]
    lib.t4c_scene_destroy.restype = None
    lib.t4c_scene_set_active.argtypes = [VP, C.POINTER(I32)]
    lib.t4c_scene_set_enabled.argtypes = [VP, I32, I32]
    lib.t4c_scene_set_acm.argtypes = [VP, I32, C.POINTER(I32)]
    lib.t4c_scene_set_margins.argtypes = [VP, C.c_double, I32, C.POINTER(I32), DP]
    lib.t4c_scene_set_link_pose.argtypes = [VP, I32, DP]
    lib.t4c_scene_set_link_poses.argtypes = [VP, I32, C.POINTER(I32), DP]
    lib.t4c_scene_num_active.argtypes = [VP]
    lib.t4c_scene_add_object.argtypes = [VP, C.POINTER(Link), I32, C.POINTER(Shape), I32, DP, C.POINTER(I32)]
    lib.t4c_scene_remove_object.argtypes = [VP, I32]
    lib.t4c_fk_model_create.argtypes = [C.POINTER(FkModelDesc), VP, C.POINTER(VP)]
    lib.t4c_fk_model_destroy.argtypes = [VP]
    lib.t4c_fk_model_destroy.restype = None
    lib.t4c_fk.argtypes = [VP, VP, DP, I64, DP]
    lib.t4c_fk_noalloc.argtypes = [VP, VP, DP, I64, DP]
    lib.t4c_lvs_count.argtypes = [DP, I64, I32, I32, C.c_double, I32, C.POINTER(I64)]
    lib.t4c_lvs_items.argtypes = [DP, I64, I32, I32, C.c_double, I32, I64, C.POINTER(I32), C.POINTER(I32),
                                  C.POINTER(I32), C.POINTER(I64)]
    lib.t4c_lvs_item_states.argtypes = [DP, I64, I32, I32, C.c_double, I32, I64, C.POINTER(I64), DP, DP]
    lib.t4c_check_discrete.argtypes = [VP, VP, DP, I64, C.POINTER(Request), C.POINTER(Result)]
    lib.t4c_check_link_poses_discrete.argtypes = [VP, DP, I64, C.POINTER(Request), C.POINTER(Result)]
    lib.t4c_check_cast.argtypes = [VP, VP, DP, DP, I64, C.POINTER(Request), C.POINTER(Result)]
    lib.t4c_check_link_poses_cast.argtypes = [VP, DP, DP, I64, C.POINTER(Request), C.POINTER(Result)]
    lib.t4c_check_trajectories_discrete.argtypes = [VP, VP, DP, I64, I32, C.c_double, C.POINTER(Request),
                                                    C.POINTER(Result), C.POINTER(I64)]
    lib.t4c_check_trajectories_cast.argtypes = [VP, VP, DP, I64, I32, C.c_double, C.POINTER(Request),
                                                C.POINTER(Result), C.POINTER(I64)]
    lib.t4c_check_discrete_device.argtypes = [VP, VP, I64, C.POINTER(Request), C.POINTER(DeviceIO)]
    lib.t4c_check_cast_device.argtypes = [VP, VP, I64, VP, C.POINTER(Request), C.POINTER(DeviceIO)]
    U8P = C.POINTER(C.c_uint8)
    lib.t4c_comm_unique_id.argtypes = [U8P]
    lib.t4c_comm_create.argtypes = [U8P, I32, I32, I32, C.POINTER(VP)]
    lib.t4c_comm_create_local.argtypes = [C.POINTER(I32), I32, C.POINTER(VP)]
    lib.t4c_comm_destroy.argtypes = [VP]
    lib.t4c_comm_destroy.restype = None
    lib.t4c_check_discrete_sharded.argtypes = [VP, VP, VP, I32, I64, I64, C.POINTER(Request), C.POINTER(DeviceIO),
                                               C.POINTER(Gathered)]
    lib.t4c_check_cast_sharded.argtypes = [VP, VP, VP, I32, I64, I64, VP, C.POINTER(Request), C.POINTER(DeviceIO),
                                           C.POINTER(Gathered)]
    lib.t4c_check_discrete_multi.argtypes = [C.POINTER(VP), VP, C.POINTER(VP), I32, DP, I64, C.POINTER(Request),
                                             C.POINTER(Result)]
    lib.t4c_jacobian.argtypes = [VP, VP, DP, I64, I32, DP, DP]
    lib.t4c_contact_gradients.argtypes = [VP, VP, DP, I64, VP, I64, DP]
    lib.t4c_contact_gradients_device.argtypes = [VP, VP, VP, I32, I64, VP, I64, VP]
    lib.t4c_kernel_launch_count.restype = C.c_uint64
    lib.t4c_profile_enable.argtypes = [I32]
    lib.t4c_profile_enable.restype = None
    lib.t4c_profile_reset.restype = None
    lib.t4c_profile_get.argtypes = [DP, C.POINTER(C.c_uint64), I32]
    lib.t4c_profile_get.restype = I32
    lib.t4c_measure_fp32_peak.argtypes = [I32, DP]
    lib.t4c_measure_fp64_peak.argtypes = [I32, DP]
    lib.t4c_debug_sincos.argtypes = [I32, DP, I64, DP, DP]
    lib.t4c_debug_hull_support.argtypes = [VP, I32, DP, I64, C.POINTER(I32), DP]
    for name in ("t4c_scene_create", "t4c_scene_clone", "t4c_scene_set_active", "t4c_scene_set_enabled",
                 "t4c_scene_set_acm", "t4c_scene_set_margins", "t4c_scene_set_link_pose", "t4c_scene_num_active",
                 "t4c_fk_model_create", "t4c_fk", "t4c_fk_noalloc", "t4c_check_discrete", "t4c_check_link_poses_discrete",
                 "t4c_check_cast", "t4c_check_link_poses_cast", "t4c_check_trajectories_discrete",
                 "t4c_check_trajectories_cast", "t4c_check_discrete_device", "t4c_check_cast_device",
                 "t4c_measure_fp32_peak", "t4c_lvs_count", "t4c_lvs_items", "t4c_lvs_item_states",
                 "t4c_debug_hull_support", "t4c_measure_fp64_peak", "t4c_debug_sincos", "t4c_comm_unique_id",
                 "t4c_comm_create", "t4c_comm_create_local", "t4c_check_discrete_sharded", "t4c_check_cast_sharded",
                 "t4c_check_discrete_multi", "t4c_jacobian", "t4c_contact_gradients",
                 "t4c_contact_gradients_device", "t4c_scene_add_object", "t4c_scene_remove_object",
                 "t4c_scene_set_link_poses"):
        getattr(lib, name).restype = I32
    if path is None:
        _lib = lib
    return lib


KERNEL_NAMES = ("fk", "convert_poses", "broad_static", "broad_active", "cull_fp32", "gjk_fp64", "epa_fp64",
                "closest_insert", "closest_emit", "count_hit")


def profile_get(lib):
    ms = (C.c_double * 16)()
    n = (C.c_uint64 * 16)()
    k = lib.t4c_profile_get(ms, n, 16)
    return {KERNEL_NAMES[i]: (float(ms[i]), int(n[i])) for i in range(min(k, len(KERNEL_NAMES)))}


class T4CError(RuntimeError):
    pass


def check(status, lib=None):
    if status != T4C_OK:
        lib = lib or load_library()
        msg = lib.t4c_last_error()
        raise T4CError(f"t4c status {status}: {msg.decode() if msg else ''}")
