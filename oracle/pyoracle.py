"""ctypes loader for the CPU ORACLE (test infrastructure).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  It never touches /root/reference at run time.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from tesseract_b200 import capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libt4oracle.so")

CONTACT_DTYPE = np.dtype([("state", "<i8"), ("link", "<i4", 2), ("shape_id", "<i4", 2), ("subshape_id", "<i4", 2),
                          ("cc_type", "<i4", 2), ("distance", "<f8"), ("point", "<f8", (2, 3)),
                          ("point_local", "<f8", (2, 3)), ("normal", "<f8", 3), ("cc_time", "<f8", 2),
                          ("transform", "<f8", (2, 12)), ("cc_transform", "<f8", (2, 12))], align=True)

_lib = None


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in ("t4o_narrow.cpp", "t4o_manager.cpp", "t4o.h")]
    srcs.append(os.path.join(HERE, "..", "include", "t4c.h"))
    stale = force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", HERE, "-B", "libt4oracle.so"], stdout=subprocess.DEVNULL)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        VP, I32, I64, DP = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_double)
        L.t4o_manager_create.restype = VP
        L.t4o_manager_create.argtypes = [C.POINTER(capi.SceneDesc), C.c_int]
        L.t4o_manager_clone.restype = VP
        L.t4o_manager_clone.argtypes = [VP]
        L.t4o_manager_destroy.argtypes = [VP]
        L.t4o_manager_destroy.restype = None
        L.t4o_manager_set_active.argtypes = [VP, C.POINTER(I32)]
        L.t4o_manager_set_enabled.argtypes = [VP, I32, I32]
        L.t4o_manager_set_acm.argtypes = [VP, I32, C.POINTER(I32)]
        L.t4o_manager_set_margins.argtypes = [VP, C.c_double, I32, C.POINTER(I32), DP]
        L.t4o_manager_set_pose.argtypes = [VP, I32, DP]
        L.t4o_manager_set_pose_cast.argtypes = [VP, I32, DP, DP]
        L.t4o_manager_contact_test.restype = I64
        L.t4o_manager_contact_test.argtypes = [VP, C.POINTER(capi.Request), I64, C.c_void_p, I64]
        L.t4o_manager_get_counters.argtypes = [VP, C.POINTER(C.c_uint64)]
        L.t4o_fk.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.FkModelDesc), DP, I64, DP]
        L.t4o_fk.restype = None
        for fn in (L.t4o_check_discrete,):
            fn.restype = I64
            fn.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.FkModelDesc), DP, I64, C.POINTER(capi.Request),
                           C.c_void_p, C.c_void_p, I64, C.c_int, C.POINTER(C.c_uint64)]
        L.t4o_check_cast.restype = I64
        L.t4o_check_cast.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.FkModelDesc), DP, DP, I64,
                                     C.POINTER(capi.Request), C.c_void_p, C.c_void_p, I64, C.c_int,
                                     C.POINTER(C.c_uint64)]
        L.t4o_lvs_expand.restype = I64
        L.t4o_lvs_expand.argtypes = [DP, I64, I32, I32, C.c_double, C.c_int, DP, DP, C.POINTER(I64),
                                     C.POINTER(I32), I64]
        L.t4o_pair_distance.restype = C.c_int
        L.t4o_pair_distance.argtypes = [C.POINTER(capi.Shape), DP, C.POINTER(capi.Shape), DP, DP, C.c_double,
                                        C.c_int, C.c_int, DP, DP, DP]  # noqa
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


COUNTER_NAMES = ("candidates", "gjk_pairs", "gjk_iterations", "epa_pairs", "epa_iterations", "contacts")


class OracleManager:
    """Restated BulletDiscreteBVHManager (cast=False) / BulletCastBVHManager (cast=True) driven by link index
    or name, mirroring the reference manager API used by its test-suite."""

    def __init__(self, builder, cast=False, _handle=None):
        self.b = builder
        self.cast = cast
        self.L = lib()
        self.desc = builder.scene_desc()
        self.h = _handle or self.L.t4o_manager_create(C.byref(self.desc), 1 if cast else 0)

    def clone(self):
        return OracleManager(self.b, self.cast, self.L.t4o_manager_clone(self.h))

    def __del__(self):
        try:
            self.L.t4o_manager_destroy(self.h)
        except Exception:
            pass

    def _idx(self, link):
        return link if isinstance(link, int) else self.b.link_index[link]

    def set_active(self, names):
        flags = (C.c_int32 * len(self.b.link_names))(*[1 if (not names or n in names) else 0
                                                      for n in self.b.link_names])
        self.L.t4o_manager_set_active(self.h, flags)

    def set_enabled(self, link, enabled):
        self.L.t4o_manager_set_enabled(self.h, self._idx(link), 1 if enabled else 0)

    def set_margins(self, default, pairs=None):
        pairs = pairs or {}
        keys = list(pairs.keys())
        pi = np.array([[self._idx(a), self._idx(b)] for a, b in keys], dtype=np.int32).reshape(-1, 2)
        pv = np.array([pairs[k] for k in keys], dtype=np.float64)
        self.L.t4o_manager_set_margins(self.h, float(default), len(keys), pi.ctypes.data_as(C.POINTER(C.c_int32)),
                                       _dp(pv))

    def set_acm(self, pairs):
        pi = np.array([[self._idx(a), self._idx(b)] for a, b in pairs], dtype=np.int32).reshape(-1, 2)
        self.L.t4o_manager_set_acm(self.h, len(pairs), pi.ctypes.data_as(C.POINTER(C.c_int32)))

    def set_pose(self, link, P, P1=None):
        P = np.ascontiguousarray(P, dtype=np.float64).reshape(12)
        if P1 is None:
            self.L.t4o_manager_set_pose(self.h, self._idx(link), _dp(P))
        else:
            P1 = np.ascontiguousarray(P1, dtype=np.float64).reshape(12)
            self.L.t4o_manager_set_pose_cast(self.h, self._idx(link), _dp(P), _dp(P1))

    def contact_test(self, test_type=capi.TEST_ALL, calculate_penetration=True, calculate_distance=True, tag=0):
        req = capi.Request(test_type, int(calculate_penetration), int(calculate_distance))
        cap = 4096
        while True:
            out = np.zeros(cap, dtype=CONTACT_DTYPE)
            n = self.L.t4o_manager_contact_test(self.h, C.byref(req), tag, out.ctypes.data, cap)
            if n <= cap:
                return out[:n]
            cap = int(n)

    def counters(self):
        c = (C.c_uint64 * 6)()
        self.L.t4o_manager_get_counters(self.h, c)
        return dict(zip(COUNTER_NAMES, [int(x) for x in c]))


def pair_distance(builder, shape_a, pose_a, shape_b, pose_b, threshold, calculate_penetration=True,
                  calculate_distance=True):
    """One convex-convex query of the restated detector (collideConvexConvex): shapes are indices into builder.shapes,
    poses 3x4 WORLD poses of the shapes.  Returns (distance, normal_on_b, point_on_b) or None."""
    L = lib()
    sd = builder.scene_desc()
    pa = np.ascontiguousarray(pose_a, dtype=np.float64).reshape(12)
    pb = np.ascontiguousarray(pose_b, dtype=np.float64).reshape(12)
    d = C.c_double()
    n = np.zeros(3)
    p = np.zeros(3)
    ok = L.t4o_pair_distance(C.byref(sd.shapes[shape_a]), _dp(pa), C.byref(sd.shapes[shape_b]), _dp(pb), sd.vertices,
                             float(threshold), int(calculate_penetration), int(calculate_distance), C.byref(d), _dp(n),
                             _dp(p))
    return (float(d.value), n, p) if ok else None


def fk(builder, q):
    L = lib()
    sd, fd = builder.scene_desc(), builder.fk_desc()
    q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, fd.n_q)
    out = np.zeros((len(q), fd.n_links, 3, 4))
    L.t4o_fk(C.byref(sd), C.byref(fd), _dp(q), len(q), _dp(out))
    return out


def check_discrete(builder, q, test_type=capi.TEST_CLOSEST, calculate_penetration=True, calculate_distance=True,
                   want_contacts=True, capacity=None, n_threads=1):
    """Per-state FK -> setCollisionObjectsTransform -> contactTest loop (core/utils.h:129-160) on the oracle."""
    L = lib()
    sd, fd = builder.scene_desc(), builder.fk_desc()
    q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, fd.n_q)
    n = len(q)
    req = capi.Request(test_type, int(calculate_penetration), int(calculate_distance))
    hit = np.zeros(n, dtype=np.uint8)
    cap = capacity or max(1024, 64 * n)
    out = np.zeros(cap if want_contacts else 0, dtype=CONTACT_DTYPE)
    cnt = (C.c_uint64 * 6)()
    total = L.t4o_check_discrete(C.byref(sd), C.byref(fd), _dp(q), n, C.byref(req), hit.ctypes.data,
                                 out.ctypes.data if want_contacts else None, cap, n_threads, cnt)
    if want_contacts and total > cap:
        return check_discrete(builder, q, test_type, calculate_penetration, calculate_distance, True, int(total),
                              n_threads)
    return hit, out[:total] if want_contacts else out, dict(zip(COUNTER_NAMES, [int(x) for x in cnt]))


def check_cast(builder, q0, q1, test_type=capi.TEST_CLOSEST, want_contacts=True, capacity=None, n_threads=1):
    L = lib()
    sd, fd = builder.scene_desc(), builder.fk_desc()
    q0 = np.ascontiguousarray(q0, dtype=np.float64).reshape(-1, fd.n_q)
    q1 = np.ascontiguousarray(q1, dtype=np.float64).reshape(-1, fd.n_q)
    n = len(q0)
    req = capi.Request(test_type, 1, 1)
    hit = np.zeros(n, dtype=np.uint8)
    cap = capacity or max(1024, 64 * n)
    out = np.zeros(cap if want_contacts else 0, dtype=CONTACT_DTYPE)
    cnt = (C.c_uint64 * 6)()
    total = L.t4o_check_cast(C.byref(sd), C.byref(fd), _dp(q0), _dp(q1), n, C.byref(req), hit.ctypes.data,
                             out.ctypes.data if want_contacts else None, cap, n_threads, cnt)
    if want_contacts and total > cap:
        return check_cast(builder, q0, q1, test_type, True, int(total), n_threads)
    return hit, out[:total] if want_contacts else out, dict(zip(COUNTER_NAMES, [int(x) for x in cnt]))


def lvs_expand(traj, lvs, cast):
    L = lib()
    traj = np.ascontiguousarray(traj, dtype=np.float64)
    n_traj, n_wp, n_q = traj.shape
    n = L.t4o_lvs_expand(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), None, None, None, None, 0)
    q0 = np.zeros((n, n_q))
    q1 = np.zeros((n, n_q))
    it = np.zeros(n, dtype=np.int64)
    st = np.zeros(n, dtype=np.int32)
    L.t4o_lvs_expand(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), _dp(q0), _dp(q1),
                     it.ctypes.data_as(C.POINTER(C.c_int64)), st.ctypes.data_as(C.POINTER(C.c_int32)), n)
    return q0, (q1 if cast else None), it, st
