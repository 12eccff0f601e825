"""Host-side mirror of the reference's contact-manager interface on top of the C ABI (include/t4c.h).

`B200ContactManager` plays the role of tesseract_collision::DiscreteContactManager (cast=False,
tesseract_collision/include/tesseract_collision/core/discrete_contact_manager.h:44-208) and
ContinuousContactManager (cast=True, continuous_contact_manager.h:44-257) for the Python test tooling: same call
names in snake case, same argument meaning, same error behaviour (unknown link => silent no-op for transforms,
False for enable/disable).  The C++ twin with the exact reference signatures is tesseract_b200/csrc/b200_contact_managers.h.

`BatchChecker` is the batched entry point (FK + check of many joint vectors / segments / trajectories).

There is no CPU path here: every call goes to libt4c.so (CUDA, sm_100a) and raises if it is missing.
"""
import ctypes as C

import numpy as np

from . import capi

RESULT_DTYPE = np.dtype([("state", "<i8"), ("link", "<i4", 2), ("shape_id", "<i4", 2), ("subshape_id", "<i4", 2),
                         ("cc_type", "<i4", 2), ("distance", "<f8"), ("point", "<f8", (2, 3)),
                         ("point_local", "<f8", (2, 3)), ("normal", "<f8", 3), ("cc_time", "<f8", 2),
                         ("transform", "<f8", (2, 12)), ("cc_transform", "<f8", (2, 12))])


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _inv_apply(T, p):
    T = np.asarray(T).reshape(3, 4)
    return T[:, :3].T @ (np.asarray(p) - T[:, 3])


class _SceneHandle:
    def __init__(self, lib, handle):
        self.lib, self.h = lib, handle

    def __del__(self):
        try:
            self.lib.t4c_scene_destroy(self.h)
        except Exception:
            pass


def create_scene(lib, builder, device=0):
    desc = builder.scene_desc()
    h = C.c_void_p()
    capi.check(lib.t4c_scene_create(C.byref(desc), device, C.byref(h)), lib)
    return _SceneHandle(lib, h)


class B200ContactManager:
    def __init__(self, builder, cast=False, lib=None, device=0, _scene=None, _state=None):
        self.lib = lib or capi.load_library()
        self.b = builder
        self.cast = cast
        self.device = device
        self.scene = _scene or create_scene(self.lib, builder, device)
        n = len(builder.link_names)
        if _state is None:
            self.active = np.array(builder.link_active, dtype=np.int32)
            self.pose0 = np.array([p.reshape(12) for p in builder.link_pose], dtype=np.float64).reshape(n, 12)
            self.pose1 = self.pose0.copy()
        else:
            self.active, self.pose0, self.pose1 = (x.copy() for x in _state)

    # -- identity / cloning (discrete_contact_manager.h:65-73)
    def clone(self):
        h = C.c_void_p()
        capi.check(self.lib.t4c_scene_clone(self.scene.h, C.byref(h)), self.lib)
        return B200ContactManager(self.b, self.cast, self.lib, self.device, _SceneHandle(self.lib, h),
                                  (self.active, self.pose0, self.pose1))

    def _idx(self, link):
        if isinstance(link, (int, np.integer)):
            return int(link) if 0 <= link < len(self.b.link_names) else None
        return self.b.link_index.get(link)

    # -- setActiveCollisionObjects: empty list => everything active (core/common.h:66-69)
    def set_active(self, names):
        self.active = np.array([1 if (not names or n in names) else 0 for n in self.b.link_names], dtype=np.int32)
        capi.check(self.lib.t4c_scene_set_active(self.scene.h, self.active.ctypes.data_as(C.POINTER(C.c_int32))),
                   self.lib)

    def set_enabled(self, link, enabled):
        i = self._idx(link)
        if i is None:
            return False
        capi.check(self.lib.t4c_scene_set_enabled(self.scene.h, i, 1 if enabled else 0), self.lib)
        return True

    def set_margins(self, default, pairs=None):
        pairs = pairs or {}
        keys = list(pairs.keys())
        pi = np.array([[self._idx(a), self._idx(b)] for a, b in keys], dtype=np.int32).reshape(-1, 2)
        pv = np.array([pairs[k] for k in keys], dtype=np.float64)
        capi.check(self.lib.t4c_scene_set_margins(self.scene.h, float(default), len(keys),
                                                  pi.ctypes.data_as(C.POINTER(C.c_int32)), _dp(pv)), self.lib)

    def set_acm(self, pairs):
        pi = np.array([[self._idx(a), self._idx(b)] for a, b in pairs], dtype=np.int32).reshape(-1, 2)
        capi.check(self.lib.t4c_scene_set_acm(self.scene.h, len(pairs), pi.ctypes.data_as(C.POINTER(C.c_int32))),
                   self.lib)

    # -- setCollisionObjectsTransform(name, T[, T2]): unknown name is a silent no-op
    #    (bullet_discrete_bvh_manager.cpp:198-199)
    def set_pose(self, link, P, P1=None):
        i = self._idx(link)
        if i is None:
            return
        P = np.ascontiguousarray(P, dtype=np.float64).reshape(12)
        self.pose0[i] = P
        self.pose1[i] = P if P1 is None else np.ascontiguousarray(P1, dtype=np.float64).reshape(12)
        if not self.active[i]:
            capi.check(self.lib.t4c_scene_set_link_pose(self.scene.h, i, _dp(self.pose0[i])), self.lib)

    # -- contactTest(ContactResultMap&, ContactRequest)
    def contact_test(self, test_type=capi.TEST_ALL, calculate_penetration=True, calculate_distance=True, tag=0):
        act = np.flatnonzero(self.active)
        p0 = np.ascontiguousarray(self.pose0[act].reshape(1, -1, 12))
        p1 = np.ascontiguousarray(self.pose1[act].reshape(1, -1, 12))
        req = capi.Request(test_type, int(calculate_penetration), int(calculate_distance))
        cap = 4096
        while True:
            out = np.zeros(cap, dtype=capi.CONTACT_DTYPE)
            loc = np.zeros((cap, 2, 3), dtype=np.float32)
            hit = np.zeros(1, dtype=np.uint8)
            res = capi.Result()
            res.hit = hit.ctypes.data_as(C.POINTER(C.c_uint8))
            res.contacts = out.ctypes.data
            res.capacity = cap
            res.cc_local = loc.ctypes.data_as(C.POINTER(C.c_float))
            if self.cast:
                st = self.lib.t4c_check_link_poses_cast(self.scene.h, _dp(p0), _dp(p1), 1, C.byref(req), C.byref(res))
            else:
                st = self.lib.t4c_check_link_poses_discrete(self.scene.h, _dp(p0), 1, C.byref(req), C.byref(res))
            if st == capi.T4C_ERR_CAPACITY:
                cap = int(res.n_contacts)
                continue
            capi.check(st, self.lib)
            break
        self.last_stats = res.stats.as_dict()
        return self._expand(out[:res.n_contacts], loc[:res.n_contacts], tag)

    def _expand(self, raw, loc, tag):
        """t4c_contact PODs -> ContactResult-like records (names / transforms are host knowledge,
        bullet_utils.h:388-441)."""
        r = np.zeros(len(raw), dtype=RESULT_DTYPE)
        ident = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float64)
        for k, c in enumerate(raw):
            o = r[k]
            o["state"] = tag
            o["link"] = c["link"]
            o["shape_id"] = c["shape_id"]
            o["subshape_id"] = c["subshape_id"]
            o["cc_type"] = c["cc_type"]
            o["distance"] = c["distance"]
            o["point"] = c["point"]
            o["normal"] = c["normal"]
            o["cc_time"] = c["cc_time"]
            for j in range(2):
                li = int(c["link"][j])
                o["transform"][j] = self.pose0[li]
                o["cc_transform"][j] = ident
                o["point_local"][j] = _inv_apply(self.pose0[li], c["point"][j])
                if self.cast and int(c["cc_type"][j]) != capi.CC_NONE:
                    o["cc_transform"][j] = self.pose1[li]
                    o["point_local"][j] = loc[k][j]
        return r


class BatchChecker:
    """Batched entry point: FK + discrete / cast contact tests of many joint vectors (host buffers)."""

    def __init__(self, builder, lib=None, device=0):
        self.lib = lib or capi.load_library()
        self.b = builder
        self.scene = create_scene(self.lib, builder, device)
        self.fk_desc = builder.fk_desc()
        self.model = C.c_void_p()
        capi.check(self.lib.t4c_fk_model_create(C.byref(self.fk_desc), self.scene.h, C.byref(self.model)), self.lib)

    def __del__(self):
        try:
            self.lib.t4c_fk_model_destroy(self.model)
        except Exception:
            pass

    def fk(self, q):
        q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, self.fk_desc.n_q)
        out = np.zeros((len(q), len(self.b.link_names), 3, 4))
        capi.check(self.lib.t4c_fk(self.model, self.scene.h, _dp(q), len(q), _dp(out)), self.lib)
        return out

    def _run(self, fn, args, n, test_type, pen, dist, want_contacts, capacity, cast):
        req = capi.Request(test_type, int(pen), int(dist))
        cap = capacity or max(1024, 16 * n)
        while True:
            hit = np.zeros(n, dtype=np.uint8)
            out = np.zeros(cap if want_contacts else 0, dtype=capi.CONTACT_DTYPE)
            loc = np.zeros((cap if (want_contacts and cast) else 0, 2, 3), dtype=np.float32)
            res = capi.Result()
            res.hit = hit.ctypes.data_as(C.POINTER(C.c_uint8))
            res.contacts = out.ctypes.data if want_contacts else None
            res.capacity = cap if want_contacts else 0
            res.cc_local = loc.ctypes.data_as(C.POINTER(C.c_float)) if (want_contacts and cast) else None
            st = fn(*args, C.byref(req), C.byref(res))
            if st == capi.T4C_ERR_CAPACITY:
                cap = int(res.n_contacts)
                continue
            capi.check(st, self.lib)
            k = int(res.n_contacts) if want_contacts else 0
            return hit, out[:k], loc[:k], res.stats.as_dict()

    def check_discrete(self, q, test_type=capi.TEST_CLOSEST, calculate_penetration=True, calculate_distance=True,
                       want_contacts=True, capacity=None):
        q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, self.fk_desc.n_q)
        hit, out, _, stats = self._run(self.lib.t4c_check_discrete, (self.scene.h, self.model, _dp(q), len(q)), len(q),
                                       test_type, calculate_penetration, calculate_distance, want_contacts, capacity,
                                       False)
        return hit, out, stats

    def check_link_poses(self, T0, T1=None, test_type=capi.TEST_CLOSEST, want_contacts=True, capacity=None):
        """FK bypass: T0 (and T1 for cast) are (n, n_links, 3, 4) world poses of ALL links; the kinematic links'
        rows are handed to t4c_check_link_poses_{discrete,cast}."""
        act = self.b.active_link_indices()
        p0 = np.ascontiguousarray(np.asarray(T0, dtype=np.float64)[:, act].reshape(len(T0), len(act), 12))
        if T1 is None:
            return self._run(self.lib.t4c_check_link_poses_discrete, (self.scene.h, _dp(p0), len(p0)), len(p0),
                             test_type, True, True, want_contacts, capacity, False)
        p1 = np.ascontiguousarray(np.asarray(T1, dtype=np.float64)[:, act].reshape(len(T1), len(act), 12))
        return self._run(self.lib.t4c_check_link_poses_cast, (self.scene.h, _dp(p0), _dp(p1), len(p0)), len(p0),
                         test_type, True, True, want_contacts, capacity, True)

    def check_cast(self, q0, q1, test_type=capi.TEST_CLOSEST, want_contacts=True, capacity=None):
        q0 = np.ascontiguousarray(q0, dtype=np.float64).reshape(-1, self.fk_desc.n_q)
        q1 = np.ascontiguousarray(q1, dtype=np.float64).reshape(-1, self.fk_desc.n_q)
        return self._run(self.lib.t4c_check_cast, (self.scene.h, self.model, _dp(q0), _dp(q1), len(q0)), len(q0),
                         test_type, True, True, want_contacts, capacity, True)

    def check_trajectories(self, traj, lvs, cast, test_type=capi.TEST_CLOSEST, want_contacts=True, capacity=None):
        traj = np.ascontiguousarray(traj, dtype=np.float64)
        n_traj, n_wp, n_q = traj.shape
        n_items = C.c_int64()
        capi.check(self.lib.t4c_lvs_count(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), C.byref(n_items)),
                   self.lib)
        fn = self.lib.t4c_check_trajectories_cast if cast else self.lib.t4c_check_trajectories_discrete
        req_items = C.c_int64()

        def call(scene, model, tp, nt, nw, l, req, res):
            return fn(scene, model, tp, nt, nw, l, req, res, C.byref(req_items))

        return self._run(call, (self.scene.h, self.model, _dp(traj), n_traj, n_wp, float(lvs)), int(n_items.value),
                         test_type, True, True, want_contacts, capacity, cast)

    def lvs_items(self, traj, lvs, cast):
        """(item_traj, item_step, item_sub) of every LVS work item, in the reference's visiting order
        (tesseract_environment/core/utils.h:189-205, 328-347)"""
        traj = np.ascontiguousarray(traj, dtype=np.float64)
        n_traj, n_wp, n_q = traj.shape
        n = C.c_int64()
        capi.check(self.lib.t4c_lvs_count(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), C.byref(n)), self.lib)
        out = [np.zeros(max(1, n.value), dtype=np.int32) for _ in range(3)]
        ip = [o.ctypes.data_as(C.POINTER(C.c_int32)) for o in out]
        capi.check(self.lib.t4c_lvs_items(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), n.value, ip[0], ip[1],
                                          ip[2], C.byref(n)), self.lib)
        return tuple(o[:n.value] for o in out)

    def lvs_item_states(self, traj, lvs, cast, items):
        """joint vectors (q0, q1 or None) of the selected work items (strictly ascending flat indices)"""
        traj = np.ascontiguousarray(traj, dtype=np.float64)
        n_traj, n_wp, n_q = traj.shape
        items = np.ascontiguousarray(items, dtype=np.int64)
        q0 = np.zeros((len(items), n_q))
        q1 = np.zeros((len(items), n_q)) if cast else None
        capi.check(self.lib.t4c_lvs_item_states(_dp(traj), n_traj, n_wp, n_q, float(lvs), int(cast), len(items),
                                                items.ctypes.data_as(C.POINTER(C.c_int64)), _dp(q0),
                                                _dp(q1) if cast else None), self.lib)
        return q0, q1

