"""Live scene edits through the C ABI: addCollisionObject / removeCollisionObject on an existing
scene, a static object moved with a BVH refit instead of a rebuild, clones that share the compiled image.  Every edited
scene must give exactly the contacts of a scene built from scratch in its final state (host simulation of the kernels on
the CPU tier; test_gpu_scene_edit runs the same body on the device)."""
import ctypes as C
import time

import numpy as np
import pytest

from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker, create_scene


def run(lib, scene_h, model, q, tt=capi.TEST_CLOSEST):
    n = len(q)
    hit = np.zeros(n, dtype=np.uint8)
    con = np.zeros(64 * n, dtype=capi.CONTACT_DTYPE)
    res = capi.Result()
    res.hit = hit.ctypes.data_as(C.POINTER(C.c_uint8))
    res.contacts, res.capacity = con.ctypes.data, len(con)
    req = capi.Request(tt, 1, 1)
    capi.check(lib.t4c_check_discrete(scene_h, model, q.ctypes.data_as(C.POINTER(C.c_double)), n, C.byref(req),
                                      C.byref(res)), lib)
    return hit, con[:int(res.n_contacts)]


def keyed(con, names):
    return {(int(c["state"]), frozenset((names[c["link"][0]], names[c["link"][1]]))): float(c["distance"]) for c in con}


def scene_edit_case(lib, n_states=300, n_prims=1500):
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    q = S.random_states(b, n_states, seed=31, dtype=np.float64)
    sc = create_scene(lib, b)
    fkd = b.fk_desc()
    model = C.c_void_p()
    capi.check(lib.t4c_fk_model_create(C.byref(fkd), sc.h, C.byref(model)), lib)
    names = list(b.link_names)
    hit0, con0 = run(lib, sc.h, model, q)
    # ---- clone: shares the image; must give identical results, and cost far less than a build
    t = time.perf_counter()
    clone = C.c_void_p()
    capi.check(lib.t4c_scene_clone(sc.h, C.byref(clone)), lib)
    clone_ms = 1e3 * (time.perf_counter() - t)
    hit_c, con_c = run(lib, clone, model, q)
    assert np.array_equal(hit_c, hit0) and keyed(con_c, names) == keyed(con0, names)
    # ---- move a static primitive into the robot's reach: refit on the clone, the original is untouched
    moved = b.link_index["env_00007"]
    P = S.pose((0.35, 0.0, 0.6))
    capi.check(lib.t4c_scene_set_link_pose(clone, moved, P.ctypes.data_as(C.POINTER(C.c_double))), lib)
    hit_m, con_m = run(lib, clone, model, q)
    hit_again, con_again = run(lib, sc.h, model, q)
    assert keyed(con_again, names) == keyed(con0, names)
    b2 = S.build_config2(n_prims=n_prims, margin=0.05)
    b2.set_link_pose("env_00007", P)
    hit_r, con_r, _ = BatchChecker(b2, lib=lib).check_discrete(q, capi.TEST_CLOSEST)
    assert np.array_equal(hit_m, hit_r) and keyed(con_m, names) == keyed(con_r, names)
    assert any("env_00007" in k[1] for k in keyed(con_m, names)), "the moved object should now be in contact"
    # ---- add an object (a box in front of the robot), then remove another one
    link = capi.Link()
    link.name, link.type_id, link.enabled, link.active = b"added_box", 0, 1, 0
    link.pose[:] = S.pose((0.0, 0.45, 0.5)).reshape(-1).tolist()
    shp = capi.Shape()
    shp.type, shp.shape_id, shp.subshape_id = capi.BOX, 0, -1
    shp.params[:] = [0.05, 0.05, 0.05]
    shp.pose[:] = S.identity().reshape(-1).tolist()
    li = C.c_int32()
    capi.check(lib.t4c_scene_add_object(clone, C.byref(link), 1, C.byref(shp), 0, None, C.byref(li)), lib)
    assert li.value == len(names)
    gone = b.link_index["env_00011"]
    capi.check(lib.t4c_scene_remove_object(clone, gone), lib)
    hit_e, con_e = run(lib, clone, model, q)
    b3 = S.build_config2(n_prims=n_prims, margin=0.05)
    b3.set_link_pose("env_00007", P)
    b3.add_collision_object("added_box", 0, [S.Box(0.1, 0.1, 0.1)], [S.identity()], pose_=S.pose((0.0, 0.45, 0.5)))
    b3.link_enabled[b3.link_index["env_00011"]] = 0
    hit_f, con_f, _ = BatchChecker(b3, lib=lib).check_discrete(q, capi.TEST_CLOSEST)
    names3 = list(b3.link_names)
    assert np.array_equal(hit_e, hit_f) and keyed(con_e, names + ["added_box"]) == keyed(con_f, names3)
    assert any("added_box" in k[1] for k in keyed(con_e, names + ["added_box"]))
    lib.t4c_scene_destroy(clone)
    lib.t4c_fk_model_destroy(model)
    return dict(clone_ms=clone_ms, contacts=len(con_e))


def test_scene_edit_hostsim():
    from hostsim import build as hostsim
    print(scene_edit_case(hostsim.lib()))


@pytest.mark.gpu
def test_gpu_scene_edit():
    print(scene_edit_case(capi.load_library(), n_states=20000, n_prims=10000))
