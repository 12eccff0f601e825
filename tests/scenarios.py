"""Known-answer scenarios restated from the reference's manager-parametrised test suite
(tesseract_collision/include/tesseract_collision/test_suite/*.hpp).

Every function takes `make(builder, cast) -> manager` so the SAME bodies run against the CPU oracle
(tests/test_oracle_golden.py) and the CUDA managers (tests/test_gpu_golden.py), the way the reference
instantiates one body per manager class.  A manager is duck-typed:
    set_active(names) set_margins(default, pairs) set_pose(link, P[, P1]) set_enabled(link, bool)
    contact_test(test_type, ...) -> structured array (link, distance, point, normal, shape_id, subshape_id, cc_*)
    clone()
"""
import json
import os

import numpy as np

from tesseract_b200 import capi
from tesseract_b200 import scene as S

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_cache = {}


def meshes():
    if "m" not in _cache:
        with open(os.path.join(GOLD, "meshes.json")) as f:
            _cache["m"] = json.load(f)
    return _cache["m"]


def sphere_hull():
    return S.ConvexMesh(np.array(meshes()["sphere_p25m.ply"]["vertices"]))


def sphere_mesh():
    m = meshes()["sphere_p25m.ply"]
    return S.Mesh(np.array(m["vertices"]), np.array(m["faces"]))


def box2_hull():
    return S.ConvexMesh(np.array(meshes()["box2_2m.ply"]["vertices"]))


def octree(name, sub_type):
    d = np.load(os.path.join(GOLD, name + "_octree.npz"))
    return S.Octree(d["leaves"].astype(np.float64), sub_type)


def plane_mesh():
    d = np.load(os.path.join(GOLD, "plane_4m.npz"))
    return S.Mesh(d["vertices"], d["faces"])


I = S.identity


def tr(x=0.0, y=0.0, z=0.0):
    return S.pose((x, y, z))


def near(a, b, tol):
    assert abs(a - b) <= tol, f"{a} vs {b} (tol {tol})"


def oriented(res, b, first_name):
    """idx = {0,1,1} / {1,0,-1} convention of the reference tests (e.g. collision_box_box_unit.hpp:159-161)."""
    r = res[0]
    if b.link_names[int(r["link"][0])] == first_name:
        return r, 0, 1, 1.0
    return r, 1, 0, -1.0


# ------------------------------------------------------------------------------------------------ discrete
def box_box(make, use_convex_mesh, test_type):
    """collision_box_box_unit.hpp:13-302"""
    b = S.SceneBuilder()
    b.add_collision_object("box_link", 0, [S.Box(1, 1, 1)], [I()], enabled=False)
    b.link_enabled[b.link_index["box_link"]] = 1  # enableCollisionObject
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()])
    b.link_enabled[b.link_index["thin_box_link"]] = 0  # disableCollisionObject
    second = box2_hull() if use_convex_mesh else S.Box(2, 2, 2)
    b.add_collision_object("second_box_link", 0, [second], [I()])
    assert b.add_collision_object("empty_link", 0, [], []) is False
    b.set_active(["box_link", "second_box_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("box_link", tr(0.2, 0.1, 0))
    m.set_pose("second_box_link", I())
    res = m.contact_test(test_type)
    assert len(res) > 0
    r, i0, i1, sg = oriented(res, b, "box_link")
    near(r["distance"], -1.30, 1e-3)
    near(r["point"][0][1], r["point"][1][1], 1e-3)
    near(r["point"][0][2], r["point"][1][2], 1e-3)
    near(r["point"][i0][0], -0.3, 1e-3)
    near(r["point"][i1][0], 1.0, 1e-3)
    near(r["normal"][0], sg * -1.0, 1e-3)
    near(r["normal"][1], 0.0, 1e-3)
    near(r["normal"][2], 0.0, 1e-3)
    # outside the contact distance
    m.set_pose("box_link", tr(1.60, 0, 0))
    assert len(m.contact_test(test_type)) == 0
    # unrelated pair margin raises the max margin to 1.7 but this pair keeps 0.1 (:199-218)
    m.set_margins(0.1, {("box_link", "thin_box_link"): 1.7})
    assert len(m.contact_test(test_type)) == 0
    # pair margin 0.25 for this pair
    m.set_margins(0.1, {("box_link", "second_box_link"): 0.25})
    for _ in range(2):
        res = m.contact_test(test_type)
        assert len(res) > 0
        r, i0, i1, sg = oriented(res, b, "box_link")
        near(r["distance"], 0.1, 1e-3)
        near(r["point"][0][1], r["point"][1][1], 1e-3)
        near(r["point"][0][2], r["point"][1][2], 1e-3)
        near(r["point"][i0][0], 1.1, 1e-3)
        near(r["point"][i1][0], 1.0, 1e-3)
        near(r["normal"][0], sg * -1.0, 1e-3)
        near(r["normal"][1], 0.0, 1e-3)
        near(r["normal"][2], 0.0, 1e-3)
        m.set_margins(0.25)  # second pass: default margin 0.25


def sphere_sphere_primitive(make):
    """collision_sphere_sphere_unit.hpp:132-243"""
    b = S.SceneBuilder()
    b.add_collision_object("sphere_link", 0, [S.Sphere(0.25)], [I()])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object("sphere1_link", 0, [S.Sphere(0.25)], [I()])
    b.set_active(["sphere_link", "sphere1_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("sphere_link", I())
    m.set_pose("sphere1_link", tr(0.2))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], -0.30, 1e-4)
    for k, v in enumerate((0.25, 0.0, 0.0)):
        near(r["point"][i0][k], v, 1e-3)
    for k, v in enumerate((-0.05, 0.0, 0.0)):
        near(r["point"][i1][k], v, 1e-3)
    near(r["normal"][0], sg * 1.0, 1e-3)
    m.set_pose("sphere1_link", tr(1, 0, 0))
    assert len(m.contact_test(capi.TEST_CLOSEST)) == 0
    m.set_margins(0.52)
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], 0.5, 1e-4)
    near(r["point"][i0][0], 0.25, 1e-3)
    near(r["point"][i1][0], 0.75, 1e-3)
    near(r["normal"][0], sg * 1.0, 1e-3)
    near(r["normal"][1], 0.0, 1e-3)
    near(r["normal"][2], 0.0, 1e-3)


def sphere_sphere_convex(make):
    """collision_sphere_sphere_unit.hpp:245-413 (runTestConvex1/2/3)"""
    b = S.SceneBuilder()
    b.add_collision_object("sphere_link", 0, [sphere_hull()], [I()])
    b.add_collision_object("sphere1_link", 0, [sphere_hull()], [I()])
    b.set_active(["sphere_link", "sphere1_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("sphere_link", I())
    m.set_pose("sphere1_link", tr(0.2))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    # Bullet (with btConvexHullComputer snapping): -0.270548; exact hull of the ply vertices: -0.270636
    near(r["distance"], -0.270548, 1e-3)
    # face-face contact: the witness is unique only along the normal; the reference's y = 0 is one point of the
    # contact polygon, so y is not compared
    for k, v in ((0, 0.232874), (2, -0.025368)):
        near(r["point"][i0][k], v, 1e-3)
    for k, v in ((0, -0.032874), (2, 0.025368)):
        near(r["point"][i1][k], v, 1e-3)
    m.set_pose("sphere1_link", tr(1, 0, 0))
    assert len(m.contact_test(capi.TEST_CLOSEST)) == 0
    m.set_margins(0.55)
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], 0.52448, 1e-3)
    near(r["point"][i0][0], 0.23776, 1e-3)
    near(r["point"][i1][0], 0.76224, 1e-3)
    near(r["point"][0][1], r["point"][1][1], 1e-3)
    near(r["point"][0][2], r["point"][1][2], 1e-3)
    n = sg * np.asarray(r["normal"], dtype=np.float64)
    assert n[0] > 0 and abs(np.arccos(min(1.0, n[0]))) < 1e-5
    # runTestConvex3: shift in y
    m.set_margins(0.1)
    m.set_pose("sphere1_link", tr(0, 0.2, 0))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], -0.280223, 1e-3)
    n = sg * np.asarray(r["normal"], dtype=np.float64)
    near(abs(n[0]), 0.3037316, 1e-3)
    near(n[1], 0.9340783, 1e-3)
    near(abs(n[2]), 0.1877358, 1e-3)


def box_sphere(make, use_convex_mesh):
    """collision_box_sphere_unit.hpp:132-330"""
    b = S.SceneBuilder()
    b.add_collision_object("box_link", 0, [S.Box(1, 1, 1)], [I()])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object("sphere_link", 0, [sphere_hull() if use_convex_mesh else S.Sphere(0.25)], [I()])
    b.set_active(["box_link", "sphere_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("box_link", I())
    m.set_pose("sphere_link", tr(0.2))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "box_link")
    if use_convex_mesh:
        near(r["distance"], -0.53776, 1e-3)
        near(r["point"][i0][0], 0.5, 1e-3)
        near(r["point"][i1][0], -0.03776, 1e-3)
    else:
        near(r["distance"], -0.55, 1e-4)
        near(r["point"][i0][0], 0.5, 1e-3)
        near(r["point"][i1][0], -0.05, 1e-3)
    near(r["point"][0][1], r["point"][1][1], 1e-3)
    near(r["point"][0][2], r["point"][1][2], 1e-3)
    near(r["normal"][0], sg * 1.0, 1e-3)
    m.set_pose("sphere_link", tr(1, 0, 0))
    assert len(m.contact_test(capi.TEST_CLOSEST)) == 0
    m.set_margins(0.27 if use_convex_mesh else 0.251)
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "box_link")
    near(r["distance"], 0.26224 if use_convex_mesh else 0.25, 1e-3)
    near(r["point"][i0][0], 0.5, 1e-3)
    near(r["point"][i1][0], 0.76224 if use_convex_mesh else 0.75, 1e-3)
    near(r["normal"][0], sg * 1.0, 1e-3)
    near(r["normal"][1], 0.0, 1e-3)
    near(r["normal"][2], 0.0, 1e-3)


def box_cylinder_cone_capsule(make, kind):
    """collision_box_cylinder_unit.hpp:113-208, collision_box_cone_unit.hpp:115-220,
    collision_box_capsule_unit.hpp:112-204"""
    geom = {"cylinder": S.Cylinder(0.25, 0.25), "cone": S.Cone(0.25, 0.25), "capsule": S.Capsule(0.25, 0.25)}[kind]
    other = kind + "_link"
    b = S.SceneBuilder()
    b.add_collision_object("box_link", 0, [S.Box(1, 1, 1)], [I()])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object(other, 0, [geom], [I()])
    b.set_active(["box_link", other])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("box_link", I())
    m.set_pose(other, tr(0.2))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "box_link")
    near(r["distance"], -0.55, 1e-4)
    near(r["point"][i0][0], 0.5, 1e-3)
    near(r["point"][i1][0], -0.05, 1e-3)
    near(r["point"][0][1], r["point"][1][1], 1e-3)
    near(r["point"][0][2], r["point"][1][2], 1e-3)
    if kind == "cone":
        near(r["point"][i0][2], -0.125, 1e-3)
        near(r["point"][i1][2], -0.125, 1e-3)
    near(r["normal"][0], sg * 1.0, 1e-3)
    near(r["normal"][1], 0.0, 1e-3)
    near(r["normal"][2], 0.0, 1e-3)
    if kind == "capsule":
        m.set_pose(other, tr(0, 0, 1))
    else:
        m.set_pose(other, tr(1, 0, 0))
    assert len(m.contact_test(capi.TEST_CLOSEST)) == 0
    m.set_margins(0.251)
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "box_link")
    if kind == "capsule":
        near(r["distance"], 0.125, 1e-3)
        near(r["point"][i0][2], 0.5, 1e-3)
        near(r["point"][i1][2], 0.625, 1e-3)
        near(r["normal"][2], sg * 1.0, 1.1e-3)
    else:
        near(r["distance"], 0.25, 1e-3)
        near(r["point"][i0][0], 0.5, 1e-3)
        near(r["point"][i1][0], 0.75, 1e-3)
        near(r["normal"][0], sg * 1.0, 1e-3)
        if kind == "cone":
            near(r["point"][i0][2], -0.125, 1e-3)
            near(r["point"][i1][2], -0.125, 1e-3)


def mesh_mesh(make):
    """collision_mesh_mesh_unit.hpp:126-214 (no setActiveCollisionObjects => all active)"""
    b = S.SceneBuilder()
    b.add_collision_object("sphere_link", 0, [sphere_mesh()], [I()])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object("sphere1_link", 0, [sphere_mesh()], [I()])
    b.set_active([])
    b.set_margin(0.0)
    m = make(b, False)
    m.set_pose("sphere_link", I())
    m.set_pose("sphere1_link", tr(0.2))
    res = m.contact_test(capi.TEST_ALL)
    assert len(res) >= 37
    m.set_pose("sphere1_link", tr(1, 0, 0))
    assert len(m.contact_test(capi.TEST_ALL)) == 0
    m.set_margins(0.55)
    res = m.contact_test(capi.TEST_CLOSEST)
    assert len(res) == 1
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], 0.52448, 1e-3)
    near(r["point"][i0][0], 0.23776, 1e-3)
    near(r["point"][i1][0], 0.76224, 1e-3)
    n = sg * np.asarray(r["normal"], dtype=np.float64)
    assert n[0] > 0 and abs(np.arccos(min(1.0, n[0]))) < 1e-5
    m.set_pose("sphere1_link", tr(0, 1, 0))
    res = m.contact_test(capi.TEST_CLOSEST)
    r, i0, i1, sg = oriented(res, b, "sphere_link")
    near(r["distance"], 0.5, 1e-3)
    near(r["point"][i0][1], 0.25, 1e-3)
    near(r["point"][i1][1], 0.75, 1e-3)
    n = sg * np.asarray(r["normal"], dtype=np.float64)
    assert n[1] > 0 and abs(np.arccos(min(1.0, n[1]))) < 1e-5


def octomap_sphere(make, use_convex_mesh, tol):
    """collision_octomap_sphere_unit.hpp:91-124 (blender_monkey.bt, 414 194 BOX voxels)"""
    b = S.SceneBuilder()
    b.add_collision_object("octomap_link", 0, [octree("blender_monkey", S.Octree.BOX)], [I()])
    b.add_collision_object("sphere_link", 0, [sphere_hull() if use_convex_mesh else S.Sphere(0.25)], [I()])
    b.set_active(["octomap_link", "sphere_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("octomap_link", I())
    m.set_pose("sphere_link", tr(0, 0, 1))
    res = m.contact_test(capi.TEST_FIRST)
    assert len(res) == 1 and res[0]["distance"] < 0.1
    res = m.contact_test(capi.TEST_CLOSEST)
    assert len(res) == 1
    near(res[0]["distance"], -0.25, tol)
    res = m.contact_test(capi.TEST_ALL)
    assert len(res) > 0 and res[0]["distance"] < 0.1


def octomap_mesh(make):
    """collision_octomap_mesh_unit.hpp:76-119: exactly 2712 contacts"""
    b = S.SceneBuilder()
    b.add_collision_object("octomap_link", 0, [octree("box_2m", S.Octree.BOX)], [I()])
    b.add_collision_object("plane_link", 0, [plane_mesh()], [I()])
    b.set_active(["octomap_link", "plane_link"])
    b.set_margin(0.1)
    m = make(b, False)
    m.set_pose("octomap_link", I())
    m.set_pose("plane_link", I())
    res = m.contact_test(capi.TEST_ALL)
    assert len(res) == 2712


def _two_octomaps(sub0, sub1):
    b = S.SceneBuilder()
    b.add_collision_object("octomap1_link", 0, [octree("box_2m", sub0)], [tr(1.1, 0, 0)])
    b.add_collision_object("octomap2_link", 0, [octree("box_2m", sub1)], [tr(-1.1, 0, 0)])
    b.set_margin(0.25)
    return b


def octomap_octomap(make, cast):
    """collision_octomap_octomap_unit.hpp:80-131"""
    b = _two_octomaps(S.Octree.SPHERE_OUTSIDE, S.Octree.SPHERE_INSIDE)
    b.set_active(["octomap1_link"] if cast else ["octomap1_link", "octomap2_link"])
    m = make(b, cast)

    def run(mm, types):
        for t in types:
            if cast:
                mm.set_pose("octomap1_link", tr(0, -2.0, 0), tr(0, 2.0, 0))
            else:
                mm.set_pose("octomap1_link", I())
                mm.set_pose("octomap2_link", I())
            res = mm.contact_test(t)
            assert len(res) > 0
            for r in res:
                near(r["distance"], -0.0071, 1e-3)

    run(m, (capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL))
    run(m.clone(), (capi.TEST_FIRST,))


def compound_compound(make, cast):
    """collision_compound_compound_unit.hpp:79-130"""
    b = _two_octomaps(S.Octree.BOX, S.Octree.BOX)
    b.set_active(["octomap1_link"] if cast else ["octomap1_link", "octomap2_link"])
    m = make(b, cast)
    for mm in (m, m.clone()):
        if cast:
            mm.set_pose("octomap1_link", tr(0, -2.0, 0), tr(0, 2.0, 0))
        else:
            mm.set_pose("octomap1_link", I())
            mm.set_pose("octomap2_link", I())
        res = mm.contact_test(capi.TEST_ALL)
        assert len(res) > 0
        for r in res:
            near(r["distance"], 0.20, 1e-3)


def large_dataset(make, use_convex_mesh, edge=5, expected=300):
    """collision_large_dataset_unit.hpp:18-97: edge^3 spheres, pitch 0.55, margin 0.1, ALL => 300 link pairs"""
    b = S.SceneBuilder()
    g = sphere_hull() if use_convex_mesh else S.Sphere(0.25)
    names = []
    delta = 0.55
    for x in range(edge):
        for y in range(edge):
            for z in range(edge):
                nm = f"sphere_link_{x}{y}{z}"
                b.add_collision_object(nm, 0, [g], [I()], True, tr(x * delta, y * delta, z * delta))
                names.append(nm)
    b.set_active(names)
    b.set_margin(0.1)
    m = make(b, False)
    res = m.contact_test(capi.TEST_ALL)
    keys = {(int(r["link"][0]), int(r["link"][1])) for r in res}
    assert len(keys) == expected, len(keys)
    return m, b, names


def multi_clone(make, edge=10, expected=2700, clones=4):
    """collision_multi_threaded_unit.hpp:18-136: 10^3 spheres, 4 clones, each shifts all objects, 2700 pairs each"""
    b = S.SceneBuilder()
    names = []
    delta = 0.55
    poses = {}
    for x in range(edge):
        for y in range(edge):
            for z in range(edge):
                nm = f"sphere_link_{x}{y}{z}"
                poses[nm] = tr(x * delta, y * delta, z * delta)
                b.add_collision_object(nm, 0, [S.Sphere(0.25)], [I()], True, poses[nm])
                names.append(nm)
    b.set_active(names)
    b.set_margin(0.1)
    m = make(b, False)
    for t in range(clones):
        c = m.clone()
        shift = 0.25 if t == 0 else (-0.25 if t == 1 else (0.5 if t == 2 else -0.5))
        for nm in names:
            P = poses[nm].copy()
            P[0, 3] += shift
            c.set_pose(nm, P)
        res = c.contact_test(capi.TEST_ALL)
        keys = {(int(r["link"][0]), int(r["link"][1])) for r in res}
        assert len(keys) == expected, len(keys)


# ------------------------------------------------------------------------------------------------ cast
def box_box_cast(make):
    """collision_box_box_cast_unit.hpp:113-174"""
    b = S.SceneBuilder()
    b.add_collision_object("static_box_link", 0, [S.Box(1, 1, 1)], [I()])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object("moving_box_link", 0, [S.Box(0.25, 0.25, 0.25)], [tr(0.5, -0.5, 0)])
    b.set_active(["moving_box_link"])
    b.set_margin(0.1)
    m = make(b, True)
    m.set_pose("static_box_link", I())
    m.set_pose("moving_box_link", tr(-1.9, 0.0, 0), tr(1.9, 3.8, 0))
    for t in (capi.TEST_ALL, capi.TEST_CLOSEST, capi.TEST_FIRST):
        res = m.contact_test(t)
        assert len(res) > 0
        r = res[0]
        assert b.link_names[int(r["link"][1])] == "moving_box_link"  # cast link is index 1 (bullet_utils.h:595-607)
        near(r["distance"], -0.2475, 1e-3)
        near(r["cc_time"][0], -1.0, 1e-3)
        near(r["cc_time"][1], 0.25, 1e-3)
        assert int(r["cc_type"][0]) == capi.CC_NONE and int(r["cc_type"][1]) == capi.CC_BETWEEN
        for k, v in enumerate((-0.5, 0.5, 0.0)):
            near(r["point"][0][k], v, 1e-3)
        for k, v in enumerate((-0.325, 0.325, 0.0)):
            near(r["point"][1][k], v, 1e-3)
        if "point_local" in r.dtype.names:
            T = np.asarray(r["transform"][1]).reshape(3, 4)
            p0 = T[:, :3] @ r["point_local"][1] + T[:, 3]
            for k, v in enumerate((-1.275, -0.625, 0.0)):
                near(p0[k], v, 1e-3)
            T = np.asarray(r["cc_transform"][1]).reshape(3, 4)
            p1 = T[:, :3] @ r["point_local"][1] + T[:, 3]
            for k, v in enumerate((2.525, 3.175, 0.0)):
                near(p1[k], v, 1e-3)


def sphere_sphere_cast(make, use_convex_mesh):
    """collision_sphere_sphere_cast_unit.hpp:136-435"""
    b = S.SceneBuilder()
    g = (lambda: sphere_hull()) if use_convex_mesh else (lambda: S.Sphere(0.25))
    b.add_collision_object("sphere_link", 0, [g()], [tr(0, 0, 0.25)])
    b.add_collision_object("thin_box_link", 0, [S.Box(0.1, 1, 1)], [I()], enabled=False)
    b.add_collision_object("sphere1_link", 0, [g()], [tr(0, 0, 0.25)])
    b.set_active(["sphere_link", "sphere1_link"])
    b.set_margin(0.1)
    m = make(b, True)
    expect = [(-1.0, -0.1, 0.5, 0.5), (-0.5, -0.1, 0.3333, 0.5)]
    if use_convex_mesh:
        expect = [(-1.0, -0.0754, 0.5, 0.5), (-0.5, -0.0754, 0.3848, 0.5)]
    for y0, d, t0, t1 in expect:
        m.set_pose("sphere_link", tr(-0.2, y0, 0), tr(-0.2, 1.0, 0))
        m.set_pose("sphere1_link", tr(0.2, 0, -1.0), tr(0.2, 0, 1.0))
        res = m.contact_test(capi.TEST_CLOSEST)
        assert len(res) > 0
        r, i0, i1, sg = oriented(res, b, "sphere_link")
        near(r["distance"], d, 1e-3)
        if use_convex_mesh and y0 == -0.5:
            # The two swept hulls touch along a flat strip |y| <= 0.077253 (edge of the icosphere swept along z), so
            # the witness y -- and with it cc_time = (y - y0)/(y1 - y0) (bullet_utils.h:509-522) -- is not unique.
            # The reference pins Bullet's EPA pick y = +0.0772 (cc_time 0.3848, collision_sphere_sphere_cast_unit.hpp
            # :402-412); any point of the strip is an equally valid answer, so check consistency.
            y = float(r["point"][i0][1])
            assert -0.0773 - 1e-3 <= y <= 0.0773 + 1e-3
            near(r["cc_time"][i0], (y - y0) / (1.0 - y0), 1e-3)
            assert (0.0772 - y0) / (1.0 - y0) >= r["cc_time"][i0] - 1e-3  # reference value is the strip's end
        else:
            near(r["cc_time"][i0], t0, 1e-3)
        near(r["cc_time"][i1], t1, 1e-3)
        assert int(r["cc_type"][i0]) == capi.CC_BETWEEN and int(r["cc_type"][i1]) == capi.CC_BETWEEN
        if not use_convex_mesh:
            for k, v in enumerate((0.05, 0.0, 0.25)):
                near(r["point"][i0][k], v, 1e-3)
            for k, v in enumerate((-0.05, 0.0, 0.25)):
                near(r["point"][i1][k], v, 1e-3)
            if "point_local" in r.dtype.names:
                for k, v in enumerate((0.25, 0.0, 0.25)):
                    near(r["point_local"][i0][k], v, 1e-3)
                for k, v in enumerate((-0.25, 0.0, 0.25)):
                    near(r["point_local"][i1][k], v, 1e-3)
        else:
            near(r["point"][i0][0], 0.0377, 1e-3)
        near(r["normal"][0], sg * 1.0, 1e-3)
        near(r["normal"][1], 0.0, 1e-3)
        near(r["normal"][2], 0.0, 1e-3)
