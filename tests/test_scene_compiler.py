"""C++ scene compiler (include/t4c_compiler.h) against independent references:
  convex hull     scipy / Qhull vertex sets on random clouds and on the robots' link meshes
  octomap .bt     the leaf tables under tests/golden (written by the Python fixture tool) -- values AND order
  URDF / SRDF     the JSON robot fixtures (links, joints, limits, ACM, hull vertex sets), then the compiled scene runs
                  through the host simulation of the kernels and must give the contacts of the Python-built scene.
The file-based cases read tesseract_support under /root/reference and are skipped where it does not exist (GPU box)."""
import ctypes as C
import os

import numpy as np
import pytest
from scipy.spatial import ConvexHull

from tesseract_b200 import capi, compiler, scene as S

REF = "/root/reference/tesseract_support"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
needs_ref = pytest.mark.skipif(not os.path.isdir(REF), reason="reference data not present")


def test_convex_hull_matches_qhull():
    rng = np.random.default_rng(11)
    for n in (4, 8, 30, 200, 2000):
        for kind in range(3):
            p = rng.normal(size=(n, 3))
            if kind == 1:
                p /= np.linalg.norm(p, axis=1)[:, None]  # every point on the hull
            if kind == 2:
                p = np.round(p * 3) / 3  # many coplanar / duplicate points
            mine = set(compiler.convex_hull_vertices(p).tolist())
            ref = set(ConvexHull(p).vertices.tolist())
            # duplicates / points ON a facet may be kept by one side only: compare the geometry, not the index sets
            hull_mine = ConvexHull(p[sorted(mine)])
            assert abs(hull_mine.volume - ConvexHull(p).volume) <= 1e-9 * max(1.0, hull_mine.volume)
            extra = mine - ref
            eq = ConvexHull(p).equations
            for i in extra:  # an extra vertex must lie on the hull's boundary
                assert np.max(eq[:, :3] @ p[i] + eq[:, 3]) > -1e-9


def test_robot_link_hulls():
    for robot in ("iiwa14", "abb_irb2400"):
        r = S.load_robot(robot)
        for l in r["links"]:
            if l["hull"] is None:
                continue
            v = np.asarray(l["mesh_vertices"], dtype=np.float64)
            mine = v[compiler.convex_hull_vertices(v)]
            ref = np.asarray(l["hull"], dtype=np.float64)
            assert abs(ConvexHull(mine).volume - ConvexHull(ref).volume) <= 1e-12
            assert len(mine) >= len(ref)


@needs_ref
@pytest.mark.parametrize("name", ["box_2m", "blender_monkey"])
def test_octomap_reader(name):
    mine = compiler.read_octomap_bt(os.path.join(REF, "meshes", name + ".bt"))
    ref = np.load(os.path.join(GOLD, name + "_octree.npz"))["leaves"].astype(np.float64)
    assert mine.shape == ref.shape
    assert np.abs(mine - ref).max() <= 1e-6  # the fixture stores float32


def _by_name(desc):
    return {desc.links[i].name.decode(): i for i in range(desc.n_links)}


@needs_ref
@pytest.mark.parametrize("robot,urdf", [("iiwa14", "lbr_iiwa_14_r820"), ("abb_irb2400", "abb_irb2400")])
def test_urdf_compiler_matches_fixture(robot, urdf):
    cs = compiler.CompiledScene(os.path.join(REF, "urdf", urdf + ".urdf"), os.path.join(REF, "urdf", urdf + ".srdf"),
                                package_root=os.path.dirname(REF), convex=True, margin=0.05)
    sd, fd = cs.scene_desc(), cs.fk_desc()
    r = S.load_robot(robot)
    names = _by_name(sd)
    assert set(names) == {l["name"] for l in r["links"]}
    # joints: same origins / axes / limits, keyed by child link
    jr = {j["child"]: j for j in r["joints"]}
    lo, hi = cs.limits()
    for k in range(fd.n_joints):
        j = fd.joints[k]
        child = sd.links[j.child_link].name.decode()
        ref = jr[child]
        assert sd.links[j.parent_link].name.decode() == ref["parent"]
        assert np.allclose(np.array(j.origin[:]).reshape(3, 4), np.asarray(ref["origin"]), atol=1e-12)
        if ref["type"] != "fixed":
            assert np.allclose(j.axis[:], ref["axis"], atol=1e-12)
            assert abs(lo[j.q_index] - ref["lower"]) < 1e-12 and abs(hi[j.q_index] - ref["upper"]) < 1e-12
    # ACM
    acm = {frozenset((sd.links[sd.acm_pairs[2 * i]].name.decode(), sd.links[sd.acm_pairs[2 * i + 1]].name.decode()))
           for i in range(sd.n_acm_pairs)}
    assert acm == {frozenset(p) for p in r["acm"]}
    # hull vertex sets
    V = np.ctypeslib.as_array(sd.vertices, shape=(sd.n_vertices, 3))
    for i in range(sd.n_shapes):
        s = sd.shapes[i]
        link = sd.links[s.link].name.decode()
        ref = np.asarray([l for l in r["links"] if l["name"] == link][0]["hull"], dtype=np.float64)
        mine = V[s.vertex_offset:s.vertex_offset + s.vertex_count]
        assert s.type == capi.HULL
        assert abs(ConvexHull(mine).volume - ConvexHull(ref).volume) <= 1e-12
    # active links: below a moving joint
    active = {sd.links[i].name.decode() for i in range(sd.n_links) if sd.links[i].active}
    assert "base_link" not in active and "link_1" in active and "tool0" in active


@needs_ref
def test_compiled_scene_gives_the_same_contacts():
    """iiwa from the URDF by the C++ compiler + the config-2 primitives added as objects == the Python-built scene"""
    from hostsim import build as hostsim
    from tesseract_b200.managers import BatchChecker
    lib = hostsim.lib()
    b = S.build_config2(n_prims=2000, margin=0.05)
    cs = compiler.CompiledScene(os.path.join(REF, "urdf", "lbr_iiwa_14_r820.urdf"),
                                os.path.join(REF, "urdf", "lbr_iiwa_14_r820.srdf"),
                                package_root=os.path.dirname(REF), convex=True, margin=0.05)
    for s in b.shapes:
        name = b.link_names[s["link"]]
        if name.startswith("env_"):
            cs.add_object(name, [s], pose=b.link_pose[s["link"]])
    sd, fd = cs.scene_desc(), cs.fk_desc()
    scene, model = C.c_void_p(), C.c_void_p()
    capi.check(lib.t4c_scene_create(C.byref(sd), 0, C.byref(scene)), lib)
    capi.check(lib.t4c_fk_model_create(C.byref(fd), scene, C.byref(model)), lib)
    # the same joint vectors in both column orders (URDF order on both sides for this robot)
    q = S.random_states(b, 300, seed=8, dtype=np.float64)
    n = len(q)
    hit = np.zeros(n, dtype=np.uint8)
    con = np.zeros(64 * n, dtype=capi.CONTACT_DTYPE)
    res = capi.Result()
    res.hit = hit.ctypes.data_as(C.POINTER(C.c_uint8))
    res.contacts, res.capacity = con.ctypes.data, len(con)
    req = capi.Request(capi.TEST_CLOSEST, 1, 1)
    capi.check(lib.t4c_check_discrete(scene, model, q.ctypes.data_as(C.POINTER(C.c_double)), n, C.byref(req),
                                      C.byref(res)), lib)
    con = con[:int(res.n_contacts)]
    hit_p, con_p, _ = BatchChecker(b, lib=lib).check_discrete(q, capi.TEST_CLOSEST)
    assert np.array_equal(hit, hit_p)
    nm = {i: sd.links[i].name.decode() for i in range(sd.n_links)}
    mine = {(int(c["state"]), frozenset((nm[c["link"][0]], nm[c["link"][1]]))): float(c["distance"]) for c in con}
    ref = {(int(c["state"]), frozenset((b.link_names[c["link"][0]], b.link_names[c["link"][1]]))): float(c["distance"])
           for c in con_p}
    assert set(mine) == set(ref)
    assert max(abs(mine[k] - ref[k]) for k in mine) <= 1e-6  # hull vertex ORDER may differ (support ties), not the hull
    lib.t4c_fk_model_destroy(model)
    lib.t4c_scene_destroy(scene)


def _write_bt(path, occupied, res=0.1, depth=16):
    """minimal writer of the octomap binary format: a tree with the given occupied depth-16 voxels
    (integer keys per axis, 0 .. 2^16 - 1), everything else unknown; children of a node in index order 0..7,
    two bits per child at (2c, 2c + 1): (0, 1) occupied leaf, (1, 1) inner node, (0, 0) unknown"""
    tree = {}
    for k in occupied:
        node = tree
        for d in range(depth - 1, -1, -1):
            c = (((k[0] >> d) & 1) << 0) | (((k[1] >> d) & 1) << 1) | (((k[2] >> d) & 1) << 2)
            node = node.setdefault(c, {})
    out = bytearray()
    n_nodes = [1]

    def emit(node):
        bits = 0
        for c in range(8):
            if c in node:
                bits |= (0b11 if node[c] else 0b10) << (2 * c)
                n_nodes[0] += 1
        out.extend(int(bits).to_bytes(2, "little"))
        for c in range(8):
            if c in node and node[c]:
                emit(node[c])

    emit(tree)
    with open(path, "wb") as f:
        f.write(b"# Octomap OcTree binary file\nid OcTree\nsize %d\nres %g\ndata\n" % (n_nodes[0], res))
        f.write(bytes(out))


def test_urdf_octomap_element(tmp_path):
    """<octomap shape_type=...><octree filename=.bt/></octomap> in a URDF collision (tesseract_urdf/src/octomap.cpp):
    one box / sphere per occupied leaf in the collision's frame; self-contained (writes its own .bt)"""
    half = 1 << 15
    keys = [(half + 3, half + 1, half), (half - 2, half + 7, half + 4), (half, half, half), (half + 3, half + 1, half + 1)]
    bt = tmp_path / "tiny.bt"
    _write_bt(str(bt), keys, res=0.1)
    leaves = compiler.read_octomap_bt(str(bt))
    assert leaves.shape == (4, 4)
    want = sorted(tuple((np.array(k) - half + 0.5) * 0.1) for k in keys)
    assert np.abs(np.array(sorted(map(tuple, leaves[:, :3]))) - np.array(want)).max() < 1e-9
    assert np.allclose(leaves[:, 3], 0.1)
    for shape_type, t4_type, r in (("box", capi.BOX, 0.05), ("sphere_inside", capi.SPHERE, 0.05),
                                   ("sphere_outside", capi.SPHERE, np.sqrt(2 * 0.05 ** 2))):
        urdf = tmp_path / f"map_{shape_type}.urdf"
        urdf.write_text(f"""<robot name="m"><link name="world"/><link name="map">
  <collision><origin xyz="1 2 3" rpy="0 0 1.5707963267948966"/>
    <geometry><octomap shape_type="{shape_type}"><octree filename="{bt}"/></octomap></geometry></collision></link>
  <joint name="j" type="fixed"><parent link="world"/><child link="map"/></joint></robot>""")
        cs = compiler.CompiledScene(str(urdf), None, convex=True)
        sd = cs.scene_desc()
        assert sd.n_shapes == 4
        li = cs.link_index("map")
        for i in range(4):
            sh = sd.shapes[i]
            assert sh.link == li and sh.type == t4_type and sh.subshape_id == i
            assert abs(sh.params[0] - r) < 1e-12
            # leaf centre (x, y, z) in the collision frame, rotated by 90 degrees about z and shifted by (1, 2, 3)
            x, y, z = leaves[i, :3]
            p = np.array([sh.pose[3], sh.pose[7], sh.pose[11]])
            assert np.abs(p - np.array([1 - y, 2 + x, 3 + z])).max() < 1e-9
    bad = tmp_path / "bad.urdf"
    bad.write_text(f"""<robot name="m"><link name="map"><collision><geometry><octomap shape_type="cube">
      <octree filename="{bt}"/></octomap></geometry></collision></link></robot>""")
    with pytest.raises(compiler.CompilerError):
        compiler.CompiledScene(str(bad), None)
