"""CPU tier: the hull support tables (t4_host_scene.h buildSupportTable / t4_shapes.h coreSupport) return exactly the
vertex of the plain arg-max scan (btConvexHullShape::localGetSupportVertexWithoutMarginNonVirtual semantics: first
index with the maximum projection), for random, axis-aligned, cell-boundary and degenerate directions."""
import ctypes as C

import numpy as np
import pytest

from tesseract_b200 import scene as S
from tesseract_b200.managers import create_scene

from hostsim import build as hostsim


def directions(rng, n):
    d = rng.normal(size=(n, 3))
    d *= 10.0 ** rng.uniform(-6, 3, size=(n, 1))  # GJK directions are not normalised
    R = 16
    edges = np.arange(-R // 2, R // 2 + 1) * (2.0 / R)
    grid = []
    for m in range(3):  # cube-map cell boundaries and corners, both signs, and one ulp either side
        for sgn in (-1.0, 1.0):
            for u in edges:
                for v in edges[::3]:
                    x = np.empty(3)
                    x[m] = sgn
                    x[[i for i in range(3) if i != m]] = (u, v)
                    grid += [x, np.nextafter(x, 2.0), np.nextafter(x, -2.0)]
    special = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (-1, 0, 0), (0, -1, 0), (0, 0, -1), (1, 1, 1), (-1, -1, -1),
               (1, 1, 0), (1, -1, 0), (0, 0, 0), (1e-300, 0, 0), (0, -1e-310, 1e-310), (1e30, 1e29, -1e30)]
    return np.ascontiguousarray(np.vstack([d, np.array(grid), np.array(special, dtype=np.float64)]))


@pytest.mark.parametrize("robot", ["iiwa14", "abb_irb2400"])
def test_table_equals_scan(robot):
    lib = hostsim.lib()
    b = S.SceneBuilder()
    S.add_robot(b, S.load_robot(robot), convex=True)
    b.set_margin(0.05)
    h = create_scene(lib, b)
    d = directions(np.random.default_rng(5), 20000)
    desc = b.scene_desc()
    n_hulls = 0
    for sh in range(desc.n_shapes):
        if desc.shapes[sh].type != 5:
            continue
        out = np.zeros((len(d), 4), dtype=np.int32)
        mean = C.c_double()
        rc = lib.t4c_debug_hull_support(h.h, sh, d.ctypes.data_as(C.POINTER(C.c_double)), len(d),
                                        out.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(mean))
        assert rc == 0, lib.t4c_last_error()
        assert mean.value > 0, "robot hulls must carry a support table"
        assert mean.value < 0.25 * desc.shapes[sh].vertex_count
        assert (out >= 0).all()
        assert np.array_equal(out[:, 0], out[:, 1]), f"FP64 table != scan, shape {sh}"
        # FP32: the scan may round two near-equal projections differently from the exact order only at ~1e-7 relative;
        # the table must still return a vertex whose projection equals the scan's maximum to FP32 accuracy
        bad = np.nonzero(out[:, 2] != out[:, 3])[0]
        assert len(bad) < 1e-3 * len(d)
        n_hulls += 1
        print(robot, sh, desc.shapes[sh].vertex_count, "verts, mean cell length", round(mean.value, 2),
              "fp32 index mismatches", len(bad))
    assert n_hulls >= 7
