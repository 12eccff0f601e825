"""Independent second opinion for the GJK / EPA restatements: exact distance / penetration depth of random
polytope pairs from the convex hull of their Minkowski difference (scipy / qhull), no iteration tolerances.

  origin outside conv{a - b}: distance = min over the hull's triangles of the point-triangle distance;
  origin inside:              depth = min facet offset, normal = that facet's normal (minimum translation).

Spheres enter as a point core plus radius (distance = core distance - r), the way the reference treats them
(bullet_utils.cpp:256-307).  Used against BOTH the CPU oracle and the device path (tests/test_second_opinion.py), so
the two restatements -- written by one author from one source -- are checked against arithmetic neither of them shares.
"""
import numpy as np
from scipy.spatial import ConvexHull

from tesseract_b200 import scene as S


def random_rotation(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    return S.quat_to_R(q)


def random_polytope(rng, n, size):
    """n points on a randomly stretched sphere: every one is a hull vertex, spacing >> the detector's 1e-6 merge radius"""
    p = rng.normal(size=(n, 3))
    p /= np.linalg.norm(p, axis=1)[:, None]
    return p * rng.uniform(0.4, 1.0, size=3) * size


def box_vertices(h):
    return np.array([[sx * h[0], sy * h[1], sz * h[2]] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)])


def _tri_dist_origin(a, b, c):
    """distance of the origin from triangles (vectorised over the leading axis); Ericson 5.1.5"""
    ab, ac, ap = b - a, c - a, -a
    d1, d2 = np.einsum("ij,ij->i", ab, ap), np.einsum("ij,ij->i", ac, ap)
    bp = -b
    d3, d4 = np.einsum("ij,ij->i", ab, bp), np.einsum("ij,ij->i", ac, bp)
    cp = -c
    d5, d6 = np.einsum("ij,ij->i", ab, cp), np.einsum("ij,ij->i", ac, cp)
    vc, vb, va = d1 * d4 - d3 * d2, d5 * d2 - d1 * d6, d3 * d6 - d5 * d4
    out = np.empty(len(a))
    pt = np.empty_like(a)
    done = np.zeros(len(a), dtype=bool)

    def put(mask, p):
        m = mask & ~done
        pt[m] = p[m]
        done[m] = True

    put((d1 <= 0) & (d2 <= 0), a)
    put((d3 >= 0) & (d4 <= d3), b)
    with np.errstate(divide="ignore", invalid="ignore"):
        put((vc <= 0) & (d1 >= 0) & (d3 <= 0), a + ab * (d1 / (d1 - d3))[:, None])
        put((d6 >= 0) & (d5 <= d6), c)
        put((vb <= 0) & (d2 >= 0) & (d6 <= 0), a + ac * (d2 / (d2 - d6))[:, None])
        put((va <= 0) & (d4 - d3 >= 0) & (d5 - d6 >= 0), b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)))[:, None])
        den = 1.0 / (va + vb + vc)
        put(np.ones(len(a), dtype=bool), a + ab * (vb * den)[:, None] + ac * (vc * den)[:, None])
    out[:] = np.linalg.norm(pt, axis=1)
    return out, pt


def exact_pair(va_world, vb_world, ra=0.0, rb=0.0):
    """(signed distance, unit normal from B towards A or None when not unique enough to compare)"""
    if len(va_world) == 1 and len(vb_world) == 1:
        v = va_world[0] - vb_world[0]
        d = np.linalg.norm(v)
        return d - ra - rb, v / d
    m = (va_world[:, None, :] - vb_world[None, :, :]).reshape(-1, 3)
    hull = ConvexHull(m)
    eq = hull.equations
    if np.all(eq[:, 3] <= 0):  # origin inside: penetration
        k = int(np.argmax(eq[:, 3]))
        depth = -eq[k, 3]
        # minimum translation that separates A from B moves A along +n: the contact normal (B -> A) is -n ... the
        # reference reports normalOnB pointing from B to A, and for penetration that is the facet normal reversed
        second = np.partition(eq[:, 3], -2)[-2] if len(eq) > 1 else -np.inf
        unique = (-second - depth) > 1e-6 or np.allclose(eq[np.argsort(eq[:, 3])[-2], :3], eq[k, :3], atol=1e-9)
        return -(depth + ra + rb), (-eq[k, :3] if unique else None)
    tri = hull.points[hull.simplices]
    d, pt = _tri_dist_origin(tri[:, 0], tri[:, 1], tri[:, 2])
    k = int(np.argmin(d))
    return d[k] - ra - rb, (pt[k] / d[k] if d[k] > 1e-9 else None)


class PairFamily:
    """one (shape A on an active link, shape B on a static link) scene + n random relative poses"""

    def __init__(self, kind_a, kind_b, rng, size=0.08):
        self.b = S.SceneBuilder()
        self.ra = self.rb = 0.0
        self.va, ga = self._make(kind_a, rng, size)
        self.vb, gb = self._make(kind_b, rng, size)
        if kind_a == "sphere":
            self.ra = ga.r
        if kind_b == "sphere":
            self.rb = gb.r
        self.b.add_collision_object("A", 0, [ga], [S.identity()])
        self.b.add_collision_object("B", 0, [gb], [S.identity()])
        self.b.set_active(["A"])
        self.b.set_margin(0.25)
        self.size = size

    @staticmethod
    def _make(kind, rng, size):
        if kind == "hull":
            v = random_polytope(rng, int(rng.integers(8, 40)), size)
            return v, S.ConvexMesh(v)
        if kind == "box":
            h = rng.uniform(0.3, 1.0, size=3) * size
            return box_vertices(h), S.Box(*(2 * h))
        if kind == "sphere":
            r = float(rng.uniform(0.3, 1.0) * size)
            return np.zeros((1, 3)), S.Sphere(r)
        raise ValueError(kind)

    def poses(self, rng, n):
        """world poses of link A: centre offsets from deep penetration to clearly separated"""
        T = np.zeros((n, 2, 3, 4))
        T[:, 1] = S.identity()
        for i in range(n):
            T[i, 0, :, :3] = random_rotation(rng)
            d = rng.normal(size=3)
            d /= np.linalg.norm(d)
            T[i, 0, :, 3] = d * rng.uniform(0.0, 2.6) * self.size
        return T

    def truth(self, T):
        out = []
        for t in T:
            va = self.va @ t[0, :, :3].T + t[0, :, 3]
            out.append(exact_pair(va, self.vb, self.ra, self.rb))
        return out
