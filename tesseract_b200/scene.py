"""Host-side scene assembly: the Python mirror of how Tesseract feeds a contact manager.

Mirrors (names and argument meaning) the calls an Environment makes when it populates a
manager clone (tesseract_environment/src/core/environment.cpp:613-640):
    addCollisionObject(name, mask_id, shapes, shape_poses, enabled)
    setActiveCollisionObjects(names)
    setCollisionMarginData(default, pair overrides)
    setIsContactAllowedFn  (here: explicit allowed-pair list = AllowedCollisionMatrix entries)
and flattens the geometry with the reference's conversion rules
(tesseract_collision/src/bullet/bullet_utils.cpp:59-365) into the t4c_scene_desc PODs of
include/t4c.h.  Used by tests, bench.py and smoke(); pure numpy + ctypes.
"""
import ctypes as C
import json
import math
import os

import numpy as np

from . import capi

DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


# --------------------------------------------------------------------------- geometry PODs
SHAPE_DTYPE = np.dtype([("type", "<i4"), ("link", "<i4"), ("shape_id", "<i4"), ("subshape_id", "<i4"),
                        ("params", "<f8", 3), ("vertex_offset", "<i4"), ("vertex_count", "<i4"), ("pose", "<f8", 12)])
assert SHAPE_DTYPE.itemsize == C.sizeof(capi.Shape)


class Geometry:
    pass


class Box(Geometry):
    def __init__(self, x, y, z):  # full edge lengths (tesseract_geometry::Box)
        self.x, self.y, self.z = float(x), float(y), float(z)


class Sphere(Geometry):
    def __init__(self, r):
        self.r = float(r)


class Cylinder(Geometry):
    def __init__(self, r, l):
        self.r, self.l = float(r), float(l)


class Cone(Geometry):
    def __init__(self, r, l):
        self.r, self.l = float(r), float(l)


class Capsule(Geometry):
    def __init__(self, r, l):
        self.r, self.l = float(r), float(l)


class ConvexMesh(Geometry):
    def __init__(self, vertices):
        self.vertices = np.ascontiguousarray(vertices, dtype=np.float64).reshape(-1, 3)


class Mesh(Geometry):
    def __init__(self, vertices, faces):
        self.vertices = np.ascontiguousarray(vertices, dtype=np.float64).reshape(-1, 3)
        self.faces = np.ascontiguousarray(faces, dtype=np.int64).reshape(-1, 3)


class Octree(Geometry):
    BOX, SPHERE_INSIDE, SPHERE_OUTSIDE = 0, 1, 2

    def __init__(self, leaves, sub_type=0):
        """leaves: (n,4) = centre xyz + edge size, in octomap leaf-iterator order"""
        self.leaves = np.ascontiguousarray(leaves, dtype=np.float64).reshape(-1, 4)
        self.sub_type = sub_type


def identity():
    return np.hstack([np.eye(3), np.zeros((3, 1))])


def pose(xyz=(0, 0, 0), R=None):
    T = identity()
    if R is not None:
        T[:, :3] = np.asarray(R, dtype=np.float64)
    T[:, 3] = xyz
    return T


def compose(A, B):
    A = np.asarray(A, dtype=np.float64).reshape(3, 4)
    B = np.asarray(B, dtype=np.float64).reshape(3, 4)
    out = np.empty((3, 4))
    out[:, :3] = A[:, :3] @ B[:, :3]
    out[:, 3] = A[:, :3] @ B[:, 3] + A[:, 3]
    return out


def quat_to_R(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def rot_z(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1.0]])


# --------------------------------------------------------------------------- builder
class SceneBuilder:
    """Accumulates links / shapes / joints and emits ctypes descriptions (keeps buffers alive)."""

    def __init__(self):
        self.link_names = []
        self.link_index = {}
        self.link_type_id = []
        self.link_enabled = []
        self.link_active = []
        self.link_pose = []
        self.shapes = []      # dicts
        self.vertices = []    # list of (k,3) arrays
        self.n_vertices = 0
        self.acm = set()
        self.default_margin = 0.0
        self.pair_margins = {}
        self.joints = []      # dicts
        self.n_q = 0
        self.joint_names = []
        self.joint_limits = []
        self._keep = []

    # -- manager-like API
    def add_link(self, name, pose_=None):
        """Declare a link without geometry (e.g. tool0)."""
        if name in self.link_index:
            return self.link_index[name]
        idx = len(self.link_names)
        self.link_index[name] = idx
        self.link_names.append(name)
        self.link_type_id.append(0)
        self.link_enabled.append(1)
        self.link_active.append(0)
        self.link_pose.append(identity() if pose_ is None else np.asarray(pose_, dtype=np.float64).reshape(3, 4))
        return idx

    def add_collision_object(self, name, mask_id, shapes, shape_poses, enabled=True, pose_=None):
        """DiscreteContactManager::addCollisionObject (discrete_contact_manager.h:75-79)."""
        if not shapes or len(shapes) != len(shape_poses):
            return False  # bullet_utils.h:926-930
        li = self.add_link(name, pose_)
        if pose_ is not None:
            self.link_pose[li] = np.asarray(pose_, dtype=np.float64).reshape(3, 4)
        self.link_type_id[li] = int(mask_id)
        self.link_enabled[li] = 1 if enabled else 0
        self.shapes = [s for s in self.shapes if s["link"] != li]  # re-adding replaces
        single = len(shapes) == 1 and np.array_equal(np.asarray(shape_poses[0]).reshape(3, 4), identity())
        for j, (g, P) in enumerate(zip(shapes, shape_poses)):
            P = np.asarray(P, dtype=np.float64).reshape(3, 4)
            plain_sub = -1 if single else j  # compound child index becomes subshape_id (bullet_utils.h:422-423)
            if isinstance(g, Box):
                self._shape(capi.BOX, li, j, plain_sub, (g.x / 2, g.y / 2, g.z / 2), P)
            elif isinstance(g, Sphere):
                self._shape(capi.SPHERE, li, j, plain_sub, (g.r, 0, 0), P)
            elif isinstance(g, Cylinder):
                self._shape(capi.CYLINDER, li, j, plain_sub, (g.r, g.l / 2, 0), P)
            elif isinstance(g, Cone):
                self._shape(capi.CONE, li, j, plain_sub, (g.r, g.l, 0), P)
            elif isinstance(g, Capsule):
                self._shape(capi.CAPSULE, li, j, plain_sub, (g.r, g.l / 2, 0), P)
            elif isinstance(g, ConvexMesh):
                off = self._verts(g.vertices)
                self._shape(capi.HULL, li, j, plain_sub, (0, 0, 0), P, off, len(g.vertices))
            elif isinstance(g, Mesh):
                tri = g.vertices[g.faces.reshape(-1)]  # one triangle shape per face (bullet_utils.cpp:111-132)
                off = self._verts(tri)
                for f in range(len(g.faces)):
                    self._shape(capi.TRIANGLE, li, j, f, (0, 0, 0), P, off + 3 * f, 3)
            elif isinstance(g, Octree):
                # bullet_utils.cpp:160-243: one box (half = size/2) / inscribed sphere / circumscribed-in-plane sphere
                # (r = sqrt(2) size/2) per occupied leaf; child pose = P * Translation(centre)
                centres = g.leaves[:, :3] @ P[:, :3].T + P[:, 3]
                half = g.leaves[:, 3] * 0.5
                for v in range(len(g.leaves)):
                    Pv = P.copy()
                    Pv[:, 3] = centres[v]
                    h = float(half[v])
                    if g.sub_type == Octree.BOX:
                        self._shape(capi.BOX, li, j, v, (h, h, h), Pv)
                    elif g.sub_type == Octree.SPHERE_INSIDE:
                        self._shape(capi.SPHERE, li, j, v, (h, 0, 0), Pv)
                    else:
                        self._shape(capi.SPHERE, li, j, v, (math.sqrt(2 * (h * h)), 0, 0), Pv)
            else:
                raise TypeError(f"unsupported geometry {type(g)}")
        return True

    def _verts(self, v):
        off = self.n_vertices
        self.vertices.append(np.asarray(v, dtype=np.float64).reshape(-1, 3))
        self.n_vertices += len(self.vertices[-1])
        return off

    def _shape(self, typ, link, shape_id, sub, params, P, voff=0, vcount=0):
        self.shapes.append(dict(type=typ, link=link, shape_id=shape_id, subshape_id=sub, params=params, pose=P,
                                vertex_offset=voff, vertex_count=vcount))

    def set_active(self, names):
        """setActiveCollisionObjects: empty list => everything active (core/common.h:66-69)."""
        for i, n in enumerate(self.link_names):
            self.link_active[i] = 1 if (not names or n in names) else 0

    def set_link_pose(self, name, P):
        self.link_pose[self.link_index[name]] = np.asarray(P, dtype=np.float64).reshape(3, 4)

    def allow_collision(self, a, b):
        ia, ib = self.link_index[a], self.link_index[b]
        self.acm.add((min(ia, ib), max(ia, ib)))

    def set_margin(self, default, pairs=None):
        self.default_margin = float(default)
        self.pair_margins = {}
        for (a, b), v in (pairs or {}).items():
            ia, ib = self.link_index[a], self.link_index[b]
            self.pair_margins[(min(ia, ib), max(ia, ib))] = float(v)

    def add_joint(self, name, jtype, parent, child, origin, axis=(0, 0, 1), lower=0.0, upper=0.0):
        q_index = -1
        if jtype in ("revolute", "continuous", "prismatic"):
            q_index = self.n_q
            self.n_q += 1
            self.joint_names.append(name)
            self.joint_limits.append((lower, upper))
        t = {"fixed": capi.JOINT_FIXED, "revolute": capi.JOINT_REVOLUTE, "continuous": capi.JOINT_REVOLUTE,
             "prismatic": capi.JOINT_PRISMATIC}[jtype]
        self.joints.append(dict(type=t, parent=self.add_link(parent), child=self.add_link(child), q_index=q_index,
                                origin=np.asarray(origin, dtype=np.float64).reshape(3, 4), axis=axis))

    # -- emit ctypes
    def scene_desc(self):
        n_links = len(self.link_names)
        links = (capi.Link * max(n_links, 1))()
        names = [n.encode() for n in self.link_names]
        for i in range(n_links):
            links[i].name = names[i]
            links[i].type_id = self.link_type_id[i]
            links[i].enabled = self.link_enabled[i]
            links[i].active = self.link_active[i]
            links[i].pose[:] = self.link_pose[i].reshape(-1).tolist()
        # t4c_shape records through a numpy mirror of the struct (a million voxels fill in seconds, not minutes)
        ns = len(self.shapes)
        arr = np.zeros(max(ns, 1), dtype=SHAPE_DTYPE)
        if ns:
            for f in ("type", "link", "shape_id", "subshape_id", "vertex_offset", "vertex_count"):
                arr[f][:ns] = np.fromiter((s[f] for s in self.shapes), dtype=np.int32, count=ns)
            arr["params"][:ns] = np.array([s["params"] for s in self.shapes], dtype=np.float64).reshape(ns, 3)
            arr["pose"][:ns] = np.array([s["pose"] for s in self.shapes], dtype=np.float64).reshape(ns, 12)
        shapes = (capi.Shape * max(ns, 1)).from_buffer(arr)
        verts = np.ascontiguousarray(np.vstack(self.vertices) if self.vertices else np.zeros((0, 3)))
        acm = np.ascontiguousarray(np.array(sorted(self.acm), dtype=np.int32).reshape(-1, 2))
        mp = np.ascontiguousarray(np.array(sorted(self.pair_margins.keys()), dtype=np.int32).reshape(-1, 2))
        mv = np.ascontiguousarray(np.array([self.pair_margins[k] for k in sorted(self.pair_margins.keys())],
                                           dtype=np.float64))
        d = capi.SceneDesc()
        d.n_links, d.links = n_links, C.cast(links, C.POINTER(capi.Link))
        d.n_shapes, d.shapes = len(self.shapes), C.cast(shapes, C.POINTER(capi.Shape))
        d.n_vertices, d.vertices = len(verts), verts.ctypes.data_as(C.POINTER(C.c_double))
        d.n_acm_pairs, d.acm_pairs = len(acm), acm.ctypes.data_as(C.POINTER(C.c_int32))
        d.default_margin = self.default_margin
        d.n_margin_pairs = len(mp)
        d.margin_pairs = mp.ctypes.data_as(C.POINTER(C.c_int32))
        d.margin_values = mv.ctypes.data_as(C.POINTER(C.c_double))
        self._keep.append((links, names, shapes, arr, verts, acm, mp, mv))
        return d

    def fk_desc(self, root=None):
        joints = (capi.Joint * max(len(self.joints), 1))()
        # parent-before-child order
        order, placed = [], set()
        children = {j["child"] for j in self.joints}
        roots = [i for i in range(len(self.link_names)) if i not in children]
        placed.update(roots)
        pending = list(range(len(self.joints)))
        while pending:
            progressed = False
            for k in list(pending):
                if self.joints[k]["parent"] in placed:
                    order.append(k)
                    placed.add(self.joints[k]["child"])
                    pending.remove(k)
                    progressed = True
            if not progressed:
                raise ValueError("joint graph is not a forest")
        for n, k in enumerate(order):
            j = self.joints[k]
            joints[n].type, joints[n].parent_link, joints[n].child_link = j["type"], j["parent"], j["child"]
            joints[n].q_index = j["q_index"]
            joints[n].origin[:] = j["origin"].reshape(-1).tolist()
            joints[n].axis[:] = [float(x) for x in j["axis"]]
        d = capi.FkModelDesc()
        d.n_links, d.n_joints, d.joints = len(self.link_names), len(self.joints), C.cast(joints, C.POINTER(capi.Joint))
        d.n_q = self.n_q
        d.root_link = self.link_index[root] if root else (roots[0] if roots else 0)
        self._keep.append(joints)
        return d

    def active_link_indices(self):
        return [i for i in range(len(self.link_names)) if self.link_active[i]]


# --------------------------------------------------------------------------- robots
def load_robot(name):
    """name: 'iiwa14' or 'abb_irb2400' (fixtures made by tools/make_fixtures.py from tesseract_support)."""
    with open(os.path.join(DATA_DIR, name + ".json")) as f:
        return json.load(f)


def add_robot(b, robot, prefix="", base_pose=None, convex=True, active=True):
    """Add a robot the way Environment does: one collision object per link with collision geometry, ACM from
    the SRDF, joints for FK.  convex=True loads meshes as ConvexMesh (tesseract_urdf convex_mesh convert),
    else as triangle Mesh.  Active links = links below a moving joint (core/utils.h:55-76)."""
    root = None
    children = {j["child"] for j in robot["joints"]}
    for l in robot["links"]:
        nm = prefix + l["name"]
        if l["name"] not in children:
            root = nm
        if l["hull"] is None:
            b.add_link(nm)
            continue
        P = np.asarray(l["collision_origin"], dtype=np.float64)
        g = ConvexMesh(l["hull"]) if convex else Mesh(l["mesh_vertices"], l["mesh_faces"])
        b.add_collision_object(nm, 0, [g], [P])
    if base_pose is not None:
        b.set_link_pose(root, base_pose)
    for j in robot["joints"]:
        b.add_joint(prefix + j["name"], j["type"], prefix + j["parent"], prefix + j["child"], j["origin"], j["axis"],
                    j["lower"], j["upper"])
    for a, c in robot["acm"]:
        b.allow_collision(prefix + a, prefix + c)
    # active links: everything downstream of a non-fixed joint
    moving = set()
    changed = True
    while changed:
        changed = False
        for j in robot["joints"]:
            if (j["type"] != "fixed" or j["parent"] in moving) and j["child"] not in moving:
                moving.add(j["child"])
                changed = True
    return root, [prefix + m for m in moving]


def random_environment(b, n_prims=10000, seed=42, extent=3.0, height=3.0, keepout=0.25, prefix="env_"):
    """Config-2 environment: 40 % boxes, 30 % spheres, 30 % cylinders, size U(0.01,0.05), each its own
    static link, centres uniform in [-extent,extent]^2 x [0,height] outside the cylinder r<keepout around the base,
    orientations uniform quaternions.  RNG: numpy MT19937(seed)."""
    rng = np.random.Generator(np.random.MT19937(seed))
    names = []
    i = 0
    while i < n_prims:
        c = np.array([rng.uniform(-extent, extent), rng.uniform(-extent, extent), rng.uniform(0, height)])
        if math.hypot(c[0], c[1]) < keepout:
            continue
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        kind = rng.uniform()
        s = rng.uniform(0.01, 0.05, size=3)
        if kind < 0.4:
            g = Box(2 * s[0], 2 * s[1], 2 * s[2])
        elif kind < 0.7:
            g = Sphere(s[0])
        else:
            g = Cylinder(s[0], 2 * s[1])
        nm = f"{prefix}{i:05d}"
        b.add_collision_object(nm, 0, [g], [identity()], True, pose(c, quat_to_R(q)))
        names.append(nm)
        i += 1
    return names


def random_states(b, n, seed=1234, dtype=np.float32):
    """n joint vectors uniform in the URDF limits; returned row-major (n, n_q)."""
    rng = np.random.Generator(np.random.MT19937(seed))
    lo = np.array([l for l, _ in b.joint_limits])
    hi = np.array([h for _, h in b.joint_limits])
    q = rng.uniform(lo, hi, size=(n, len(lo)))
    return np.ascontiguousarray(q.astype(dtype))


def random_trajectories(b, n_traj, n_wp, seed=7, sigma=0.02):
    """Config-3 trajectories: start uniform in limits, random walk N(0, sigma^2) per joint per step, clamped."""
    rng = np.random.Generator(np.random.MT19937(seed))
    lo = np.array([l for l, _ in b.joint_limits])
    hi = np.array([h for _, h in b.joint_limits])
    q0 = rng.uniform(lo, hi, size=(n_traj, 1, len(lo)))
    steps = rng.normal(0.0, sigma, size=(n_traj, n_wp - 1, len(lo)))
    traj = np.concatenate([q0, steps], axis=1)
    traj = np.cumsum(traj, axis=1)
    return np.ascontiguousarray(np.clip(traj, lo, hi))


def build_config2(n_prims=10000, margin=0.05, env_seed=42):
    """KUKA iiwa 14 convex-hull links vs the 10k box/sphere/cylinder environment (BASELINE.json configs[1])."""
    b = SceneBuilder()
    robot = load_robot("iiwa14")
    root, moving = add_robot(b, robot)
    random_environment(b, n_prims, env_seed)
    b.set_active(moving)
    b.set_margin(margin)
    return b


def build_config4(octree_leaves, plane=None, margin=0.10, robot_xyz=(0.0, -2.0, 0.0), sub_type=Octree.BOX):
    """BASELINE.json configs[3] style: ABB IRB2400 convex-hull links vs an octomap-derived voxel environment (one box
    or sphere per occupied leaf, bullet_utils.cpp:160-243) plus a concave workcell mesh (one triangle shape per
    face).  octree_leaves: (n, 4) centre + size; plane: optional (vertices, faces)."""
    b = SceneBuilder()
    root, moving = add_robot(b, load_robot("abb_irb2400"), base_pose=pose(robot_xyz))
    b.add_collision_object("octomap", 0, [Octree(octree_leaves, sub_type)], [identity()])
    if plane is not None:
        b.add_collision_object("workcell", 0, [Mesh(plane[0], plane[1])], [pose((0.0, 0.0, -0.05))])
    b.set_active(moving)
    b.set_margin(margin)
    return b


def synthetic_room_voxels(n_voxels=1000000, res=0.01, seed=99):
    """SURVEY 8(d) config 4 occupancy: ONE layer of voxels on the inner surfaces of a room plus the surfaces of six box
    fixtures around the robot -- what an octomap of a scanned workcell holds -- thinned at random (seed) to exactly
    n_voxels.  The survey's 4 x 4 x 2.5 m room has 180 k surface cells at its 0.02 m: a million leaves need 0.01 m
    cells and a 4.4 x 4.4 x 2.6 m room (1.04 M cells before thinning).  Returns (n, 4) centre + edge."""
    rng = np.random.Generator(np.random.MT19937(seed))
    sx, sy, sz = 4.4, 4.4, 2.6
    cells = []

    def face(u0, u1, v0, v1, fixed, axis):
        """cell centres of an axis-aligned rectangle: axis = the constant coordinate"""
        nu, nv = int(round((u1 - u0) / res)), int(round((v1 - v0) / res))
        u = u0 + (np.arange(nu) + 0.5) * res
        v = v0 + (np.arange(nv) + 0.5) * res
        U, V = np.meshgrid(u, v, indexing="ij")
        W = np.full_like(U, fixed)
        order = {0: (W, U, V), 1: (U, W, V), 2: (U, V, W)}[axis]
        cells.append(np.stack([a.reshape(-1) for a in order], axis=1))

    h = res / 2
    face(-sx / 2, sx / 2, -sy / 2, sy / 2, -h, 2)          # floor (top of the cells at z = 0)
    face(-sx / 2, sx / 2, -sy / 2, sy / 2, sz + h, 2)      # ceiling
    face(-sy / 2, sy / 2, 0.0, sz, -sx / 2 - h, 0)         # walls
    face(-sy / 2, sy / 2, 0.0, sz, sx / 2 + h, 0)
    face(-sx / 2, sx / 2, 0.0, sz, -sy / 2 - h, 1)
    face(-sx / 2, sx / 2, 0.0, sz, sy / 2 + h, 1)
    fixtures = [((1.05, 0.0), (0.6, 0.7, 0.9)), ((-1.0, 0.3), (0.7, 0.6, 0.8)), ((0.2, 1.1), (0.8, 0.5, 1.0)),
                ((0.0, -1.15), (0.7, 0.7, 0.7)), ((1.2, 1.2), (0.6, 0.6, 1.2)), ((-1.2, -1.1), (0.7, 0.7, 1.1))]
    for (cx, cy), (fx, fy, fz) in fixtures:
        face(cx - fx / 2, cx + fx / 2, cy - fy / 2, cy + fy / 2, fz - h, 2)   # top
        face(cy - fy / 2, cy + fy / 2, 0.0, fz, cx - fx / 2 + h, 0)           # four sides
        face(cy - fy / 2, cy + fy / 2, 0.0, fz, cx + fx / 2 - h, 0)
        face(cx - fx / 2, cx + fx / 2, 0.0, fz, cy - fy / 2 + h, 1)
        face(cx - fx / 2, cx + fx / 2, 0.0, fz, cy + fy / 2 - h, 1)
    allc = np.concatenate(cells)
    if len(allc) < n_voxels:
        raise ValueError(f"only {len(allc)} occupied cells")
    idx = np.sort(rng.choice(len(allc), size=n_voxels, replace=False))
    return np.concatenate([allc[idx], np.full((n_voxels, 1), res)], axis=1)


def build_config4_synthetic(n_voxels=1000000, margin=0.10, plane=None, seed=99):
    """BASELINE.json configs[3]: ABB IRB2400 hull links vs n_voxels synthetic octomap voxels -- half as boxes, half as
    SPHERE_OUTSIDE (bullet_utils.cpp:182-183,228) -- plus the concave workcell mesh (one triangle per face)."""
    leaves = synthetic_room_voxels(n_voxels, seed=seed)
    b = SceneBuilder()
    root, moving = add_robot(b, load_robot("abb_irb2400"), base_pose=pose((0.0, 0.0, 0.0)))
    b.add_collision_object("octomap_boxes", 0, [Octree(leaves[0::2], Octree.BOX)], [identity()])
    b.add_collision_object("octomap_spheres", 0, [Octree(leaves[1::2], Octree.SPHERE_OUTSIDE)], [identity()])
    if plane is not None:
        b.add_collision_object("workcell", 0, [Mesh(plane[0], plane[1])], [pose((0.0, 0.0, 0.001))])
    b.set_active(moving)
    b.set_margin(margin)
    return b


def lbracket_mesh(seed):
    """small concave fixture: an L-bracket as a closed triangle mesh (two boxes sharing an edge), ~24 triangles"""
    rng = np.random.default_rng(seed)
    a, t, w = rng.uniform(0.15, 0.3), rng.uniform(0.02, 0.04), rng.uniform(0.1, 0.2)
    boxes = [((0, 0, 0), (a, w, t)), ((0, 0, 0), (t, w, a))]
    V, F = [], []
    for (x0, y0, z0), (dx, dy, dz) in boxes:
        o = len(V)
        V += [(x0 + i * dx, y0 + j * dy, z0 + k * dz) for i in (0, 1) for j in (0, 1) for k in (0, 1)]
        quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
        for q in quads:
            F += [(o + q[0], o + q[1], o + q[2]), (o + q[0], o + q[2], o + q[3])]
    return np.array(V, dtype=np.float64), np.array(F, dtype=np.int64)


def config5_fixtures():
    """the four concave fixtures of configs[4] (seed 5.., between the two arms)"""
    out = []
    for i in range(4):
        v, f = lbracket_mesh(5 + i)
        out.append((f"fixture_{i}", v, f, pose((0.45 * (1 if i % 2 else -1), 0.3 * (1 if i < 2 else -1), 0.35))))
    return out


def build_config5(n_prims=2000, margin=0.05, fixtures=None):
    """BASELINE.json configs[4] style: two iiwa (bases at y = -/+0.45 m, the second yawed by pi), 14 joints, cross-arm
    pairs enabled, optional concave fixture meshes [(name, vertices, faces, pose)], plus the config-2 primitives."""
    b = SceneBuilder()
    robot = load_robot("iiwa14")
    _, m0 = add_robot(b, robot, prefix="a_", base_pose=pose((0.0, -0.45, 0.0)))
    _, m1 = add_robot(b, robot, prefix="b_", base_pose=pose((0.0, 0.45, 0.0), rot_z(math.pi)))
    for name, v, f, P in fixtures or []:
        b.add_collision_object(name, 0, [Mesh(v, f)], [identity()], pose_=P)
    random_environment(b, n_prims, 42, keepout=0.8)
    b.set_active(m0 + m1)
    b.set_margin(margin)
    return b
