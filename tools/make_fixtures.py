#!/usr/bin/env python3
"""Generate the committed data fixtures from the read-only reference checkout.

Runs ONLY in the build container (needs /root/reference); the GPU box never
sees /root/reference, so everything the tests / bench need is written as
small data files under tesseract_b200/data/ (robots) and tests/golden/
(meshes and octomaps used by the reference's known-answer tests).

What it reads (data, not code):
  tesseract_support/urdf/{lbr_iiwa_14_r820,abb_irb2400}.{urdf,srdf} + collision STLs
  tesseract_support/meshes/{sphere_p25m.ply,box2_2m.ply,box_2m.bt,blender_monkey.bt,plane_4m.stl}

Robot fixture format (JSON, see tesseract_b200/scene.py:load_robot):
  links[]  : name, parent joint index (-1 root), hull vertices (float32-exact
             STL coordinates of the convex-hull vertex subset), triangle mesh
             (vertices + faces, deduplicated) for the "Mesh" variant
  joints[] : name, type, parent/child link, origin 3x4 (row-major), axis, limits
  acm[]    : SRDF disable_collisions pairs

"Convex-hull links" follow tesseract_urdf/src/convex_mesh.cpp:80-100 /
tesseract_collision/core/common.h:258-264 (makeConvexMesh): the hull vertex
set of the mesh.  We take the hull with scipy (Qhull) on the exact STL
vertices -- Bullet's btConvexHullComputer grid snapping is NOT reproduced
(SURVEY Appendix A.8); every engine in this repo receives these vertices.
"""
import json
import math
import os
import struct
import sys
import xml.etree.ElementTree as ET

import numpy as np
from scipy.spatial import ConvexHull

REF = "/root/reference/tesseract_support"
OUT_ROBOT = os.path.join(os.path.dirname(__file__), "..", "tesseract_b200", "data")
OUT_GOLD = os.path.join(os.path.dirname(__file__), "..", "tests", "golden")


def read_stl(path):
    with open(path, "rb") as f:
        data = f.read()
    if data[:5] == b"solid" and b"facet" in data[:400]:
        verts = []
        for line in data.decode("ascii", "ignore").splitlines():
            t = line.split()
            if len(t) == 4 and t[0] == "vertex":
                verts.append([float(t[1]), float(t[2]), float(t[3])])
        tri = np.asarray(verts, dtype=np.float32).reshape(-1, 3, 3)
    else:
        n = struct.unpack_from("<I", data, 80)[0]
        rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n, offset=84)
        tri = rec["v"].copy()
    flat = tri.reshape(-1, 3)
    uniq, inv = np.unique(flat, axis=0, return_inverse=True)
    faces = inv.reshape(-1, 3)
    return uniq.astype(np.float32), faces.astype(np.int32)


def read_ply_ascii(path):
    with open(path) as f:
        lines = f.read().splitlines()
    nv = nf = 0
    i = 0
    while lines[i] != "end_header":
        t = lines[i].split()
        if t[:2] == ["element", "vertex"]:
            nv = int(t[2])
        if t[:2] == ["element", "face"]:
            nf = int(t[2])
        i += 1
    i += 1
    v = np.array([[float(x) for x in lines[i + k].split()[:3]] for k in range(nv)], dtype=np.float64)
    faces = []
    for k in range(nf):
        t = [int(x) for x in lines[i + nv + k].split()]
        faces.append(t[1:1 + t[0]])
    return v, faces


def rpy_to_R(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    # URDF fixed-axis roll-pitch-yaw: R = Rz(y) Ry(p) Rx(r)
    return [[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
            [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
            [-sp, cp * sr, cp * cr]]


def origin_3x4(el):
    xyz = [0.0, 0.0, 0.0]
    rpy = [0.0, 0.0, 0.0]
    if el is not None:
        if el.get("xyz"):
            xyz = [float(x) for x in el.get("xyz").split()]
        if el.get("rpy"):
            rpy = [float(x) for x in el.get("rpy").split()]
    R = rpy_to_R(*rpy)
    return [R[0] + [xyz[0]], R[1] + [xyz[1]], R[2] + [xyz[2]]]


def make_robot(urdf_name, srdf_name, out_name):
    urdf = ET.parse(os.path.join(REF, "urdf", urdf_name)).getroot()
    srdf = ET.parse(os.path.join(REF, "urdf", srdf_name)).getroot()
    links = []
    index = {}
    for l in urdf.findall("link"):
        e = {"name": l.get("name"), "hull": None, "mesh_vertices": None, "mesh_faces": None,
             "collision_origin": origin_3x4(None)}
        col = l.find("collision")
        if col is not None:
            mesh = col.find("geometry").find("mesh")
            fn = mesh.get("filename").replace("package://tesseract_support/", "")
            v, f = read_stl(os.path.join(REF, fn))
            hull = ConvexHull(v.astype(np.float64))
            hv = v[np.sort(hull.vertices)]
            e["hull"] = [[float(x) for x in p] for p in hv]
            e["mesh_vertices"] = [[float(x) for x in p] for p in v]
            e["mesh_faces"] = f.tolist()
            e["collision_origin"] = origin_3x4(col.find("origin"))
            e["mesh_file"] = fn
        index[e["name"]] = len(links)
        links.append(e)
    joints = []
    for j in urdf.findall("joint"):
        ax = j.find("axis")
        axis = [float(x) for x in ax.get("xyz").split()] if ax is not None else [1.0, 0.0, 0.0]
        lim = j.find("limit")
        joints.append({
            "name": j.get("name"), "type": j.get("type"),
            "parent": j.find("parent").get("link"), "child": j.find("child").get("link"),
            "origin": origin_3x4(j.find("origin")), "axis": axis,
            "lower": float(lim.get("lower")) if lim is not None and lim.get("lower") else 0.0,
            "upper": float(lim.get("upper")) if lim is not None and lim.get("upper") else 0.0,
        })
    acm = [[d.get("link1"), d.get("link2")] for d in srdf.findall("disable_collisions")]
    robot = {"name": urdf.get("name"), "source": [urdf_name, srdf_name], "links": links, "joints": joints, "acm": acm}
    os.makedirs(OUT_ROBOT, exist_ok=True)
    with open(os.path.join(OUT_ROBOT, out_name), "w") as f:
        json.dump(robot, f, separators=(",", ":"))
    print(out_name, "links", len(links), "joints", len(joints), "acm", len(acm),
          "hull verts", [len(l["hull"]) for l in links if l["hull"]])


def parse_bt(path):
    """Octomap binary tree reader, format per SURVEY Appendix A.9. Returns res, leaves[(x,y,z,size)] in
    depth-first child order 0..7 (octomap leaf-iterator order)."""
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    res = None
    size = None
    while True:
        end = data.index(b"\n", pos)
        line = data[pos:end].decode("ascii")
        pos = end + 1
        if line.startswith("res "):
            res = float(line.split()[1])
        if line.startswith("size "):
            size = int(line.split()[1])
        if line.strip() == "data":
            break
    leaves = []
    count = 0
    depth_max = 16

    sys.setrecursionlimit(10000)

    def rec(pos, key, depth):
        nonlocal count
        count += 1
        bits = data[pos] | (data[pos + 1] << 8)
        pos += 2
        for c in range(8):
            lo = (bits >> (2 * c)) & 1
            hi = (bits >> (2 * c + 1)) & 1
            step = 1 << (depth_max - depth - 1)
            ck = (key[0] + (step if c & 1 else 0), key[1] + (step if c & 2 else 0), key[2] + (step if c & 4 else 0))
            if lo == 1 and hi == 1:
                pos = rec(pos, ck, depth + 1)
            elif lo == 0 and hi == 1:
                count += 1
                edge = res * step
                leaves.append(((ck[0] - 32768) * res + edge / 2, (ck[1] - 32768) * res + edge / 2,
                               (ck[2] - 32768) * res + edge / 2, edge))
            elif lo == 1 and hi == 0:
                count += 1
        return pos

    endpos = rec(pos, (0, 0, 0), 0)
    assert endpos == len(data), (endpos, len(data))
    assert count == size, (count, size)
    return res, np.asarray(leaves, dtype=np.float64)


def make_golden_meshes():
    os.makedirs(OUT_GOLD, exist_ok=True)
    out = {}
    for name in ("sphere_p25m.ply", "box2_2m.ply"):
        v, faces = read_ply_ascii(os.path.join(REF, "meshes", name))
        out[name] = {"vertices": v.tolist(), "faces": faces}
    v, f = read_stl(os.path.join(REF, "meshes", "plane_4m.stl"))
    # plane_4m.stl: keep the ORIGINAL triangle order (subshape ids) -> re-read without dedup of faces
    out_plane = {"vertices": v.astype(np.float64).tolist(), "faces": f.tolist()}
    with open(os.path.join(OUT_GOLD, "meshes.json"), "w") as fo:
        json.dump(out, fo, separators=(",", ":"))
    np.savez_compressed(os.path.join(OUT_GOLD, "plane_4m.npz"), vertices=v.astype(np.float64), faces=f)
    for name in ("box_2m.bt", "blender_monkey.bt"):
        res, leaves = parse_bt(os.path.join(REF, "meshes", name))
        np.savez_compressed(os.path.join(OUT_GOLD, name.replace(".bt", "_octree.npz")), res=res,
                            leaves=leaves.astype(np.float32) if "monkey" in name else leaves)
        print(name, "res", res, "occupied leaves", len(leaves))
    print("plane_4m", v.shape, f.shape, "ply", {k: len(o["vertices"]) for k, o in out.items()})
    del out_plane


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("reference checkout not present; fixtures are already committed")
    make_robot("lbr_iiwa_14_r820.urdf", "lbr_iiwa_14_r820.srdf", "iiwa14.json")
    make_robot("abb_irb2400.urdf", "abb_irb2400.srdf", "abb_irb2400.json")
    make_golden_meshes()
