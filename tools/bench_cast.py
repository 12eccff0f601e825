#!/usr/bin/env python
"""BASELINE.json configs[2]: KUKA iiwa continuous cast-hull checks, n_traj trajectories x 100 waypoints with LVS
subdivision (tesseract_environment/core/utils.h:172-297), through the C ABI.  Unit of work = LVS sub-segment.

  value : t4c_check_cast_device on pre-expanded sub-segments already in HBM (FP64 row-major q0 / q1)
  e2e   : t4c_check_trajectories_cast on host trajectories (H2D + LVS subdivision kernels + check + D2H)
Not part of the driver's bench contract (bench.py is); prints one JSON line."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def lvs_expand(traj, lvs):
    """sub-segments (q0, q1) of every waypoint segment: cnt = ceil(dist / lvs) + 1 points when dist > lvs"""
    a, b = traj[:, :-1].reshape(-1, traj.shape[2]), traj[:, 1:].reshape(-1, traj.shape[2])
    dist = np.linalg.norm(b - a, axis=1)
    sub = np.where(dist > lvs, np.ceil(dist / lvs).astype(np.int64), 1)
    idx = np.repeat(np.arange(len(a)), sub)
    k = np.arange(len(idx)) - np.repeat(np.cumsum(sub) - sub, sub)
    n = sub[idx].astype(np.float64)
    d = b[idx] - a[idx]
    q0 = a[idx] + d * (k / n)[:, None]
    q1 = a[idx] + d * ((k + 1) / n)[:, None]
    return np.ascontiguousarray(q0), np.ascontiguousarray(q1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--traj", type=int, default=16384)
    ap.add_argument("--wp", type=int, default=100)
    ap.add_argument("--lvs", type=float, default=0.05)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--mode", default="first", choices=["first", "closest"])
    args = ap.parse_args()
    import torch
    from tesseract_b200 import capi, scene as S
    lib = capi.load_library()
    b = S.build_config2(n_prims=10000, margin=0.05)
    desc, fkd = b.scene_desc(), b.fk_desc()
    scene, model = C.c_void_p(), C.c_void_p()
    capi.check(lib.t4c_scene_create(C.byref(desc), 0, C.byref(scene)), lib)
    capi.check(lib.t4c_fk_model_create(C.byref(fkd), scene, C.byref(model)), lib)
    traj = S.random_trajectories(b, args.traj, args.wp, seed=7, sigma=0.02)
    q0, q1 = lvs_expand(traj, args.lvs)
    n = len(q0)
    tt = 0 if args.mode == "first" else 1
    req = capi.Request(tt, 1, 1)
    dev = torch.device("cuda", 0)
    q0d, q1d = torch.from_numpy(q0).to(dev), torch.from_numpy(q1).to(dev)
    hit = torch.zeros(n, dtype=torch.uint8, device=dev)
    cap = 2 * n if tt == 0 else 16 * n
    con = torch.empty(cap * 80, dtype=torch.uint8, device=dev)

    def step_device():
        io = capi.DeviceIO()
        io.q, io.q_layout = q0d.data_ptr(), capi.Q_F64_ROWMAJOR
        io.hit, io.contacts, io.capacity = hit.data_ptr(), con.data_ptr(), cap
        capi.check(lib.t4c_check_cast_device(scene, model, n, q1d.data_ptr(), C.byref(req), C.byref(io)), lib)
        return io

    hit_h = np.zeros(n, dtype=np.uint8)
    con_h = np.zeros(cap, dtype=capi.CONTACT_DTYPE)

    def step_host():
        res = capi.Result()
        res.hit = hit_h.ctypes.data_as(C.POINTER(C.c_uint8))
        res.contacts, res.capacity = con_h.ctypes.data, cap
        items = C.c_int64()
        capi.check(lib.t4c_check_trajectories_cast(scene, model, traj.ctypes.data_as(C.POINTER(C.c_double)), args.traj,
                                                   args.wp, args.lvs, C.byref(req), C.byref(res), C.byref(items)), lib)
        return res, items.value

    lib.t4c_profile_enable(1)
    for _ in range(args.warmup):
        step_device()
    lib.t4c_profile_reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        io = step_device()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    prof = {k: v[0] / args.steps for k, v in capi.profile_get(lib).items() if v[1]}
    lib.t4c_profile_enable(0)
    step_host()
    t = time.perf_counter()
    res, items = step_host()
    ms_host = 1e3 * (time.perf_counter() - t)
    print(json.dumps({"metric": "cast (continuous) LVS sub-segments checked per second", "value": n / (ms * 1e-3),
                      "unit": "sub-segments/s", "trajectories_per_s": args.traj / (ms * 1e-3), "ms_per_step": ms,
                      "config": {"workload": f"configs[2]: iiwa cast-hull checks, {args.traj} trajectories x {args.wp} "
                                             f"waypoints, LVS {args.lvs} rad -> {n} sub-segments, 10k-primitive "
                                             "environment, margin 0.05 m", "request": args.mode},
                      "e2e": {"value": items / (ms_host * 1e-3), "unit": "sub-segments/s", "ms_per_step": ms_host,
                              "items": items, "note": "H2D of the trajectories + LVS subdivision kernels + check + D2H"},
                      "kernel_ms_per_step": prof, "stats": io.stats.as_dict(), "contacts": int(io.n_contacts)}))


if __name__ == "__main__":
    main()
