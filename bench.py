#!/usr/bin/env python
"""Benchmark of the collision-query hot path on BASELINE.json's configurations.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--mode first|closest|first_flags]
                    [--no-secondary] [--no-cpu-baseline]

Headline line (`metric`, `value`, `e2e`, `roofline`, `cpu_baseline`): configs[1] -- KUKA iiwa 14 convex-hull links vs a
10k box / sphere / cylinder environment, 1 Mi random joint states per GPU, discrete contactTest, margin 0.05 m, FIRST
with one contact record per colliding state.  One "step" = one pass of the hot path (FK + broadphase + narrowphase +
result bridging) over the whole batch.  `value` times the pass with the joint vectors already resident in HBM; `e2e`
times the host-buffer C-ABI call (pinned host q -> H2D -> check -> D2H hit flags + contact records).

`secondary` (same JSON line, short step counts; every entry has its own ms_per_step, e2e, stats and parity_sample):
    config1_closest   configs[1], ContactTestType::CLOSEST with full records (SURVEY 8(d) mode B)
    config2_cast      configs[2]: 65 536 trajectories x 100 waypoints, LVS_CONTINUOUS 0.05 rad, cast hulls, CLOSEST
    config2_cast_fine configs[2] on a 4 096-trajectory subset at the reference's default LVS 0.005
    config3_voxels    configs[3]: ABB IRB2400 vs 1 M synthetic octomap voxels + concave mesh, CLOSEST, margin 0.10
    config4_dual_arm  configs[4]: two iiwa + concave fixtures + primitives, 2 Mi states per GPU (16 Mi on 8 GPUs), FIRST
                      and CLOSEST -- multi-GPU runs only (the single-GPU line carries the four entries above)

Multi-GPU: one process per GPU (torchrun), states sharded by contiguous ranges, environment replicated, per-rank
results gathered to rank 0 once per step; weak scaling (fixed states per GPU).

`parity_sample`: the device results of the timed batch against the CPU oracle on a subsample of the same inputs, by the
strict rule of tests/parity.py (computed in the cpu_baseline leg: the oracle is the checker, never the thing timed as
"ours").
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "collision-checked states/sec (discrete, iiwa hull links vs 10k-primitive environment, margin 0.05 m)"
UNIT = "states/s"
MODES = {"first": ("ContactTestType::FIRST, one contact record per colliding state", 0, True),
         "first_flags": ("ContactTestType::FIRST, per-state collision flag only", 0, False),
         "closest": ("ContactTestType::CLOSEST, full contact records", 1, True)}

# SURVEY 8(d) constants of the algorithmic-work formula (flop per event)
F_SIMPLEX, F_SUP_HULL_FIXED, F_SUP_PRIM, F_VERTEX, F_EPA, F_NODE, F_CAND, F_OUT = 100.0, 33.0, 26.0, 6.0, 400.0, 12.0, 30.0, 60.0
FLOP_FORMULA = ("fk: states*(80*moving_joints+63*fixed_joints); broad_static: 12*bvh_nodes+30*candidates; "
                "cull_fp32: cull_iterations*(100+33+26)+6*(support_vertices-support_vertices_fp64); "
                "gjk_fp64: (phase1_iterations+phase2_iterations)*(100+33+26)+6*(support_vertices_fp64-support_vertices_epa); "
                "epa_fp64: epa_iterations*(400+33+26)+epa_gjk_iterations*(100+33+26)+6*support_vertices_epa; "
                "33 = hull-side support without its vertex scan, 26 = mean primitive support (0.4 box 24 + 0.3 sphere 18 + "
                "0.3 cylinder 36), 6 per hull vertex actually evaluated (support tables: a handful, not V)")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="first", choices=sorted(MODES))
    ap.add_argument("--states", type=int, default=1 << 20, help="joint states per GPU")
    ap.add_argument("--prims", type=int, default=10000)
    ap.add_argument("--cpu-sample", type=int, default=0, help="states in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--secondary", default="", help="comma list of secondary entries to run (default: all that apply)")
    ap.add_argument("--cast-traj", type=int, default=65536)
    ap.add_argument("--voxels", type=int, default=1000000)
    ap.add_argument("--voxel-states", type=int, default=1 << 17,
                    help="configs[3] states per GPU (the configuration names 1 Mi; a state has ~1 600 voxel pairs inside "
                         "the 0.10 m margin, so the default bench line times a 128 Ki-state slice of it)")
    ap.add_argument("--dual-states", type=int, default=1 << 21, help="configs[4] states per GPU")
    return ap.parse_args()


def workload(args):
    return {"workload": f"configs[1]: KUKA iiwa 14 convex-hull links vs {args.prims} box/sphere/cylinder environment, "
                        f"{args.states} random joint states per GPU (seed 1234 + rank), discrete, margin 0.05 m",
            "request": MODES[args.mode][0], "states_per_gpu": args.states, "environment_primitives": args.prims,
            "parallelism": f"dp{args.gpus} (states sharded, environment BVH replicated)",
            "l2": "L2 flushed between timed iterations (512 MiB device memset)"}


class ClockSampler:
    """SM clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows, self.stop_flag, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        # in-process NVML first: spawning nvidia-smi every 200 ms initialises NVML on every GPU of the box each time
        # and was measured to skew the ranks of a multi-GPU run by ~5 ms per step (profiles/README, r2f)
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                    ("sw_power_cap", 0x4))
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                nv.nvmlDeviceGetCurrentClocksThrottleReasons
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            while not self.stop_flag:
                r = int(get_reasons(h))
                self.rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(mx), "0"] +
                                 ["Active" if r & b else "Not Active" for _, b in bits])
                time.sleep(0.05)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                parts = [p.strip() for p in out.stdout.strip().split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.2)

    def start(self):
        self.t.start()

    def stop(self):
        self.stop_flag = True
        self.t.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  The reference itself
    cannot be built here (Bullet / Eigen / KDL are not installed, DESIGN.md "Oracle"), so this is the FP64 CPU
    restatement (oracle/), one manager clone per thread -- kind "port"."""
    if rank != 0:
        return
    from oracle import pyoracle
    from tesseract_b200 import scene as S
    mode_name, tt, want = MODES[args.mode]
    b = S.build_config2(n_prims=args.prims, margin=0.05)
    cores = os.cpu_count() or 1
    n = args.cpu_sample or max(cores * 250, 2000)
    q = S.random_states(b, n, seed=1234, dtype=np.float64)
    times = []
    for it in range(args.warmup + args.steps):
        t = time.perf_counter()
        pyoracle.check_discrete(b, q, tt, want_contacts=want, n_threads=cores)
        dt = time.perf_counter() - t
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    val = n / (ms * 1e-3)
    sample = (f"first {n} states of the workload, {cores} threads, one manager clone per thread; CPU restatement "
              "(oracle/), not Bullet: Bullet / FCL are not installed on this box")
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": workload(args),
                      "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---------------------------------------------------------------------------------------------------------------------
# one scene + model on one device, timed through the C ABI
# ---------------------------------------------------------------------------------------------------------------------
class Engine:
    def __init__(self, lib, builder, device, dist_ctx):
        import torch
        from tesseract_b200 import capi
        self.torch, self.capi, self.lib, self.b, self.device = torch, capi, lib, builder, device
        self.rank, self.world, self.dist = dist_ctx
        self.dev = torch.device("cuda", device)
        desc = builder.scene_desc()
        self.scene = C.c_void_p()
        capi.check(lib.t4c_scene_create(C.byref(desc), device, C.byref(self.scene)), lib)
        self.fkd = builder.fk_desc()
        self.model = C.c_void_p()
        capi.check(lib.t4c_fk_model_create(C.byref(self.fkd), self.scene, C.byref(self.model)), lib)
        self.nq = self.fkd.n_q
        self.flush = torch.empty(512 << 20, dtype=torch.uint8, device=self.dev)
        # multi-GPU: the library's own communicator (t4_multi.h); torch.distributed only carries the 128-byte id
        self.comm = None
        if self.world > 1:
            idt = torch.zeros(128, dtype=torch.uint8, device=self.dev)
            if self.rank == 0:
                buf = (C.c_uint8 * 128)()
                capi.check(lib.t4c_comm_unique_id(buf), lib)
                idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
            self.dist.broadcast(idt, src=0)
            buf = (C.c_uint8 * 128)(*idt.cpu().tolist())
            self.comm = C.c_void_p()
            capi.check(lib.t4c_comm_create(buf, self.rank, self.world, device, C.byref(self.comm)), lib)

    def close(self):
        if self.comm is not None:
            self.lib.t4c_comm_destroy(self.comm)
        self.lib.t4c_fk_model_destroy(self.model)
        self.lib.t4c_scene_destroy(self.scene)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps, warmup, after=None):
        """max-over-ranks device time per step of fn() [+ after()], L2 flushed and ranks aligned before every step"""
        torch = self.torch
        for _ in range(warmup):
            fn()
            if after:
                after()
        total, tail, last = 0.0, 0.0, None
        for _ in range(steps):
            self.flush.zero_()
            self.barrier()
            e0, em, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            last = fn()  # synchronous C-ABI call: kernels run on the library's stream and are joined inside
            em.record()
            if after:
                after()
            e1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1), em.elapsed_time(e1), e0.elapsed_time(em)], device=self.dev)
            if self.world > 1:
                self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
            total += float(ms[0].item())
            tail += float(ms[1].item())
        return total / steps, tail / steps, last

    # ---- discrete states -------------------------------------------------------------------------------------------
    def discrete(self, q_host, tt, want, steps, warmup, e2e_steps, contacts_per_state, gather=None, chunk=None):
        """q_host: (n, nq) float64.  Returns timings / stats / a results handle for the parity sample."""
        torch, capi, lib = self.torch, self.capi, self.lib
        n = len(q_host)
        chunk = chunk or n
        cap = int(contacts_per_state * min(chunk, n))
        q_soa = torch.from_numpy(np.ascontiguousarray(q_host.T.astype(np.float32))).to(self.dev)  # FP32 SoA q[j][i]
        q_pin = torch.from_numpy(q_host).pin_memory()
        hit_dev = torch.zeros(n, dtype=torch.uint8, device=self.dev)
        con_dev = torch.empty((cap if want else 1) * 80, dtype=torch.uint8, device=self.dev)
        hit_pin = torch.zeros(n, dtype=torch.uint8).pin_memory()
        con_pin = torch.empty((cap if want else 1) * 80, dtype=torch.uint8).pin_memory()
        req = capi.Request(tt, 1, 1)
        acc = {}
        sharded = self.comm is not None
        root = self.rank == 0
        g_hit = g_con = g_hit_pin = g_con_pin = None
        if sharded and root:
            # the root's gathered buffers: flags of the whole batch, records of every shard
            g_hit = torch.zeros(self.world * n, dtype=torch.uint8, device=self.dev)
            g_con = torch.empty(self.world * (cap if want else 1) * 80, dtype=torch.uint8, device=self.dev)
            g_hit_pin = torch.zeros(self.world * n, dtype=torch.uint8).pin_memory()
            g_con_pin = torch.empty(self.world * (cap if want else 1) * 80, dtype=torch.uint8).pin_memory()
        q_dev64 = torch.empty((n, self.nq), dtype=torch.float64, device=self.dev) if sharded else None
        gathered_n = [0]

        def add_stats(st):
            for k, v in st.as_dict().items():
                acc[k] = acc.get(k, 0) + v

        def call_sharded(io, m, a):
            """this rank's shard [a, a + m) of its n states; one gather to rank 0 inside the entry"""
            g = capi.Gathered()
            if root:
                g.hit = g_hit.data_ptr()
                g.contacts = g_con.data_ptr() if want else None
                g.capacity = self.world * cap if want else 0
            capi.check(lib.t4c_check_discrete_sharded(self.scene, self.model, self.comm, 0, m, self.rank * n + a,
                                                      C.byref(req), C.byref(io), C.byref(g) if root else None), lib)
            if root:
                gathered_n[0] = int(g.n_contacts)

        def step_device():
            acc.clear()
            ncon = 0
            for a in range(0, n, chunk):
                m = min(chunk, n - a)
                io = capi.DeviceIO()
                # SoA column offset: q[j * n + i] -> the chunk reads columns [a, a + m) of every row, so hand it the
                # row-major FP32 slice instead when the batch is cut into chunks
                if chunk >= n:
                    io.q, io.q_layout = q_soa.data_ptr(), capi.Q_F32_SOA
                else:
                    io.q, io.q_layout = q_rm.data_ptr() + a * self.nq * 4, capi.Q_F32_ROWMAJOR
                io.hit = hit_dev.data_ptr() + a
                io.contacts = con_dev.data_ptr() if want else None
                io.capacity = cap if want else 0
                if sharded:
                    call_sharded(io, m, a)
                else:
                    capi.check(lib.t4c_check_discrete_device(self.scene, self.model, m, C.byref(req), C.byref(io)), lib)
                add_stats(io.stats)
                ncon += int(io.n_contacts)
            return ncon

        def step_host_sharded():
            """end to end at N GPUs: pinned host q -> H2D, sharded check with its gather, D2H of the gathered flags
            and records on rank 0"""
            q_dev64.copy_(q_pin, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            ncon = 0
            for a in range(0, n, chunk):
                m = min(chunk, n - a)
                io = capi.DeviceIO()
                io.q, io.q_layout = q_dev64.data_ptr() + a * self.nq * 8, capi.Q_F64_ROWMAJOR
                io.hit = hit_dev.data_ptr() + a
                io.contacts = con_dev.data_ptr() if want else None
                io.capacity = cap if want else 0
                call_sharded(io, m, a)
                ncon += int(io.n_contacts)
                if root:
                    g_hit_pin.copy_(g_hit, non_blocking=True)
                    if want and gathered_n[0] > 0:
                        g_con_pin[:gathered_n[0] * 80].copy_(g_con[:gathered_n[0] * 80], non_blocking=True)
                    torch.cuda.current_stream().synchronize()
            return ncon

        def step_host():
            if sharded:
                return step_host_sharded()
            ncon = 0
            for a in range(0, n, chunk):
                m = min(chunk, n - a)
                res = capi.Result()
                res.hit = C.cast(hit_pin.data_ptr() + a, C.POINTER(C.c_uint8))
                res.contacts = con_pin.data_ptr() if want else None
                res.capacity = cap if want else 0
                capi.check(lib.t4c_check_discrete(self.scene, self.model,
                                                  C.cast(q_pin.data_ptr() + a * self.nq * 8, C.POINTER(C.c_double)), m,
                                                  C.byref(req), C.byref(res)), lib)
                ncon += int(res.n_contacts)
            return ncon

        q_rm = None
        if chunk < n:
            q_rm = torch.from_numpy(np.ascontiguousarray(q_host.astype(np.float32))).to(self.dev)
        after = (lambda: gather(hit_dev)) if (gather and not sharded) else None
        lib.t4c_profile_enable(1)
        for _ in range(warmup):
            step_device()
            if after:
                after()
        lib.t4c_profile_reset()
        launches0 = lib.t4c_kernel_launch_count()
        ms_dev, ms_tail, ncon = self.timed(step_device, steps, 0, after)
        launches = int(lib.t4c_kernel_launch_count() - launches0)
        prof = capi.profile_get(lib)
        lib.t4c_profile_enable(0)
        stats = dict(acc)
        ms_e2e, _, ncon_h = self.timed(step_host, e2e_steps, 1, after)
        # results of the last host call stay in the pinned buffers: the parity sample reads them
        con_view = None
        if want and chunk >= n:
            con_view = np.frombuffer(con_pin.numpy(), dtype=capi.CONTACT_DTYPE)[:ncon_h]
        if sharded and root:
            hit_pin, want_view = g_hit_pin, False  # gathered flags of the whole batch
        return dict(ms=ms_dev, ms_e2e=ms_e2e, exchange_ms=ms_tail, stats=stats, n_contacts=ncon, launches=launches,
                    gathered_contacts=gathered_n[0],
                    kernels={k: v[0] / max(1, steps) for k, v in prof.items() if v[1]},
                    kernel_launches={k: v[1] // max(1, steps) for k, v in prof.items() if v[1]},
                    hit=hit_pin.numpy().copy(), contacts=con_view, h2d=n * self.nq * 8,
                    d2h=(self.world * n + (gathered_n[0] * 80 if want else 0)) if sharded
                    else n + (ncon_h * 80 if want else 0))

    # ---- cast trajectories -----------------------------------------------------------------------------------------
    def cast(self, traj, lvs, tt, steps, warmup, e2e_steps, contacts_per_item, traj_chunk):
        """traj: (n_traj, n_wp, nq) float64.  value: sub-segments already expanded and resident in HBM
        (t4c_check_cast_device); e2e: t4c_check_trajectories_cast with host waypoints (H2D, LVS expansion on the device,
        check, D2H)."""
        torch, capi, lib = self.torch, self.capi, self.lib
        n_traj, n_wp, nq = traj.shape
        DP = C.POINTER(C.c_double)
        req = capi.Request(tt, 1, 1)
        # LVS items per chunk of trajectories (host-only helpers of the library; same visiting order as the reference)
        chunks = []
        total_items = 0
        for a in range(0, n_traj, traj_chunk):
            t = np.ascontiguousarray(traj[a:a + traj_chunk])
            cnt = C.c_int64()
            capi.check(lib.t4c_lvs_count(t.ctypes.data_as(DP), len(t), n_wp, nq, float(lvs), 1, C.byref(cnt)), lib)
            m = int(cnt.value)
            items = np.arange(m, dtype=np.int64)
            q0, q1 = np.empty((m, nq)), np.empty((m, nq))
            capi.check(lib.t4c_lvs_item_states(t.ctypes.data_as(DP), len(t), n_wp, nq, float(lvs), 1, m,
                                               items.ctypes.data_as(C.POINTER(C.c_int64)), q0.ctypes.data_as(DP),
                                               q1.ctypes.data_as(DP)), lib)
            chunks.append(dict(traj=torch.from_numpy(t).pin_memory(), n=m, q0=torch.from_numpy(q0).to(self.dev),
                               q1=torch.from_numpy(q1).to(self.dev), n_traj=len(t)))
            total_items += m
        max_items = max(c["n"] for c in chunks)
        cap = int(contacts_per_item * max_items)
        hit_dev = torch.zeros(max_items, dtype=torch.uint8, device=self.dev)
        con_dev = torch.empty(cap * 80, dtype=torch.uint8, device=self.dev)
        hit_pin = torch.zeros(max_items, dtype=torch.uint8).pin_memory()
        con_pin = torch.empty(cap * 80, dtype=torch.uint8).pin_memory()
        acc = {}
        first = {}

        def step_device():
            acc.clear()
            ncon = 0
            for c in chunks:
                io = capi.DeviceIO()
                io.q, io.q_layout = c["q0"].data_ptr(), capi.Q_F64_ROWMAJOR
                io.hit, io.contacts, io.capacity = hit_dev.data_ptr(), con_dev.data_ptr(), cap
                capi.check(lib.t4c_check_cast_device(self.scene, self.model, c["n"], c["q1"].data_ptr(), C.byref(req),
                                                     C.byref(io)), lib)
                for k, v in io.stats.as_dict().items():
                    acc[k] = acc.get(k, 0) + v
                ncon += int(io.n_contacts)
            return ncon

        def step_host():
            ncon = 0
            for i, c in enumerate(chunks):
                res = capi.Result()
                res.hit = C.cast(hit_pin.data_ptr(), C.POINTER(C.c_uint8))
                res.contacts, res.capacity = con_pin.data_ptr(), cap
                items = C.c_int64()
                capi.check(lib.t4c_check_trajectories_cast(self.scene, self.model,
                                                           C.cast(c["traj"].data_ptr(), DP), c["n_traj"], n_wp,
                                                           float(lvs), C.byref(req), C.byref(res), C.byref(items)), lib)
                if i == 0:  # the parity sample reads the first chunk's results
                    first["hit"] = hit_pin.numpy()[:int(items.value)].copy()
                    first["contacts"] = np.frombuffer(con_pin.numpy(), dtype=capi.CONTACT_DTYPE)[:int(res.n_contacts)].copy()
                ncon += int(res.n_contacts)
            return ncon

        lib.t4c_profile_enable(1)
        for _ in range(warmup):
            step_device()
        lib.t4c_profile_reset()
        launches0 = lib.t4c_kernel_launch_count()
        ms_dev, _, ncon = self.timed(step_device, steps, 0)
        launches = int(lib.t4c_kernel_launch_count() - launches0)
        prof = capi.profile_get(lib)
        lib.t4c_profile_enable(0)
        stats = dict(acc)
        ms_e2e, _, ncon_h = self.timed(step_host, e2e_steps, 0)
        return dict(ms=ms_dev, ms_e2e=ms_e2e, stats=stats, n_contacts=ncon, launches=launches, items=total_items,
                    kernels={k: v[0] / max(1, steps) for k, v in prof.items() if v[1]},
                    hit=first.get("hit"), contacts=first.get("contacts"), first_traj=chunks[0]["n_traj"],
                    h2d=n_traj * n_wp * nq * 8, d2h=total_items + ncon_h * 80)


# ---------------------------------------------------------------------------------------------------------------------
# roofline accounting (counted events x fixed constants; FP64 kernels against a measured DFMA peak)
# ---------------------------------------------------------------------------------------------------------------------
def kernel_flops(stats, n_moving, n_fixed):
    s = stats
    sv32 = s["support_vertices"] - s["support_vertices_fp64"]
    sv_epa = s.get("support_vertices_epa", 0)
    pair = F_SIMPLEX + F_SUP_HULL_FIXED + F_SUP_PRIM
    return {"fk": s["states"] * (80.0 * n_moving + 63.0 * n_fixed),
            "broad_static": F_NODE * s["bvh_nodes"] + F_CAND * s["candidates"],
            "cull_fp32": s["cull_iterations"] * pair + F_VERTEX * sv32,
            "gjk_fp64": (s["phase1_iterations"] + s["phase2_iterations"]) * pair +
                        F_VERTEX * (s["support_vertices_fp64"] - sv_epa),
            "epa_fp64": s["epa_iterations"] * (F_EPA + F_SUP_HULL_FIXED + F_SUP_PRIM) + s["epa_gjk_iterations"] * pair +
                        F_VERTEX * sv_epa}


def roofline(lib, capi, device, r, n_states, nq, want, n_moving, n_fixed, traffic_key):
    kern = r["kernels"]
    if not kern:
        return None
    top = max(kern, key=lambda k: kern[k])
    f32, f64 = C.c_double(), C.c_double()
    capi.check(lib.t4c_measure_fp32_peak(device, C.byref(f32)), lib)
    capi.check(lib.t4c_measure_fp64_peak(device, C.byref(f64)), lib)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        if tj.get("_config") == traffic_key:
            traffic = {k: v for k, v in tj.items() if not k.startswith("_")}
    except Exception:
        pass
    flops = kernel_flops(r["stats"], n_moving, n_fixed)
    fp64_kernels = ("fk", "gjk_fp64", "epa_fp64")
    per = {}
    for k, ms in kern.items():
        if k not in flops or ms <= 0:
            continue
        peak = f64.value if k in fp64_kernels else f32.value
        ach = flops[k] / (ms * 1e-3) / 1e12
        per[k] = {"ms": ms, "gflop": flops[k] / 1e9, "achieved_tflops": ach,
                  "peak_tflops": peak, "precision": "fp64" if k in fp64_kernels else "fp32",
                  "frac": ach / peak if peak else None, "traffic": traffic.get(k)}
    st = r["stats"]
    bytes_state = 4 * nq + 1 + (80.0 * st["contacts"] / max(1, st["states"]) if want else 0.0)
    t = per.get(top, {})
    return {"bound": t.get("precision", "fp32"), "kernel": top, "achieved": t.get("achieved_tflops"),
            "peak": t.get("peak_tflops"), "unit": "TFLOP/s", "frac": t.get("frac"), "traffic": t.get("traffic"),
            "peak_source": "measured in this run: FFMA loop (fp32 kernels) / DFMA loop (fp64 kernels) on every SM, "
                           "t4c_measure_fp32_peak / t4c_measure_fp64_peak; the FP64 kernels are built without FMA "
                           "contraction (one flop per instruction), so 0.5 is their ceiling",
            "flop_formula": FLOP_FORMULA, "fp32_peak_tflops": f32.value, "fp64_peak_tflops": f64.value,
            "kernels": per, "kernel_ms_per_step": kern,
            "traffic_note": "DRAM bytes per step per kernel from this round's ncu --set full capture "
                            "(profiles/traffic.json), null when the capture is of another configuration",
            "hbm": {"achieved": bytes_state * n_states / (r["ms"] * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "algorithmic_bytes_per_state": bytes_state,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"}}


# ---------------------------------------------------------------------------------------------------------------------
# parity samples (rank 0, single-GPU runs): the oracle as the checker, timed as the CPU baseline of the same work
# ---------------------------------------------------------------------------------------------------------------------
def parity_discrete(b, q, tt, r, n_sample, margin, cores):
    """device results of the timed batch vs the oracle on its first n_sample states"""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import parity
    from oracle import pyoracle
    from tesseract_b200 import capi
    ns = min(n_sample, len(q))
    t = time.perf_counter()
    hit_o, con_o, _ = pyoracle.check_discrete(b, q[:ns], tt, n_threads=cores)
    dt = time.perf_counter() - t
    hit_g = r["hit"][:ns]
    out = {"n": ns, "flag_mismatch": int((hit_g != hit_o).sum()), "rule": "tests/parity.py compare()"}
    cpu = {"value": ns / dt, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"first {ns} states, {cores} threads, one manager clone per thread (oracle/, not Bullet)"}
    con = r["contacts"]
    if con is None:
        return out, cpu
    con_g = con[con["state"] < ns]
    try:
        if tt == capi.TEST_FIRST:
            # FIRST is order dependent in the reference: every device record must be a contact the oracle reports
            # for that (state, link pair, shape pair) in ALL mode
            _, con_a, _ = pyoracle.check_discrete(b, q[:ns], capi.TEST_ALL, n_threads=cores)
            A = parity.index(con_a, True)
            G = parity.index(con_g, True)
            dd, miss = 0.0, 0
            for k, v in G.items():
                if k not in A:
                    miss += 0 if abs(v[0] - margin) <= 1e-5 + 4e-7 else 1
                    continue
                dd = max(dd, abs(v[0] - A[k][0]))
            out.update(records=len(G), records_not_in_oracle=miss, max_dd=dd)
            out["ok"] = bool(miss == 0 and dd <= 1e-5 + 4e-7 and out["flag_mismatch"] == 0)
        else:
            T = pyoracle.fk(b, q[:ns])
            rep = parity.compare(b, con_g, con_o, hit_g, hit_o, lambda a, c: margin, link_poses=T)
            out.update({k: rep[k] for k in ("n_gpu", "n_oracle", "band_exempt", "max_dd", "max_dp", "max_dn",
                                            "nonunique", "curved_pen", "curved_max_dn", "curved_dn_hist")})
            out["ok"] = True
    except AssertionError as e:
        out["ok"] = False
        out["violation"] = str(e)[:300]
    return out, cpu


def parity_cast(b, traj, lvs, r, n_traj, margin, cores):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import parity
    from oracle import pyoracle
    from tesseract_b200 import capi
    n_traj = min(n_traj, r["first_traj"])
    q0, q1, _, _ = pyoracle.lvs_expand(traj[:n_traj], lvs, cast=True)
    ns = len(q0)
    t = time.perf_counter()
    hit_o, con_o, _ = pyoracle.check_cast(b, q0, q1, capi.TEST_CLOSEST, n_threads=cores)
    dt = time.perf_counter() - t
    out = {"n": ns, "trajectories": n_traj}
    cpu = {"value": ns / dt, "unit": "sub-segments/s", "cores": cores, "kind": "port",
           "sample": f"the {ns} LVS sub-segments of the first {n_traj} trajectories, {cores} threads (oracle/, not Bullet)"}
    if r["contacts"] is None:
        return out, cpu
    # work items of a chunk are numbered trajectory by trajectory: the first n_traj trajectories are items [0, ns)
    hit_g = r["hit"][:ns]
    con_g = r["contacts"][r["contacts"]["state"] < ns]
    out["flag_mismatch"] = int((hit_g != hit_o).sum())
    try:
        rep = parity.compare(b, con_g, con_o, hit_g, hit_o, lambda a, c: margin, check_cc=True)
        out.update({k: rep[k] for k in ("n_gpu", "n_oracle", "band_exempt", "max_dd", "max_dp", "max_dn", "max_dt",
                                        "nonunique", "curved_pen", "curved_max_dn")})
        out["ok"] = True
    except AssertionError as e:
        out["ok"] = False
        out["violation"] = str(e)[:300]
    return out, cpu


def joints_of(b):
    moving = sum(1 for j in b.joints if j["type"] != 0)
    return moving, len(b.joints) - moving


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch
    import torch.distributed as dist
    from tesseract_b200 import capi, scene as S
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = capi.load_library()
    dev = torch.device("cuda", local_rank)
    cores = os.cpu_count() or 1
    mode_name, tt, want = MODES[args.mode]
    n = args.states
    ctx = (rank, world, dist)
    seed_off = int(os.environ.get("T4_BENCH_SEED_OFFSET", "0"))

    # ---- result exchange of a multi-GPU step: per-rank hit flags to rank 0, one collective
    gathered = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

    def gather_hits(hit_dev):
        if world > 1:
            dist.gather(hit_dev, gathered, dst=0)

    # ================================================================================================================
    # headline: configs[1]
    # ================================================================================================================
    b = S.build_config2(n_prims=args.prims, margin=0.05)
    eng = Engine(lib, b, local_rank, ctx)
    q_host = S.random_states(b, n, seed=1234 + rank + seed_off, dtype=np.float64)
    sampler = ClockSampler(local_rank)
    # NCCL communicator set-up happens in the warm-up, outside the timed region
    sampler_started = [False]

    def run_primary():
        sampler.start()
        sampler_started[0] = True
        return eng.discrete(q_host, tt, want, args.steps, args.warmup, max(1, min(args.steps, 3)),
                            2 if tt == 0 else 12, gather=gather_hits if world > 1 else None)

    r = run_primary()
    clocks = sampler.stop()
    collective_ms = None
    if world > 1:
        hit_tmp = torch.zeros(n, dtype=torch.uint8, device=dev)
        gather_hits(hit_tmp)  # connection set-up of the yardstick collective stays outside its timing
        eng.barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(10):
            gather_hits(hit_tmp)
        c1.record()
        torch.cuda.synchronize()
        collective_ms = c0.elapsed_time(c1) / 10
    nm, nf = joints_of(b)
    roof = roofline(lib, capi, local_rank, r, n, eng.nq, want, nm, nf,
                    {"mode": args.mode, "states": n, "prims": args.prims})
    # per-rank kernel times: which kernel of which rank sets the step time of a multi-GPU run
    per_rank = None
    if world > 1:
        names = list(capi.KERNEL_NAMES)
        mine = torch.tensor([r["kernels"].get(k, 0.0) for k in names] + [r["ms"]], device=dev, dtype=torch.float64)
        allk = [torch.zeros_like(mine) for _ in range(world)] if rank == 0 else None
        dist.gather(mine, allk, dst=0)
        if rank == 0:
            M = torch.stack(allk).cpu().numpy()
            per_rank = {k: {"max_ms": float(M[:, i].max()), "argmax_rank": int(M[:, i].argmax()),
                            "min_ms": float(M[:, i].min())} for i, k in enumerate(names) if M[:, i].max() > 0}
            per_rank["kernels_total_ms_by_rank"] = [float(x) for x in M[:, :len(names)].sum(axis=1)]
    cpu = None
    psample = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ns = args.cpu_sample or max(cores * 250, 2000)
        psample, cpu = parity_discrete(b, q_host, tt, r, ns, 0.05, cores)
    eng.close()

    # ================================================================================================================
    # secondary entries
    # ================================================================================================================
    secondary = {}
    wanted = [s for s in args.secondary.split(",") if s]

    def on(name, default):
        return (name in wanted) if wanted else (default and not args.no_secondary)

    def finish(name, r2, builder, unit, units, want2, note, ps=None, cpu2=None, n_for_roof=None, nq2=7):
        nm2, nf2 = joints_of(builder)
        ms = r2["ms"]
        if world > 1:
            tms = torch.tensor([ms, r2["ms_e2e"]], device=dev, dtype=torch.float64)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms, r2["ms_e2e"] = float(tms[0].item()), float(tms[1].item())
        e = {"workload": note, "value": world * units / (ms * 1e-3), "unit": unit, "ms_per_step": ms,
             "e2e": {"value": world * units / (r2["ms_e2e"] * 1e-3), "unit": unit, "ms_per_step": r2["ms_e2e"],
                     "h2d_bytes_per_step": r2["h2d"], "d2h_bytes_per_step": r2["d2h"]},
             "stats_per_step": r2["stats"], "contacts_per_step": r2["n_contacts"], "gpu_launches": r2["launches"],
             "roofline": roofline(lib, capi, local_rank, r2, n_for_roof or units, nq2, want2, nm2, nf2, None),
             "parity_sample": ps, "cpu_baseline": cpu2}
        secondary[name] = e

    single = world == 1
    check = rank == 0 and single and not args.no_cpu_baseline
    if on("config1_closest", single):
        eng = Engine(lib, b, local_rank, ctx)
        r2 = eng.discrete(q_host, 1, True, 2, 1, 1, 12)
        ps, c2 = parity_discrete(b, q_host, 1, r2, 2048, 0.05, cores) if check else (None, None)
        finish("config1_closest", r2, b, UNIT, n, True,
               f"configs[1] mode B: ContactTestType::CLOSEST with full records, {n} states, margin 0.05", ps, c2)
        eng.close()
    for name, n_traj, lvs, chunk_traj in (("config2_cast", args.cast_traj, 0.05, 8192),
                                          ("config2_cast_fine", min(4096, args.cast_traj), 0.005, 1024)):
        if not on(name, single):
            continue
        eng = Engine(lib, b, local_rank, ctx)
        traj = S.random_trajectories(b, n_traj, 100, seed=7)
        r2 = eng.cast(traj, lvs, 1, 2, 1, 1, 14, chunk_traj)
        ps, c2 = parity_cast(b, traj, lvs, r2, 16 if lvs > 0.01 else 2, 0.05, cores) if check else (None, None)
        finish(name, r2, b, "sub-segments/s", r2["items"], True,
               f"configs[2]: {n_traj} trajectories x 100 waypoints (random walk sigma 0.02 rad, seed 7), LVS_CONTINUOUS "
               f"{lvs} rad -> {r2['items']} cast sub-segments, CLOSEST, margin 0.05; run as calls of {chunk_traj} "
               f"trajectories; also {n_traj / (r2['ms'] * 1e-3):.0f} trajectories/s", ps, c2)
        secondary[name]["trajectories_per_s"] = world * n_traj / (r2["ms"] * 1e-3)
        eng.close()
    if on("config3_voxels", single):
        pl = np.load(os.path.join(ROOT, "tests", "golden", "plane_4m.npz"))
        b4 = S.build_config4_synthetic(args.voxels, margin=0.10, plane=(pl["vertices"], pl["faces"]))
        eng = Engine(lib, b4, local_rank, ctx)
        n4 = args.voxel_states
        q4 = S.random_states(b4, n4, seed=99 + rank, dtype=np.float64)
        r2 = eng.discrete(q4, 1, True, 1, 1, 1, 24, chunk=1 << 16)
        ps = c2 = None
        if check:
            # chunked batches keep no single contact buffer: run the sample through the host entry once more
            from tesseract_b200.managers import BatchChecker
            bc = BatchChecker(b4)
            hit_s, con_s, _ = bc.check_discrete(q4[:256], 1)
            ps, c2 = parity_discrete(b4, q4, 1, dict(hit=hit_s, contacts=con_s), 256, 0.10, cores)
            del bc
        finish("config3_voxels", r2, b4, UNIT, n4, True,
               f"configs[3]: ABB IRB2400 hull links vs {args.voxels} synthetic octomap voxels (one layer of 0.01 m cells "
               f"on the walls of a 4.4 x 4.4 x 2.6 m room and on six box fixtures; half boxes, half SPHERE_OUTSIDE) + "
               f"plane_4m.stl as a triangle mesh, {n4} states (slice of the 1 Mi-state configuration), CLOSEST, "
               f"margin 0.10", ps, c2, nq2=6)
        eng.close()
    if on("config4_dual_arm", not single):
        b5 = S.build_config5(n_prims=args.prims, margin=0.05, fixtures=S.config5_fixtures())
        n5 = args.dual_states
        q5 = S.random_states(b5, n5, seed=2024 + rank, dtype=np.float64)
        for sub, tt5, per in (("first", 0, 2), ("closest", 1, 24)):
            eng = Engine(lib, b5, local_rank, ctx)
            r2 = eng.discrete(q5, tt5, True, 2, 1, 1, per, chunk=1 << 19 if tt5 else 1 << 20)
            ps = c2 = None
            if check:
                from tesseract_b200.managers import BatchChecker
                bc = BatchChecker(b5)
                hit_s, con_s, _ = bc.check_discrete(q5[:512], tt5)
                ps, c2 = parity_discrete(b5, q5, tt5, dict(hit=hit_s, contacts=con_s), 512, 0.05, cores)
                del bc
            finish(f"config4_dual_arm_{sub}", r2, b5, UNIT, n5, True,
                   f"configs[4]: two iiwa (14 joints, cross-arm pairs on) + 4 concave fixtures + {args.prims} "
                   f"primitives, {n5} states per GPU ({world * n5} in total), "
                   f"{'FIRST with records' if tt5 == 0 else 'CLOSEST'}, margin 0.05", ps, c2, nq2=14)
            eng.close()

    if rank == 0:
        out = {"metric": METRIC, "value": world * n / (r["ms"] * 1e-3), "unit": UNIT, "n_gpus": world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms"], "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "f32 cull + f64 exact pass", "data": "synthetic",
               "config": workload(args), "clocks": clocks, "gpu_launches": r["launches"],
               "e2e": {"value": world * n / (r["ms_e2e"] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": r["h2d"],
                       "d2h_bytes_per_step": r["d2h"], "ms_per_step": r["ms_e2e"]},
               "roofline": roof, "cpu_baseline": cpu, "parity_sample": psample, "stats_per_step": r["stats"],
               "mode": args.mode,
               "exchange": {"kind": ("none" if world == 1 else "t4c_check_discrete_sharded: one grouped ncclSend/ncclRecv "
                                     "of hit flags + contact records + counters to rank 0, inside the C-ABI entry"),
                            "gathered_contacts_per_step": r.get("gathered_contacts"),
                            "collective_only_ms": collective_ms,
                            "note": "the exchange is part of the timed entry (value and e2e)"},
               "per_rank": per_rank, "secondary": secondary}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
