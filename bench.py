#!/usr/bin/env python
"""Headline benchmark: collision-checked states/sec on BASELINE.json configs[1]
(KUKA iiwa 14 convex-hull links vs a 10k box/sphere/cylinder environment, 1M random joint states per GPU, discrete
contactTest, 0.05 m margin).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--mode first|closest|first_flags]

One "step" = one pass of the hot path (FK + broadphase + narrowphase + result bridging) over the whole batch of
joint states.  `value` times the pass with the joint vectors already resident in HBM; `e2e` times the host-buffer
C-ABI call (t4c_check_discrete: pinned host q -> H2D -> check -> D2H hit flags / contacts).
Multi-GPU: one process per GPU (torchrun), states sharded by contiguous ranges, environment replicated, per-rank hit
flags gathered to rank 0 with a single NCCL gather; weak scaling (1M states per GPU).
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
    return ap.parse_args()


def workload(args):
    return {"workload": f"configs[1]: KUKA iiwa 14 convex-hull links vs {args.prims} box/sphere/cylinder environment, "
                        f"{args.states} random joint states per GPU (seed 1234 + rank), discrete, margin 0.05 m",
            "request": MODES[args.mode][0], "states_per_gpu": args.states, "environment_primitives": args.prims,
            "parallelism": f"dp{args.gpus} (states sharded, environment BVH replicated)",
            "l2": "L2 flushed between timed iterations (512 MiB device memset)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

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
    sample = f"first {n} states of the workload, {cores} threads, one manager clone per thread"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": workload(args),
                      "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


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
    mode_name, tt, want = MODES[args.mode]
    n = args.states
    b = S.build_config2(n_prims=args.prims, margin=0.05)
    desc = b.scene_desc()
    scene = C.c_void_p()
    capi.check(lib.t4c_scene_create(C.byref(desc), local_rank, C.byref(scene)), lib)
    fkd = b.fk_desc()
    model = C.c_void_p()
    capi.check(lib.t4c_fk_model_create(C.byref(fkd), scene, C.byref(model)), lib)
    nq = fkd.n_q
    # ---- synthetic inputs: this rank's shard of the joint states
    q_host = S.random_states(b, n, seed=1234 + rank + int(os.environ.get("T4_BENCH_SEED_OFFSET", "0")), dtype=np.float64)  # (n, nq)
    dev = torch.device("cuda", local_rank)
    q_soa = torch.from_numpy(np.ascontiguousarray(q_host.T.astype(np.float32))).to(dev)  # FP32 SoA q[j][i] in HBM
    q_pin = torch.from_numpy(q_host).pin_memory()
    hit_dev = torch.zeros(n, dtype=torch.uint8, device=dev)
    cap = 2 * n if tt == 0 else 12 * n
    con_dev = torch.empty((cap if want else 1) * 80, dtype=torch.uint8, device=dev)
    hit_pin = torch.zeros(n, dtype=torch.uint8).pin_memory()
    con_pin = torch.empty((cap if want else 1) * 80, dtype=torch.uint8).pin_memory()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    req = capi.Request(tt, 1, 1)

    def step_device():
        io = capi.DeviceIO()
        io.q, io.q_layout = q_soa.data_ptr(), capi.Q_F32_SOA
        io.hit = hit_dev.data_ptr()
        io.contacts = con_dev.data_ptr() if want else None
        io.capacity = cap if want else 0
        capi.check(lib.t4c_check_discrete_device(scene, model, n, C.byref(req), C.byref(io)), lib)
        return io

    def step_host():
        res = capi.Result()
        res.hit = C.cast(hit_pin.data_ptr(), C.POINTER(C.c_uint8))
        res.contacts = con_pin.data_ptr() if want else None
        res.capacity = cap if want else 0
        capi.check(lib.t4c_check_discrete(scene, model, C.cast(q_pin.data_ptr(), C.POINTER(C.c_double)), n,
                                          C.byref(req), C.byref(res)), lib)
        return res

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, gather):
        for _ in range(warmup):
            fn()
            if gather:
                gather()
        total_ms, last = 0.0, None
        gather_ms[0] = 0.0
        for _ in range(steps):
            flush.zero_()
            barrier()
            e0, em, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            last = fn()          # synchronous C-ABI call: kernels run on the library's stream and are joined here
            em.record()
            if gather:
                gather()
            e1.record()
            torch.cuda.synchronize()
            if os.environ.get("T4_BENCH_DEBUG"):
                print(f"[rank {rank}] fn {e0.elapsed_time(em):.2f} ms, exchange {em.elapsed_time(e1):.2f} ms, "
                      f"kernels {sum(v[0] for v in capi.profile_get(lib).values()):.2f} ms (cumulative)",
                      file=sys.stderr, flush=True)
            ms = torch.tensor([e0.elapsed_time(e1), em.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            total_ms += float(ms[0].item())
            gather_ms[0] += float(ms[1].item()) / steps  # result exchange incl. waiting for the slowest rank
        return total_ms / steps, last

    gather_ms = [0.0]
    gather_kind = os.environ.get("T4_BENCH_GATHER", "gather")
    gathered = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None
    gathered_all = torch.empty(world * n, dtype=torch.uint8, device=dev) if (world > 1 and gather_kind == "allgather") else None

    def gather_hits():
        if world > 1 and gather_kind == "allgather":
            dist.all_gather_into_tensor(gathered_all, hit_dev)
        elif world > 1:
            dist.gather(hit_dev, gathered, dst=0)  # the single result exchange (north_star): flags to rank 0 over NVLink

    sampler = ClockSampler(local_rank)
    launches0 = lib.t4c_kernel_launch_count()
    lib.t4c_profile_enable(1)
    for _ in range(args.warmup):
        step_device()
        gather_hits()  # NCCL communicator set-up happens here, outside the timed region
    lib.t4c_profile_reset()
    launches0 = lib.t4c_kernel_launch_count()
    sampler.start()
    ms_dev, io = timed(step_device, args.steps, 0, gather_hits)
    exchange_ms = gather_ms[0]
    collective_ms = None
    if world > 1:
        # the collective alone, ranks aligned by a barrier: separates its own cost from the wait for the slowest rank
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(10):
            gather_hits()
        c1.record()
        torch.cuda.synchronize()
        collective_ms = c0.elapsed_time(c1) / 10
    clocks = sampler.stop()
    launches = int(lib.t4c_kernel_launch_count() - launches0)
    prof = capi.profile_get(lib)
    lib.t4c_profile_enable(0)
    ms_e2e, res = timed(step_host, max(1, min(args.steps, 3)), 1, gather_hits)
    stats = io.stats.as_dict()
    # ---- roofline of the dominant kernel (FP32/FP64 CUDA-core issue bound, not HBM and not tensor: DESIGN.md)
    kern = {k: v for k, v in prof.items() if v[1]}
    top = max(kern, key=lambda k: kern[k][0]) if kern else None
    roof = None
    if top:
        tf = C.c_double()
        capi.check(lib.t4c_measure_fp32_peak(local_rank, C.byref(tf)), lib)
        ms_top = kern[top][0] / max(1, args.steps)
        # ALGORITHMIC work (SURVEY 8(d) constants): see DESIGN.md "Roofline accounting"
        V = 190.0  # mean hull vertices per robot link
        per_state = {"fk": 9 * 80 + 63 * 2.0,
                     "broad_static": (stats["bvh_nodes"] * 12 + stats["candidates"] * 30) / max(1, stats["states"]),
                     "cull_fp32": (stats["gjk_iterations"] * (6 * V + 33 + 36 + 100)) / max(1, stats["states"]),
                     "gjk_fp64": (stats["gjk_iterations"] * (6 * V + 33 + 36 + 100)) / max(1, stats["states"]),
                     "epa_fp64": (stats["epa_iterations"] * (400 + 2 * (6 * V + 33))) / max(1, stats["states"])}
        flops = per_state.get(top, 0.0) * n
        ach = flops / (ms_top * 1e-3) / 1e12 if ms_top > 0 else 0.0
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        bytes_state = 4 * nq + 1 + (80 * stats["contacts"] / max(1, stats["states"]) if want else 0)
        # DRAM bytes per launch of that kernel from the committed ncu --set full capture (same workload only)
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            if tj["_config"] == {"mode": args.mode, "states": n, "prims": args.prims}:
                traffic = tj.get(top)
        except Exception:
            pass
        roof = {"bound": "fp32", "kernel": top, "achieved": ach, "peak": tf.value, "unit": "TFLOP/s",
                "frac": ach / tf.value if tf.value else None, "traffic": traffic,
                "peak_source": "measured in this run by t4c_measure_fp32_peak (FFMA loop on every SM)",
                "kernel_ms_per_step": {k: v[0] / max(1, args.steps) for k, v in kern.items()},
                "hbm": {"achieved": bytes_state * n / (ms_dev * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"}}
    # ---- CPU baseline on a bounded sample (rank 0, N == 1)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pyoracle
        cores = os.cpu_count() or 1
        ns = args.cpu_sample or max(cores * 250, 2000)
        t = time.perf_counter()
        pyoracle.check_discrete(b, q_host[:ns], tt, want_contacts=want, n_threads=cores)
        dt = time.perf_counter() - t
        cpu = {"value": ns / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {ns} states of the workload, {cores} threads, one manager clone per thread; "
                         "CPU restatement (oracle/), not Bullet: Bullet/FCL are not installed on this box"}
    if rank == 0:
        h2d = n * nq * 8
        d2h = n + (int(res.n_contacts) * 80 if want else 0)
        out = {"metric": METRIC, "value": world * n / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev, "higher_is_better": True,
               "scaling": "weak", "vs_baseline": None, "dtype": "f32 cull + f64 exact pass", "data": "synthetic",
               "config": workload(args), "clocks": clocks, "gpu_launches": launches,
               "e2e": {"value": world * n / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                       "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e},
               "roofline": roof, "cpu_baseline": cpu, "stats_per_step": stats, "mode": args.mode,
               "exchange": {"kind": ("none" if world == 1 else gather_kind), "ms_per_step": exchange_ms,
                            "collective_only_ms": collective_ms,
                            "note": "hit flags to rank 0; includes waiting for the slowest rank"}}
        print(json.dumps(out))
    lib.t4c_fk_model_destroy(model)
    lib.t4c_scene_destroy(scene)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
