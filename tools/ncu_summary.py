#!/usr/bin/env python
"""Summarise an `ncu --set full --page raw --csv` export of ONE bench step: per kernel the launches, time, registers,
occupancy, active threads per instruction, FP64 / FMA pipe use, L2 hit rate and DRAM bytes; writes profiles/traffic.json
(DRAM bytes per step per kernel, read by bench.py for roofline.traffic) when --traffic is given.

    python tools/ncu_summary.py profiles/r3e_full_raw.csv --traffic profiles/traffic.json --mode first
"""
import argparse
import csv
import json
import sys
from collections import OrderedDict

NAMES = OrderedDict([("fkKernelT", "fk"), ("broadStaticKernel", "broad_static"), ("liveKeyKernel", "broad_static"), ("broadActiveKernel", "broad_active"),
                     ("cullKernel", "cull_fp32"), ("pruneKernel", "cull_fp32"), ("gjkKernelR128", "gjk_fp64"),
                     ("epaKernelR128", "epa_fp64")])


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return 0.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--traffic")
    ap.add_argument("--mode", default="first")
    ap.add_argument("--states", type=int, default=1 << 20)
    ap.add_argument("--prims", type=int, default=10000)
    a = ap.parse_args()
    rows = [r for r in csv.reader(open(a.csv)) if r]
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def get(r, name, scale_unit=None):
        i = col.get(name)
        if i is None:
            return 0.0
        v = num(r[i])
        u = units[i]
        if scale_unit == "bytes":
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
        if scale_unit == "ms":
            v *= {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1, "second": 1e3}.get(u, 1)
        return v

    agg = OrderedDict()
    for r in data:
        kname = r[col["Kernel Name"]]
        key = next((v for k, v in NAMES.items() if k in kname), None)
        if key is None:
            continue
        g = agg.setdefault(key, dict(launches=0, ms=0.0, dram=0.0, inst=0.0, w=[], regs=0))
        ms = get(r, "gpu__time_duration.sum", "ms")
        g["launches"] += 1
        g["ms"] += ms
        g["dram"] += get(r, "dram__bytes_read.sum", "bytes") + get(r, "dram__bytes_write.sum", "bytes")
        g["inst"] += get(r, "smsp__inst_executed.sum")
        g["regs"] = max(g["regs"], int(get(r, "launch__registers_per_thread")))
        g["w"].append((ms, get(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
                       get(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
                       get(r, "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
                       get(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                       get(r, "lts__t_sector_hit_rate.pct"),
                       get(r, "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
                       get(r, "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"),
                       get(r, "smsp__inst_executed_pipe_fp64.sum") or get(r, "sm__inst_executed_pipe_fp64.sum")))
    print("| kernel | launches | ms | regs | warps active % | thr/inst | FP64 pipe % | FMA pipe % | L2 hit % | long-scoreboard | no-instruction | warp inst (M) | FP64 warp inst (M) | DRAM (GB) |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for k, g in agg.items():
        t = sum(w[0] for w in g["w"]) or 1.0
        avg = [sum(w[0] * w[i] for w in g["w"]) / t for i in range(1, 8)]
        fp64_inst = sum(w[8] for w in g["w"])
        print(f"| {k} | {g['launches']} | {g['ms']:.2f} | {g['regs']} | {avg[0]:.0f} | {avg[1]:.1f} | {avg[2]:.1f} | "
              f"{avg[3]:.1f} | {avg[4]:.0f} | {avg[5]:.1f} | {avg[6]:.1f} | {g['inst'] / 1e6:.0f} | {fp64_inst / 1e6:.0f} | "
              f"{g['dram'] / 1e9:.2f} |")
    if a.traffic:
        out = {"_source": f"{a.csv}: ncu --set full --clock-control none, bench.py --steps 1 --warmup 1 (mode {a.mode}, "
                          f"{a.states} states, {a.prims} primitives); dram__bytes_read.sum + dram__bytes_write.sum summed "
                          "over the launches of one step",
               "_config": {"mode": a.mode, "states": a.states, "prims": a.prims}}
        for k, g in agg.items():
            out[k] = int(g["dram"])
        json.dump(out, open(a.traffic, "w"), indent=1)


if __name__ == "__main__":
    main()
