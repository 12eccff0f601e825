#!/usr/bin/env python3
"""Join an `ncu --page source --csv --print-source sass` export with the line table of the same cubin.

The SASS view of ncu's CSV export has the per-instruction counters (stall samples, executed instructions, active
threads) but no file / line columns; `nvdisasm -g -c` of the cubin that was profiled gives the line of every
instruction at the same section offset.  Prints, for one kernel: stall-reason totals, opcode mix, and samples /
executed warp instructions / mean active threads by source file, by out-of-line routine and by (file, line).

    cuobjdump -xelf all libt4c.so ; nvdisasm -g -c t4c_exact.sm_100a.cubin > exact.sass
    python tools/ncu_source_join.py source_sass.csv exact.sass gjkKernelR128 [launch index]
"""
import csv
import re
import sys
from collections import defaultdict


def read_sass_lines(path, kernel):
    """section offset -> (file, line, routine, instruction text) for the section of `kernel`"""
    out, in_sec, cur, routine = {}, False, ("?", 0), kernel
    sec_re = re.compile(r"^\s*\.section\s+\.text\.(\S+?),")
    line_re = re.compile(r'//## File "([^"]+)", line (\d+)')
    ins_re = re.compile(r"^\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);")
    lab_re = re.compile(r"^(\$\S+|\.text\.\S+|_Z\S+):")
    for l in open(path, errors="replace"):
        m = sec_re.match(l)
        if m:
            in_sec, routine = kernel in m.group(1), kernel
            continue
        if not in_sec:
            continue
        m = line_re.search(l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = lab_re.match(l)
        if m and "$" in m.group(1):
            routine = m.group(1).split("$")[-1]
            continue
        m = ins_re.match(l)
        if m:
            out[int(m.group(1), 16)] = (cur[0], cur[1], routine, m.group(2))
    return out


def read_csv_kernels(path, kernel):
    """list of launches of `kernel`; each a list of dict rows (one per SASS instruction)"""
    launches, hdr, cur, take = [], None, None, False
    for row in csv.reader(open(path, errors="replace")):
        if not row:
            continue
        if row[0] == "Kernel Name":
            take, cur, hdr = kernel in row[1], [], None
            if take:
                launches.append(cur)
        elif row[0] == "Address":
            hdr = row
        elif take and hdr and row[0].startswith("0x"):
            cur.append(dict(zip(hdr, row)))
    return launches


def num(s):
    try:
        return float(s.replace(",", ""))
    except (ValueError, AttributeError):
        return 0.0


def main():
    csv_path, sass_path, kernel = sys.argv[1:4]
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    lines = read_sass_lines(sass_path, kernel)
    launches = read_csv_kernels(csv_path, kernel)
    rows = launches[which]
    base = int(rows[0]["Address"], 16)
    tot_s = tot_i = tot_t = 0.0
    by_file, by_line, by_rout = (defaultdict(lambda: [0.0, 0.0, 0.0]) for _ in range(3))
    stalls, ops = defaultdict(float), defaultdict(float)
    for r in rows:
        off = int(r["Address"], 16) - base
        s, i, t = num(r.get("# Samples")), num(r.get("Instructions Executed")), num(r.get("Thread Instructions Executed"))
        tot_s, tot_i, tot_t = tot_s + s, tot_i + i, tot_t + t
        f, ln, rt, ins = lines.get(off, ("?", 0, "?", "?"))
        w = ins.split()
        ops[(w[1] if w[0].startswith("@") and len(w) > 1 else w[0]).split(".")[0]] += i
        for d, k in ((by_file, f), (by_line, (f, ln)), (by_rout, rt[-60:])):
            d[k][0] += s
            d[k][1] += i
            d[k][2] += t
        for k, v in r.items():
            if k.startswith("stall_") and "Not Issued" not in k:
                stalls[k] += num(v)
    executed = sum(1 for r in rows if num(r.get("Instructions Executed")) > 0)
    print(f"# {kernel}: launch {which} of {len(launches)}; {len(rows)} SASS instructions ({len(rows) * 16} B), "
          f"{executed} executed at least once ({executed * 16} B)")
    print(f"# samples {tot_s:.0f}, warp instructions {tot_i:.0f}, mean active threads {tot_t / max(tot_i, 1):.2f}")
    print("# stall samples by reason:")
    for k, v in sorted(stalls.items(), key=lambda kv: -kv[1])[:8]:
        print(f"  {k:28s} {v:10.0f} {100 * v / max(tot_s, 1):5.1f}%")
    print("# opcode mix (executed warp instructions):")
    print("  " + ", ".join(f"{k} {100 * v / max(tot_i, 1):.1f}%" for k, v in sorted(ops.items(), key=lambda kv: -kv[1])[:16]))

    def show(title, d, n):
        print(f"# {title}: samples %, warp instructions %, active threads")
        for k, (s, i, t) in sorted(d.items(), key=lambda kv: -kv[1][0])[:n]:
            print(f"  {str(k):62s} {100 * s / max(tot_s, 1):5.1f}% {100 * i / max(tot_i, 1):5.1f}% {t / max(i, 1):5.1f}")

    show("by file", by_file, 12)
    show("by routine", by_rout, 16)
    show("by line", by_line, 40)


if __name__ == "__main__":
    main()
