#!/usr/bin/env python
"""Regenerates the table block of tesseract_b200/csrc/t4_sincos.h.

The device FK has to produce the SAME bits as the host libm the reference (Eigen::AngleAxisd -> std::sin / std::cos) runs
on: glibc >= 2.28 on x86-64 (IBM Accurate Mathematical Library, sysdeps/ieee754/dbl-64/s_sin.c, FMA multiarch variant).
Its 110 x 4 table (sin(k/128), low part, cos(k/128), low part) is NOT the correctly rounded split of those values in 18 of
440 entries, so it is read out of the libm this container ships (located by the correctly rounded high parts, which do
match) instead of being recomputed.  Run once; the output is committed inside the header.
"""
import math
import struct
import sys


def main(path="/lib/x86_64-linux-gnu/libm.so.6"):
    blob = open(path, "rb").read()
    key = b"".join(struct.pack("<d", v) for v in (0.0, 0.0, 1.0, 0.0)) + struct.pack("<d", math.sin(1 / 128))
    at = blob.find(key)
    if at < 0:
        sys.exit("table not found")
    tab = struct.unpack("<440d", blob[at:at + 440 * 8])
    for k in range(110):  # sanity: high parts are the correctly rounded sin / cos of k/128
        assert abs(tab[4 * k] - math.sin(k / 128)) <= 2e-16 and abs(tab[4 * k + 2] - math.cos(k / 128)) <= 2e-16
    out = []
    for k in range(110):
        out.append("  " + ", ".join(float.hex(v) for v in tab[4 * k:4 * k + 4]) + ",")
    print("\n".join(out))


if __name__ == "__main__":
    main(*sys.argv[1:])
