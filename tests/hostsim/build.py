"""TEST-ONLY host simulator of the CUDA kernels.

Compiles tesseract_b200/csrc/t4c.cu with g++ -DT4C_HOSTSIM: the very same kernel bodies (FK, broadphase, FP32 cull,
FP64 GJK/EPA narrowphase, result bridging) run serially on the host, so their numerics and the host-side scene
compiler / C ABI plumbing can be checked against the oracle in the CPU-only test tier (`-m "not gpu"`).
This library is NEVER loaded by the tesseract_b200 package (capi.load_library only opens csrc/libt4c.so, built by nvcc,
and raises when it is missing); it is not a fallback and is not shipped.
"""
import glob
import os
import subprocess

import ctypes as C

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "..", "tesseract_b200", "csrc")
LIB = os.path.join(HERE, "libt4c_hostsim.so")


def build(force=False):
    srcs = glob.glob(os.path.join(CSRC, "*.h")) + [os.path.join(CSRC, "t4c.cu"),
                                                   os.path.join(HERE, "..", "..", "include", "t4c.h")]
    stale = force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs)
    if stale:
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DT4C_HOSTSIM", "-ffp-contract=off",
                               "-x", "c++", os.path.join(CSRC, "t4c.cu"), "-o", LIB])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        from tesseract_b200 import capi
        try:
            _lib = capi.load_library(build())
        except OSError:
            _lib = capi.load_library(build(force=True))
    return _lib
