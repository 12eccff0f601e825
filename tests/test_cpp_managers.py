"""The C++ contact-manager plugin (tesseract_b200/csrc/plugin) run through the reference's own manager tests
(tests/cpp/test_managers.cpp).  CPU tier: linked against the host simulator of the kernels (host logic + numerics);
GPU tier: linked against libt4c.so, i.e. the shipped plugin path."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = [os.path.join(ROOT, "tests", "cpp", "test_managers.cpp"),
       os.path.join(ROOT, "tesseract_b200", "csrc", "plugin", "b200_contact_managers.cpp")]


def build_and_run(libdir, libname, exe):
    out = os.path.join(ROOT, "tests", "cpp", exe)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror"] + SRC +
                          ["-o", out, "-L" + libdir, "-l" + libname, "-Wl,-rpath," + libdir])
    r = subprocess.run([out], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout


def test_cpp_managers_hostsim():
    from hostsim import build as hostsim
    hostsim.build()
    build_and_run(os.path.join(ROOT, "tests", "hostsim"), "t4c_hostsim", "test_managers_hostsim.bin")


@pytest.mark.gpu
def test_cpp_managers_gpu():
    import __graft_entry__ as g
    g.build()
    build_and_run(os.path.join(ROOT, "tesseract_b200", "csrc"), "t4c", "test_managers_gpu.bin")


def test_plugin_library_exports():
    """libtesseract_b200_managers.so is built by __graft_entry__.build() and exports the factory symbols."""
    import ctypes
    import __graft_entry__ as g
    g.build()
    lib = ctypes.CDLL(os.path.join(ROOT, "tesseract_b200", "csrc", "libtesseract_b200_managers.so"))
    for sym in ("tesseract_b200_create_discrete_manager", "tesseract_b200_create_cast_manager"):
        assert hasattr(lib, sym)
