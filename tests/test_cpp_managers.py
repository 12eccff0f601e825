"""The C++ contact-manager plugin (tesseract_b200/csrc/plugin) run through the reference's own manager tests
(tests/cpp/test_managers.cpp).  CPU tier: linked against the host simulator of the kernels (host logic + numerics);
GPU tier: linked against libt4c.so, i.e. the shipped plugin path."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = [os.path.join(ROOT, "tesseract_b200", "csrc", "plugin", f)
          for f in ("b200_contact_managers.cpp", "b200_trajectory_check.cpp")]


def build_and_run(libdir, libname, exe, test="test_managers.cpp"):
    out = os.path.join(ROOT, "tests", "cpp", exe)
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror",
                           os.path.join(ROOT, "tests", "cpp", test)] + PLUGIN +
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


def test_cpp_trajectory_check_hostsim():
    """checkTrajectory / checkTrajectories / isValid against the reference's per-state loop"""
    from hostsim import build as hostsim
    hostsim.build()
    build_and_run(os.path.join(ROOT, "tests", "hostsim"), "t4c_hostsim", "test_trajectory_check_hostsim.bin",
                  "test_trajectory_check.cpp")


@pytest.mark.gpu
def test_cpp_trajectory_check_gpu():
    import __graft_entry__ as g
    g.build()
    build_and_run(os.path.join(ROOT, "tesseract_b200", "csrc"), "t4c", "test_trajectory_check_gpu.bin",
                  "test_trajectory_check.cpp")


def test_plugin_library_exports():
    """libtesseract_b200_managers.so is built by __graft_entry__.build() and exports the factory symbols."""
    import ctypes
    import __graft_entry__ as g
    g.build()
    lib = ctypes.CDLL(os.path.join(ROOT, "tesseract_b200", "csrc", "libtesseract_b200_managers.so"))
    for sym in ("tesseract_b200_create_discrete_manager", "tesseract_b200_create_cast_manager"):
        assert hasattr(lib, sym)


def test_cpp_margin_data():
    """CollisionMarginData (a14) with the numbers of the reference's collision_margin_data_unit.cpp; host only"""
    out = os.path.join(ROOT, "tests", "cpp", "test_margin_data.bin")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-Wextra", "-Werror",
                           os.path.join(ROOT, "tests", "cpp", "test_margin_data.cpp"), "-I" + os.path.join(ROOT, "include"),
                           "-o", out])
    r = subprocess.run([out], capture_output=True, text=True, timeout=60)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and "0 failures" in r.stdout, r.stdout + r.stderr
