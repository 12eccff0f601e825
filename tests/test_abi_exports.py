"""CPU tier: libt4c.so (built by nvcc for sm_100a) loads without a GPU and exports every function include/t4c.h
declares; the compute entries fail loudly (T4C_ERR_CUDA) instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "t4c.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = set(re.findall(r"\b(t4c_[a-z0-9_]+)\s*\(", src))
    assert len(names) >= 25
    return sorted(names)


def test_every_declared_symbol_is_exported():
    import __graft_entry__ as g
    g.build()
    from tesseract_b200 import capi
    lib = capi.load_library()
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.t4c_abi_version() == 2


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return  # exercised on the CPU-only tier
    from tesseract_b200 import capi, scene as S
    lib = capi.load_library()
    assert lib.t4c_device_count() == 0
    b = S.SceneBuilder()
    b.add_collision_object("a", 0, [S.Box(1, 1, 1)], [S.identity()])
    desc = b.scene_desc()
    h = C.c_void_p()
    st = lib.t4c_scene_create(C.byref(desc), 0, C.byref(h))
    assert st == capi.T4C_ERR_CUDA
    assert b"no CPU path" in lib.t4c_last_error()


def test_product_binding_refuses_the_host_simulator(monkeypatch):
    """T4C_LIB may point the binding at another CUDA build of the library, never at the CPU-tier host simulator"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from hostsim import build as hostsim
    from tesseract_b200 import capi
    assert capi.load_library().t4c_build_info() == b"cuda sm_100a"
    assert hostsim.lib().t4c_build_info() == b"hostsim"  # explicit path: the tests' own use
    monkeypatch.setenv("T4C_LIB", hostsim.LIB)
    monkeypatch.setattr(capi, "_lib", None)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.load_library()
