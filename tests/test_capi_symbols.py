"""CPU: the C-ABI library loads, exports every symbol include/ekf_b200.h declares, and the ctypes table
matches the header. No compute entry point is called (no GPU here) except the host-only odometry helpers."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from oracle import ekf_oracle as eo


@pytest.fixture(scope="module")
def capi():
    import __graft_entry__ as g
    from slam_duckietown_b200 import capi
    if not os.path.isfile(capi.LIB_PATH):
        g.build()
    return capi


def _header_functions():
    src = open(os.path.join(ROOT, "include", "ekf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(int|double|const char\*)\s+(ekf3?_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = [a.strip() for a in m.group(3).split(",") if a.strip() and a.strip() != "void"]
        out[m.group(2)] = (m.group(1), args)
    return out


def test_every_declared_symbol_is_exported_and_bound(capi):
    decl = _header_functions()
    assert len(decl) >= 30
    lib = capi.lib()
    assert set(decl) == set(capi.SIGNATURES), set(decl) ^ set(capi.SIGNATURES)
    for name, (ret, args) in decl.items():
        assert hasattr(lib, name), f"{name} not exported"
        res, argtypes = capi.SIGNATURES[name]
        assert len(argtypes) == len(args), f"{name}: header has {len(args)} args, ctypes table {len(argtypes)}"
        for a, t in zip(args, argtypes):
            if "*" in a:
                assert t in (C.c_void_p, C.c_char_p) or hasattr(t, "contents") or issubclass(t, C._Pointer), (name, a)
            elif a.startswith("double"):
                assert t is C.c_double, (name, a)
            elif a.startswith("int64_t"):
                assert t is C.c_int64, (name, a)
            elif a.startswith("int"):
                assert t is C.c_int, (name, a)


def test_no_torch_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "ekf_b200.h")).read()
    assert "torch" not in src.lower().replace("torch tensor storage", "") and "at::" not in src


def test_host_odometry_helpers_bit_exact(capi):
    lib = capi.lib()
    rng = np.random.default_rng(0)
    for _ in range(200):
        t, p = int(rng.integers(0, 100000)), int(rng.integers(0, 100000))
        assert lib.ekf_delta_phi(t, p, 135) == eo.delta_phi(t, p, 135)
    a, l = C.c_double(), C.c_double()
    for _ in range(200):
        dl, dr = rng.normal(0, 0.3, 2)
        capi.call("ekf_displacement", 0.0318, 0.1, float(dl), float(dr), C.byref(a), C.byref(l))
        assert (a.value, l.value) == eo.displacement(0.0318, 0.1, float(dl), float(dr))


def test_errors_are_codes_not_exceptions(capi):
    lib = capi.lib()
    assert lib.ekf_displacement(1.0, 1.0, 0.0, 0.0, None, None) == -1
    assert b"null" in lib.ekf_last_error()
    n = C.c_int()
    assert lib.ekf_batch_layout(20, C.byref(n), None, None, None) == 0 and n.value == 43
    ld = C.c_int64()
    assert lib.ekf_lm_layout(4096, C.byref(n), C.byref(ld)) == 0 and (n.value, ld.value) == (8195, 8208)
    with pytest.raises(capi.EkfError):
        capi.call("ekf_batch_layout", -1, None, None, None, None)


def test_compute_fails_loudly_without_gpu(capi):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from slam_duckietown_b200 import ekf_bindings as eb
    with pytest.raises(capi.EkfLibraryError):
        eb.BatchedEKF(4, 8)
    with pytest.raises(capi.EkfLibraryError):
        eb.EKF_pose_estimation(0.0, 0.0, np.zeros(3), np.eye(3), 0.7, [], {})
