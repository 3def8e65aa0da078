"""The drop-in boundary on a box WITHOUT a GPU: libepsgpu.so loads, exports every symbol
include/epsgpu.h declares, the ctypes table covers exactly those symbols, and the product
fails loudly (no CPU fallback) when no B200 is present.  No compute calls."""
import ctypes
import os
import re

import pytest

from epsilon_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "epsgpu.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(eps_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_something():
    syms = declared_symbols()
    assert "eps_solver_solve" in syms and "eps_map_apply" in syms and len(syms) >= 35


def test_library_exports_every_declared_symbol():
    assert os.path.exists(capi.LIB_PATH), "run `python -m epsilon_b200.build`"
    lib = ctypes.CDLL(capi.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_ctypes_table_matches_header():
    assert sorted(capi.SIGNATURES) == declared_symbols()


def test_every_entry_point_cites_the_reference():
    """Each group of entry points names the reference interface (file:line) it replaces."""
    src = open(HEADER).read()
    for needle in ("linear_map.h:33-57", "prox_admm.cc:131-169", "solvemodule.cc:189-242",
                   "vector_util.cc:261-281", "solver_params.proto", "solver.proto", "affine.cc:22-129"):
        assert needle in src, needle


def test_params_default_and_struct_layout():
    lib = capi.load()
    p = capi.Params()
    lib.eps_params_default(ctypes.byref(p))
    # SolverParams defaults, proto/epsilon/solver_params.proto
    assert (p.max_iterations, p.rho, p.abs_tol, p.rel_tol, p.epoch_iterations) == (10000, 1.0, 1e-4, 1e-2, 10)
    assert p.warm_start == 0 and p.kkt_mode == 0 and p.cg_tol > 0


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_device_is_a_loud_error_not_a_fallback():
    lib = capi.load()
    h = ctypes.c_void_p()
    rc = lib.eps_ctx_create(0, None, ctypes.byref(h))
    assert rc != 0 and not h.value
    assert lib.eps_last_error()  # message set
    from epsilon_b200.solver import Context
    with pytest.raises(capi.EpsError):
        Context(0)


def test_product_does_not_import_the_oracle():
    """Nothing under epsilon_b200/ may import, load or execute oracle/ (or the reference)."""
    pkg = os.path.join(ROOT, "epsilon_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if not fn.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")):
                continue
            text = open(os.path.join(dirpath, fn), errors="replace").read()
            for line in text.splitlines():
                code = line.split("#")[0] if fn.endswith(".py") else line
                assert not re.search(r"\b(from|import)\s+oracle\b", code), (fn, line)
                assert "oracle/_ref" not in code and "epsilon_ref" not in code, (fn, line)
                assert "/root/reference" not in code, (fn, line)
