"""CPU suite: the C-ABI library loads, exports every symbol include/hsfm_ba.h declares, its PODs
match the ctypes mirrors, it refuses to compute without a GPU (no CPU fallback), and nothing in the
product tree touches the oracle."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from sfm_b200 import api, build

    build.build()
    return api.load_library()


def header_symbols():
    src = open(os.path.join(ROOT, "include", "hsfm_ba.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hsfm_ba_[a-z_0-9]+)\s*\(", src)))


def test_exports_every_declared_symbol(lib):
    from sfm_b200 import api

    names = header_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hsfm_ba.h but not exported"
    assert set(names) == set(api.SYMBOLS), set(names) ^ set(api.SYMBOLS)
    assert lib.hsfm_ba_abi_version() == 2


def test_struct_layout_and_defaults(lib):
    from sfm_b200 import api
    from sfm_b200._cdefs import Iteration, Options, Summary

    assert C.sizeof(Options) == 8 * 4 + 10 * 8
    assert C.sizeof(Iteration) == 7 * 8 + 4 * 4
    assert C.sizeof(Summary) == 3 * 8 + 6 * 4 + 8 + 4 * 8 + 8 + 2 * 4 + 8 * 8 + 8 + 4 * 4 + 8
    o = api.default_options()
    # ceres::Solver::Options defaults; the reference overrides only the linear solver
    assert (o.max_num_iterations, o.max_consecutive_invalid_steps, o.jacobi_scaling) == (50, 5, 1)
    assert (o.initial_trust_region_radius, o.max_trust_region_radius, o.min_trust_region_radius) == (1e4, 1e16, 1e-32)
    assert (o.min_relative_decrease, o.min_lm_diagonal, o.max_lm_diagonal) == (1e-3, 1e-6, 1e32)
    assert (o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance) == (1e-6, 1e-10, 1e-8)


def test_oracle_options_share_the_layout():
    from oracle import pyoracle as po
    from sfm_b200._cdefs import Iteration, Options

    assert C.sizeof(po.OOptions) == C.sizeof(Options) and C.sizeof(po.OIteration) == C.sizeof(Iteration)
    assert [f for f, _ in po.OOptions._fields_][8:] == [f for f, _ in Options._fields_][8:]


def test_no_cpu_fallback_without_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.hsfm_ba_create(C.byref(h), 0)
    assert rc == -2 and not h.value  # HSFM_BA_ERR_NO_DEVICE: fails loudly, computes nothing
    assert b"no CPU path" in lib.hsfm_ba_error_string(rc)
    from sfm_b200 import api

    with pytest.raises(api.BAError):
        api.Context(0)


def test_product_never_references_the_oracle():
    bad = []
    for d, _, files in os.walk(os.path.join(ROOT, "sfm_b200")):
        if "build" in d.split(os.sep)[-1:]:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                if re.search(r"(^|\n)\s*(from|import)\s+oracle|pyoracle|libba_oracle|ba_oracle", txt):
                    bad.append(os.path.join(d, f))
    assert not bad, bad
    # and the shared library has no undefined oracle symbols / dependency
    import subprocess

    out = subprocess.run(["ldd", os.path.join(ROOT, "sfm_b200", "libhsfm_ba.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out
