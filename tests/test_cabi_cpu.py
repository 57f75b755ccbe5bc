"""The C-ABI library loads on a CPU-only machine and exports every symbol include/bsnls.h declares; compute calls
fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

from beam_slam_b200 import cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bsnls.h")).read()
    return sorted(set(re.findall(r"\b(bsnls_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    lib = cabi.load_library()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), n


def test_default_options_match_ceres_defaults():
    lib = cabi.load_library()
    o = cabi.Options()
    lib.bsnls_default_options(C.byref(o))
    d = cabi.Options.defaults()
    for f, _ in cabi.Options._fields_:
        assert getattr(o, f) == getattr(d, f), f
    assert o.max_num_iterations == 50 and o.function_tolerance == 1e-6 and o.initial_trust_region_radius == 1e4


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_create_fails_loudly_without_a_gpu():
    lib = cabi.load_library()
    h = C.c_void_p()
    rc = lib.bsnls_create(None, C.byref(h))
    assert rc == cabi.ERR_CUDA and not h.value
    assert b"no CUDA device" in lib.bsnls_last_error(None)


def test_invalid_options_are_rejected_before_touching_the_device():
    lib = cabi.load_library()
    h = C.c_void_p()
    o = cabi.Options.defaults(max_num_iterations=-1)
    assert lib.bsnls_create(C.byref(o), C.byref(h)) == cabi.ERR_INVALID


def test_product_package_never_references_the_oracle():
    """The oracle is test infrastructure: nothing under beam_slam_b200/ may import, link or call it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "beam_slam_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in txt and "oracle_" not in txt and "oracle/" not in txt, os.path.join(dirpath, f)
