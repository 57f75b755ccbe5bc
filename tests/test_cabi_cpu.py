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


def test_snapshot_round_trip(tmp_path):
    """bsnls_problem_save / bsnls_problem_load: every array of the flattened problem comes back bit for bit (host only),
    the CPU checker gives the same answer on the copy, and a damaged file is refused."""
    import numpy as np
    import oracle_lib as orc
    from beam_slam_b200 import host
    p = host.HashGraph.synthetic(2, scale=0.1).flatten()
    path = tmp_path / "c2.bsnls"
    p.save(path)
    q = cabi.FlatProblem.load(path)
    for name in ("block_size", "block_manifold", "block_constant", "block_group", "block_values", "cameras"):
        assert np.array_equal(getattr(p, name), getattr(q, name)), name
    assert sorted(p.groups) == sorted(q.groups)
    for t in p.groups:
        for k, v in p.groups[t].items():
            w = q.groups[t][k]
            assert (v is None and w is None) or np.array_equal(v, w), (t, k)
    assert orc.evaluate(p)[0] == orc.evaluate(q)[0]
    q2 = cabi.FlatProblem.load_numpy(path)       # the numpy-only reader of the same format
    assert np.array_equal(q2.block_values, p.block_values) and sorted(q2.groups) == sorted(p.groups)
    assert all(np.array_equal(q2.groups[t]["blocks"], p.groups[t]["blocks"]) and np.array_equal(q2.groups[t]["consts"], p.groups[t]["consts"]) for t in p.groups)
    raw = path.read_bytes()
    bad = tmp_path / "bad.bsnls"
    bad.write_bytes(b"XSNLS" + raw[5:])
    with pytest.raises(OSError):
        cabi.FlatProblem.load(bad)
    bad.write_bytes(raw[:len(raw) // 2])     # truncated
    with pytest.raises(OSError):
        cabi.FlatProblem.load(bad)
