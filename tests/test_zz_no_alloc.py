"""bsnls_reprojection_cost_noalloc: the allocation-free entry point.

It serves callers that cannot ask the driver for device memory — a real-time optimisation thread (cudaMalloc and
cudaFree synchronise the whole device), a process confined to a pre-sized memory arena, code recording a stream
capture. They pass buffers they already own; the library's only working memory is a fixed 128 KiB array inside its
image, so the number of observations per call is capped.

The test drives the entry point through the C-ABI with caller-owned device buffers (torch tensors stand in for the
caller's arena) and compares the cost with what the handle-based entry point bsnls_evaluate returns for the same
observations (and with the CPU oracle's cost): two calls in a row, the second on a smaller set so stale scratch from
the first would show.
Collected last on purpose (file name), and independent of what ran before it.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as orc
from beam_slam_b200 import cabi, host
from beam_slam_b200.solver import Solver

pytestmark = pytest.mark.gpu


def _reprojection_only(p, keep):
    """The same graph with only its first `keep` reprojection factors (createProblem order kept)."""
    g = p.groups[cabi.FACTOR_REPROJECTION]
    order = np.argsort(g["res_index"])[:keep]
    order.sort()
    sub = dict(blocks=g["blocks"][order], consts=g["consts"][order], aux=g["aux"][order], loss_type=None, loss_param=None,
               res_index=np.argsort(np.argsort(g["res_index"][order])).astype(np.int32))
    return cabi.FlatProblem(p.block_size, p.block_manifold, p.block_constant, p.block_group, p.block_values, p.cameras,
                            {cabi.FACTOR_REPROJECTION: sub})


def _noalloc_cost(lib, q, torch):
    g = q.groups[cabi.FACTOR_REPROJECTION]
    dev = torch.device("cuda", 0)
    off = q.offsets.astype(np.int32)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    values, cams, pixel = t(q.block_values), t(q.cameras.reshape(-1)), t(g["consts"].reshape(-1))
    qo, to, po = t(off[g["blocks"][:, 0]]), t(off[g["blocks"][:, 1]]), t(off[g["blocks"][:, 2]])
    cam = t(g["aux"])
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    msg = C.create_string_buffer(256)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.bsnls_reprojection_cost_noalloc(values.data_ptr(), qo.data_ptr(), to.data_ptr(), po.data_ptr(), cam.data_ptr(),
                                             pixel.data_ptr(), cams.data_ptr(), len(g["res_index"]), out.data_ptr(),
                                             C.c_void_p(stream), msg, 256)
    assert rc == 0, msg.value
    torch.cuda.synchronize(dev)
    return float(out.item())


def test_noalloc_cost_matches_evaluate_on_two_calls_in_a_row():
    import torch
    torch.cuda.set_device(0)
    lib = cabi.load_library()
    p = host.HashGraph.synthetic(3, scale=0.05).flatten()
    n = p.count(cabi.FACTOR_REPROJECTION)
    for keep in (n, n // 3):                       # second call: fewer blocks of partial sums than the first
        q = _reprojection_only(p, keep)
        s = Solver()
        s.set_problem(q)
        expect, _, _ = s.evaluate()                # existing entry point (allocates through its handle)
        got = _noalloc_cost(lib, q, torch)
        assert abs(got - expect) <= 1e-12 * expect, (keep, got, expect)
        assert abs(got - orc.evaluate(q)[0]) <= 1e-12 * expect   # and against the CPU checker, not only the sibling entry point
        s.close()
    # beyond the fixed scratch: refused with a reason, nothing launched
    msg = C.create_string_buffer(256)
    one = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = lib.bsnls_reprojection_cost_noalloc(one.data_ptr(), one.data_ptr(), one.data_ptr(), one.data_ptr(), one.data_ptr(),
                                             one.data_ptr(), one.data_ptr(), cabi.NOALLOC_MAX_OBSERVATIONS + 1, one.data_ptr(),
                                             None, msg, 256)
    assert rc == cabi.ERR_INVALID and b"too many observations" in msg.value
