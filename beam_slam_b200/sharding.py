"""Host-side view of the landmark sharding used by the multi-GPU solve: which landmarks a
rank keeps, and the sub-problem a rank effectively evaluates. The partition itself is computed by the C library
(bsnls_shard_range); this module wraps it and is what the CPU (gloo) tests and bench.py use."""
import ctypes as C

import numpy as np

from . import cabi


def elimination_order(problem):
    """Landmark blocks (group 0, non-constant, referenced) in elimination order and the observations per landmark."""
    g = problem.groups.get(cabi.FACTOR_REPROJECTION)
    if g is None:
        return np.zeros(0, np.int64), np.zeros(1, np.int32)
    lm = g["blocks"][:, 2]
    blocks = np.unique(lm)                         # ascending block index == program order
    pos = np.searchsorted(blocks, lm)
    counts = np.bincount(pos, minlength=len(blocks))
    return blocks, np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)


def shard_range(lm_ptr, rank, world):
    lib = cabi.load_library()
    lo, hi = C.c_int32(), C.c_int32()
    lm_ptr = np.ascontiguousarray(lm_ptr, np.int32)
    rc = lib.bsnls_shard_range(lm_ptr.ctypes.data_as(C.POINTER(C.c_int32)), len(lm_ptr) - 1, rank, world, C.byref(lo), C.byref(hi))
    if rc != 0:
        raise ValueError("bsnls_shard_range rejected its arguments")
    return lo.value, hi.value


def shard_problem(problem, rank, world):
    """The factors rank `rank` evaluates: reprojection factors of its landmarks; every other factor on rank 0 only
    (in the library they are evaluated redundantly on every rank and counted once)."""
    blocks, lm_ptr = elimination_order(problem)
    lo, hi = shard_range(lm_ptr, rank, world)
    groups = {}
    n = 0
    for t, g in problem.groups.items():
        if t == cabi.FACTOR_REPROJECTION:
            keep = np.isin(g["blocks"][:, 2], blocks[lo:hi])
        else:
            keep = np.full(len(g["res_index"]), rank == 0)
        if not keep.any():
            continue
        sub = {k: (None if v is None else v[keep]) for k, v in g.items()}
        groups[t] = sub
        n += int(keep.sum())
    # res_index must be a permutation of 0..n-1: keep relative program order
    allri = np.concatenate([g["res_index"] for g in groups.values()]) if groups else np.zeros(0, np.int32)
    rank_of = {int(v): i for i, v in enumerate(np.sort(allri))}
    for g in groups.values():
        g["res_index"] = np.array([rank_of[int(v)] for v in g["res_index"]], np.int32)
    return cabi.FlatProblem(problem.block_size, problem.block_manifold, problem.block_constant, problem.block_group,
                            problem.block_values, problem.cameras, groups), (lo, hi)


def _point_at_loaded_nccl():
    """The library resolves NCCL by soname at run time; tell it (BSNLS_NCCL_LIB) which instance this process already
    mapped — the one torch.distributed uses — so that both share one NCCL."""
    import os
    if os.environ.get("BSNLS_NCCL_LIB"):
        return
    try:
        with open("/proc/self/maps") as f:
            for line in f:
                if "libnccl.so" in line:
                    os.environ["BSNLS_NCCL_LIB"] = line.split()[-1]
                    return
    except OSError:
        pass


def create_sharded_solver(options=None):
    """One Solver per rank of an initialised torch.distributed job (NCCL or gloo for the id exchange)."""
    import torch
    import torch.distributed as dist
    from .solver import Solver
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return Solver(options)
    lib = cabi.load_library()
    _point_at_loaded_nccl()
    buf = (C.c_char * 128)()
    if rank == 0 and lib.bsnls_nccl_unique_id(buf) != 0:
        raise RuntimeError(lib.bsnls_last_error(None).decode())
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.broadcast(t, 0)
    return Solver(options, sharded=(rank, world, bytes(t.cpu().tolist())))
