"""N>1 host logic on CPU: world_size-2 gloo job. Each rank takes its landmark shard (bsnls_shard_range), evaluates
it with the oracle, and the all-reduced sums reproduce the single-process result — the linearity the multi-GPU solve
relies on (cost and pose gradient are sums over landmarks)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_lib as orc
from beam_slam_b200 import cabi, host, sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = host.HashGraph.synthetic(2, scale=0.2).flatten()
    sub, (lo, hi) = sharding.shard_problem(p, rank, world)
    cost, _, grad = orc.evaluate(sub)
    # scatter the shard's tangent gradient into the full problem's program-order gradient, then sum over ranks
    full_cost, _, full_grad = orc.evaluate(p)
    active_full = _active(p)
    active_sub = _active(sub)
    g = np.zeros(len(full_grad))
    pos = {b: i for i, b in enumerate(active_full)}
    for i, b in enumerate(active_sub):
        g[3 * pos[b]:3 * pos[b] + 3] = grad[3 * i:3 * i + 3]
    t = torch.from_numpy(np.concatenate([[cost], g]))
    dist.all_reduce(t)
    ranges = [None] * world
    dist.all_gather_object(ranges, (lo, hi))
    if rank == 0:
        out["cost"] = (float(t[0]), full_cost)
        out["grad"] = float(np.abs(t[1:].numpy() - full_grad).max() / np.abs(full_grad).max())
        out["ranges"] = ranges
        blocks, lm_ptr = sharding.elimination_order(p)
        out["obs"] = [int(lm_ptr[h] - lm_ptr[l]) for l, h in ranges]
        out["total"] = (len(blocks), int(lm_ptr[-1]))
    dist.destroy_process_group()


def _active(p):
    used = set()
    for t, g in p.groups.items():
        for row in g["blocks"]:
            used.update(int(b) for b in row if not p.block_constant[int(b)])
    return sorted(used)


def test_two_rank_landmark_sharding_is_a_partition_and_sums_back():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    (lo0, hi0), (lo1, hi1) = out["ranges"]
    nlm, nobs = out["total"]
    assert lo0 == 0 and hi0 == lo1 and hi1 == nlm                    # contiguous partition of the elimination order
    assert sum(out["obs"]) == nobs and abs(out["obs"][0] - out["obs"][1]) <= 0.05 * nobs  # balanced by observations
    c, cf = out["cost"]
    assert abs(c - cf) <= 1e-12 * cf
    assert out["grad"] < 1e-12


def test_shard_range_edge_cases():
    lm_ptr = np.array([0, 3, 3, 10, 11], np.int32)
    assert sharding.shard_range(lm_ptr, 0, 1) == (0, 4)
    parts = [sharding.shard_range(lm_ptr, r, 3) for r in range(3)]
    assert parts[0][0] == 0 and parts[-1][1] == 4 and all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
    # more ranks than landmarks: empty shards are allowed, still a partition
    parts = [sharding.shard_range(lm_ptr, r, 8) for r in range(8)]
    assert parts[0][0] == 0 and parts[-1][1] == 4 and all(a[1] == b[0] and a[0] <= a[1] for a, b in zip(parts, parts[1:]))


def _pcg_worker(rank, world, port, out):
    """The sharded solve's conjugate gradients (csrc/bsnls.cu RunPartitionedPcg) restated with numpy + gloo: block rows of S
    by bsnls_shard_range on the non-zero-block prefix, own x / r / z, whole p, one scalar all-reduce and one all-reduce of
    the zero-padded z slice with the two dot products per iteration."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = host.HashGraph.synthetic(3, scale=0.05).flatten()
    S, b = orc.reduced_system(p, 1e4)
    n = len(b); nf = n // 3
    blk = np.abs(S.reshape(nf, 3, nf, 3)).max(axis=(1, 3)) > 0
    rowptr = np.concatenate([[0], np.cumsum(blk.sum(1))]).astype(np.int32)
    rb0, rb1 = sharding.shard_range(rowptr, rank, world)
    lo, hi = 3 * rb0, 3 * rb1
    Minv = [np.linalg.inv(S[3 * i:3 * i + 3, 3 * i:3 * i + 3]) for i in range(nf)]
    prec = lambda r, i0, i1: np.concatenate([Minv[i] @ r[3 * (i - i0):3 * (i - i0) + 3] for i in range(i0, i1)]) if i1 > i0 else np.zeros(0)
    x = np.zeros(hi - lo); r = b[lo:hi].copy(); z = prec(r, rb0, rb1)
    send = np.zeros(n + 2); send[lo:hi] = z; send[n] = r @ r; send[n + 1] = r @ z
    t = torch.from_numpy(send.copy()); dist.all_reduce(t); full = t.numpy()
    pvec = full[:n].copy(); bb, rz = full[n], full[n + 1]
    iters = 0
    for iters in range(1, 501):
        Ap = S[lo:hi] @ pvec
        t = torch.tensor([float(Ap @ pvec[lo:hi])], dtype=torch.float64); dist.all_reduce(t); pAp = float(t[0])
        alpha = rz / pAp
        x += alpha * pvec[lo:hi]; r -= alpha * Ap; z = prec(r, rb0, rb1)
        send = np.zeros(n + 2); send[lo:hi] = z; send[n] = r @ r; send[n + 1] = r @ z
        t = torch.from_numpy(send.copy()); dist.all_reduce(t); full = t.numpy()
        if full[n] <= 1e-26 * bb:
            break
        pvec = full[:n] + (full[n + 1] / rz) * pvec; rz = full[n + 1]
    xs = np.zeros(n); xs[lo:hi] = x
    t = torch.from_numpy(xs); dist.all_reduce(t)
    ranges = [None] * world
    dist.all_gather_object(ranges, (rb0, rb1))
    if rank == 0:
        out["err"] = float(np.abs(t.numpy() - np.linalg.solve(S, b)).max() / np.abs(np.linalg.solve(S, b)).max())
        out["iters"] = iters
        out["ranges"] = ranges
        out["nf"] = nf
    dist.destroy_process_group()


def test_two_rank_partitioned_conjugate_gradients():
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_pcg_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    (a0, a1), (b0, b1) = out["ranges"]
    assert a0 == 0 and a1 == b0 and b1 == out["nf"] and a1 > a0 and b1 > b0     # the block rows are partitioned
    assert out["err"] < 1e-9 and out["iters"] < 500
