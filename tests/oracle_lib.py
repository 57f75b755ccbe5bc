"""Loader for oracle/liboracle.so — the CPU checker. Test infrastructure: imported only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
import ctypes as C
import os
import subprocess

import numpy as np

from beam_slam_b200 import cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_lib = None
_pf64 = C.POINTER(C.c_double)
_pi32 = C.POINTER(C.c_int32)


def load():
    global _lib
    if _lib is not None:
        return _lib
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(so):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    lib = C.CDLL(so)
    lib.oracle_solve.argtypes = [C.POINTER(cabi.Problem), C.POINTER(cabi.Options), C.POINTER(cabi.Summary),
                                 C.POINTER(cabi.Iteration), C.c_int32, _pf64]
    lib.oracle_evaluate.argtypes = [C.POINTER(cabi.Problem), _pf64, _pf64, C.POINTER(C.c_int64), _pf64,
                                    C.POINTER(C.c_int64)]
    lib.oracle_cost_function.argtypes = [C.POINTER(cabi.Problem), C.c_int32, C.c_int32, _pf64, _pf64]
    lib.oracle_linearize.argtypes = [C.POINTER(cabi.Problem), C.c_int32, C.c_int32, _pf64, _pf64, _pi32]
    lib.oracle_reduced_system.argtypes = [C.POINTER(cabi.Problem), C.POINTER(cabi.Options), C.c_double, _pf64,
                                          _pf64, _pi32]
    lib.oracle_get_ordering.argtypes = [C.POINTER(cabi.Problem), _pi32, _pi32, _pi32, _pi32]
    lib.oracle_quat_plus.argtypes = [_pf64, _pf64, _pf64]
    lib.oracle_quat_minus.argtypes = [_pf64, _pf64, _pf64]
    lib.oracle_quat_plus_jacobian.argtypes = [_pf64, _pf64]
    lib.oracle_point_transformation.argtypes = [_pf64, _pf64, _pf64, _pf64]
    lib.oracle_time_eval_schur.argtypes = [C.POINTER(cabi.Problem), C.POINTER(cabi.Options), C.c_double, C.c_int32,
                                           _pi32]
    lib.oracle_time_eval_schur.restype = C.c_double
    lib.oracle_set_linear_mode.argtypes = [C.c_int32]
    lib.oracle_loss.argtypes = [C.c_int32, C.c_double, C.c_double, _pf64]
    lib.oracle_set_num_threads.argtypes = [C.c_int32]
    lib.oracle_reduced_blocks.argtypes = [C.POINTER(cabi.Problem), C.POINTER(cabi.Options), C.c_double, _pi32, _pi32, _pf64,
                                          C.POINTER(C.c_int64), _pf64]
    lib.oracle_solve_trace.argtypes = [C.POINTER(cabi.Problem), C.POINTER(cabi.Options), C.POINTER(cabi.Summary),
                                       C.POINTER(cabi.Iteration), C.c_int32, _pf64, _pf64, C.c_int32, _pi32]
    _lib = lib
    return lib


def _p(a):
    return a.ctypes.data_as(_pf64)


def solve(problem, options=None, max_iters=256):
    lib = load()
    options = options or cabi.Options.defaults()
    s = cabi.Summary()
    its = (cabi.Iteration * max_iters)()
    x = np.zeros(problem.num_values)
    rc = lib.oracle_solve(C.byref(problem.c), C.byref(options), C.byref(s), its, max_iters, _p(x))
    assert rc == 0, s.message
    return s, [its[i] for i in range(min(s.num_iterations, max_iters))], x


LINEAR_AUTO, LINEAR_DENSE, LINEAR_SPARSE = 0, 1, 2


def loss(loss_type, a, s):
    """rho(s), rho'(s), rho''(s) of one loss class."""
    rho = np.zeros(3)
    load().oracle_loss(loss_type, a, s, _p(rho))
    return rho


def set_num_threads(n):
    """OpenMP threads of the oracle (torchrun sets OMP_NUM_THREADS=1 for its workers)."""
    load().oracle_set_num_threads(int(n))


def set_linear_mode(mode):
    """Reduced-system storage / Cholesky of the oracle: automatic (dense up to 2400 unknowns), dense, block-sparse."""
    load().oracle_set_linear_mode(mode)


def solve_trace(problem, options=None, max_iters=256):
    """solve() + the linearisation point of every attempted step: trace[k-1] is where iteration k linearised."""
    lib = load()
    options = options or cabi.Options.defaults()
    s = cabi.Summary()
    its = (cabi.Iteration * max_iters)()
    x = np.zeros(problem.num_values)
    trace = np.zeros((max_iters, problem.num_values))
    n = C.c_int32()
    rc = lib.oracle_solve_trace(C.byref(problem.c), C.byref(options), C.byref(s), its, max_iters, _p(x), _p(trace), max_iters, C.byref(n))
    assert rc == 0, s.message
    return s, [its[i] for i in range(min(s.num_iterations, max_iters))], x, trace[:min(n.value, max_iters)].copy()


def reduced_blocks(problem, radius, options=None):
    """Lower-triangular 3x3 blocks of the reduced system on its structural pattern: rows, cols (f-block positions in
    program order, sorted by (row, col)), vals (n, 3, 3), rhs."""
    lib = load()
    options = options or cabi.Options.defaults()
    n = C.c_int64()
    assert lib.oracle_reduced_blocks(C.byref(problem.c), C.byref(options), radius, None, None, None, C.byref(n), None) == 0
    rows, cols, vals = np.zeros(n.value, np.int32), np.zeros(n.value, np.int32), np.zeros((n.value, 3, 3))
    _, f = ordering(problem)
    rhs = np.zeros(3 * len(f))
    assert lib.oracle_reduced_blocks(C.byref(problem.c), C.byref(options), radius, rows.ctypes.data_as(_pi32),
                                     cols.ctypes.data_as(_pi32), _p(vals), C.byref(n), _p(rhs)) == 0
    return rows, cols, vals, rhs


def evaluate(problem):
    lib = load()
    cost = C.c_double()
    nres, ngrad = C.c_int64(), C.c_int64()
    rc = lib.oracle_evaluate(C.byref(problem.c), C.byref(cost), None, C.byref(nres), None, C.byref(ngrad))
    assert rc == 0
    res, grad = np.zeros(nres.value), np.zeros(ngrad.value)
    lib.oracle_evaluate(C.byref(problem.c), C.byref(cost), _p(res), C.byref(nres), _p(grad), C.byref(ngrad))
    return cost.value, res, grad


def cost_function(problem, ftype, index, sizes=None):
    """Raw CostFunction::Evaluate: residual, list of ambient row-major Jacobians per block (sizes: the ambient block sizes of
    a factor type with a variable shape -- the marginal prior)."""
    lib = load()
    nr = cabi.FACTOR_NRES[ftype]
    sizes = sizes if sizes is not None else cabi.FACTOR_BLOCK_SIZES[ftype]
    res = np.zeros(nr)
    jac = np.zeros(nr * sum(sizes))
    assert lib.oracle_cost_function(C.byref(problem.c), ftype, index, _p(res), _p(jac)) == 0
    out, o = [], 0
    for s in sizes:
        out.append(jac[o:o + nr * s].reshape(nr, s).copy())
        o += nr * s
    return res, out


def linearize(problem, ftype, index):
    lib = load()
    nr = cabi.FACTOR_NRES[ftype]
    res = np.zeros(nr)
    jac = np.zeros(nr * 32)
    nc = C.c_int32()
    assert lib.oracle_linearize(C.byref(problem.c), ftype, index, _p(res), _p(jac), C.byref(nc)) == 0
    return res, jac[:nr * nc.value].reshape(nr, nc.value).copy()


def reduced_system(problem, radius, options=None):
    lib = load()
    options = options or cabi.Options.defaults()
    dim = C.c_int32()
    assert lib.oracle_reduced_system(C.byref(problem.c), C.byref(options), radius, None, None, C.byref(dim)) == 0
    S, rhs = np.zeros((dim.value, dim.value)), np.zeros(dim.value)
    assert lib.oracle_reduced_system(C.byref(problem.c), C.byref(options), radius, _p(S), _p(rhs), C.byref(dim)) == 0
    return S, rhs


def ordering(problem):
    lib = load()
    ne, nf = C.c_int32(), C.c_int32()
    e = np.zeros(len(problem.block_size), np.int32)
    f = np.zeros(len(problem.block_size), np.int32)
    assert lib.oracle_get_ordering(C.byref(problem.c), e.ctypes.data_as(_pi32), C.byref(ne),
                                   f.ctypes.data_as(_pi32), C.byref(nf)) == 0
    return e[:ne.value].copy(), f[:nf.value].copy()


def quat_plus(x, d):
    lib = load()
    x, d, out = np.asarray(x, float), np.asarray(d, float), np.zeros(4)
    lib.oracle_quat_plus(_p(x), _p(d), _p(out))
    return out


def quat_minus(x1, x2):
    lib = load()
    x1, x2, out = np.asarray(x1, float), np.asarray(x2, float), np.zeros(3)
    lib.oracle_quat_minus(_p(x1), _p(x2), _p(out))
    return out


def time_eval_schur(problem, radius=1e4, reps=3, options=None):
    lib = load()
    options = options or cabi.Options.defaults()
    th = C.c_int32()
    sec = lib.oracle_time_eval_schur(C.byref(problem.c), C.byref(options), radius, reps, C.byref(th))
    return sec, th.value
