"""Thin object wrapper over the C-ABI (include/bsnls.h). Every method is one C call; errors raise with the
library's message. There is no fallback: without libbsnls.so or without a CUDA device these calls fail."""
import ctypes as C

import numpy as np

from . import cabi

_pf64 = C.POINTER(C.c_double)
_pi32 = C.POINTER(C.c_int32)


class BsnlsError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"bsnls error {code}: {message}")
        self.code = code


class Solver:
    def __init__(self, options=None, sharded=None):
        """sharded = (rank, world_size, nccl_id bytes) for the landmark-sharded multi-GPU solve."""
        self._lib = cabi.load_library()
        self._h = C.c_void_p()
        self.options = options or cabi.Options.defaults()
        if sharded is None:
            rc = self._lib.bsnls_create(C.byref(self.options), C.byref(self._h))
        else:
            rank, world, nccl_id = sharded
            buf = (C.c_char * 128).from_buffer_copy(nccl_id)
            rc = self._lib.bsnls_create_sharded(C.byref(self.options), rank, world, buf, C.byref(self._h))
        if rc != 0:
            msg = self._lib.bsnls_last_error(None)
            raise BsnlsError(rc, msg.decode() if msg else "?")
        self.problem = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.bsnls_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise BsnlsError(rc, self._lib.bsnls_last_error(self._h).decode())

    def set_options(self, options):
        self.options = options
        self._check(self._lib.bsnls_set_options(self._h, C.byref(options)))

    def set_problem(self, problem):
        self.problem = problem
        self._check(self._lib.bsnls_set_problem(self._h, C.byref(problem.c)))

    def set_values(self, values):
        v = np.ascontiguousarray(values, np.float64)
        assert v.size == self.problem.num_values
        self._check(self._lib.bsnls_set_values(self._h, v.ctypes.data_as(_pf64)))

    def get_values(self, out=None):
        v = out if out is not None else np.zeros(self.problem.num_values)
        self._check(self._lib.bsnls_get_values(self._h, v.ctypes.data_as(_pf64)))
        return v

    def solve(self, max_iters=None):
        n = (max_iters or self.options.max_num_iterations) + 2
        s = cabi.Summary()
        its = (cabi.Iteration * n)()
        self._check(self._lib.bsnls_solve(self._h, C.byref(s), its, n))
        return s, [its[i] for i in range(min(s.num_iterations, n))]

    def evaluate(self):
        cost = C.c_double()
        nr, ng = C.c_int64(), C.c_int64()
        self._check(self._lib.bsnls_evaluate(self._h, C.byref(cost), None, C.byref(nr), None, C.byref(ng)))
        res, grad = np.zeros(nr.value), np.zeros(ng.value)
        self._check(self._lib.bsnls_evaluate(self._h, C.byref(cost), res.ctypes.data_as(_pf64), C.byref(nr),
                                             grad.ctypes.data_as(_pf64), C.byref(ng)))
        return cost.value, res, grad

    def linearize(self, residual_blocks):
        """bsnls_linearize: per residual block (createProblem number) the loss-corrected residual (15, zero padded) and the
        tangent Jacobian (15 x 30: three columns per parameter block in the factor's block order)."""
        ids = np.ascontiguousarray(residual_blocks, np.int32)
        res, jac = np.zeros((len(ids), 15)), np.zeros((len(ids), 15, 30))
        self._check(self._lib.bsnls_linearize(self._h, len(ids), ids.ctypes.data_as(_pi32), res.ctypes.data_as(_pf64), jac.ctypes.data_as(_pf64)))
        return res, jac

    def ordering(self):
        nb = len(self.problem.block_size)
        e, f = np.zeros(nb, np.int32), np.zeros(nb, np.int32)
        ne, nf = C.c_int32(), C.c_int32()
        self._check(self._lib.bsnls_get_ordering(self._h, e.ctypes.data_as(_pi32), C.byref(ne), f.ctypes.data_as(_pi32), C.byref(nf)))
        return e[:ne.value].copy(), f[:nf.value].copy()

    def shard(self):
        """Positions (in the elimination order) of the landmarks this rank keeps."""
        n = C.c_int32(0)
        self._check(self._lib.bsnls_get_shard(self._h, None, C.byref(n)))
        out = np.zeros(n.value, np.int32)
        self._check(self._lib.bsnls_get_shard(self._h, out.ctypes.data_as(_pi32), C.byref(n)))
        return out

    def reduced_pattern(self):
        n = C.c_int64(0)
        self._check(self._lib.bsnls_get_reduced_pattern(self._h, None, None, C.byref(n)))
        r, c = np.zeros(n.value, np.int32), np.zeros(n.value, np.int32)
        self._check(self._lib.bsnls_get_reduced_pattern(self._h, r.ctypes.data_as(_pi32), c.ctypes.data_as(_pi32), C.byref(n)))
        return r, c

    def reduced_system(self, radius):
        _, f = self.ordering()
        dim = 3 * len(f)
        S, rhs = np.zeros((dim, dim)), np.zeros(dim)
        self._check(self._lib.bsnls_get_reduced_system(self._h, radius, S.ctypes.data_as(_pf64), rhs.ctypes.data_as(_pf64), dim))
        return S, rhs  # S is symmetric, so the column-major buffer reads the same row-major

    def covariance(self, pairs, tangent_space=True, sizes=None):
        """HashGraph::getCovariance: list of covariance blocks for (block_i, block_j) pairs at the current values."""
        pr = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        out = np.zeros((len(pr), 16))
        self._check(self._lib.bsnls_get_covariance(self._h, pr.ctypes.data_as(_pi32), len(pr), 1 if tangent_space else 0, out.ctypes.data_as(_pf64)))
        if tangent_space:
            return [out[k, :9].reshape(3, 3).copy() for k in range(len(pr))]
        return [out[k, :sizes[i] * sizes[j]].reshape(sizes[i], sizes[j]).copy() for k, (i, j) in enumerate(pr)]

    def reduced_blocks(self, radius):
        """Lower-triangular 3x3 blocks of the reduced system on its structural pattern: rows, cols, vals (n,3,3), rhs."""
        n = C.c_int64(0)
        self._check(self._lib.bsnls_get_reduced_blocks(self._h, radius, None, None, None, C.byref(n), None))
        r, c, v = np.zeros(n.value, np.int32), np.zeros(n.value, np.int32), np.zeros((n.value, 3, 3))
        _, f = self.ordering()
        rhs = np.zeros(3 * len(f))
        self._check(self._lib.bsnls_get_reduced_blocks(self._h, radius, r.ctypes.data_as(_pi32), c.ctypes.data_as(_pi32),
                                                       v.ctypes.data_as(_pf64), C.byref(n), rhs.ctypes.data_as(_pf64)))
        return r, c, v, rhs

    def trial_step(self, values, radius, keep_jacobi_scale):
        """One LM step from `values` with the given trust-region radius: list of 1 or 2 iteration records."""
        out = (cabi.Iteration * 2)()
        n = C.c_int32(0)
        v = None if values is None else np.ascontiguousarray(values, np.float64)
        self._check(self._lib.bsnls_trial_step(self._h, None if v is None else v.ctypes.data_as(_pf64), radius,
                                               1 if keep_jacobi_scale else 0, out, C.byref(n)))
        return [out[i] for i in range(n.value)]

    def run_eval(self):
        c = C.c_double()
        self._check(self._lib.bsnls_run_eval(self._h, C.byref(c)))
        return c.value

    def run_schur(self, radius=1e4):
        self._check(self._lib.bsnls_run_schur(self._h, radius))

    def run_pass(self, radius=1e4):
        """One complete linearisation (residuals, Jacobians, landmark elimination, reduced system); returns the cost."""
        c = C.c_double()
        self._check(self._lib.bsnls_run_pass(self._h, radius, C.byref(c)))
        return c.value

    def enable_breakdown(self, on=True):
        """run_pass with CUDA events between its kernels (direct launches) instead of one graph launch."""
        self._check(self._lib.bsnls_enable_breakdown(self._h, 1 if on else 0))

    def last_kernel_ms(self):
        a, b = C.c_float(), C.c_float()
        self._check(self._lib.bsnls_last_kernel_ms(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    BREAKDOWN = ("eval_reproj", "eval_generic", "landmark", "gradient", "clear_S", "pair", "generic_rows", "exchange")

    def last_breakdown_ms(self):
        out = (C.c_float * 8)()
        self._check(self._lib.bsnls_last_breakdown_ms(self._h, out))
        return dict(zip(self.BREAKDOWN, [float(v) for v in out]))

    def launch_count(self):
        return int(self._lib.bsnls_launch_count(self._h))
