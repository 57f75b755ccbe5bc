"""ctypes mirror of include/bsnls.h (the C-ABI boundary) and a numpy-backed holder for a flattened problem.

Plumbing only: structs, enums and the loader for beam_slam_b200/lib/libbsnls.so. The loader raises when the
CUDA library is missing; there is no CPU fallback anywhere in this package.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libbsnls.so")

# ---- enums (include/bsnls.h) ----
MANIFOLD_EUCLIDEAN, MANIFOLD_QUATERNION = 0, 1
(FACTOR_REPROJECTION, FACTOR_IMU_DELTA, FACTOR_IMU_PRIOR, FACTOR_RELATIVE_POSE, FACTOR_RELATIVE_POSE_EXTRINSICS,
 FACTOR_ABSOLUTE_POSE, FACTOR_RELATIVE_ORIENTATION, FACTOR_ABSOLUTE_VEC3, FACTOR_RELATIVE_VEC3, FACTOR_MARGINAL,
 FACTOR_GRAVITY_ALIGNMENT, FACTOR_ABSOLUTE_ORIENTATION) = range(12)
NUM_FACTOR_TYPES = 12
MARGINAL_MAX_BLOCKS, MARGINAL_MAX_ROWS = 10, 15
FACTOR_NBLOCKS = [3, 10, 5, 4, 6, 2, 2, 1, 2, 10, 1, 1]   # block slots per factor (the marginal prior pads unused ones with -1)
FACTOR_NRES = [2, 15, 15, 6, 6, 6, 3, 3, 3, 15, 2, 3]
FACTOR_STRIDE = [2, 288, 241, 43, 43, 43, 13, 12, 12, 508, 7, 13]
FACTOR_BLOCK_SIZES = [[4, 3, 3], [4, 3, 3, 3, 3, 4, 3, 3, 3, 3], [4, 3, 3, 3, 3], [3, 4, 3, 4],
                      [3, 4, 3, 4, 3, 4], [3, 4], [4, 4], [3], [3, 3], None, [4], [4]]
LOSS_NONE, LOSS_CAUCHY, LOSS_HUBER, LOSS_SOFTLONE, LOSS_ARCTAN, LOSS_TUKEY, LOSS_DCS, LOSS_FAIR, LOSS_GEMAN_MCCLURE, LOSS_WELSCH = range(10)
NOALLOC_MAX_OBSERVATIONS = 4194304
CAMERA_STRIDE = 20
LINEAR_DENSE_SCHUR, LINEAR_ITERATIVE_SCHUR, LINEAR_AUTO = 0, 1, 2
CONVERGENCE, NO_CONVERGENCE, FAILURE = 0, 1, 2
OK, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_STATE = 0, -1, -2, -3, -4

_pi32 = C.POINTER(C.c_int32)
_pf64 = C.POINTER(C.c_double)
_pu8 = C.POINTER(C.c_uint8)


class FactorGroup(C.Structure):
    _fields_ = [("type", C.c_int32), ("count", C.c_int32), ("blocks", _pi32), ("consts", _pf64),
                ("aux", _pi32), ("loss_type", _pi32), ("loss_param", _pf64), ("res_index", _pi32)]


class Problem(C.Structure):
    _fields_ = [("num_blocks", C.c_int32), ("block_size", _pi32), ("block_manifold", _pi32),
                ("block_constant", _pu8), ("block_group", _pi32), ("block_values", _pf64),
                ("num_cameras", C.c_int32), ("cameras", _pf64), ("num_groups", C.c_int32),
                ("groups", C.POINTER(FactorGroup))]


class Options(C.Structure):
    _fields_ = [("max_num_iterations", C.c_int32), ("max_solver_time_in_seconds", C.c_double),
                ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
                ("parameter_tolerance", C.c_double), ("initial_trust_region_radius", C.c_double),
                ("max_trust_region_radius", C.c_double), ("min_trust_region_radius", C.c_double),
                ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double),
                ("max_lm_diagonal", C.c_double), ("max_num_consecutive_invalid_steps", C.c_int32),
                ("jacobi_scaling", C.c_int32), ("linear_solver", C.c_int32), ("pcg_max_iterations", C.c_int32),
                ("pcg_tolerance", C.c_double), ("device", C.c_int32)]

    @staticmethod
    def defaults(**kw):
        """ceres::Solver::Options defaults (the values fuse_core/src/ceres_options.cpp falls back to)."""
        o = Options(50, 1e9, 1e-6, 1e-10, 1e-8, 1e4, 1e16, 1e-32, 1e-3, 1e-6, 1e32, 5, 1, LINEAR_AUTO, 500,
                    1e-13, -1)
        for k, v in kw.items():
            if not hasattr(o, k):
                raise AttributeError(k)
            setattr(o, k, v)
        return o


class Iteration(C.Structure):
    _fields_ = [("iteration", C.c_int32), ("step_is_valid", C.c_int32), ("step_is_successful", C.c_int32),
                ("linear_solver_iterations", C.c_int32), ("cost", C.c_double), ("cost_change", C.c_double),
                ("gradient_max_norm", C.c_double), ("gradient_norm", C.c_double), ("step_norm", C.c_double),
                ("relative_decrease", C.c_double), ("trust_region_radius", C.c_double)]


class Summary(C.Structure):
    _fields_ = [("termination_type", C.c_int32), ("num_iterations", C.c_int32),
                ("num_successful_steps", C.c_int32), ("num_unsuccessful_steps", C.c_int32),
                ("initial_cost", C.c_double), ("final_cost", C.c_double), ("fixed_cost", C.c_double),
                ("total_time_in_seconds", C.c_double), ("device_time_in_seconds", C.c_double),
                ("num_parameter_blocks_reduced", C.c_int32), ("num_parameters_reduced", C.c_int32),
                ("num_effective_parameters_reduced", C.c_int32), ("num_residual_blocks_reduced", C.c_int32),
                ("num_residuals_reduced", C.c_int32), ("num_e_blocks", C.c_int32), ("num_f_blocks", C.c_int32),
                ("reduced_system_dim", C.c_int32), ("linear_solver_used", C.c_int32),
                ("reduced_system_nnz_blocks", C.c_int64), ("message", C.c_char * 160)]

    def is_solution_usable(self):
        """ceres::Solver::Summary::IsSolutionUsable (the only check at fixed_lag_smoother.cpp:286)."""
        return self.termination_type in (CONVERGENCE, NO_CONVERGENCE)


def _ptr(a, ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


class FlatProblem:
    """Owns the numpy arrays of one flattened graph and exposes them as a bsnls_problem.

    groups: dict factor_type -> dict(blocks [n,nb] i32, consts [n,stride] f64, aux [n] i32 | None,
                                     loss_type [n] i32 | None, loss_param [n] f64 | None, res_index [n] i32)
    """

    def __init__(self, block_size, block_manifold, block_constant, block_group, block_values, cameras, groups):
        self.block_size = np.ascontiguousarray(block_size, np.int32)
        self.block_manifold = np.ascontiguousarray(block_manifold, np.int32)
        self.block_constant = np.ascontiguousarray(block_constant, np.uint8)
        self.block_group = np.ascontiguousarray(block_group, np.int32)
        self.block_values = np.ascontiguousarray(block_values, np.float64)
        self.cameras = np.ascontiguousarray(cameras, np.float64).reshape(-1, CAMERA_STRIDE)
        self.offsets = np.concatenate([[0], np.cumsum(self.block_size)]).astype(np.int64)
        assert self.block_values.size == self.offsets[-1]
        self.groups = {}
        for t, g in sorted(groups.items()):
            n = len(g["res_index"])
            d = dict(blocks=np.ascontiguousarray(g["blocks"], np.int32).reshape(n, FACTOR_NBLOCKS[t]),
                     consts=np.ascontiguousarray(g["consts"], np.float64).reshape(n, FACTOR_STRIDE[t]),
                     aux=None if g.get("aux") is None else np.ascontiguousarray(g["aux"], np.int32),
                     loss_type=None if g.get("loss_type") is None else np.ascontiguousarray(g["loss_type"], np.int32),
                     loss_param=None if g.get("loss_param") is None else np.ascontiguousarray(g["loss_param"], np.float64),
                     res_index=np.ascontiguousarray(g["res_index"], np.int32))
            self.groups[t] = d
        self._build()

    def _build(self):
        self._garr = (FactorGroup * max(1, len(self.groups)))()
        for i, (t, g) in enumerate(self.groups.items()):
            self._garr[i] = FactorGroup(t, len(g["res_index"]), _ptr(g["blocks"], _pi32), _ptr(g["consts"], _pf64),
                                        _ptr(g["aux"], _pi32), _ptr(g["loss_type"], _pi32),
                                        _ptr(g["loss_param"], _pf64), _ptr(g["res_index"], _pi32))
        self.c = Problem(len(self.block_size), _ptr(self.block_size, _pi32), _ptr(self.block_manifold, _pi32),
                         _ptr(self.block_constant, _pu8), _ptr(self.block_group, _pi32),
                         _ptr(self.block_values, _pf64), len(self.cameras), _ptr(self.cameras, _pf64),
                         len(self.groups), self._garr)

    @property
    def num_values(self):
        return int(self.offsets[-1])

    @property
    def num_residual_blocks(self):
        return sum(len(g["res_index"]) for g in self.groups.values())

    def count(self, t):
        return len(self.groups[t]["res_index"]) if t in self.groups else 0

    def value(self, block):
        return self.block_values[self.offsets[block]:self.offsets[block + 1]]

    def save(self, path):
        """Write the snapshot file (bsnls_problem_save: the SoA buffers in createProblem order)."""
        rc = load_library().bsnls_problem_save(C.byref(self.c), str(path).encode())
        if rc != 0:
            raise OSError(f"bsnls_problem_save({path}) failed: {rc}")

    @classmethod
    def load(cls, path):
        """Read a snapshot file back (bsnls_problem_load) into numpy-owned arrays."""
        lib = load_library()
        pp = C.POINTER(Problem)()
        rc = lib.bsnls_problem_load(str(path).encode(), C.byref(pp))
        if rc != 0:
            raise OSError(f"bsnls_problem_load({path}) failed: {rc}")
        try:
            cp = pp.contents
            arr = lambda ptr, n, dt: np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dt)
            nb = cp.num_blocks
            size = arr(cp.block_size, nb, np.int32)
            groups = {}
            for gi in range(cp.num_groups):
                g = cp.groups[gi]
                t, n = g.type, g.count
                groups[t] = dict(blocks=arr(g.blocks, n * FACTOR_NBLOCKS[t], np.int32), consts=arr(g.consts, n * FACTOR_STRIDE[t], np.float64),
                                 aux=arr(g.aux, n, np.int32) if g.aux else None, loss_type=arr(g.loss_type, n, np.int32) if g.loss_type else None,
                                 loss_param=arr(g.loss_param, n, np.float64) if g.loss_param else None, res_index=arr(g.res_index, n, np.int32))
            return cls(size, arr(cp.block_manifold, nb, np.int32), arr(cp.block_constant, nb, np.uint8), arr(cp.block_group, nb, np.int32),
                       arr(cp.block_values, int(size.sum()), np.float64), arr(cp.cameras, cp.num_cameras * CAMERA_STRIDE, np.float64), groups)
        finally:
            lib.bsnls_problem_free(pp)

    @classmethod
    def load_numpy(cls, path):
        """The same snapshot read with numpy alone (format: csrc/snapshot.cpp) — for consumers that must not load the
        CUDA library, e.g. the CPU reference arm of bench.py."""
        with open(path, "rb") as f:
            if f.read(8) != b"BSNLS\x00\x01\x00":
                raise OSError(f"{path}: not a bsnls snapshot (format 1)")
            rd = lambda dt, n: np.fromfile(f, dtype=dt, count=n)
            nb, ncam, ngroups, _ = (int(v) for v in rd(np.int32, 4))
            size, manifold, group = rd(np.int32, nb), rd(np.int32, nb), rd(np.int32, nb)
            constant = rd(np.uint8, nb)
            f.read((8 - nb % 8) % 8)
            values = rd(np.float64, int(size.sum()))
            cameras = rd(np.float64, ncam * CAMERA_STRIDE)
            groups = {}
            for _ in range(ngroups):
                t, n, has_aux, has_loss = (int(v) for v in rd(np.int32, 4))
                g = dict(blocks=rd(np.int32, n * FACTOR_NBLOCKS[t]), consts=rd(np.float64, n * FACTOR_STRIDE[t]), res_index=rd(np.int32, n),
                         aux=None, loss_type=None, loss_param=None)
                if has_aux:
                    g["aux"] = rd(np.int32, n)
                if has_loss:
                    g["loss_type"], g["loss_param"] = rd(np.int32, n), rd(np.float64, n)
                if len(g["res_index"]) != n:
                    raise OSError(f"{path}: truncated snapshot")
                groups[t] = g
        return cls(size, manifold, constant, group, values, cameras, groups)

    def with_values(self, values):
        p = FlatProblem.__new__(FlatProblem)
        p.__dict__.update(self.__dict__)
        p.block_values = np.ascontiguousarray(values, np.float64).copy()
        p._build()
        return p


_lib = None


def load_library():
    """Load the CUDA library. Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                           "g.build()'` (nvcc, sm_100a). beam_slam_b200 has no CPU path.")
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    lib.bsnls_default_options.argtypes = [C.POINTER(Options)]
    lib.bsnls_default_options.restype = None
    lib.bsnls_create.argtypes = [C.POINTER(Options), C.POINTER(H)]
    lib.bsnls_destroy.argtypes = [H]
    lib.bsnls_destroy.restype = None
    lib.bsnls_last_error.argtypes = [H]
    lib.bsnls_last_error.restype = C.c_char_p
    lib.bsnls_set_problem.argtypes = [H, C.POINTER(Problem)]
    lib.bsnls_set_values.argtypes = [H, _pf64]
    lib.bsnls_set_options.argtypes = [H, C.POINTER(Options)]
    lib.bsnls_solve.argtypes = [H, C.POINTER(Summary), C.POINTER(Iteration), C.c_int32]
    lib.bsnls_get_values.argtypes = [H, _pf64]
    lib.bsnls_evaluate.argtypes = [H, _pf64, _pf64, C.POINTER(C.c_int64), _pf64, C.POINTER(C.c_int64)]
    lib.bsnls_get_ordering.argtypes = [H, _pi32, _pi32, _pi32, _pi32]
    lib.bsnls_get_reduced_pattern.argtypes = [H, _pi32, _pi32, C.POINTER(C.c_int64)]
    lib.bsnls_get_reduced_system.argtypes = [H, C.c_double, _pf64, _pf64, C.c_int32]
    lib.bsnls_get_covariance.argtypes = [H, _pi32, C.c_int32, C.c_int32, _pf64]
    lib.bsnls_get_covariance.restype = C.c_int
    lib.bsnls_get_reduced_blocks.argtypes = [H, C.c_double, _pi32, _pi32, _pf64, C.POINTER(C.c_int64), _pf64]
    lib.bsnls_trial_step.argtypes = [H, _pf64, C.c_double, C.c_int32, C.POINTER(Iteration), _pi32]
    lib.bsnls_run_eval.argtypes = [H, _pf64]
    lib.bsnls_run_schur.argtypes = [H, C.c_double]
    lib.bsnls_run_pass.argtypes = [H, C.c_double, _pf64]
    lib.bsnls_enable_breakdown.argtypes = [H, C.c_int32]
    lib.bsnls_enable_breakdown.restype = C.c_int
    lib.bsnls_last_kernel_ms.argtypes = [H, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.bsnls_last_breakdown_ms.argtypes = [H, C.POINTER(C.c_float)]
    lib.bsnls_last_breakdown_ms.restype = C.c_int
    lib.bsnls_launch_count.argtypes = [H]
    lib.bsnls_launch_count.restype = C.c_int64
    lib.bsnls_reprojection_cost_noalloc.argtypes = [C.c_void_p] * 7 + [C.c_int32, C.c_void_p, C.c_void_p, C.c_char_p, C.c_int32]
    lib.bsnls_reprojection_cost_noalloc.restype = C.c_int
    lib.bsnls_shard_range.argtypes = [_pi32, C.c_int32, C.c_int32, C.c_int32, _pi32, _pi32]
    lib.bsnls_shard_range.restype = C.c_int
    lib.bsnls_problem_save.argtypes = [C.POINTER(Problem), C.c_char_p]
    lib.bsnls_problem_save.restype = C.c_int
    lib.bsnls_problem_load.argtypes = [C.c_char_p, C.POINTER(C.POINTER(Problem))]
    lib.bsnls_problem_load.restype = C.c_int
    lib.bsnls_problem_free.argtypes = [C.POINTER(Problem)]
    lib.bsnls_problem_free.restype = None
    lib.bsnls_get_shard.argtypes = [H, _pi32, _pi32]
    lib.bsnls_get_shard.restype = C.c_int
    lib.bsnls_linearize.argtypes = [H, C.c_int32, _pi32, _pf64, _pf64]
    lib.bsnls_linearize.restype = C.c_int
    lib.bsnls_nccl_unique_id.argtypes = [C.c_char * 128]
    lib.bsnls_create_sharded.argtypes = [C.POINTER(Options), C.c_int32, C.c_int32, C.c_char * 128, C.POINTER(H)]
    for name in ("bsnls_create", "bsnls_set_problem", "bsnls_set_values", "bsnls_set_options", "bsnls_solve",
                 "bsnls_get_values", "bsnls_evaluate", "bsnls_get_ordering", "bsnls_get_reduced_pattern",
                 "bsnls_get_reduced_system", "bsnls_get_reduced_blocks", "bsnls_trial_step", "bsnls_run_eval", "bsnls_run_schur", "bsnls_run_pass", "bsnls_last_kernel_ms",
                 "bsnls_nccl_unique_id", "bsnls_create_sharded"):
        getattr(lib, name).restype = C.c_int
    _lib = lib
    return lib
