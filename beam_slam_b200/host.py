"""ctypes plumbing over lib/libbsfuse.so: the C++ host mirror of the fuse graph interface (host/), its
createProblem-order flattener and the synthetic-graph generator. `HashGraph.optimize` is the drop-in call
(GpuHashGraph::optimize, host/gpu_hash_graph.cpp) and runs on the GPU only."""
import ctypes as C
import os

import numpy as np

from . import cabi

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libbsfuse.so")
_lib = None


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
    lib = C.CDLL(LIB_PATH)
    G = C.c_void_p
    lib.bsfuse_graph_new.restype = G
    lib.bsfuse_graph_free.argtypes = [G]
    lib.bsfuse_last_error.argtypes = [G]
    lib.bsfuse_last_error.restype = C.c_char_p
    lib.bsfuse_build_synthetic.argtypes = [G, C.c_int, C.c_double, C.c_uint64, C.POINTER(C.c_longlong)]
    lib.bsfuse_synthetic_options.argtypes = [C.c_int, C.POINTER(cabi.Options)]
    lib.bsfuse_flatten.argtypes = [G]
    lib.bsfuse_flatten.restype = C.POINTER(cabi.Problem)
    lib.bsfuse_num_values.argtypes = [G]
    lib.bsfuse_num_values.restype = C.c_longlong
    lib.bsfuse_block_uuid.argtypes = [G, C.c_int, C.c_char_p]
    lib.bsfuse_truth_values.argtypes = [G, C.POINTER(C.c_double)]
    lib.bsfuse_current_values.argtypes = [G, C.POINTER(C.c_double)]
    lib.bsfuse_set_values.argtypes = [G, C.POINTER(C.c_double)]
    lib.bsfuse_optimize.argtypes = [G, C.POINTER(cabi.Options), C.POINTER(cabi.Summary), C.POINTER(cabi.Iteration), C.c_int]
    lib.bsfuse_optimize_for.argtypes = [G, C.c_double, C.POINTER(cabi.Options), C.POINTER(cabi.Summary), C.POINTER(cabi.Iteration), C.c_int]
    lib.bsfuse_last_timing.argtypes = [G, C.POINTER(C.c_double)]
    lib.bsfuse_evaluate.argtypes = [G, C.POINTER(C.c_double)]
    lib.bsfuse_get_covariance.argtypes = [G, C.POINTER(C.c_int), C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.bsfuse_build_pose_chain.argtypes = [G, C.c_int, C.c_uint64]
    lib.bsfuse_build_orientation_chain.argtypes = [G]
    lib.bsfuse_marginalize.argtypes = [G, C.c_char_p, C.c_int, C.POINTER(C.c_int)]
    _pi = C.POINTER(C.c_int)
    lib.bsfuse_marginalize_next.argtypes = [C.c_int, _pi, _pi, _pi, C.POINTER(C.c_double), C.POINTER(C.c_double), _pi, _pi,
                                            C.POINTER(C.c_double), C.POINTER(C.c_double)]
    _pd = C.POINTER(C.c_double)
    lib.bsfuse_preintegrate.argtypes = [_pd, C.c_int, C.c_double, _pd, _pd, _pd, C.c_double, _pd, _pd]
    R = C.c_void_p
    lib.bsfuse_replay_new.argtypes = [C.c_int, C.c_int, C.c_uint64]
    lib.bsfuse_replay_new.restype = R
    lib.bsfuse_replay_free.argtypes = [R]
    lib.bsfuse_replay_error.argtypes = [R]
    lib.bsfuse_replay_error.restype = C.c_char_p
    lib.bsfuse_replay_done.argtypes = [R]
    lib.bsfuse_replay_advance.argtypes = [R]
    lib.bsfuse_replay_advance.restype = C.c_double
    lib.bsfuse_replay_flatten.argtypes = [R]
    lib.bsfuse_replay_flatten.restype = C.POINTER(cabi.Problem)
    lib.bsfuse_replay_observations.argtypes = [R]
    lib.bsfuse_replay_observations.restype = C.c_longlong
    lib.bsfuse_replay_optimize.argtypes = [R, C.POINTER(cabi.Options), C.POINTER(cabi.Summary), C.POINTER(cabi.Iteration), C.c_int, C.POINTER(C.c_double)]
    lib.bsfuse_replay_current_values.argtypes = [R, C.POINTER(C.c_double)]
    lib.bsfuse_uuid_name.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.c_char_p]
    lib.bsfuse_uuid_stamped.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_char_p]
    lib.bsfuse_uuid_landmark.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p]
    lib.bsfuse_uuid_hash.argtypes = [C.c_char_p]
    lib.bsfuse_uuid_hash.restype = C.c_uint64
    lib.bsfuse_selftest_containers.restype = C.c_int
    _lib = lib
    return lib


def _np(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def problem_to_numpy(cp):
    """Copy a C bsnls_problem (owned by C++) into a numpy-backed cabi.FlatProblem."""
    nb = cp.num_blocks
    size = _np(cp.block_size, nb, np.int32)
    nv = int(size.sum())
    groups = {}
    for gi in range(cp.num_groups):
        g = cp.groups[gi]
        t, n = g.type, g.count
        groups[t] = dict(blocks=_np(g.blocks, n * cabi.FACTOR_NBLOCKS[t], np.int32),
                         consts=_np(g.consts, n * cabi.FACTOR_STRIDE[t], np.float64),
                         aux=_np(g.aux, n, np.int32) if g.aux else None,
                         loss_type=_np(g.loss_type, n, np.int32) if g.loss_type else None,
                         loss_param=_np(g.loss_param, n, np.float64) if g.loss_param else None,
                         res_index=_np(g.res_index, n, np.int32))
    return cabi.FlatProblem(size, _np(cp.block_manifold, nb, np.int32), _np(cp.block_constant, nb, np.uint8),
                            _np(cp.block_group, nb, np.int32), _np(cp.block_values, nv, np.float64),
                            _np(cp.cameras, cp.num_cameras * cabi.CAMERA_STRIDE, np.float64), groups)


class HashGraph:
    """Python face of bsfuse::GpuHashGraph (mirrors fuse_graphs::HashGraph's optimize/evaluate)."""

    INFO = ("num_keyframes", "num_landmarks", "num_observations", "num_imu", "num_relative_pose", "num_priors",
            "num_variables", "num_constraints")

    def __init__(self):
        self._lib = load_library()
        self._g = self._lib.bsfuse_graph_new()
        self.info = {}

    def __del__(self):
        if getattr(self, "_g", None):
            self._lib.bsfuse_graph_free(self._g)
            self._g = None

    @classmethod
    def synthetic(cls, config, scale=1.0, seed=0):
        g = cls()
        out = (C.c_longlong * 8)()
        if g._lib.bsfuse_build_synthetic(g._g, config, scale, seed, out) != 0:
            raise ValueError(g._lib.bsfuse_last_error(g._g).decode())
        g.info = dict(zip(cls.INFO, [int(v) for v in out]))
        g.config = config
        return g

    @classmethod
    def pose_chain(cls, n, seed=0):
        """n stamped poses, a prior on the first, odometry between neighbours and the links 0-2 and 1-3 (no landmarks)."""
        g = cls()
        if g._lib.bsfuse_build_pose_chain(g._g, n, seed) != 0:
            raise ValueError(g._lib.bsfuse_last_error(g._g).decode())
        g.config = 3
        return g

    def marginalize(self, uuids):
        """fuse_constraints::marginalizeVariables (host/marginalize.h) for the variables with these UUIDs, then
        Graph::update with the transaction. Returns (constraints removed, marginal constraints added, variables removed)."""
        out = (C.c_int * 3)()
        rc = self._lib.bsfuse_marginalize(self._g, b"".join(uuids), len(uuids), out)
        if rc == -1:
            raise KeyError(self._lib.bsfuse_last_error(self._g).decode())
        if rc == -2:
            raise ValueError(self._lib.bsfuse_last_error(self._g).decode())
        if rc != 0:
            raise RuntimeError(self._lib.bsfuse_last_error(self._g).decode())
        return tuple(int(v) for v in out)

    @staticmethod
    def synthetic_options(config):
        o = cabi.Options()
        load_library().bsfuse_synthetic_options(config, C.byref(o))
        return o

    def flatten(self):
        cp = self._lib.bsfuse_flatten(self._g)
        if not cp:
            raise ValueError(self._lib.bsfuse_last_error(self._g).decode())
        return problem_to_numpy(cp.contents)

    @classmethod
    def orientation_chain(cls):
        """The four orientations and four constraints of test_marginalize_variables.cpp:416-470, literal values."""
        g = cls()
        if g._lib.bsfuse_build_orientation_chain(g._g) != 0:
            raise ValueError(g._lib.bsfuse_last_error(g._g).decode())
        g.config = 3
        return g

    def block_uuids(self, n=None):
        """UUID (16 bytes) of the first n blocks (default: all), in flatten() order."""
        n = len(self.flatten().block_size) if n is None else n
        out = []
        buf = C.create_string_buffer(16)
        for i in range(n):
            self._lib.bsfuse_block_uuid(self._g, i, buf)
            out.append(buf.raw)
        return out

    def truth_values(self):
        v = np.zeros(self._lib.bsfuse_num_values(self._g))
        self._lib.bsfuse_truth_values(self._g, v.ctypes.data_as(C.POINTER(C.c_double)))
        return v

    def current_values(self):
        v = np.zeros(self._lib.bsfuse_num_values(self._g))
        self._lib.bsfuse_current_values(self._g, v.ctypes.data_as(C.POINTER(C.c_double)))
        return v

    def set_values(self, values):
        v = np.ascontiguousarray(values, np.float64)
        self._lib.bsfuse_set_values(self._g, v.ctypes.data_as(C.POINTER(C.c_double)))

    TIMING = ("flatten_s", "upload_s", "solve_s", "download_s", "structure_reused")

    def last_timing(self):
        """Host wall-clock split of the last optimize(): flatten (or value gather), bsnls_set_problem (or
        bsnls_set_values), bsnls_solve, bsnls_get_values + scatter; structure_reused = 1 when the device problem was kept."""
        t = (C.c_double * 5)()
        self._lib.bsfuse_last_timing(self._g, t)
        return dict(zip(self.TIMING, [float(v) for v in t]))

    def optimize(self, options=None, max_iters=256, max_time_s=None):
        """GpuHashGraph::optimize (optimizeFor when max_time_s is given): flatten in createProblem order (or just gather
        the values when the graph's structure is unchanged), solve on the GPU, write variables back."""
        options = options or cabi.Options.defaults()
        s = cabi.Summary()
        its = (cabi.Iteration * max_iters)()
        rc = self._lib.bsfuse_optimize_for(self._g, -1.0 if max_time_s is None else float(max_time_s), C.byref(options), C.byref(s), its, max_iters)
        if rc == -1:
            raise ValueError(self._lib.bsfuse_last_error(self._g).decode())
        if rc != 0:
            raise RuntimeError(self._lib.bsfuse_last_error(self._g).decode())
        return s, [its[i] for i in range(min(s.num_iterations, max_iters))]

    def covariance(self, block_pairs, tangent_space=True):
        """GpuHashGraph::getCovariance for pairs of block indices (positions in flatten() order): 3x3 tangent blocks, or
        flat ambient blocks."""
        pr = np.ascontiguousarray(block_pairs, np.int32).reshape(-1, 2)
        out = np.zeros((len(pr), 16))
        rc = self._lib.bsfuse_get_covariance(self._g, pr.ctypes.data_as(C.POINTER(C.c_int)), len(pr), 1 if tangent_space else 0,
                                             out.ctypes.data_as(C.POINTER(C.c_double)))
        if rc == -1:
            raise KeyError(self._lib.bsfuse_last_error(self._g).decode())
        if rc != 0:
            raise RuntimeError(self._lib.bsfuse_last_error(self._g).decode())
        return [out[k, :9].reshape(3, 3).copy() for k in range(len(pr))] if tangent_space else [out[k].copy() for k in range(len(pr))]

    def evaluate(self):
        c = C.c_double()
        if self._lib.bsfuse_evaluate(self._g, C.byref(c)) != 0:
            raise RuntimeError(self._lib.bsfuse_last_error(self._g).decode())
        return c.value


def marginalize_next(terms):
    """fuse_constraints::detail::marginalizeNext (host/marginalize.cpp). terms: list of (variables, [A_i rows x 3], b).
    Returns (variables, [A_i], b) of the marginal on the variables other than the lowest-ordered one."""
    lib = load_library()
    rows = np.array([len(t[2]) for t in terms], np.int32); nv = np.array([len(t[0]) for t in terms], np.int32)
    vs = np.array([v for t in terms for v in t[0]], np.int32)
    A = np.concatenate([np.asarray(a, np.float64).reshape(-1) for t in terms for a in t[1]])
    b = np.concatenate([np.asarray(t[2], np.float64) for t in terms])
    onv = C.c_int(); ov = np.zeros(16, np.int32); oA = np.zeros(16 * 64 * 3); ob = np.zeros(64)
    pi, pd = C.POINTER(C.c_int), C.POINTER(C.c_double)
    r = lib.bsfuse_marginalize_next(len(terms), rows.ctypes.data_as(pi), nv.ctypes.data_as(pi), vs.ctypes.data_as(pi), A.ctypes.data_as(pd),
                                    b.ctypes.data_as(pd), C.byref(onv), ov.ctypes.data_as(pi), oA.ctypes.data_as(pd), ob.ctypes.data_as(pd))
    n = onv.value
    return [int(v) for v in ov[:n]], [oA[i * r * 3:(i + 1) * r * 3].reshape(r, 3).copy() for i in range(n)], ob[:r].copy()


def preintegrate(imu, t_end, bg, ba, noise=(0.00057, 0.00094, 0.0000037, 0.0000024), info_weight=1.0):
    """bs_common::PreIntegrator::Integrate on an IMU stream (rows t, w, a with t <= t_end): returns the delta
    (t, q wxyz, p, v) and the constant block of a BSNLS_FACTOR_IMU_DELTA factor."""
    lib = load_library()
    imu = np.ascontiguousarray(imu, np.float64)
    pd = C.POINTER(C.c_double)
    bg, ba, noise = (np.ascontiguousarray(v, np.float64) for v in (bg, ba, noise))
    delta, consts = np.zeros(11), np.zeros(cabi.FACTOR_STRIDE[cabi.FACTOR_IMU_DELTA])
    rc = lib.bsfuse_preintegrate(imu.ctypes.data_as(pd), len(imu), float(t_end), bg.ctypes.data_as(pd), ba.ctypes.data_as(pd),
                                 noise.ctypes.data_as(pd), float(info_weight), delta.ctypes.data_as(pd), consts.ctypes.data_as(pd))
    if rc != 0:
        raise ValueError("preintegration failed")
    return dict(t=delta[0], q=delta[1:5].copy(), p=delta[5:8].copy(), v=delta[8:11].copy()), consts


class SmootherReplay:
    """Python face of bsfuse::SmootherReplay (host/smoother_replay.h): the fixed-lag smoother's cycle — new keyframe in,
    oldest keyframe pseudo-marginalised, optimize — on a synthetic VIO trajectory."""

    def __init__(self, total_keyframes, window_keyframes=10, seed=0):
        self._lib = load_library()
        self._r = self._lib.bsfuse_replay_new(total_keyframes, window_keyframes, seed)
        err = self._lib.bsfuse_replay_error(self._r)
        if err:
            raise ValueError(err.decode())

    def __del__(self):
        if getattr(self, "_r", None):
            self._lib.bsfuse_replay_free(self._r)
            self._r = None

    def done(self):
        return bool(self._lib.bsfuse_replay_done(self._r))

    def advance(self):
        t = self._lib.bsfuse_replay_advance(self._r)
        if t < 0:
            raise RuntimeError(self._lib.bsfuse_replay_error(self._r).decode())
        return t

    def flatten(self):
        cp = self._lib.bsfuse_replay_flatten(self._r)
        if not cp:
            raise ValueError(self._lib.bsfuse_replay_error(self._r).decode())
        return problem_to_numpy(cp.contents)

    def observations(self):
        return int(self._lib.bsfuse_replay_observations(self._r))

    def optimize(self, options=None, max_iters=64):
        options = options or cabi.Options.defaults()
        s = cabi.Summary()
        its = (cabi.Iteration * max_iters)()
        t = (C.c_double * 5)()
        rc = self._lib.bsfuse_replay_optimize(self._r, C.byref(options), C.byref(s), its, max_iters, t)
        if rc != 0:
            raise RuntimeError(self._lib.bsfuse_replay_error(self._r).decode())
        return s, [its[i] for i in range(min(s.num_iterations, max_iters))], dict(zip(HashGraph.TIMING, [float(v) for v in t]))

    def current_values(self, n):
        v = np.zeros(n)
        self._lib.bsfuse_replay_current_values(self._r, v.ctypes.data_as(C.POINTER(C.c_double)))
        return v
