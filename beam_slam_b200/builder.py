"""Incremental builder of a flattened problem (bsnls_problem) from blocks and factors given in program order.

Small, Python-side counterpart of the C++ flattener (host/): used by tests and by fixtures that restate the
reference's unit-test graphs literally. Blocks are added in createProblem variable order, factors in
createProblem constraint order (dependencies/fuse/fuse_graphs/src/hash_graph.cpp:478,507).
"""
import numpy as np

from . import cabi


class ProblemBuilder:
    def __init__(self):
        self.size, self.manifold, self.constant, self.group, self.values = [], [], [], [], []
        self.cameras = []
        self.factors = {}
        self.nres = 0

    def add_block(self, values, manifold=cabi.MANIFOLD_EUCLIDEAN, constant=False, group=1):
        values = [float(v) for v in values]
        self.size.append(len(values))
        self.manifold.append(manifold)
        self.constant.append(1 if constant else 0)
        self.group.append(group)
        self.values.extend(values)
        return len(self.size) - 1

    def add_quaternion(self, wxyz, constant=False):
        return self.add_block(wxyz, cabi.MANIFOLD_QUATERNION, constant, 1)

    def add_landmark(self, xyz, constant=False):
        return self.add_block(xyz, cabi.MANIFOLD_EUCLIDEAN, constant, 0)

    def add_camera(self, fx, fy, cx, cy, T_cam_baselink, weight=1.0, loss_type=cabi.LOSS_NONE, loss_a=0.0):
        T = np.asarray(T_cam_baselink, np.float64)[:3, :4].reshape(-1)
        self.cameras.append([fx, fy, cx, cy, *T, weight, float(loss_type), loss_a, 0.0])
        return len(self.cameras) - 1

    def add_factor(self, ftype, blocks, consts, aux=0, loss_type=cabi.LOSS_NONE, loss_param=0.0):
        assert len(blocks) == cabi.FACTOR_NBLOCKS[ftype]
        c = np.zeros(cabi.FACTOR_STRIDE[ftype])
        consts = np.asarray(consts, np.float64).reshape(-1)
        c[:consts.size] = consts
        g = self.factors.setdefault(ftype, dict(blocks=[], consts=[], aux=[], loss_type=[], loss_param=[],
                                                res_index=[]))
        g["blocks"].append(list(blocks))
        g["consts"].append(c)
        g["aux"].append(aux)
        g["loss_type"].append(loss_type)
        g["loss_param"].append(loss_param)
        g["res_index"].append(self.nres)
        self.nres += 1
        return len(g["res_index"]) - 1

    def add_marginal(self, blocks, A, b, x_bar):
        """fuse_constraints::MarginalConstraint (marginal_constraint.h:85-96): blocks in order, A[i] = rows x 3 per block
        (tangent columns), b = rows, x_bar[i] = the block's linearisation point (ambient). Rows <= 15, blocks <= 10."""
        n, rows = len(blocks), len(b)
        assert 1 <= n <= cabi.MARGINAL_MAX_BLOCKS and rows <= cabi.MARGINAL_MAX_ROWS
        c = np.zeros(cabi.FACTOR_STRIDE[cabi.FACTOR_MARGINAL])
        c[0] = n
        off = 0
        for i, blk in enumerate(blocks):
            xb = np.asarray(x_bar[i], np.float64).reshape(-1)
            assert xb.size == self.size[blk]
            c[1 + i] = xb.size
            c[26 + off:26 + off + xb.size] = xb
            off += xb.size
            Ai = np.asarray(A[i], np.float64).reshape(rows, 3)
            for r in range(rows):
                c[58 + r * 30 + 3 * i:58 + r * 30 + 3 * i + 3] = Ai[r]
        c[11:11 + rows] = np.asarray(b, np.float64)
        slots = list(blocks) + [-1] * (cabi.MARGINAL_MAX_BLOCKS - n)
        return self.add_factor(cabi.FACTOR_MARGINAL, slots, c)

    def build(self):
        cams = np.asarray(self.cameras, np.float64) if self.cameras else np.zeros((0, cabi.CAMERA_STRIDE))
        groups = {t: {k: np.asarray(v) for k, v in g.items()} for t, g in self.factors.items()}
        return cabi.FlatProblem(self.size, self.manifold, self.constant, self.group, self.values, cams, groups)
