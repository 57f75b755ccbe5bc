"""Host-side logic: UUID scheme, hash-map (createProblem) ordering, flattener, synthetic generator. CPU only."""
import ctypes as C
import struct
import uuid as pyuuid

import numpy as np

import oracle_lib as orc
from beam_slam_b200 import cabi, host


def _name(ns_bytes, data):
    lib = host.load_library()
    out = C.create_string_buffer(16)
    lib.bsfuse_uuid_name(ns_bytes, data, len(data), out)
    return out.raw


def test_name_uuid_is_rfc4122_v5():
    """boost::uuids::name_generator == RFC 4122 v5 (uuid.h:85-88): check against python's uuid5."""
    nil = bytes(16)
    for s in [b"", b"fuse_variables::Orientation3DStamped", b"x" * 200]:
        assert _name(nil, s) == pyuuid.uuid5(pyuuid.UUID(bytes=nil), s.decode()).bytes
    ns = pyuuid.uuid5(pyuuid.UUID(bytes=nil), "bs_variables::Point3DLandmark")
    data = struct.pack("<Q", 12345)
    # uuid5 takes str; do the sha1 by hand for raw bytes
    import hashlib
    d = bytearray(hashlib.sha1(ns.bytes + data).digest()[:16])
    d[6] = (d[6] & 0x0F) | 0x50
    d[8] = (d[8] & 0x3F) | 0x80
    lib = host.load_library()
    out = C.create_string_buffer(16)
    lib.bsfuse_uuid_landmark(b"bs_variables::Point3DLandmark", 12345, out)
    assert out.raw == bytes(d)
    # stamped: {uint32 sec, uint32 nsec, 16-byte device id} (uuid.cpp:82-97)
    ns2 = pyuuid.uuid5(pyuuid.UUID(bytes=nil), "fuse_variables::Position3DStamped")
    d2 = bytearray(hashlib.sha1(ns2.bytes + struct.pack("<II", 7, 500000000) + nil).digest()[:16])
    d2[6] = (d2[6] & 0x0F) | 0x50
    d2[8] = (d2[8] & 0x3F) | 0x80
    lib.bsfuse_uuid_stamped(b"fuse_variables::Position3DStamped", 7, 500000000, out)
    assert out.raw == bytes(d2)


def test_uuid_hash_is_boost_hash_combine():
    u = bytes(range(16))
    seed = 0
    for b in u:
        seed ^= (b + 0x9e3779b9 + ((seed << 6) & 0xFFFFFFFFFFFFFFFF) + (seed >> 2)) & 0xFFFFFFFFFFFFFFFF
    assert host.load_library().bsfuse_uuid_hash(u) == seed


def test_hash_graph_container_semantics():
    """std::logic_error / std::out_of_range / invalid_argument behaviour of hash_graph.cpp:100-301"""
    assert host.load_library().bsfuse_selftest_containers() == 0


def test_flatten_order_is_deterministic_and_hash_ordered():
    g1 = host.HashGraph.synthetic(1, scale=0.1)
    g2 = host.HashGraph.synthetic(1, scale=0.1)
    p1, p2 = g1.flatten(), g2.flatten()
    n = len(p1.block_size)
    assert g1.block_uuids(n) == g2.block_uuids(n)
    assert np.array_equal(p1.block_values, p2.block_values)
    for t in p1.groups:
        assert np.array_equal(p1.groups[t]["blocks"], p2.groups[t]["blocks"])
    # hash-map order is not insertion order: landmarks and stamped variables interleave
    grp = p1.block_group
    assert len(set(np.diff(np.flatnonzero(grp == 0)))) > 1
    # res_index is a permutation over all factor groups
    ri = np.concatenate([g["res_index"] for g in p1.groups.values()])
    assert sorted(ri) == list(range(len(ri)))


def test_synthetic_config1_shape():
    g = host.HashGraph.synthetic(1)
    assert g.info["num_keyframes"] == 10 and g.info["num_imu"] == 9 and g.info["num_priors"] == 1
    assert g.info["num_landmarks"] == 300 and 2500 <= g.info["num_observations"] <= 3000
    p = g.flatten()
    assert p.count(cabi.FACTOR_REPROJECTION) == g.info["num_observations"]
    assert p.count(cabi.FACTOR_IMU_DELTA) == 9 and p.count(cabi.FACTOR_IMU_PRIOR) == 1
    assert len(p.cameras) == 1 and p.cameras[0, 17] == cabi.LOSS_CAUCHY and p.cameras[0, 18] == 5.0
    # pixels are integers (visual_map.cpp:188-193)
    px = p.groups[cabi.FACTOR_REPROJECTION]["consts"]
    assert np.array_equal(px, np.round(px))


def test_oracle_solves_config1_towards_truth():
    g = host.HashGraph.synthetic(1, scale=0.5)
    p = g.flatten()
    truth = g.truth_values()
    opts = host.HashGraph.synthetic_options(1)
    s, its, x = orc.solve(p, opts)
    assert s.is_solution_usable()
    assert s.final_cost < 0.05 * s.initial_cost
    off = p.offsets
    poses = np.concatenate([np.arange(off[b], off[b + 1]) for b in np.flatnonzero(p.block_group == 1)])
    lms = np.concatenate([np.arange(off[b], off[b + 1]) for b in np.flatnonzero(p.block_group == 0)])
    err = lambda v, idx: np.nanmedian(np.abs(v[idx] - truth[idx]))
    assert err(x, poses) < 0.5 * err(p.block_values, poses)     # keyframe states move towards the truth
    assert err(x, lms) < err(p.block_values, lms)               # landmarks (depth-limited) do not get worse
    # costs are monotone over successful steps
    succ = [i.cost for i in its if i.step_is_successful]
    assert all(b <= a for a, b in zip(succ, succ[1:]))


def test_oracle_solves_lvio_and_refinement_configs():
    for cfg, scale in ((2, 0.1), (3, 0.01)):
        g = host.HashGraph.synthetic(cfg, scale=scale)
        p = g.flatten()
        s, its, x = orc.solve(p, host.HashGraph.synthetic_options(cfg))
        assert s.is_solution_usable(), s.message
        assert s.final_cost < 0.2 * s.initial_cost
