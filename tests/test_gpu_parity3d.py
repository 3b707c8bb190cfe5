"""GPU parity tests of the 3D path: csm_stack3d_build / csm_match3d_batch through the
reference-shaped host class against the oracle (oracle/csm_oracle3d.h)."""
import numpy as np
import pytest

import oracle_binding as ob
import synth3d

pytestmark = pytest.mark.gpu


def _embed(lo, dense, box_lo, box_hi):
    """Places a level ([z,y,x] bytes at cell offset lo) into the common box."""
    shape = tuple(int(v) for v in (box_hi - box_lo + 1)[::-1])
    out = np.zeros(shape, np.uint8)
    if dense.size:
        o = lo - box_lo
        nz, ny, nx = dense.shape
        out[o[2]:o[2] + nz, o[1]:o[1] + ny, o[0]:o[0] + nx] = dense
    return out


def _assert_levels_identical(oracle_matcher, gpu_matcher, depth):
    for lvl in range(depth):
        lo_o, d_o = oracle_matcher.level(lvl)
        lo_g, d_g = gpu_matcher.level(lvl)
        # the device box must contain every voxel the reference creates
        nz = np.argwhere(d_o)
        if len(nz):
            lo_nz = lo_o + nz.min(0)[::-1]
            hi_nz = lo_o + nz.max(0)[::-1]
            assert np.all(lo_nz >= lo_g) and np.all(hi_nz <= lo_g + np.array(d_g.shape[::-1]) - 1), lvl
        hi_o = lo_o + np.array(d_o.shape[::-1]) - 1
        hi_g = lo_g + np.array(d_g.shape[::-1]) - 1
        box_lo, box_hi = np.minimum(lo_o, lo_g), np.maximum(hi_o, hi_g)
        assert np.array_equal(_embed(lo_o, d_o, box_lo, box_hi), _embed(lo_g, d_g, box_lo, box_hi)), \
            "level %d differs" % lvl


def _random_grid(seed, n=1000, lo=-50, hi=49):
    rng = np.random.default_rng(seed)
    xyz = np.unique(rng.integers(lo, hi + 1, size=(n, 3)), axis=0).astype(np.int32)
    values = synth3d.probability_to_value(rng.uniform(0.1, 0.9, len(xyz)))
    return xyz, values


def _pair(csm, gpu_ctx, res, xyz, val, low_res, lxyz, lval, hist, opt):
    og, olg = ob.OracleGrid3D(res, xyz, val), ob.OracleGrid3D(low_res, lxyz, lval)
    om = ob.OracleFCSM3D(og, olg, hist, opt)
    gm = csm.Matcher3D(gpu_ctx, res, xyz, val, low_res, lxyz, lval, hist, opt)
    return om, gm


@pytest.mark.parametrize("depth,frd,seed,n", [(4, 4, 1, 1000), (6, 3, 2, 1000), (5, 1, 3, 300),
                                              (8, 3, 4, 2000), (3, 2, 5, 1), (7, 9, 6, 200)])
def test_stack3d_bytes_identical(csm, gpu_ctx, depth, frd, seed, n):
    """PrecomputedGridGenerator3DTest's random grid (precomputation_grid_3d_test.cc:33-69)
    through the whole stack, incl. half-resolution levels and negative indices."""
    xyz, val = _random_grid(seed, n)
    opt = ob.options3d(depth, frd)
    om, gm = _pair(csm, gpu_ctx, 2.0, xyz, val, 4.0, xyz[:10], val[:10], np.zeros(10, np.float32), opt)
    assert gm.grid_size == om.grid.grid_size
    _assert_levels_identical(om, gm, depth)


def test_stack3d_world_and_empty(csm, gpu_ctx):
    w = synth3d.World3D(11, size_xy=8.0)
    xyz, val = w.voxels(w.res)
    lxyz, lval = w.voxels(w.low_res)
    opt = ob.options3d(8, 3)
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
    _assert_levels_identical(om, gm, 8)
    empty = np.zeros((0, 3), np.int32), np.zeros(0, np.uint16)
    om, gm = _pair(csm, gpu_ctx, 0.1, empty[0], empty[1], 0.45, empty[0], empty[1], np.zeros(4, np.float32),
                   ob.options3d(3, 2))
    _assert_levels_identical(om, gm, 3)


def _compare(r_o, r_g, check_scans=True):
    assert bool(r_g.found) == bool(r_o.found)
    if check_scans:
        assert r_g.num_scans == r_o.num_scans
    if not r_o.found:
        return
    assert (r_g.scan_index, tuple(r_g.offset)) == (r_o.scan_index, tuple(r_o.offset))
    for name in ("score", "low_resolution_score", "rotational_score"):
        assert np.float32(getattr(r_g, name)).tobytes() == np.float32(getattr(r_o, name)).tobytes(), name
    # the reference widens a float pose to double; the values are float-exact
    assert np.array_equal(np.array(r_g.pose, np.float64).astype(np.float32), np.array(r_o.pose, np.float32))


AXES_CLOUD = np.array([[d if a == 0 else 0, d if a == 1 else 0, d if a == 2 else 0]
                       for a in range(3) for d in (4.0, 4.5, 5.0, 5.5)], np.float32)


def _fixture_matchers(csm, gpu_ctx, pose_xyz, theta):
    """fast_correlative_scan_matcher_3d_test.cc:104-118: the axes cloud inserted at a pose."""
    q = synth3d.yaw_quat(theta)
    world = synth3d.quat_rotate(q, AXES_CLOUD.astype(np.float64)) + pose_xyz
    grid = ob.OracleGrid3D(0.05)
    grid.insert(np.asarray(pose_xyz, np.float32), world.astype(np.float32))
    xyz, val = grid.cells()
    opt = ob.options3d(6, 6, 0.1, 0.15, 0.8, 0.8, 0.3)
    hist = np.zeros(10, np.float32)
    om = ob.OracleFCSM3D(grid, grid, hist, opt)
    gm = csm.Matcher3D(gpu_ctx, 0.05, xyz, val, 0.05, xyz, val, hist, opt)
    return om, gm


def test_match3d_reference_fixture(csm, gpu_ctx):
    """FastCorrelativeScanMatcher3DTest.CorrectPoseForMatch (:146-178), bit-exact
    against the oracle, including the negative low-resolution case."""
    rng = np.random.default_rng(42)
    hist = np.zeros(10, np.float32)
    far = np.array([[42.0, 42.0, 42.0]], np.float32)
    for _ in range(8):
        t = 0.7 * rng.uniform(-1, 1, 3)
        theta = 0.2 * rng.uniform(-1, 1)
        om, gm = _fixture_matchers(csm, gpu_ctx, t, theta)
        node = csm.Node3D(AXES_CLOUD, AXES_CLOUD, hist)
        node_far = csm.Node3D(AXES_CLOUD, far, hist)
        r_o = om.match(AXES_CLOUD, AXES_CLOUD, hist, 0.1)
        r_g = csm.match3d_batch(gpu_ctx, [gm, gm], [node, node_far], [0.1, 0.1])
        assert r_o.found and r_g[0].found
        _compare(r_o, r_g[0])
        assert np.linalg.norm(np.array(r_g[0].pose[:3]) - t) < 0.05
        assert abs(2 * np.arctan2(r_g[0].pose[6], r_g[0].pose[3]) - theta) < 0.05
        r_o_far = om.match(AXES_CLOUD, far, hist, 0.1)
        assert not r_o_far.found and not r_g[1].found
        # the reference's single-search entry point returns nullptr / a Result
        single = csm.match3d(gm, node, 0.1)
        assert single is not None and np.float32(single.score).tobytes() == np.float32(r_o.score).tobytes()
        assert csm.match3d(gm, node_far, 0.1) is None


def test_match3d_full_submap_fixture(csm, gpu_ctx):
    """FastCorrelativeScanMatcher3DTest.CorrectPoseForMatchFullSubmap (:180-204)."""
    t, theta = np.array([0.31, -0.22, 0.4]), 0.12
    om, gm = _fixture_matchers(csm, gpu_ctx, t, theta)
    hist = np.zeros(10, np.float32)
    node = csm.Node3D(AXES_CLOUD, AXES_CLOUD, hist)
    r_o = om.match(AXES_CLOUD, AXES_CLOUD, hist, 0.1, full=True)
    r_g = csm.match3d_batch(gpu_ctx, [gm], [node], [0.1], full=[1])[0]
    assert r_o.found
    _compare(r_o, r_g)
    assert np.linalg.norm(np.array(r_g.pose[:3]) - t) < 0.05


def _world_case(seed, depth, frd, **kw):
    w = synth3d.World3D(seed, **kw)
    xyz, val = w.voxels(w.res)
    lxyz, lval = w.voxels(w.low_res)
    return w, xyz, val, lxyz, lval


def test_discretize3d_cells_identical(csm, gpu_ctx):
    w, xyz, val, lxyz, lval = _world_case(21, 6, 3, size_xy=8.0)
    opt = ob.options3d(6, 3, 0.0, 0.4, 1.0, 0.5, np.radians(12))
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
    hi, lo, hist, grav, true = w.node([0.7, -1.1, 0.3], 0.4, gravity_yaw=0.15, tilt=(0.02, -0.03))
    init = true.copy()
    init[:3] += [0.2, -0.3, 0.1]
    submap = synth3d.pose7([0.5, 0.25, -0.125], synth3d.yaw_quat(0.3))
    p_o, rot_o, cells_o = om.discrete_scans(hi, hist, init.astype(np.float32), submap.astype(np.float32), grav)
    node = csm.Node3D(hi, lo, hist, grav)
    # the oracle takes float poses: feed the device path the same float-representable values
    p_g, rot_g, cells_g = csm.discrete_scans3d(gpu_ctx, gm, node, init.astype(np.float32).astype(np.float64),
                                               submap.astype(np.float32).astype(np.float64))
    assert len(rot_o) > 3
    assert np.array_equal(p_o, p_g) and np.array_equal(rot_o, rot_g)
    assert np.array_equal(cells_o, cells_g)


@pytest.mark.parametrize("seed,depth,frd,min_rot", [(31, 6, 3, 0.3), (32, 7, 2, 0.0), (33, 5, 5, 0.3),
                                                    (34, 8, 3, 0.05)])
def test_match3d_world_bit_exact(csm, gpu_ctx, seed, depth, frd, min_rot):
    w, xyz, val, lxyz, lval = _world_case(seed, depth, frd, size_xy=10.0)
    opt = ob.options3d(depth, frd, min_rot, 0.4, 1.5, 0.5, np.radians(10))
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
    rng = np.random.default_rng(seed)
    nodes, inits, oracle_results = [], [], []
    submap = synth3d.pose7([0.0, 0.0, 0.0], synth3d.yaw_quat(0.0)).astype(np.float32)
    for k in range(6):
        pos = np.append(rng.uniform(-3, 3, 2), rng.uniform(-0.2, 0.6))
        yaw = rng.uniform(-1, 1)
        hi, lo, hist, grav, true = w.node(pos, yaw, n_high=260 + 17 * k, n_low=60 + k,
                                          gravity_yaw=rng.uniform(-0.2, 0.2))
        init = true.copy()
        init[:3] += rng.uniform(-0.6, 0.6, 3) * [1, 1, 0.3]
        init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.1, 0.1)), true[3:])
        if k == 5:  # a node that does not belong here: must agree on "not found" too
            init[:3] += [40.0, 0, 0]
        init = init.astype(np.float32)
        min_score = 0.45
        oracle_results.append(om.match(hi, lo, hist, min_score, init, submap, grav))
        nodes.append(csm.Node3D(hi, lo, hist, grav))
        inits.append(init.astype(np.float64))
    res = csm.match3d_batch(gpu_ctx, [gm] * len(nodes), nodes, [0.45] * len(nodes), np.array(inits),
                            np.tile(submap.astype(np.float64), (len(nodes), 1)))
    assert sum(1 for r in oracle_results if r.found) >= 3
    for r_o, r_g in zip(oracle_results, res):
        _compare(r_o, r_g)


def test_match3d_full_submap_world(csm, gpu_ctx):
    w, xyz, val, lxyz, lval = _world_case(41, 7, 3, size_xy=6.0)
    opt = ob.options3d(7, 3, 0.5, 0.4, 1.5, 0.5, np.radians(10))
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
    hi, lo, hist, grav, true = w.node([0.8, -0.4, 0.1], 0.5, n_high=200, n_low=50, max_range=4.0)
    rot_only = synth3d.pose7([0, 0, 0], true[3:]).astype(np.float32)
    r_o = om.match(hi, lo, hist, 0.45, rot_only, ob.IDENTITY_POSE7, grav, full=True)
    node = csm.Node3D(hi, lo, hist, grav)
    r_g = csm.match3d_batch(gpu_ctx, [gm], [node], [0.45], np.array([rot_only.astype(np.float64)]), None, [1])[0]
    assert r_o.found
    _compare(r_o, r_g)


@pytest.mark.parametrize("window,mode", [(1.0, 2), (3.0, 1)])
def test_match3d_tie_order(csm, gpu_ctx, window, mode):
    """Unknown space everywhere: every candidate scores kMinProbability, so the winner is
    decided purely by the reference's traversal order (std::sort of equal lowest-
    resolution scores, then stable child order).  The small window is resolved from the
    deferred equal-bound nodes (tie_resolution 2), the large one defers too many and
    takes the ordered-walk fallback (tie_resolution 1)."""
    xyz = np.array([[500, 500, 500]], np.int32)  # one voxel far away from every lookup
    val = synth3d.probability_to_value([0.9])
    opt = ob.options3d(4, 2, 0.0, 0.05, window, 0.4, np.radians(20))
    hist = np.ones(12, np.float32)
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, xyz, val, hist, opt)
    rng = np.random.default_rng(5)
    hi = rng.uniform(-2, 2, (40, 3)).astype(np.float32)
    lo = rng.uniform(-2, 2, (9, 3)).astype(np.float32)
    r_o = om.match(hi, lo, hist, 0.05)
    node = csm.Node3D(hi, lo, hist)
    r_g = csm.match3d_batch(gpu_ctx, [gm], [node], [0.05])[0]
    assert r_o.found and r_o.num_scans > 16
    _compare(r_o, r_g)
    assert r_g.tie_resolution == mode


def test_match3d_empty_clouds(csm, gpu_ctx):
    w, xyz, val, lxyz, lval = _world_case(51, 5, 3, size_xy=6.0)
    opt = ob.options3d(5, 3, 0.3, 0.4, 1.0, 0.5, np.radians(10))
    gm = csm.Matcher3D(gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
    hi, lo, hist, grav, true = w.node([0.3, 0.2, 0.1], 0.1, n_high=100, n_low=30, max_range=4.0)
    empty = np.zeros((0, 3), np.float32)
    nodes = [csm.Node3D(empty, lo, hist, grav), csm.Node3D(hi, empty, hist, grav), csm.Node3D(hi, lo, hist, grav)]
    res = csm.match3d_batch(gpu_ctx, [gm] * 3, nodes, [0.4] * 3, np.tile(true, (3, 1)))
    assert not res[0].found and not res[1].found and res[2].found


def test_match3d_periodic_world_top_level_ties(csm, gpu_ctx):
    """A world that repeats every 16 voxels: leaves 16 voxels apart score identically and
    sit below different lowest-resolution candidates with identical scores, so the
    winner depends on the reference's std::sort order of equal lowest-resolution scores."""
    rng = np.random.default_rng(77)
    tile = np.unique(rng.integers(0, 16, size=(70, 3)) % [16, 16, 8], axis=0)
    tiles = np.array([[i, j, 0] for i in range(-8, 8) for j in range(-8, 8)]) * [16, 16, 0]
    xyz = (tile[None, :, :] + tiles[:, None, :]).reshape(-1, 3).astype(np.int32)
    prob = np.tile(rng.uniform(0.55, 0.9, len(tile)), len(tiles))
    val = synth3d.probability_to_value(prob)
    lxyz = np.unique(rng.integers(-12, 12, size=(400, 3)), axis=0).astype(np.int32)
    lval = synth3d.probability_to_value(rng.uniform(0.3, 0.9, len(lxyz)))
    hist = np.ones(8, np.float32)
    opt = ob.options3d(4, 2, 0.0, 0.1, 2.0, 0.4, np.radians(3))
    om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, hist, opt)
    centre = xyz[(np.abs(xyz[:, 0]) < 30) & (np.abs(xyz[:, 1]) < 30)]
    pick = centre[rng.choice(len(centre), 60, replace=False)]
    true = np.array([0.3, -0.2, 0.1])
    hi = (pick * 0.1 - true).astype(np.float32)
    lo = rng.uniform(-1, 1, (20, 3)).astype(np.float32)
    r_o = om.match(hi, lo, hist, 0.3)
    node = csm.Node3D(hi, lo, hist)
    r_g = csm.match3d_batch(gpu_ctx, [gm], [node], [0.3])[0]
    assert r_o.found
    _compare(r_o, r_g)
    assert r_g.tie_resolution in (1, 2)


def test_constraint_builder3d_protocol_and_results(csm, gpu_ctx):
    """ConstraintBuilder3D (constraint_builder_3d_test.cc:61-115 protocol): every search
    answered, non-null constraints in submission order, transforms identical to the
    oracle's matcher on the same (node, submap) pairs, distance gate and sampler applied."""
    worlds = [synth3d.World3D(61 + k, size_xy=8.0) for k in range(2)]
    opt = ob.options3d(6, 3, 0.3, 0.4, 1.2, 0.4, np.radians(10))
    oracle_m, gpu_m = [], []
    for w in worlds:
        xyz, val = w.voxels(w.res)
        lxyz, lval = w.voxels(w.low_res)
        om, gm = _pair(csm, gpu_ctx, 0.1, xyz, val, 0.45, lxyz, lval, w.submap_histogram(), opt)
        oracle_m.append(om)
        gpu_m.append(gm)
    rng = np.random.default_rng(8)
    nodes, raw, node_poses, search_submap, search_full = [], [], [], [], []
    for i in range(12):
        s = i % 2
        hi, lo, hist, grav, true = worlds[s].node(np.append(rng.uniform(-2, 2, 2), 0.2), rng.uniform(-1, 1),
                                                  n_high=200, n_low=50, max_range=5.0)
        init = true.copy()
        init[:3] += rng.uniform(-0.4, 0.4, 3) * [1, 1, 0.3]
        if i == 7:
            init[:3] += [100.0, 0, 0]  # beyond max_constraint_distance: never searched
        full = 1 if i in (3, 8) else 0
        if full:
            init[:3] = 0.0
        nodes.append(csm.Node3D(hi, lo, hist, grav))
        raw.append((hi, lo, hist, grav))
        node_poses.append(init.astype(np.float32).astype(np.float64))
        search_submap.append(s)
        search_full.append(full)
    ident = np.tile(ob.IDENTITY_POSE7.astype(np.float64), (12, 1))
    found, transform, stats, _ = csm.constraint_builder3d_run(
        gpu_m, search_submap, search_full, nodes, np.array(node_poses), ident, opt, min_score=0.45,
        global_localization_min_score=0.5, pool_threads=3, flush_threshold=5)
    assert stats["searches"] == 11 and stats["flushes"] >= 2
    for i in range(12):
        hi, lo, hist, grav = raw[i]
        if i == 7:
            assert not found[i]
            continue
        r = oracle_m[search_submap[i]].match(hi, lo, hist, 0.5 if search_full[i] else 0.45,
                                             node_poses[i].astype(np.float32), ob.IDENTITY_POSE7, grav,
                                             bool(search_full[i]))
        assert bool(found[i]) == bool(r.found)
        if r.found:
            assert np.array_equal(transform[i].astype(np.float32), np.array(r.pose, np.float32))
    assert found.sum() == stats["found"] and found.sum() >= 6
    # the same searches entering as whole-node bursts (MaybeAddConstraints): identical answers
    found_b, transform_b, stats_b, _ = csm.constraint_builder3d_run(
        gpu_m, search_submap, search_full, nodes, np.array(node_poses), ident, opt, min_score=0.45,
        global_localization_min_score=0.5, pool_threads=3, flush_threshold=5, burst=4)
    assert stats_b["searches"] == 11
    assert np.array_equal(found, found_b) and transform.tobytes() == transform_b.tobytes()


RT_CLOUD = np.array([[-3, 2, 0], [-4, 2, 0], [-5, 2, 0], [-6, 2, 0], [-6, 3, 1], [-6, 4, 2], [-7, 3, 1]],
                    np.float32)


def _rt_fixture_grid():
    """real_time_correlative_scan_matcher_3d_test.cc:36-48: the cloud painted at (-1, 0, 0)."""
    cells = np.round((RT_CLOUD + np.array([-1, 0, 0], np.float32)) / np.float32(0.1)).astype(np.int32)
    return cells, synth3d.probability_to_value(np.ones(len(cells)))


def test_rt_match3d_reference_fixture(csm, gpu_ctx):
    """RealTimeCorrelativeScanMatcher3DTest.* (:80-117): every initial pose returns to the
    expected pose within the reference's 1e-3 isApprox bound, bit-identical to the oracle."""
    xyz, val = _rt_fixture_grid()
    og = ob.OracleGrid3D(0.1, xyz, val)
    opt = csm.rt_options(0.3, np.radians(1.0), 1e-1, 1.0)
    a = 0.8 / 180.0 * np.pi
    s, c = np.sin(0.5 * a), np.cos(0.5 * a)
    cases = [([-1, 0, 0], [1, 0, 0, 0]), ([-0.8, 0, 0], [1, 0, 0, 0]), ([-1, 0, -0.2], [1, 0, 0, 0]),
             ([-0.9, -0.2, 0.2], [1, 0, 0, 0]), ([-1, 0, 0], [c, s, 0, 0]), ([-1, 0, 0], [c, 0, s, 0]),
             ([-1, 0, 0], [c, 0, s, s])]
    for t, q in cases:
        init = np.array(list(t) + list(q), np.float64)
        score_g, pose_g = csm.rt_match3d(0.1, xyz, val, opt, init, RT_CLOUD)
        score_o, pose_o, best_o, num_o = ob.rt_match3d(og, opt, init.astype(np.float32), RT_CLOUD)
        assert np.float32(score_g).tobytes() == np.float32(score_o).tobytes()
        assert np.array_equal(pose_g.astype(np.float32), pose_o)
        assert np.linalg.norm(pose_g[:3] - [-1, 0, 0]) < 2e-3 and abs(abs(pose_g[3]) - 1) < 1e-5


def test_rt_match3d_world_bit_exact(csm, gpu_ctx):
    w = synth3d.World3D(71, size_xy=8.0)
    xyz, val = w.voxels(w.res)
    og = ob.OracleGrid3D(0.1, xyz, val)
    opt = csm.rt_options(0.15, np.radians(1.0), 1e-1, 1e-1)
    rng = np.random.default_rng(72)
    clouds, inits = [], []
    for k in range(5):
        hi, lo, hist, grav, true = w.node(np.append(rng.uniform(-2, 2, 2), 0.2), rng.uniform(-1, 1),
                                          n_high=120 + 13 * k, n_low=10, max_range=4.0, tilt=(0.01, -0.02))
        init = true.copy()
        init[:3] += rng.uniform(-0.12, 0.12, 3)
        init[3:] = synth3d.quat_mul(synth3d.yaw_quat(rng.uniform(-0.01, 0.01)), true[3:])
        clouds.append(hi)
        inits.append(init.astype(np.float32).astype(np.float64))
    scores, poses, best, num = csm.rt_match3d_batch(gpu_ctx, 0.1, xyz, val, opt, np.array(inits), clouds)
    for k in range(5):
        score_o, pose_o, best_o, num_o = ob.rt_match3d(og, opt, inits[k].astype(np.float32), clouds[k])
        assert num[k] == num_o and num_o > 100
        assert best[k] == best_o
        assert np.float32(scores[k]).tobytes() == np.float32(score_o).tobytes()
        assert np.array_equal(poses[k].astype(np.float32), pose_o)
