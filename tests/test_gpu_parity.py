"""GPU parity tests proper: the CUDA path through the C ABI against the oracle."""
import numpy as np
import pytest

import oracle_binding as ob
import synth

pytestmark = pytest.mark.gpu


def _tiny_grid(seed, nx, ny):
    rng = np.random.default_rng(seed)
    cells = rng.integers(1, 32768, size=(ny, nx)).astype(np.uint16)
    cells[rng.random((ny, nx)) < 0.25] = 0
    return cells


@pytest.mark.parametrize("nx,ny,depth", [(250, 250, 4), (4, 4, 8), (1, 1, 3), (37, 5, 7),
                                         (400, 400, 7), (129, 257, 6),
                                         (1001, 777, 5), (2047, 300, 4)])  # odd widths: rows that start unaligned
def test_stack_bytes_identical(csm, gpu_ctx, nx, ny, depth):
    cells = _tiny_grid(nx * 1000 + ny, nx, ny)
    og = ob.OracleGrid(cells, 0.05, 1.0, 1.0)
    om = ob.OracleFCSM2D(og, 1.0, 0.1, depth)
    g = csm.Grid(cells, 0.05, 1.0, 1.0)
    m = csm.Matcher(gpu_ctx, g, 1.0, 0.1, depth)
    for lvl in range(depth):
        a, b = om.level(lvl), m.level(lvl)
        assert a.shape == b.shape
        assert np.array_equal(a, b), "level %d differs" % lvl


def test_matchers_created_together_equal_matchers_created_one_by_one(csm, gpu_ctx):
    """FastCorrelativeScanMatcher2D::CreateMany / csm_stack2d_build_many: seven grids of mixed sizes
    (more than the staging slots, so slots are reused) -> the byte levels, the quad words and the kept
    uint16 cells of every stack equal those of a matcher constructed alone, and a search on each finds
    the same winner."""
    shapes = [(250, 250), (37, 5), (400, 400), (1, 1), (129, 257), (640, 480), (250, 250)]
    depth = 6
    cells = [_tiny_grid(900 + k, nx, ny) for k, (nx, ny) in enumerate(shapes)]
    grids = [csm.Grid(c, 0.05, 1.0, 1.0) for c in cells]
    for g in grids[::2]:
        g.pin()  # page-locked cells travel without the staging copy; the others through it
    many = csm.Matcher.create_many(gpu_ctx, grids, 1.0, 0.1, depth)
    assert len(many) == len(grids)
    pts = np.array([[0.3, 0.2, 0], [-0.4, 0.1, 0], [0.1, -0.35, 0], [0.5, 0.5, 0]], np.float32)
    for k, g in enumerate(grids):
        alone = csm.Matcher(gpu_ctx, g, 1.0, 0.1, depth)
        for lvl in range(depth):
            assert np.array_equal(alone.level(lvl), many[k].level(lvl)), (k, lvl)
            assert np.array_equal(alone.quad(lvl), many[k].quad(lvl)), (k, lvl)
        a = csm.match_batch(gpu_ctx, [alone], [pts], [0.0])[0]
        b = csm.match_batch(gpu_ctx, [many[k]], [pts], [0.0])[0]
        assert (a.found, a.scan_index, a.x_index_offset, a.y_index_offset, a.score) == \
            (b.found, b.scan_index, b.x_index_offset, b.y_index_offset, b.score)
    assert csm.Matcher.create_many(gpu_ctx, [], 1.0, 0.1, depth) == []


@pytest.mark.parametrize("nx,ny,depth", [(250, 250, 4), (4, 4, 8), (1, 1, 3), (37, 5, 7), (3, 2, 2),
                                         (400, 400, 7), (129, 257, 6), (1000, 1000, 7)])
def test_stack_quad_words_identical(csm, gpu_ctx, nx, ny, depth):
    """Every word of the matcher's quad-phase layout against the oracle's byte levels:
    word(u, v) = L(u-s, v-s) | L(u-s, v) << 8 | L(u, v-s) << 16 | L(u, v) << 24 in
    window-end coordinates (0 outside the level), stored at [v % 2s, u % 2s, v / 2s, u / 2s]."""
    cells = _tiny_grid(nx * 1000 + ny + 7, nx, ny)
    og = ob.OracleGrid(cells, 0.05, 1.0, 1.0)
    om = ob.OracleFCSM2D(og, 1.0, 0.1, depth)
    m = csm.Matcher(gpu_ctx, csm.Grid(cells, 0.05, 1.0, 1.0), 1.0, 0.1, depth)
    for lvl in range(depth):
        s = 1 << lvl
        L = om.level(lvl).astype(np.uint32)
        got = m.quad(lvl)
        two_s, jn, n_i = got.shape[0], got.shape[2], got.shape[3]
        U, V = n_i * two_s, jn * two_s
        big = np.zeros((V + s, U + s), dtype=np.uint32)  # big[v + s, u + s] = L(u, v)
        big[s:s + L.shape[0], s:s + L.shape[1]] = L[:V, :U]
        assert L.shape[0] <= V and L.shape[1] <= U
        want = (big[0:V, 0:U] | (big[s:V + s, 0:U] << 8) | (big[0:V, s:U + s] << 16) |
                (big[s:V + s, s:U + s] << 24))
        want = want.reshape(jn, two_s, n_i, two_s).transpose(1, 3, 0, 2)
        assert np.array_equal(want, got), "quad level %d differs" % lvl


def _compare(r_o, r_g):
    assert r_g.found == r_o.found
    if r_o.found:
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()
        assert r_g.pose_x == r_o.pose_x and r_g.pose_y == r_o.pose_y
        assert r_g.pose_theta == r_o.pose_theta


@pytest.mark.parametrize("size,depth,min_score", [(200, 5, 0.5), (200, 5, 0.1), (400, 7, 0.6),
                                                  (300, 3, 0.3)])
def test_match_full_submap_bit_exact(csm, gpu_ctx, size, depth, min_score):
    w = synth.World(size, size, seed=size + depth)
    other = synth.World(size, size, seed=size + depth + 1)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    om = ob.OracleFCSM2D(og, 7.0, np.deg2rad(30.0), depth)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    m = csm.Matcher(gpu_ctx, g, 7.0, np.deg2rad(30.0), depth)
    clouds = []
    for i in range(6):
        _, pts = synth.make_scan(w if i % 2 == 0 else other, 40 + i, max_range=8.0)
        clouds.append(pts)
    res = csm.match_batch(gpu_ctx, [m] * len(clouds), clouds, [min_score] * len(clouds))
    for pts, r_g in zip(clouds, res):
        _compare(om.match(pts, min_score), r_g)


def test_match_local_window_bit_exact(csm, gpu_ctx):
    w = synth.World(300, 300, seed=77)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    om = ob.OracleFCSM2D(og, 3.0, 0.6, 6)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    m = csm.Matcher(gpu_ctx, g, 3.0, 0.6, 6)
    rng = np.random.default_rng(5)
    clouds, inits = [], []
    for i in range(8):
        pose, pts = synth.make_scan(w, 90 + i, max_range=8.0)
        clouds.append(pts)
        inits.append((pose[0] + rng.uniform(-1, 1), pose[1] + rng.uniform(-1, 1),
                      pose[2] + rng.uniform(-0.3, 0.3)))
    res = csm.match_batch(gpu_ctx, [m] * 8, clouds, [0.4] * 8, initial=inits)
    for pts, init, r_g in zip(clouds, inits, res):
        _compare(om.match(pts, 0.4, initial=init), r_g)


def test_tie_cases_follow_reference_order(csm, gpu_ctx):
    # ConstraintBuilder2DTest shape: empty grid, min_score 0 -> every candidate ties.
    cells = np.zeros((110, 100), dtype=np.uint16)
    og = ob.OracleGrid(cells, 1.0, 2.0, 3.0)
    g = csm.Grid(cells, 1.0, 2.0, 3.0)
    pts = np.array([[0.1, 0.2, 0.3]], dtype=np.float32)
    for depth in (1, 3, 7):
        om = ob.OracleFCSM2D(og, 7.0, np.deg2rad(30.0), depth)
        m = csm.Matcher(gpu_ctx, g, 7.0, np.deg2rad(30.0), depth)
        r_g = csm.match_batch(gpu_ctx, [m, m], [pts, pts], [0.0, 0.0],
                              initial=[None, (4.0, 5.0, 0.0)])
        _compare(om.match(pts, 0.0), r_g[0])
        _compare(om.match(pts, 0.0, initial=(4.0, 5.0, 0.0)), r_g[1])
    # symmetric corridor: two parallel walls, many equal-score poses
    cells = np.zeros((200, 200), dtype=np.uint16)
    wall = synth.probability_to_value(np.float32(0.8))
    cells[60:62, :] = wall
    cells[140:142, :] = wall
    og = ob.OracleGrid(cells, 0.05, 5.0, 5.0)
    g = csm.Grid(cells, 0.05, 5.0, 5.0)
    om = ob.OracleFCSM2D(og, 7.0, np.deg2rad(30.0), 6)
    m = csm.Matcher(gpu_ctx, g, 7.0, np.deg2rad(30.0), 6)
    ys = np.linspace(-3, 3, 40, dtype=np.float32)
    pts = np.concatenate([np.stack([np.full(40, 2.0, np.float32), ys, np.zeros(40, np.float32)], 1),
                          np.stack([np.full(40, -2.0, np.float32), ys, np.zeros(40, np.float32)], 1)])
    r_g = csm.match_batch(gpu_ctx, [m], [pts], [0.3])
    _compare(om.match(pts, 0.3), r_g[0])


def test_ordered_walk_and_path_keys_agree(csm, gpu_ctx, monkeypatch):
    """The two implementations of the reference's tie order — path keys carried through the sweep
    (default) and round 1's ordered depth-first walk (CSM_PHASE2=dfs, read when a context is
    created) — return the same winners on tie-heavy and ordinary searches."""
    monkeypatch.setenv("CSM_PHASE2", "dfs")
    dfs_ctx = csm.Context(0)
    monkeypatch.delenv("CSM_PHASE2")
    try:
        cases = []
        cells = np.zeros((110, 100), dtype=np.uint16)
        cases.append((cells, 1.0, 2.0, 3.0, np.array([[0.1, 0.2, 0.3]], dtype=np.float32), 0.0, 5))
        cells = np.zeros((200, 200), dtype=np.uint16)
        wall = synth.probability_to_value(np.float32(0.8))
        cells[60:62, :] = wall
        cells[140:142, :] = wall
        ys = np.linspace(-3, 3, 40, dtype=np.float32)
        pts = np.concatenate([np.stack([np.full(40, 2.0, np.float32), ys, np.zeros(40, np.float32)], 1),
                              np.stack([np.full(40, -2.0, np.float32), ys, np.zeros(40, np.float32)], 1)])
        cases.append((cells, 0.05, 5.0, 5.0, pts, 0.3, 6))
        w = synth.World(300, 300, seed=31)
        for k in range(3):
            cases.append((w.cells, w.res, w.max_x, w.max_y, synth.make_scan(w, 60 + k, max_range=8.0)[1], 0.4, 6))
        for cells, res, max_x, max_y, pts, min_score, depth in cases:
            g = csm.Grid(cells, res, max_x, max_y)
            a = csm.match_batch(gpu_ctx, [csm.Matcher(gpu_ctx, g, 7.0, np.deg2rad(30.0), depth)], [pts], [min_score])[0]
            b = csm.match_batch(dfs_ctx, [csm.Matcher(dfs_ctx, g, 7.0, np.deg2rad(30.0), depth)], [pts], [min_score])[0]
            assert (a.found, a.scan_index, a.x_index_offset, a.y_index_offset) == \
                   (b.found, b.scan_index, b.x_index_offset, b.y_index_offset)
            assert np.float32(a.score).tobytes() == np.float32(b.score).tobytes()
    finally:
        dfs_ctx.close()


def test_empty_cloud_is_not_found(csm, gpu_ctx):
    w = synth.World(100, 100, seed=3)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    m = csm.Matcher(gpu_ctx, g, 7.0, 0.5, 4)
    _, pts = synth.make_scan(w, 1, max_range=4.0)
    res = csm.match_batch(gpu_ctx, [m, m], [np.zeros((0, 3), np.float32), pts], [0.1, 0.1])
    assert res[0].found == 0


def test_real_time_matcher_bit_exact(csm, gpu_ctx):
    # C2: live 400x400 grid, ~250 points, +-0.1 m / +-20 deg, delta weights 0.1.
    w = synth.World(400, 400, seed=21)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    opts = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(20.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    rng = np.random.default_rng(9)
    clouds, inits = [], []
    for i in range(12):
        pose, pts = synth.make_scan(w, 300 + i)
        pts = pts[::max(1, len(pts) // 250)]
        clouds.append(np.ascontiguousarray(pts))
        inits.append((pose[0] + rng.uniform(-0.08, 0.08), pose[1] + rng.uniform(-0.08, 0.08),
                      pose[2] + rng.uniform(-0.26, 0.26)))
    res = csm.rt_match_batch(gpu_ctx, [g] * len(clouds), clouds, inits, **opts)
    for pts, init, r_g in zip(clouds, inits, res):
        r_o = ob.rt_match(og, opts, init, pts)
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()
        assert r_g.pose_x == r_o.pose_x and r_g.pose_y == r_o.pose_y
        assert r_g.pose_theta == r_o.pose_theta


def test_real_time_matcher_edge_cases(csm, gpu_ctx):
    # scan partly outside the grid (out-of-limits cells read kMinProbability), a wide
    # window, zero weights, and a single-point cloud.
    w = synth.World(120, 90, seed=22)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    pose, pts = synth.make_scan(w, 7, max_range=6.0)
    cases = [
        ({"linear_search_window": 0.3, "angular_search_window": 0.5,
          "translation_delta_cost_weight": 0.0, "rotation_delta_cost_weight": 0.0},
         (pose[0] + 2.5, pose[1] - 1.0, pose[2] + 0.4), pts),
        ({"linear_search_window": 0.05, "angular_search_window": 0.05,
          "translation_delta_cost_weight": 10.0, "rotation_delta_cost_weight": 40.0},
         (pose[0], pose[1], pose[2]), pts),
        ({"linear_search_window": 0.2, "angular_search_window": 0.3,
          "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1},
         (0.3, -0.2, 1.0), pts[:1]),
    ]
    for opts, init, cloud in cases:
        r_g = csm.rt_match_batch(gpu_ctx, [g], [np.ascontiguousarray(cloud)], [init], **opts)[0]
        r_o = ob.rt_match(og, opts, init, np.ascontiguousarray(cloud))
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()
        assert (r_g.pose_x, r_g.pose_y, r_g.pose_theta) == (r_o.pose_x, r_o.pose_y, r_o.pose_theta)


def test_real_time_matcher_tsdf_bit_exact(csm, gpu_ctx):
    """TSDF2D branch (RT2D.cc:38-59): winner and score bits against the oracle, on the
    reference fixture's seven-point L, on a painted office world with random weights, with a
    scan partly outside the grid, and on a never-observed grid (every score 0)."""
    import tsdf_util
    from test_tsdf_cpu import POINTS
    cases = []
    tv, wv = tsdf_util.paint_from_points(20, 20, 0.05, 0.3, 0.5, POINTS[:, :2], (0.5, -0.5), 0.3, 1.0)
    fix = (tv, wv, 0.05, 0.3, 0.5, 0.3, 1.0)
    zero = (np.zeros_like(tv), np.zeros_like(wv), 0.05, 0.3, 0.5, 0.3, 1.0)
    o_small = {"linear_search_window": 0.1, "angular_search_window": 0.1,
               "translation_delta_cost_weight": 0.0, "rotation_delta_cost_weight": 0.0}
    cases.append((fix, o_small, (0.05, -0.05, 0.02), POINTS))
    cases.append((fix, dict(o_small, translation_delta_cost_weight=0.3, rotation_delta_cost_weight=0.7),
                  (0.0, 0.05, -0.05), POINTS))
    cases.append((zero, o_small, (0.0, 0.0, 0.0), POINTS))
    w = synth.World(300, 260, seed=31)
    tv, wv = tsdf_util.from_world(w, 0.3, 10.0, seed=5)
    world = (tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    o_c2 = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(20.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    rng = np.random.default_rng(17)
    for i in range(8):
        pose, pts = synth.make_scan(w, 500 + i, max_range=6.0)
        pts = np.ascontiguousarray(pts[::max(1, len(pts) // 250)])
        init = (pose[0] + rng.uniform(-0.08, 0.08), pose[1] + rng.uniform(-0.08, 0.08),
                pose[2] + rng.uniform(-0.26, 0.26))
        if i == 7:
            init = (pose[0] + 4.0, pose[1] - 3.0, pose[2])  # partly outside the grid
        cases.append((world, o_c2, init, pts))
    grids = {}
    for args, opts, init, pts in cases:
        key = id(args)
        if key not in grids:
            grids[key] = (csm.TsdfGrid(*args), ob.OracleTsdf(*args))
        g, og = grids[key]
        r_g = csm.rt_match_batch(gpu_ctx, [g], [np.ascontiguousarray(pts)], [init], **opts)[0]
        r_o = ob.rt_match_tsdf(og, opts, init, pts)
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()
        assert (r_g.pose_x, r_g.pose_y, r_g.pose_theta) == (r_o.pose_x, r_o.pose_y, r_o.pose_theta)
    # one mixed batch: a ProbabilityGrid job next to TSDF jobs
    pg, opg = csm.Grid(w.cells, w.res, w.max_x, w.max_y), ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g, og = grids[id(world)]
    _, _, init, pts = cases[4]
    res = csm.rt_match_batch(gpu_ctx, [g, pg, g], [pts, pts, pts], [init, init, init], **o_c2)
    for r_g, r_o in zip(res, [ob.rt_match_tsdf(og, o_c2, init, pts), ob.rt_match(opg, o_c2, init, pts),
                              ob.rt_match_tsdf(og, o_c2, init, pts)]):
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()


def _builder_case(seed, n_sub=3, n_search=24, size=240):
    rng = np.random.default_rng(seed)
    worlds = [synth.World(size, size, seed=seed * 10 + k) for k in range(n_sub)]
    poses = np.array([[rng.uniform(-2, 2), rng.uniform(-2, 2), rng.uniform(-1.5, 1.5)]
                      for _ in range(n_sub)])
    sub, full, rel, clouds = [], [], [], []
    for i in range(n_search):
        s = int(rng.integers(0, n_sub))
        src = worlds[s] if i % 3 != 2 else worlds[(s + 1) % n_sub]
        pose, pts = synth.make_scan(src, seed * 1000 + i, max_range=6.0)
        sub.append(s)
        full.append(1 if i % 2 == 0 else 0)
        # The grid is expressed in the local (map) frame, so the node's map pose is
        # `pose`; relative to the submap origin that is local_pose^-1 * pose.
        c, sn = np.cos(-poses[s][2]), np.sin(-poses[s][2])
        dx, dy = pose[0] - poses[s][0], pose[1] - poses[s][1]
        rel.append([c * dx - sn * dy + rng.uniform(-0.5, 0.5), sn * dx + c * dy + rng.uniform(-0.5, 0.5),
                    pose[2] - poses[s][2] + rng.uniform(-0.2, 0.2)])
        clouds.append(pts)
    rel[5] = [40.0, 0.0, 0.0]  # beyond max_constraint_distance: never searched
    return worlds, poses, sub, full, np.array(rel), clouds


@pytest.mark.parametrize("sampling_ratio,flush", [(1.0, 256), (0.5, 4)])
def test_constraint_builder_matches_reference_semantics(csm, sampling_ratio, flush):
    worlds, poses, sub, full, rel, clouds = _builder_case(5)
    ograds = [ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    grids = [csm.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    kw = dict(sampling_ratio=sampling_ratio, max_constraint_distance=15.0, min_score=0.45,
              global_min_score=0.5, linear_window=3.0, angular_window=0.5, depth=5)
    _, o_out, _ = ob.constraint_batch(4, ograds, poses, sub, full, rel, clouds, **kw)
    _, found, tf, stats = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds,
                                                    flush_threshold=flush, **kw)
    assert found.sum() > 0
    for i in range(len(clouds)):
        assert found[i] == o_out[i].found, i
        if found[i]:
            assert abs(tf[i][0] - o_out[i].pose_x) <= 1e-9 * max(1.0, abs(o_out[i].pose_x))
            assert abs(tf[i][1] - o_out[i].pose_y) <= 1e-9 * max(1.0, abs(o_out[i].pose_y))
            d = (tf[i][2] - o_out[i].pose_theta + np.pi) % (2 * np.pi) - np.pi
            assert abs(d) <= 1e-9
    assert stats[1] == found.sum()


def test_fast_matcher_on_tsdf_grid_bit_exact(csm, gpu_ctx):
    """FastCorrelativeScanMatcher2D takes any Grid2D: over a TSDF2D its stack is built from
    1.f - |GetCorrespondenceCost| with the base class's table (FCSM2D.cc:97-113, grid_2d.cc:59-70;
    unknown and out-of-grid cells read +truncation_distance).  Stack bytes, full-submap and
    local searches against the oracle on a painted office world."""
    import tsdf_util
    w = synth.World(300, 260, seed=31)
    tv, wv = tsdf_util.from_world(w, 0.3, 10.0, seed=5)
    args = (tv, wv, w.res, w.max_x, w.max_y, 0.3, 10.0)
    og, g = ob.OracleTsdf(*args), csm.TsdfGrid(*args)
    depth = 6
    om = ob.OracleFCSM2D(og, 3.0, 0.6, depth)
    m = csm.Matcher(gpu_ctx, g, 3.0, 0.6, depth)
    for lvl in range(depth):
        assert np.array_equal(om.level(lvl), m.level(lvl)), "level %d differs" % lvl
    assert len(np.unique(m.level(0))) > 10  # a real TSD profile, not two values
    rng = np.random.default_rng(6)
    clouds, inits = [], []
    for i in range(6):
        pose, pts = synth.make_scan(w, 700 + i, max_range=6.0)
        clouds.append(pts)
        inits.append(None if i < 3 else (pose[0] + rng.uniform(-1, 1), pose[1] + rng.uniform(-1, 1),
                                         pose[2] + rng.uniform(-0.3, 0.3)))
    # TSDF scores live in [1 - d, 1 + d]: a wall cell scores ~1, unknown space 1 - d = 0.7
    res = csm.match_batch(gpu_ctx, [m] * 6, clouds, [0.8] * 6, initial=inits)
    found = 0
    for pts, init, r_g in zip(clouds, inits, res):
        r_o = om.match(pts, 0.8, initial=init)
        _compare(r_o, r_g)
        found += r_o.found
    assert found >= 3


def test_scan_wholly_outside_a_small_grid(csm, gpu_ctx):
    """ShrinkToFit seeds its bounds at 0 (CSM2D.cc:78-79), so a short-range scan placed far
    outside a small grid keeps a window of up to linear_bound cells towards the grid: the
    lowest-resolution lattice is larger than (grid + scan extent) / step.  The reference
    reports no match; so must the device (no capacity error, no abort)."""
    w = synth.World(100, 100, seed=8)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    om = ob.OracleFCSM2D(og, 7.0, 0.3, 5)
    m = csm.Matcher(gpu_ctx, g, 7.0, 0.3, 5)
    ang = np.linspace(0, 2 * np.pi, 40, endpoint=False)
    pts = np.stack([0.5 * np.cos(ang), 0.5 * np.sin(ang), np.zeros(40)], 1).astype(np.float32)
    inits = [(20.0, 0.0, 0.0), (0.0, -20.0, 0.3), (-9.0, 9.0, 0.0), (3.2, 0.0, 0.0)]
    res = csm.match_batch(gpu_ctx, [m] * len(inits), [pts] * len(inits), [0.3] * len(inits), initial=inits)
    for init, r_g in zip(inits, res):
        _compare(om.match(pts, 0.3, initial=init), r_g)
    assert res[0].found == 0 and res[1].found == 0


def test_constraint_builder_batch_entry_and_metrics(csm):
    """MaybeAddConstraints (whole-node bursts) returns what one
    MaybeAdd*Constraint call per search returns -- same gate, same sampler sequence, same
    submission order -- and the builder reports what the reference reports: searched / found
    counters per search region, the score histograms and the queue length
    (constraint_builder_2d.cc:46-53,205-240,289-293,318-343)."""
    worlds, poses, sub, full, rel, clouds = _builder_case(7)
    ograds = [ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    grids = [csm.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    kw = dict(sampling_ratio=0.5, max_constraint_distance=15.0, min_score=0.45,
              global_min_score=0.5, linear_window=3.0, angular_window=0.5, depth=5)
    _, o_out, _ = ob.constraint_batch(4, ograds, poses, sub, full, rel, clouds, **kw)
    _, f1, t1, s1 = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, flush_threshold=6,
                                               extended=True, **kw)
    _, f2, t2, s2 = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, flush_threshold=6,
                                               burst=5, extended=True, **kw)
    assert np.array_equal(f1, f2) and np.array_equal(t1, t2)
    assert np.array_equal(f1, np.array([o.found for o in o_out], np.int32))
    for s in (s1, s2):
        searched, found = s[0], s[1]
        assert found == f1.sum() and found > 0
        assert s[4] + s[6] == searched and s[5] + s[7] == found          # local + global
        assert list(s[9:13]) == list(s[4:8])                            # the metrics families agree
        assert s[13] == 0 and s[14] == found and s[15] == found         # queue drained, scores observed
        assert s[16] == searched                                        # one device ran everything
    assert s1[0] == s2[0] < len(clouds)  # the sampler and the distance gate dropped some, identically


def _device_count(csm):
    return csm.abi().csm_device_count()


def test_constraint_builder_two_devices_partition_and_hot_submap(csm):
    """ConstraintBuilder2D(options, pool, {0, 1}): a flush is sharded by submap affinity and cost,
    both devices run their shard concurrently, results come back in submission order and equal the
    one-device run bit for bit; a burst against ONE submap is split and the submap replicated."""
    if _device_count(csm) < 2:
        pytest.skip("needs two CUDA devices")
    worlds, poses, sub, full, rel, clouds = _builder_case(9, n_sub=4, n_search=40)
    grids = [csm.Grid(w.cells, w.res, w.max_x, w.max_y) for w in worlds]
    kw = dict(sampling_ratio=1.0, max_constraint_distance=15.0, min_score=0.45,
              global_min_score=0.5, linear_window=3.0, angular_window=0.5, depth=5, extended=True)
    _, f1, t1, s1 = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, devices=(0,), **kw)
    _, f2, t2, s2 = csm.constraint_builder_run(grids, poses, sub, full, rel, clouds, devices=(0, 1), **kw)
    assert np.array_equal(f1, f2) and np.array_equal(t1, t2) and f1.sum() > 0
    assert s2[16] > 0 and s2[17] > 0 and s2[16] + s2[17] == s2[0]
    assert abs(s2[16] - s2[17]) <= 0.4 * s2[0]
    # hot submap: every search against submap 0
    hot = [0] * len(clouds)
    _, f3, t3, s3 = csm.constraint_builder_run(grids, poses, hot, full, rel, clouds, devices=(0,), **kw)
    _, f4, t4, s4 = csm.constraint_builder_run(grids, poses, hot, full, rel, clouds, devices=(0, 1), **kw)
    assert np.array_equal(f3, f4) and np.array_equal(t3, t4)
    assert s4[8] >= 1 and s4[16] > 0 and s4[17] > 0   # the stack was replicated, the node list split


def test_matcher_on_a_grid_read_from_a_proto_message(csm, gpu_ctx):
    """Grid ingestion end to end: serialized proto::Grid2D -> grid handle -> stack -> search, identical
    to the same search on the grid built directly."""
    import proto_fixtures as pf
    w = synth.World(220, 200, seed=91)
    direct = csm.Matcher(gpu_ctx, csm.Grid(w.cells, w.res, w.max_x, w.max_y), 7.0, np.deg2rad(30.0), 5)
    loaded = csm.Matcher(gpu_ctx, csm.ProtoGrid(pf.grid2d_bytes(w.cells, w.res, w.max_x, w.max_y)), 7.0,
                         np.deg2rad(30.0), 5)
    for lvl in range(5):
        assert np.array_equal(direct.level(lvl), loaded.level(lvl))
    _, pts = synth.make_scan(w, 17, max_range=8.0)
    a, b = csm.match_batch(gpu_ctx, [direct, loaded], [pts, pts], [0.4, 0.4])
    assert a.found and (a.found, a.scan_index, a.x_index_offset, a.y_index_offset, a.score) == \
        (b.found, b.scan_index, b.x_index_offset, b.y_index_offset, b.score)


def test_real_time_matcher_takes_clouds_of_any_size(csm, gpu_ctx):
    """clouds of more than 3200 points used to be rejected (the discrete scan of a warp
    did not fit shared memory); they now go through a global scratch slice.  One batch mixes a
    5000-point cloud with an ordinary one."""
    w = synth.World(300, 300, seed=23)
    og = ob.OracleGrid(w.cells, w.res, w.max_x, w.max_y)
    g = csm.Grid(w.cells, w.res, w.max_x, w.max_y)
    opts = {"linear_search_window": 0.1, "angular_search_window": float(np.deg2rad(10.0)),
            "translation_delta_cost_weight": 0.1, "rotation_delta_cost_weight": 0.1}
    rng = np.random.default_rng(4)
    pose, pts = synth.make_scan(w, 77)
    big = np.ascontiguousarray(np.concatenate([pts + rng.normal(0, 0.02, pts.shape).astype(np.float32) * [1, 1, 0]
                                               for _ in range(5000 // len(pts) + 1)])[:5000].astype(np.float32))
    clouds = [big, np.ascontiguousarray(pts)]
    inits = [(pose[0] + 0.04, pose[1] - 0.03, pose[2] + 0.05)] * 2
    res = csm.rt_match_batch(gpu_ctx, [g, g], clouds, inits, **opts)
    for c, init, r_g in zip(clouds, inits, res):
        r_o = ob.rt_match(og, opts, init, c)
        assert (r_g.scan_index, r_g.x_index_offset, r_g.y_index_offset) == \
               (r_o.scan_index, r_o.x_index_offset, r_o.y_index_offset)
        assert np.float32(r_g.score).tobytes() == np.float32(r_o.score).tobytes()
