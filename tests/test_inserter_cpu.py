"""Oracle restatement of ProbabilityGridRangeDataInserter2D / RayToPixelMask against the
reference's known answers (mapping/2d/range_data_inserter_2d_test.cc:30-137)."""
import numpy as np

import oracle_binding as ob

RETURNS = [(-3.5, 0.5, 0.0), (-2.5, 1.5, 0.0), (-1.5, 2.5, 0.0), (-0.5, 3.5, 0.0)]
ORIGIN = (-0.5, 0.5)
U, M, H = 0, 1, 2
EXPECTED = [[U, U, U, U, U], [U, H, M, M, M], [U, U, H, M, M], [U, U, U, H, M], [U, U, U, U, H]]


def fixture():
    return ob.OracleGrid(np.zeros((5, 5), np.uint16), 1.0, 1.0, 5.0)  # MapLimits(1., (1, 5), (5, 5))


def probability(grid, cell_x, cell_y):
    table = np.empty(65536, np.float32)
    ob.lib().oracle_correspondence_cost_table(table)
    return 1.0 - table[grid.cells[cell_y, cell_x]]


def test_insert_point_cloud_known_states():
    g = ob.grid_insert(fixture(), ORIGIN, RETURNS, hit_probability=0.7, miss_probability=0.4)
    assert (g.nx, g.ny) == (5, 5) and abs(g.max_x - 1.0) < 1e-9 and abs(g.max_y - 5.0) < 1e-9
    for row in range(5):          # reference :86-104: cell_index(row, column), expected[column][row]
        for column in range(5):
            state = EXPECTED[column][row]
            if state == U:
                assert g.cells[column, row] == 0
            else:
                want = 0.7 if state == H else 0.4
                assert abs(probability(g, row, column) - want) < 1e-4


def test_probability_progression():
    g = fixture()
    ob.grid_insert(g, ORIGIN, RETURNS, hit_probability=0.7, miss_probability=0.4)
    hit_cell = (int(round((5.0 - 0.5) / 1.0 - 0.5)), int(round((1.0 + 3.5) / 1.0 - 0.5)))
    miss_cell = (int(round((5.0 - 0.5) / 1.0 - 0.5)), int(round((1.0 + 2.5) / 1.0 - 0.5)))
    assert abs(probability(g, *hit_cell) - 0.7) < 1e-4 and abs(probability(g, *miss_cell) - 0.4) < 1e-4
    for _ in range(1000):
        ob.grid_insert(g, ORIGIN, RETURNS, hit_probability=0.7, miss_probability=0.4)
    assert abs(probability(g, *hit_cell) - 0.9) < 1e-3 and abs(probability(g, *miss_cell) - 0.1) < 1e-3


def test_grid_grows_around_the_range_data():
    g = ob.OracleGrid(np.zeros((4, 4), np.uint16), 0.5, 1.0, 1.0)
    ob.grid_insert(g, (0.2, 0.1), [(5.0, -3.0, 0.0)], misses=[(-4.0, 2.0, 0.0)])
    assert g.nx >= 16 and g.ny >= 16 and (g.cells > 0).sum() > 10
