"""The tall-pair GEMM walks a static schedule of 512x256 and 256x256 tiles that the host plans
(plan_tall) and every CTA pair decodes on its own (TallSched, ops/csrc/gemm_sm100.cu).  The decode
is host-callable, so its coverage invariant is checked here without a GPU: every 256x256 block of
the output is produced exactly once, whatever the shape, the number of pairs or the cost model."""
import itertools

import pytest


@pytest.fixture
def ops(dtg):
    from dtg_b200 import ops as o
    try:
        o.lib()
    except Exception as e:                      # library not built on this box
        pytest.skip("libdtg_b200.so unavailable: %s" % e)
    return o


def test_model_shapes_are_covered_once(ops):
    shapes = [(8192, 8192), (8192, 4096), (4096, 4096), (1024, 50304), (4096, 784), (8192, 1024),
              (3072, 1024), (50304, 1024), (512, 256), (256, 256), (300, 70), (8200, 1000)]
    for (M, N), clusters in itertools.product(shapes, (74, 66, 1, 7)):
        plan = ops.tall_schedule_check(M, N, clusters)
        tn = -(-N // 256)
        assert plan["big_tiles"] == plan["big_mblocks"] * tn
        assert plan["small_tiles"] == -(-(M - 512 * plan["big_mblocks"]) // 256) * tn
        assert plan["clusters"] == min(clusters, plan["big_tiles"] + plan["small_tiles"])


def test_exhaustive_small_grid(ops):
    """Every (rows, cols, pairs, group, cost) of a small grid, including ragged edges."""
    n = 0
    for M in range(1, 4200, 197):
        for N in (1, 255, 256, 257, 1000, 2048):
            for clusters in (1, 2, 3, 5, 8, 74):
                for group in (1, 2, 8, 16):
                    for cost in (0.5, 0.62, 1.0):
                        ops.tall_schedule_check(M, N, clusters, cost, group)
                        n += 1
    assert n > 6000


def test_plan_balances_the_8k_cube(ops):
    """8192^3 on 74 pairs: 16 m-blocks x 32 n-blocks.  The plan must beat both pure layouts
    (all big: 512 tiles / 74 = 7 rounds -> load 7; all small: 1024 tiles -> 14 x 0.62 = 8.68)."""
    plan = ops.tall_schedule_check(8192, 8192, 74, 0.62)
    assert plan["worst_load"] <= 7.0 and plan["big_tiles"] > 0
    assert plan["big_tiles"] + plan["small_tiles"] // 2 <= 512


def test_grouped_tile_order_is_a_bijection(ops):
    """tile_coords(): the L2-friendly grouped order all persistent GEMM kernels share."""
    lib = ops.lib()
    for tm in (1, 2, 7, 8, 9, 31, 64, 393):
        for tn in (1, 3, 16, 197):
            for group in (1, 2, 4, 8, 16, 64):
                assert lib.dtg_tile_order_check(tm, tn, group) == 0, (tm, tn, group)
    assert lib.dtg_tile_order_check(0, 3, 8) != 0


def test_bad_arguments_are_refused(ops):
    with pytest.raises(Exception):
        ops.tall_schedule_check(0, 256, 74)
