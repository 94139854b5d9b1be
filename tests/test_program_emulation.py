"""CPU: the product's shape compiler (program.hpp) and program evaluator (csg_eval.cuh, compiled for the host by
tests/emu) against the oracle's recursive tree, bit for bit."""
import json

import numpy as np
import pytest

from implisolid_b200 import shapes

SHAPES = ["c1_sphere", "c2_cube_minus_cylinder", "c3_metaballs", "c3_sum_blend", "c4_deep_chain", "demo_moon", "simple_cone", "bushy_tree",
          # SURVEY §8 f-N1: torus / heart / mushroom go through double-precision powers (glibc pow in the oracle as in the
          # reference, multiplications and sqrt in the product); on these seeded points the results are bit-identical
          "n1_torus", "n1_heart", "n1_tetrahedron", "n1_mushroom", "n1_half_plane_cut", "n1_lid_slab", "n1_medley"]


@pytest.mark.parametrize("name", SHAPES)
def test_values_and_gradients_bit_exact(emu, oracle, name):
    sj = getattr(shapes, name)()
    rng = np.random.default_rng(11)
    pts = np.concatenate([rng.uniform(-1.2, 1.2, (3000, 3)), rng.normal(0, 0.3, (1000, 3))]).astype(np.float32)
    f1, f4, g = emu.eval(sj, pts)
    fo = oracle.query_implicit_values(sj, pts)
    go = oracle.query_normals(sj, pts, normalize_and_invert=False)
    assert np.array_equal(f1.view(np.uint32), fo.view(np.uint32))
    assert np.array_equal(f4.view(np.uint32), fo.view(np.uint32))
    assert np.array_equal(g.view(np.uint32), go.view(np.uint32))


def test_inverse_matches_oracle_bitwise(emu, oracle):
    rng = np.random.default_rng(2)
    for _ in range(200):
        m = rng.normal(size=12).astype(np.float32)
        assert np.array_equal(emu.affine_inverse(m).view(np.uint32), oracle.invert_matrix(m).view(np.uint32))


def test_program_shape(emu):
    words, st = emu.compile(shapes.c4_deep_chain())
    assert st["nodes"] == 63 and st["frame_depth"] == 0 and st["value_depth"] == 32
    words, st = emu.compile(shapes.bushy_tree())
    assert st["frame_depth"] == 1
    words, st = emu.compile(shapes.c2_cube_minus_cylinder(), ignore_root_matrix=True)
    eye = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], dtype=np.float32).view(np.uint32)
    # OP_XFORM of node id 0 (root), skip word = END's pc, then the identity (copy_eye, object_factory.hpp:47)
    assert words[0] == 1 and words[1] == len(words) - 4 and np.array_equal(words[4:16], eye)
    assert len(words) % 4 == 0                                      # every instruction is 16-byte aligned


@pytest.mark.parametrize("name", SHAPES)
def test_dead_children_are_skipped_without_changing_bits(emu, oracle, name):
    """Tile specialisation, interpreter side: with every unselected child of every min/max node marked dead, the program
    evaluator must return the same bits as the full tree (skip offsets, operand bookkeeping, -f2 for a lone subtrahend)."""
    sj = getattr(shapes, name)()
    rng = np.random.default_rng(5)
    pts = np.concatenate([rng.uniform(-1.2, 1.2, (1500, 3)), rng.normal(0, 0.3, (500, 3))]).astype(np.float32)
    ft, fm, live, nodes = emu.eval_masked(sj, pts)
    fo = oracle.query_implicit_values(sj, pts)
    assert nodes > 0                                  # every synthetic shape is small enough to specialise
    assert np.array_equal(ft.view(np.uint32), fo.view(np.uint32))
    assert np.array_equal(fm.view(np.uint32), fo.view(np.uint32))
    depth_bound = int(np.ceil(np.log2(nodes + 1))) if name == "bushy_tree" else nodes
    assert live.max() <= depth_bound and live.min() >= 1


@pytest.mark.parametrize("seed", range(12))
def test_random_trees_every_interpreter_bit_exact(emu, oracle, seed):
    """Random CSG trees (every operator, n-ary unions, chains and bushy parts, random transforms on every node): the value
    interpreter (1 and 4 points), the gradient interpreter, both under per-point liveness masks and both in per-tile list
    form must return the oracle's bits."""
    sj = shapes.random_tree(seed)
    rng = np.random.default_rng(100 + seed)
    pts = np.concatenate([rng.uniform(-1.2, 1.2, (1500, 3)), rng.normal(0, 0.3, (500, 3))]).astype(np.float32)
    fo = oracle.query_implicit_values(sj, pts)
    go = oracle.query_normals(sj, pts, normalize_and_invert=False)
    f1, f4, g = emu.eval(sj, pts)
    ft, fm, live, nodes = emu.eval_masked(sj, pts)       # NaN-poisoned if any masked / list interpreter disagrees
    for got in (f1, f4, ft, fm):
        assert np.array_equal(got.view(np.uint32), fo.view(np.uint32))
    assert np.array_equal(g.view(np.uint32), go.view(np.uint32))


def _boxes_near_surface(emu, sj, rng, n, half_extent):
    """Box centres where |f| is small (the tiles K1 cannot prove inside or outside), half extents around `half_extent`."""
    pts = rng.uniform(-1.1, 1.1, (30 * n, 3)).astype(np.float32)
    f, _, _ = emu.eval(sj, pts)
    order = np.argsort(np.abs(f))[:n]
    h = (half_extent * rng.uniform(0.5, 1.5, (n, 3))).astype(np.float32)
    return np.concatenate([pts[order], h], axis=1)


@pytest.mark.parametrize("name", SHAPES + ["random_%d" % s for s in range(8)] + ["randomall_%d" % s for s in (322, 335, 339)])
def test_interval_bounds_and_tile_masks_are_sound(emu, name):
    """K1's pruning and specialisation (csg_bounds.cuh, compiled for the host): over boxes of tile and supertile size near the
    surface, the interval bounds contain the exact values at random points of the box, the liveness mask made from the
    decisions changes no bit of any interpreter's value or gradient there (mask walk, value list, gradient list), and a
    sub-box bounded under that mask (K1's second level) never gets a looser mask."""
    if name.startswith("randomall_"):      # every primitive kind, torus / heart / mushroom included; up to 187 nodes
        seed = int(name.split("_")[1])
        sj = shapes.random_tree(seed, depth=3 + seed % 4, all_prims=True)
    else:
        sj = shapes.random_tree(int(name.split("_")[1])) if name.startswith("random_") else getattr(shapes, name)()
    rng = np.random.default_rng(17)
    for half_extent in (0.012, 0.05):
        nodes, st = emu.bounds_check(sj, _boxes_near_surface(emu, sj, rng, 300, half_extent), samples=10, seed=3)
        if nodes < 0:
            pytest.skip("tree too large to specialise")
        assert st["bounds_violations"] == 0 and st["value_mismatches"] == 0 and st["gradient_mismatches"] == 0, st
        assert st["looser_second_level"] == 0, st
        assert st["live_nodes"] >= 300          # the root is always live


@pytest.mark.parametrize("name", SHAPES + ["random_%d" % s for s in range(6)] + ["randomall_%d" % s for s in (322, 335)])
def test_packed_pair_interpreters_keep_every_bit(emu, name):
    """The f32x2 interpreters (csg_eval.cuh: eval_values_list2, eval_value_gradient_list2) run two points through one instruction
    list; on the host a pair is two plain floats, so this holds their operation ORDER to the scalar interpreters' and the full
    tree's, for nearby point pairs (the GPU case: neighbouring nodes / centroids) and for unrelated ones."""
    if name.startswith("randomall_"):
        sj = shapes.random_tree(int(name.split("_")[1]), all_prims=True)
    elif name.startswith("random_"):
        sj = shapes.random_tree(int(name.split("_")[1]))
    else:
        sj = getattr(shapes, name)()
    rng = np.random.default_rng(5)
    base = rng.uniform(-1.1, 1.1, (1500, 1, 3))
    near = (base + rng.normal(0, 0.004, (1500, 4, 3))).reshape(-1, 3)
    far = rng.uniform(-1.2, 1.2, (2000, 3))
    pts = np.concatenate([near, far]).astype(np.float32)
    assert emu.eval_packed_mismatches(sj, pts) == 0
