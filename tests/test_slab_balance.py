"""z-slab boundaries (csrc/slab_balance.hpp, compiled for the host in tests/emu): the first cuts equal the Python mirror, and the
feedback from measured busy times evens out a cost profile that surface counts alone do not see."""
import numpy as np
import pytest

from implisolid_b200 import slabs


def _sphere_hist(R):
    """Surface tiles per bin of eight node planes for a sphere filling most of the box (a ring per plane)."""
    S = R + 5
    bins = (S + 7) // 8
    z = (np.arange(bins) * 8 + 4 - S / 2) / (0.45 * S)
    ring = np.sqrt(np.clip(1 - z * z, 0, None))
    return np.maximum((ring * 2 * np.pi * 0.45 * S / 8 * 1.5).astype(np.int64), 0).astype(np.uint32)


@pytest.mark.parametrize("R,world", [(256, 2), (1024, 8), (100, 3)])
def test_first_cuts_equal_the_mirror(emu, R, world):
    hist = _sphere_hist(R)
    layers = R + 2
    cuts, _, n = emu.slab_balance(hist, layers, world, np.ones(layers), 0, 0)
    assert n == 0
    assert list(cuts[0]) == slabs.balanced_cuts(hist, R, world)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_feedback_evens_out_a_hot_region(emu, world):
    """Cost per layer = surface tiles x (1 + 1.5 inside a hot band in the upper half) + a fixed cost per rank: after at most eight
    re-cuts the slowest rank is within 4 % of the mean (it starts 20-45 % above), and the cuts stay a partition."""
    R = 1024
    hist = _sphere_hist(R)
    layers = R + 2
    tiles = np.array([hist[(l + 1) >> 3] for l in range(layers)], dtype=np.float64)
    hot = np.zeros(layers)
    hot[int(0.62 * layers):int(0.80 * layers)] = 1.5
    cost = (tiles * (1.0 + hot) * 0.5).astype(np.int64) + 1
    cuts, busy, n = emu.slab_balance(hist, layers, world, cost, fixed_us=300, rounds=8)
    first = busy[0].max() * world / busy[0].sum()
    last = busy[-1].max() * world / busy[-1].sum()
    assert first > 1.15
    assert last < 1.04, (first, last, busy[-1])
    assert 1 <= n <= 8
    for c in cuts:
        assert c[0] == 0 and c[-1] == layers and (np.diff(c) >= 1).all()
    # deterministic: the same inputs give the same cuts
    cuts2, _, _ = emu.slab_balance(hist, layers, world, cost, fixed_us=300, rounds=8)
    assert np.array_equal(cuts, cuts2)


def test_no_recut_when_balanced(emu):
    R, world = 256, 4
    hist = _sphere_hist(R)
    layers = R + 2
    tiles = np.array([hist[(l + 1) >> 3] for l in range(layers)], dtype=np.int64)
    uniform = int(hist.sum()) * 8 // (4 * layers) + 1
    cuts, busy, n = emu.slab_balance(hist, layers, world, tiles + uniform, fixed_us=0, rounds=3)
    assert n == 0 and (cuts == cuts[0]).all()
