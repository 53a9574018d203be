"""Scene builders without a GPU: the Poisson-disk layers of BASELINE.json configs[0] (btest_MCORE_settling.cpp:59-131)."""
import numpy as np

from chrono_wisc_b200 import scenes


def test_poisson_disk_layers_invariants():
    r, rp = 0.02, 1.01 * 0.02
    s = scenes.settling_pd(half=(0.4, 0.3), layers=3, radius=r, seed=0)
    p = s["pos"][1:]
    assert np.array_equal(np.unique(np.round(p[:, 2] / (2 * rp))), [1, 2, 3])  # layers at 2r', 4r', 6r'
    assert np.abs(p[:, 0]).max() <= 0.4 - rp and np.abs(p[:, 1]).max() <= 0.3 - rp
    for z in np.unique(p[:, 2]):
        q = p[p[:, 2] == z][:, :2]
        d = np.sqrt(((q[:, None] - q[None]) ** 2).sum(-1)) + np.eye(q.shape[0])
        assert d.min() >= 2 * rp                      # no two closer than the minimum distance
        # maximal in practice: Bridson samples fill 0.60 .. 0.72 points per d^2 of area
        dens = q.shape[0] * (2 * rp) ** 2 / ((0.8 - 2 * rp) * (0.6 - 2 * rp))
        assert 0.58 < dens < 0.75, dens
    # same seed, same sample; the container is the reference's five boxes
    s2 = scenes.settling_pd(half=(0.4, 0.3), layers=3, radius=r, seed=0)
    assert np.array_equal(s["pos"], s2["pos"])
    assert (s["shape_type"][:5] == 2).all() and s["settings"]["bins_per_axis"] == (10, 10, 1)
