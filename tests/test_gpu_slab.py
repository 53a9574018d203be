"""Slab decomposition on the device (csrc/slab.cu + chrono_wisc_b200/slab.py): several slabs of one bed, all living
on cuda:0 in one process and exchanging their buffers directly, against the same bed in one system.

Contacts straddling a cut are evaluated on both sides from the same state but with the two shapes possibly in the
other order, and per-body sums run in a different order than in the single system, so trajectories agree to rounding,
not bit for bit: 1e-9 relative after 45 steps (stated tolerance; BASELINE.json asks for observables only).
"""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes, slab

pytestmark = pytest.mark.gpu


def moving_bed(nx=24, ny=8, nz=4, seed=11):
    s = scenes.settling_smc(nx, ny, nz, spacing_factor=1.99, jitter_factor=0.02, seed=seed, grid="bin_size",
                            bin_edge=2.5)
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.4, size=(n - 1, 3))
    s["vel"][1:, 0] += 1.0  # the whole bed drifts in +x: a layer of bodies crosses every cut plane
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    return s


def run_slabs(scene, cuts, steps, width=0.05):
    world = len(cuts) - 1
    systems = []
    for r in range(world):
        L = slab.SlabLayout(world, r, cuts, axis=0, width=width)
        systems.append(slab.SlabSystemSMC(L, device=0).LoadSlab(scene))
    def exchange():
        for s in systems:
            s.Pack()
        slab.LocalExchange.run(systems)
        for s in systems:
            s.Unpack()
    exchange()  # initial ghost layers
    for _ in range(steps):
        for s in systems:
            s.StepOnly()
        exchange()
    for s in systems:
        s.Check()
    return systems


@pytest.mark.parametrize("cuts", [(-10.0, 0.0, 10.0), (-10.0, -0.2, 0.13, 10.0)])
def test_slabs_track_single_system(cuts):
    scene = moving_bed()
    steps = 45
    one = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    one.DoSteps(steps)
    p1, q1, v1 = one.GetBodyState()
    systems = run_slabs(scene, cuts, steps)
    ids = np.concatenate([s.GetOwned()[0] for s in systems])
    pos = np.concatenate([s.GetOwned()[1] for s in systems])
    vel = np.concatenate([s.GetOwned()[3] for s in systems])
    # every dynamic body is owned by exactly one slab
    assert np.array_equal(np.sort(ids), np.arange(1, scene["mass"].size))
    order = np.argsort(ids)
    assert np.abs(pos[order] - p1[1:]).max() < 1e-9 * np.abs(p1[1:]).max()
    assert np.abs(vel[order] - v1[1:]).max() < 1e-9 * np.abs(v1[1:]).max()
    # ownership follows position, ghosts are exactly the foreign bodies inside the band
    x = p1[1:, 0]
    migrated = 0
    for s in systems:
        L = s.layout
        mine = np.nonzero(L.owns(x))[0] + 1
        assert np.array_equal(s.GetOwned()[0], mine)
        want = []
        if L.left is not None:
            want.append(np.nonzero((x < L.lo) & (x >= L.lo - L.width))[0] + 1)
        if L.right is not None:
            want.append(np.nonzero((x >= L.hi) & (x < L.hi + L.width))[0] + 1)
        assert np.array_equal(s.GetGhostIds(), np.sort(np.concatenate(want)))
        migrated += (~L.owns(scene["pos"][mine, 0])).sum()
    assert migrated > 5  # the test did exercise ownership transfer


def test_ghost_capacity_overflow_is_reported():
    scene = moving_bed(16, 6, 3)
    L = slab.SlabLayout(2, 0, (-10.0, 0.0, 10.0), width=0.05)
    s = slab.SlabSystemSMC(L, device=0, ghost_cap=4).LoadSlab(scene)
    s.Pack()
    with pytest.raises(cw._lib.Cb200Error):
        s.Check()
