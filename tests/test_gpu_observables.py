"""Ensemble observables over a long run (north_star's third correctness mode): the CUDA path and the reference's CPU path
(oracle/_ref collision + extracted force function + restated glue, ref_smc_step) advance the reference's own settling
scene (btest_MCORE_settling.cpp:59-131, Poisson-disk layers, 10 x 10 x 1 bins) for 5000 steps each.

Granular trajectories are chaotic, so nothing is compared body by body: kinetic-energy decay curve, mean height, bed
height (99th percentile of z + r) and packing fraction must agree within the stated tolerances at every checkpoint.
(Measured: 5 significant digits at step 5000 -- the two paths differ by rounding in libm calls and summation order only.)"""
import os

import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

pytestmark = [pytest.mark.gpu, pytest.mark.slow]


def test_settling_observables_match_reference_path(ref):
    r = 0.02
    scene = scenes.settling_pd(half=(0.5, 0.5), layers=8, radius=r, seed=0)
    nb = scene["mass"].size
    m, inertia = scene["mass"][1:], scene["inertia"][1:]

    def observables(p, v):
        ke = float((0.5 * m * (v[1:, :3] ** 2).sum(1) + 0.5 * (inertia * v[1:, 3:] ** 2).sum(1)).sum())
        z = p[1:, 2]
        h = float(np.percentile(z, 99) + r)
        return dict(ke=ke, mean_z=float(z.mean()), bed_height=h, packing=(nb - 1) * 4 / 3 * np.pi * r ** 3 / (1.0 * 1.0 * h))

    gpu = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    ref.set_num_threads(os.cpu_count() or 1)
    ref.smc_setup(scene)
    ke0 = None
    for k in range(10):
        gpu.DoSteps(500)
        ref.smc_step(500)
        pg, _, vg = gpu.GetBodyState()
        pr, _, vr = ref.smc_state()
        a, b = observables(pg, vg), observables(pr, vr)
        ke0 = ke0 or b["ke"]
        assert abs(a["ke"] - b["ke"]) <= 0.05 * b["ke"], (k, a, b)          # KE(t) decay curve: 5 %
        for key in ("mean_z", "bed_height", "packing"):
            assert abs(a[key] - b[key]) <= 0.01 * abs(b[key]), (k, key, a, b)  # 1 %
    assert b["ke"] < 0.5 * ke0 and 0.5 < b["packing"] < 0.64  # the bed did compact and lose energy
