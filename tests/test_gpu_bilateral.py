"""Bilateral rows + MINRES on the device against the numpy restatement of the reference's path
(oracle/np_oracle.py bilateral_solve), and a blade driven through a sphere bed by a speed motor."""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import links, scenes
from oracle import np_oracle as no

pytestmark = pytest.mark.gpu


def bed(seed=21):
    s = scenes.settling_smc(8, 8, 3, spacing_factor=1.98, jitter_factor=0.03, seed=seed)
    s["pos"][1:, 2] -= 0.025 * 0.02
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.05, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    return s


@pytest.mark.parametrize("nrows,tol,iters", [(6, 1e-12, 100), (40, 1e-10, 200), (40, 1e-4, 100), (25, 0.0, 7)])
def test_bilateral_solve_matches_restatement(nrows, tol, iters):
    scene = bed()
    nb = scene["mass"].size
    rng = np.random.default_rng(nrows)
    ids = np.stack([rng.integers(0, nb, nrows), rng.integers(0, nb, nrows)], 1).astype(np.int32)
    ids[ids[:, 0] == ids[:, 1], 1] = (ids[ids[:, 0] == ids[:, 1], 1] + 1) % nb
    ids[0] = [0, 5]  # one row against the fixed container body
    cq_a, cq_b = rng.normal(size=(nrows, 6)), rng.normal(size=(nrows, 6))
    cq_a[:, 3:] *= 0.02
    cq_b[:, 3:] *= 0.02  # angular columns act through 1/I ~ 1e4: keep the rows comparably scaled
    rhs = rng.normal(scale=0.1, size=nrows)

    free = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    free.Update(); free.RunCollision(7); free.RunSolver()
    hf = free.GetHf().reshape(nb, 6)  # gravity + gyroscopic + contact forces, before any joint
    inv_inertia = 1.0 / scene["inertia"]
    want_v, want_g, want_it, want_res = no.bilateral_solve(ids, cq_a, cq_b, rhs, scene["vel"], hf, scene["mass"],
                                                           inv_inertia, scene["active"], iters, tol)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.SetBilaterals(ids, cq_a, cq_b, rhs, max_iteration_bilateral=iters, tol_speed=tol)
    sys.Update(); sys.RunCollision(7); sys.RunSolver()
    g, it, res = sys.GetBilateralGamma(nrows)
    _, _, v = sys.GetBodyState()
    assert it == want_it
    assert np.abs(g - want_g).max() <= 1e-9 * np.abs(want_g).max()
    assert np.abs(v - want_v).max() <= 1e-9 * np.abs(want_v).max()
    if tol > 0 and tol <= 1e-10:
        viol = np.array([cq_a[i] @ v[ids[i, 0]] + cq_b[i] @ v[ids[i, 1]] + rhs[i] for i in range(nrows)])
        assert np.abs(viol).max() < 1e-7
    # rows removed: back to the unconstrained velocities
    sys2 = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys2.SetBilaterals(ids, cq_a, cq_b, rhs)
    sys2.SetBilaterals(None, None, None, None)
    sys2.Update(); sys2.RunCollision(7); sys2.RunSolver()
    _, _, v2 = sys2.GetBodyState()
    _, _, vf = free.GetBodyState()
    assert np.array_equal(v2, vf)


def test_bilaterals_refused_for_nsc():
    scene = scenes.balls_nsc(4, 4, 2)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    with pytest.raises(RuntimeError):
        sys.SetBilaterals([[0, 1]], np.zeros((1, 6)), np.zeros((1, 6)), [0.0])


def test_motor_drives_a_blade_through_the_bed():
    """Config E's mechanism (demo_MCORE_mixerSMC.cpp:80-110): a blade on a speed motor about the vertical axis."""
    scene, blade, anchor = scenes.mixer_smc()
    h = scene["settings"]["step_size"]
    omega = 4.0
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    steps = 400
    for _ in range(steps):
        pos, rot, vel = sys.GetBodyState(blade, 1)
        rows = links.motor_rotation_speed_z(0, blade, pos[0], rot[0], anchor, omega, h)
        sys.SetBilaterals(*rows, max_iteration_bilateral=100, tol_speed=1e-10)
        sys.DoSteps(1)  # whole device step: the rows are solved between force accumulation and integration
    sys.Synchronize()
    pos, rot, vel = sys.GetBodyState()
    R = links.rotmat(rot[blade])
    angle = np.arctan2(R[1, 0], R[0, 0])
    assert abs(angle - omega * h * steps) < 2e-3            # turned by omega * t
    assert np.abs(pos[blade] - anchor).max() < 1e-5         # stays on its pivot although spheres push it
    assert abs(R[2, 2] - 1) < 1e-8                          # axis stays vertical
    assert abs((R @ vel[blade, 3:])[2] - omega) < 1e-6
    g, it, res = sys.GetBilateralGamma(6)
    assert np.abs(g).max() > 0 and res < 1e-10
    # the bed felt it: spheres near the blade were swept along
    r = np.hypot(pos[1:blade, 0], pos[1:blade, 1])
    near = r < 0.3
    tang = (-pos[1:blade, 1] * vel[1:blade, 0] + pos[1:blade, 0] * vel[1:blade, 1])[near]
    assert near.sum() > 20 and tang.mean() > 0
