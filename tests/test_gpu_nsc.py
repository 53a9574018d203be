"""CUDA NSC (DVI) solve vs the CPU restatement (oracle/oracle_nsc.inc: explicit D_T / M_invD CSR products, serial
APGD exactly as the reference writes it) on identical states, through the C ABI.

Tolerances: the contact list is bit-exact (same collision path as SMC, with the NSC envelope); Jacobian rows in the
reference's arithmetic are bit-exact against the oracle (and the oracle's frame code against the reference's own
Orthogonalize / Compute_Jacobian, tests/test_oracle_cpu.py); everything that goes through a Schur product is
tolerance-based (Eigen's summation order is not pinned by any reference test):
  R_full, one Schur product, the projection      1e-12 relative
  gamma / velocities after a 30-iteration solve    1e-8 relative   (rounding differences amplified by APGD)
  residual history                                 1e-6 relative
"""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import pyoracle

pytestmark = pytest.mark.gpu


def relerr(a, b):
    scale = max(np.abs(b).max(), 1e-300)
    return np.abs(np.asarray(a) - np.asarray(b)).max() / scale


def bed(nx=8, ny=8, nz=4, seed=5, iterations=30, mode=cw.system.MODE_SLIDING, **st):
    s = scenes.balls_nsc(nx, ny, nz, seed=seed, iterations=iterations)
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    # squeeze the lattice so that many contacts are active, random velocities and orientations
    s["pos"][1:, 2] = 0.15 + (s["pos"][1:, 2] - 0.15) * 0.97
    s["vel"][1:, :3] = rng.normal(scale=0.3, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=1.0, size=(n - 1, 3))
    q = rng.normal(size=(n - 1, 4))
    s["rot"][1:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    s["settings"].update(solver_mode=mode, **st)
    return s


def both(scene):
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    orc = pyoracle.Oracle(scene)
    orc.collide()
    orc.nsc_solve()
    return sys, orc


def test_contact_list_with_envelope_is_exact():
    scene = bed()
    sys, orc = both(scene)
    g, c = sys.GetContacts(), orc.contacts()
    assert c["depth"].size > 300
    assert (c["depth"] > 0).any() and (c["depth"] < 0).any()  # envelope contacts and penetrations
    for k in ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius"):
        assert np.array_equal(g[k], c[k]), k


@pytest.mark.parametrize("mode", [cw.system.MODE_NORMAL, cw.system.MODE_SLIDING, cw.system.MODE_SPINNING])
def test_jacobian_rows_bit_exact(mode):
    scene = bed(6, 6, 3, mode=mode, iterations=2,
                **({"max_iteration_spinning": 2} if mode == cw.system.MODE_SPINNING else {}))
    sys, orc = both(scene)
    rows, bids = sys.GetJacobianRows()
    ov, ob = orc.nsc_jacobian_rows()
    # the oracle stores the block of the lower body id first (Eigen's sorted inner indices); reorder ours
    swap = bids[:, 0] > bids[:, 1]
    mine = rows.copy()
    mine[swap] = np.concatenate([rows[swap, 6:], rows[swap, :6]], axis=1)
    assert np.array_equal(np.sort(bids, axis=1), ob)
    assert np.array_equal(mine, ov)


@pytest.mark.parametrize("mode", [cw.system.MODE_NORMAL, cw.system.MODE_SLIDING, cw.system.MODE_SPINNING])
def test_rhs_schur_product_and_projection(mode):
    extra = {"max_iteration_spinning": 2} if mode == cw.system.MODE_SPINNING else {}
    if mode == cw.system.MODE_NORMAL:
        extra = {"max_iteration_normal": 2}
    scene = bed(7, 7, 3, mode=mode, iterations=2, **extra)
    scene["materials"] = np.array([scenes.nsc_material(friction=0.4, cohesion=0.02, rolling=0.01, spinning=0.005,
                                                       compliance=1e-6, complianceT=2e-6, complianceRoll=1e-6,
                                                       complianceSpin=1e-6)])
    sys, orc = both(scene)
    v = orc.nsc_vectors()
    assert relerr(sys.GetRhs(), v["R"]) < 1e-12
    rng = np.random.default_rng(0)
    x = rng.normal(size=v["R"].size)
    for lm in range(mode + 1):
        assert relerr(sys.SchurProduct(x, lm), orc.nsc_schur(x, lm)) < 1e-12, lm
        g = rng.normal(size=x.size) * 3
        assert np.array_equal(sys.Project(g, lm), orc.nsc_project(g, lm)), lm


@pytest.mark.parametrize("solver", [cw.system.SOLVER_APGD, cw.system.SOLVER_APGDREF])
def test_apgd_solve_matches_oracle(solver):
    scene = bed(8, 8, 4, iterations=30, solver_type=solver)
    sys, orc = both(scene)
    info = orc.nsc_info()
    c = sys.GetCounts()
    assert c.solver_iterations == info["iterations"] == 30
    r, o = sys.GetSolverHistory()
    ro, oo = orc.nsc_hist()
    assert r.size == ro.size
    assert relerr(r, ro) < 1e-6
    assert relerr(sys.GetGamma(), orc.nsc_vectors()["gamma"]) < 1e-8
    _, _, vel = sys.GetBodyState()
    _, _, ovel = orc.state()
    assert relerr(vel, ovel) < 1e-8
    assert abs(c.solver_residual - info["residual"]) <= 1e-6 * info["residual"]


def test_phases_normal_then_sliding_and_tolerance_stop():
    scene = bed(6, 6, 3, iterations=25, max_iteration_normal=10, tolerance=20.0)
    sys, orc = both(scene)
    info = orc.nsc_info()
    assert sys.GetCounts().solver_iterations == info["iterations"]
    assert info["iterations"] < 35  # the tolerance break fired in at least one phase
    assert relerr(sys.GetGamma(), orc.nsc_vectors()["gamma"]) < 1e-8


def test_objective_stop():
    """settings.solver.test_objective (ChSolverMulticoreAPGD.cpp:184-192): APGD leaves when the objective of its best
    iterate is at or below tolerance_objective; APGDREF keeps the residual test (APGDREF.cpp:174)."""
    full = bed(8, 8, 4, iterations=40)
    sys0, _ = both(full)
    _, o0 = sys0.GetSolverHistory()
    k = 12
    assert o0[k] < o0[k - 1]  # the objective still falls there: a threshold between the two stops at iteration k
    thr = 0.5 * (o0[k] + o0[k - 1])
    scene = bed(8, 8, 4, iterations=40, test_objective=1, tolerance_objective=float(thr))
    sys, orc = both(scene)
    info = orc.nsc_info()
    r, o = sys.GetSolverHistory()
    assert r.size == k + 1 and o[-1] <= thr < o[-2]
    assert sys.GetCounts().solver_iterations == info["iterations"] == k  # the break does not advance the counter
    assert relerr(sys.GetGamma(), orc.nsc_vectors()["gamma"]) < 1e-8
    # a tolerance on the residual that would have stopped at once is not looked at any more ...
    scene2 = bed(8, 8, 4, iterations=15, test_objective=1, tolerance_objective=-1e300, tolerance=1e9)
    sys2, _ = both(scene2)
    assert sys2.GetCounts().solver_iterations == 15
    # ... except by APGDREF
    scene3 = bed(8, 8, 4, iterations=15, solver_type=1, test_objective=1, tolerance_objective=-1e300, tolerance=1e9)
    sys3, orc3 = both(scene3)
    assert sys3.GetCounts().solver_iterations == orc3.nsc_info()["iterations"] < 15


def test_spinning_mode_solve():
    scene = bed(6, 6, 3, mode=cw.system.MODE_SPINNING, iterations=10, max_iteration_normal=5,
                max_iteration_spinning=10)
    scene["materials"] = np.array([scenes.nsc_material(friction=0.4, rolling=0.02, spinning=0.01)])
    sys, orc = both(scene)
    assert sys.GetCounts().solver_iterations == orc.nsc_info()["iterations"] == 25
    assert relerr(sys.GetGamma(), orc.nsc_vectors()["gamma"]) < 1e-8
    _, _, vel = sys.GetBodyState()
    assert relerr(vel, orc.state()[2]) < 1e-8


def test_steps_track_oracle_and_contact_forces_balance_weight():
    """utest_MCORE_contact_forces.cpp:154-177 (NSC case): once settled, the contact force on the container equals
    the weight of the balls (rel 1e-3 in the reference; the bed here is still creeping, so 5 %)."""
    scene = scenes.balls_nsc(8, 8, 3, iterations=60)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    orc = pyoracle.Oracle(scene)
    sys.DoSteps(5)
    orc.step(5)
    p, q, v = sys.GetBodyState()
    po, qo, vo = orc.state()
    assert relerr(p, po) < 1e-7
    assert relerr(v, vo) < 1e-6
    sys.DoSteps(200)
    f, t = sys.GetBodyContactForces()
    weight = 9.81 * (scene["mass"].size - 1)
    assert abs(f[0, 2] + weight) < 0.05 * weight
    assert sys.KernelLaunches() > 0


def test_slot_budget_growth():
    """A solve that back-tracks more often than the initial slot budget allows is re-run with a larger one."""
    scene = bed(5, 5, 2, iterations=4)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    sys.DoSteps(3)
    assert sys.GetCounts().solver_iterations == 4


@pytest.mark.parametrize("method", ["nsc", "smc"])
def test_reference_contact_force_acceptance(method):
    """utest_MCORE_contact_forces.cpp:154-177 on the reference's own scene and settings: from t = 0.5 s to t = 3 s the
    contact force on the container equals the total weight of the balls within rel 1e-3 (the reference's tolerance)."""
    scene = scenes.contact_force_test(method)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC() if method == "nsc" else cw.ChSystemMulticoreSMC(), scene)
    weight = 9.81 * scene["mass"][1:].sum()
    worst, t = 0.0, 0.0
    while t < 3.0 - 1e-9:
        sys.DoSteps(20)
        t += 20e-3
        if t > 0.5:
            f, _ = sys.GetBodyContactForces()
            worst = max(worst, abs(1 - (-f[0, 2]) / weight))
    assert sys.GetNumContacts() == 8
    assert worst < 1e-3, worst
