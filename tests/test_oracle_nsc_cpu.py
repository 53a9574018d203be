"""CPU checks of the NSC oracle (oracle/oracle_nsc.inc) -- no GPU.

Pins what the reference's own code can pin here: the contact frame and Jacobian rows against the reference's
Orthogonalize / Compute_Jacobian (the comparison utest_MCORE_jacobians.cpp:126-212 makes) and the cone projection
against Cone_generalized_rigid, both extracted from /root/reference into oracle/_ref.  The products themselves go
through Eigen in the reference (absent here): parity unpinned at the bit level, checked through properties.
"""
import numpy as np
import pytest

from chrono_wisc_b200 import scenes
from chrono_wisc_b200.system import MODE_NORMAL, MODE_SLIDING, MODE_SPINNING
from oracle import pyoracle


def bed(nx=5, ny=5, nz=3, seed=2, mode=MODE_SLIDING, iterations=20, **st):
    s = scenes.balls_nsc(nx, ny, nz, seed=seed, iterations=iterations)
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["pos"][1:, 2] = 0.15 + (s["pos"][1:, 2] - 0.15) * 0.97
    s["vel"][1:, :3] = rng.normal(scale=0.3, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=1.0, size=(n - 1, 3))
    q = rng.normal(size=(n - 1, 4))
    s["rot"][1:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    s["settings"].update(solver_mode=mode, **st)
    return s


def solved(scene):
    o = pyoracle.Oracle(scene)
    o.collide()
    o.nsc_solve()
    return o


def test_jacobian_rows_equal_reference_functions(ref):
    scene = bed()
    o = solved(scene)
    c = o.contacts()
    nc = c["depth"].size
    assert nc > 100
    rows, bids = o.nsc_jacobian_rows()
    for i in range(0, nc, 7):
        b1, b2 = c["body_ids"][i]
        U = c["normal"][i]
        V, W = ref.orthogonalize(U)
        TA = ref.compute_jacobian(scene["rot"][b1], U, V, W, c["pt_a"][i] - scene["pos"][b1])
        TB = ref.compute_jacobian(scene["rot"][b2], U, V, W, c["pt_b"][i] - scene["pos"][b2])
        for k, d in enumerate((U, V, W)):
            r = i if k == 0 else nc + 2 * i + (k - 1)
            blkA = np.concatenate([-d, TA[k]])
            blkB = np.concatenate([d, -TB[k]])
            want = np.concatenate([blkA, blkB]) if b1 < b2 else np.concatenate([blkB, blkA])
            assert np.array_equal(rows[r], want), (i, k)
            assert tuple(bids[r]) == (min(b1, b2), max(b1, b2))


def test_projection_equals_reference_cone(ref):
    scene = bed()
    scene["materials"] = np.array([scenes.nsc_material(friction=0.35, cohesion=0.03)])
    o = solved(scene)
    nc = o.counts()["N_c"]
    rng = np.random.default_rng(1)
    g = rng.normal(size=3 * nc) * 2
    g[:5] = 0.0  # |gamma_n| < 1e-14 branch
    p = o.nsc_project(g, MODE_SLIDING)
    mu, coh = np.float32(0.35), np.float32(0.03)
    for i in range(nc):
        _, (a, b, c) = ref.cone_generalized(g[i] + float(coh), g[nc + 2 * i], g[nc + 2 * i + 1], float(mu))
        assert (p[i], p[nc + 2 * i], p[nc + 2 * i + 1]) == (a - float(coh), b, c)
    pn = o.nsc_project(g, MODE_NORMAL)
    assert np.array_equal(pn[:nc], np.maximum(g[:nc], -float(coh)))
    assert not pn[nc:].any()


@pytest.mark.parametrize("mode", [MODE_NORMAL, MODE_SLIDING, MODE_SPINNING])
def test_schur_operator_is_symmetric_positive_semidefinite(mode):
    extra = {MODE_NORMAL: dict(max_iteration_normal=3), MODE_SLIDING: {}, MODE_SPINNING: dict(max_iteration_spinning=3)}
    o = solved(bed(4, 4, 2, mode=mode, iterations=3, **extra[mode]))
    n = o.nsc_rows() * o.counts()["N_c"]
    rng = np.random.default_rng(3)
    x, y = rng.normal(size=n), rng.normal(size=n)
    Nx, Ny = o.nsc_schur(x), o.nsc_schur(y)
    assert abs(y @ Nx - x @ Ny) < 1e-9 * abs(y @ Nx)
    assert x @ Nx >= 0


def test_apgd_residual_history_is_monotone_and_solution_feasible():
    scene = bed(6, 6, 3, iterations=60)
    o = solved(scene)
    r, obj = o.nsc_hist()
    assert r.size == 60 and (np.diff(r) <= 0).all()  # the recorded residual is the best so far
    g = o.nsc_vectors()["gamma"]
    nc = o.counts()["N_c"]
    ft = np.hypot(g[nc::2], g[nc + 1::2])
    assert (g[:nc] >= 0).all() and (ft <= float(np.float32(0.4)) * g[:nc] * (1 + 1e-9) + 1e-15).all()


def test_apgdref_equals_apgd_iterates():
    a = solved(bed(5, 5, 3, iterations=25, solver_type=0))
    b = solved(bed(5, 5, 3, iterations=25, solver_type=1))
    assert np.allclose(a.nsc_vectors()["gamma"], b.nsc_vectors()["gamma"], rtol=1e-12, atol=1e-14)


def test_nsc_steps_settle_and_hold_weight():
    """utest_MCORE_contact_forces.cpp:154-177, NSC case, on the oracle."""
    scene = scenes.balls_nsc(6, 6, 2, iterations=80)
    o = pyoracle.Oracle(scene)
    o.step(250)
    v = o.nsc_vectors()
    c = o.contacts()
    nc = c["depth"].size
    h = scene["settings"]["step_size"]
    onwall = c["body_ids"][:, 0] == 0
    F = -(v["gamma"][:nc][onwall, None] * c["normal"][onwall]).sum(0) / h
    weight = 9.81 * (scene["mass"].size - 1)
    assert abs(F[2] + weight) < 0.03 * weight


def test_omp_threads_do_not_change_nsc_results():
    scene = bed(5, 5, 3, iterations=15)
    a = pyoracle.Oracle(scene, threads=1)
    a.collide(), a.nsc_solve()
    b = pyoracle.Oracle(scene, threads=4)
    b.collide(), b.nsc_solve()
    assert np.array_equal(a.nsc_vectors()["gamma"], b.nsc_vectors()["gamma"])
    assert np.array_equal(a.state()[2], b.state()[2])
