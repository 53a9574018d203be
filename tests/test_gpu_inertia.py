"""Products of inertia (use_full_inertia_tensor, MC/ChSettings.h:107; ComputeInvMassMatrix,
MC/solver/ChIterativeSolverMulticore.cpp:61-147; ChBody::ComputeGyro) through cb200_set_inertia_full.

The check needs no second implementation: a body described in a body frame turned by r against its principal axes has
inertia R(r) J_d R(r)^T, orientation q (instead of q r) and local angular velocity R(r) w -- and is the same physical
body.  A bed of such bodies must move exactly as the bed of principal-axes bodies does (spheres: the contact geometry
does not see the frame), through contacts with friction, the gyroscopic term, the velocity update and the quaternion
step, for SMC and NSC."""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from chrono_wisc_b200._lib import Cb200Error

pytestmark = pytest.mark.gpu


def qmul(a, b):
    a0, a1, a2, a3 = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    b0, b1, b2, b3 = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([a0 * b0 - a1 * b1 - a2 * b2 - a3 * b3, a0 * b1 + a1 * b0 + a2 * b3 - a3 * b2,
                     a0 * b2 + a2 * b0 + a3 * b1 - a1 * b3, a0 * b3 + a3 * b0 + a1 * b2 - a2 * b1], -1)


def rotmat(q):
    e0, e1, e2, e3 = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    return np.stack([np.stack([1 - 2 * (e2 * e2 + e3 * e3), 2 * (e1 * e2 - e0 * e3), 2 * (e1 * e3 + e0 * e2)], -1),
                     np.stack([2 * (e1 * e2 + e0 * e3), 1 - 2 * (e1 * e1 + e3 * e3), 2 * (e2 * e3 - e0 * e1)], -1),
                     np.stack([2 * (e1 * e3 - e0 * e2), 2 * (e2 * e3 + e0 * e1), 1 - 2 * (e1 * e1 + e2 * e2)], -1)], -2)


def twin_scenes(scene, seed):
    """(principal-axes scene, turned-frame scene, r, J_full) for the dynamic bodies 1.. of `scene`."""
    rng = np.random.default_rng(seed)
    n = scene["mass"].size - 1
    A = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in scene.items()}
    I0 = A["inertia"][1:, :1]
    A["inertia"][1:] = I0 * np.array([0.6, 1.0, 1.7])  # three distinct principal moments
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    r = rng.normal(size=(n, 4))
    r /= np.linalg.norm(r, axis=1, keepdims=True)
    w = rng.normal(scale=3.0, size=(n, 3))
    A["rot"][1:] = qmul(q, r)
    A["vel"][1:, 3:] = w
    A["vel"][1:, :3] = rng.normal(scale=0.1, size=(n, 3))
    B = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in A.items()}
    R = rotmat(r)
    B["rot"][1:] = q
    B["vel"][1:, 3:] = np.einsum("nij,nj->ni", R, w)
    J = np.einsum("nij,nj,nkj->nik", R, A["inertia"][1:], R)
    J = 0.5 * (J + J.transpose(0, 2, 1))
    return A, B, r, J


def run_pair(cls, scene, seed, steps):
    A, B, r, J = twin_scenes(scene, seed)
    sa = cw.load_scene(cls(), A)
    sb = cw.load_scene(cls(), B)
    sb.SetInertiaFull(1, J)
    sa.DoSteps(steps)
    sb.DoSteps(steps)
    sa.Synchronize(), sb.Synchronize()
    pa, qa, va = sa.GetBodyState()
    pb, qb, vb = sb.GetBodyState()
    assert sa.GetNumContacts() == sb.GetNumContacts() > 0
    return (pa[1:], qa[1:], va[1:]), (pb[1:], qb[1:], vb[1:]), r


def compare(a, b, r, tol):
    (pa, qa, va), (pb, qb, vb) = a, b
    scale = np.abs(pa).max()
    assert np.abs(pa - pb).max() < tol * scale
    assert np.abs(va[:, :3] - vb[:, :3]).max() < tol * max(np.abs(va[:, :3]).max(), 1e-3)
    # orientation of the principal axes in the world, and the angular velocity in the world
    Ra, Rb = rotmat(qa), np.einsum("nij,njk->nik", rotmat(qb), rotmat(r))
    assert np.abs(Ra - Rb).max() < tol
    wa = np.einsum("nij,nj->ni", rotmat(qa), va[:, 3:])
    wb = np.einsum("nij,nj->ni", rotmat(qb), vb[:, 3:])
    assert np.abs(wa - wb).max() < tol * np.abs(wa).max()
    assert np.abs(wa).max() > 0.5  # (the bodies do spin: the test is not vacuous)


def test_smc_bed_with_turned_body_frames():
    scene = scenes.settling_smc(8, 8, 4, spacing_factor=1.98, jitter_factor=0.03, seed=21)
    a, b, r = run_pair(cw.ChSystemMulticoreSMC, scene, 3, 40)
    compare(a, b, r, 1e-9)


def test_nsc_bed_with_turned_body_frames():
    scene = scenes.balls_nsc(6, 6, 3, seed=4, iterations=30)
    a, b, r = run_pair(cw.ChSystemMulticoreNSC, scene, 5, 5)
    compare(a, b, r, 1e-7)


def test_gyroscopic_term_sees_the_products_of_inertia():
    """One free body, no gravity: after one step w' = w - h J^-1 (w x J w) (UpdateForces + ComputeGyro + the velocity
    update), evaluated here with the full matrices."""
    rng = np.random.default_rng(9)
    A = rng.normal(size=(3, 3))
    J = A @ A.T + 3 * np.eye(3)
    w = np.array([1.0, -2.0, 0.5])
    h = 1e-3
    for cls, mat in ((cw.ChSystemMulticoreSMC, scenes.smc_material()), (cw.ChSystemMulticoreNSC, scenes.nsc_material())):
        s = cls()
        s.SetStep(h)
        s.SetGravitationalAcceleration([0, 0, 0])
        s.AddBodies([2.0], [[J[0, 0], J[1, 1], J[2, 2]]], [[0, 0, 0]], vel=[[0, 0, 0, *w]])
        s.AddShapes([0], [0], [[0.1, 0, 0, 0]])
        s.SetMaterials(np.array([mat]))
        s.SetInertiaFull(0, [J])
        s.DoSteps(1)
        s.Synchronize()
        _, _, v = s.GetBodyState()
        expect = w - h * np.linalg.solve(J, np.cross(w, J @ w))
        assert np.allclose(v[0, 3:], expect, rtol=1e-12, atol=1e-15)
        assert not np.allclose(v[0, 3:], w - h * np.cross(w, np.diag(J) * w) / np.diag(J), rtol=1e-6)


def test_refusals():
    s = cw.ChSystemMulticoreSMC()
    s.AddBodies([1.0, 1.0], [[1, 1, 1], [1, 1, 1]], [[0, 0, 0], [1, 0, 0]])
    J = np.eye(3)
    J[0, 1] = 0.1  # not symmetric
    with pytest.raises(Cb200Error, match="symmetric"):
        s.SetInertiaFull(0, [J], [np.eye(3)])
    with pytest.raises(Cb200Error, match="bad argument"):
        s.SetInertiaFull(1, [np.eye(3), np.eye(3)])
