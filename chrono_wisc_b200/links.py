"""Host-side generators of bilateral rows for ``ChSystemMulticore.SetBilaterals``.

In Chrono the rows come out of the link classes (``ChLinkMotorRotationSpeed``, ``ChLinkLockRevolute`` ... fill a
``ChConstraintTwoBodies`` each; ``ChConstraintBilateral::Build_D`` copies ``Get_Cq_a / Get_Cq_b``,
src/chrono_multicore/constraints/ChConstraintBilateral.cpp:46-78).  That evaluation is per-joint host work outside the
hot path; the C++ shim forwards what Chrono computed.  This module is the stand-in used by the Python mirror, the tests
and ``bench.py --config mixed --blade``: enough to drive demo_MCORE_mixerSMC's blade (a revolute joint about a fixed
axis with an imposed angular speed, src/demos/multicore/demo_MCORE_mixerSMC.cpp:80-110).
"""
import numpy as np


def rotmat(q):
    e0, e1, e2, e3 = q
    return np.array([[(e0 * e0 + e1 * e1) * 2 - 1, (e1 * e2 - e0 * e3) * 2, (e1 * e3 + e0 * e2) * 2],
                     [(e1 * e2 + e0 * e3) * 2, (e0 * e0 + e2 * e2) * 2 - 1, (e2 * e3 - e0 * e1) * 2],
                     [(e1 * e3 - e0 * e2) * 2, (e2 * e3 + e0 * e1) * 2, (e0 * e0 + e3 * e3) * 2 - 1]])


def motor_rotation_speed_z(ground, body, pos, rot, anchor, omega, h, recovery_clamp=0.6):
    """Six rows holding `body`'s centre at `anchor` and its z axis along the world z axis while imposing the angular
    speed `omega` about it (ChLinkMotorRotationSpeed between `ground`, a fixed body, and `body`).  Velocities are
    (v in the absolute frame, w in the body frame), as host_data.v holds them; the position-level violations enter the
    right-hand side as C / h, clamped like ChConstraint's recovery speed.  Returns (ids, cq_a, cq_b, rhs)."""
    R = rotmat(rot)
    ids = np.tile([ground, body], (6, 1)).astype(np.int32)
    cq_a, cq_b, rhs = np.zeros((6, 6)), np.zeros((6, 6)), np.zeros(6)
    err = np.asarray(pos, float) - np.asarray(anchor, float)
    for k in range(3):  # v_body = 0, drift pulled back
        cq_b[k, k] = 1.0
        cq_a[k, k] = -1.0
        rhs[k] = np.clip(err[k] / h, -recovery_clamp, recovery_clamp)
    zb = R[:, 2]
    tilt = np.array([-zb[1], zb[0]])  # - (zb x ez) . (ex, ey): the rotation that brings zb back to ez, as a rate
    for k, axis in enumerate(np.eye(3)):
        cq_b[3 + k, 3:] = R.T @ axis  # world angular velocity component = (R^T a) . w_body
        cq_a[3 + k, 3:] = -axis       # the ground body is not rotated
    rhs[3:5] = np.clip(tilt / h, -recovery_clamp, recovery_clamp)
    rhs[5] = -omega
    return ids, cq_a, cq_b, rhs
