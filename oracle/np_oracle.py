"""numpy restatement of the glue around the reference's collision/force kernels -- TEST INFRASTRUCTURE.

These are the parts of the step that live in Eigen/ChBody-dependent translation units which cannot be compiled
here: composite materials, body force gathering, per-body accumulation of contact forces, the
velocity update and the position/quaternion integration.  Every function cites the reference lines it follows.
Element-wise IEEE double arithmetic in numpy rounds exactly like the scalar C++ it restates (no FMA), so results
are comparable to the last bit except where libm functions (sin/cos) enter.
"""
import numpy as np

F32 = np.float32


def composite_smc(mat_table, mat_a, mat_b):
    """ChContactMaterialCompositeSMC ctor, float arithmetic (src/chrono/physics/ChContactMaterialSMC.cpp:107-130)
    with the default combination strategy (src/chrono/physics/ChContactMaterial.h:157-169); stored widened to
    double as in ChContactContainerMulticoreSMC::AddContact (ChContactContainerMulticoreSMC.cpp:166-173)."""
    t = np.asarray(mat_table, dtype=F32)
    m1, m2 = t[np.asarray(mat_a)], t[np.asarray(mat_b)]
    one, two = F32(1), F32(2)
    E1, nu1, E2, nu2 = m1[:, 0], m1[:, 1], m2[:, 0], m2[:, 1]
    inv_E = (one - nu1 * nu1) / E1 + (one - nu2 * nu2) / E2
    inv_G = two * (two - nu1) * (one + nu1) / E1 + two * (two - nu2) * (one + nu2) / E2
    E_eff, G_eff = one / inv_E, one / inv_G
    mn = np.minimum
    half = lambda a, b: ((a + b) / two).astype(F32)
    return dict(
        fric=np.stack([mn(m1[:, 2], m2[:, 2]), mn(m1[:, 4], m2[:, 4]), mn(m1[:, 5], m2[:, 5])], 1).astype(np.float64),
        modulus=np.stack([E_eff, G_eff], 1).astype(np.float64),
        adhesion=np.stack([mn(m1[:, 7], m2[:, 7]), mn(m1[:, 8], m2[:, 8]), mn(m1[:, 9], m2[:, 9])], 1).astype(np.float64),
        cr=mn(m1[:, 6], m2[:, 6]).astype(np.float64),
        smc=np.stack([half(m1[:, 10], m2[:, 10]), half(m1[:, 11], m2[:, 11]), half(m1[:, 12], m2[:, 12]),
                      half(m1[:, 13], m2[:, 13])], 1).astype(np.float64))


def cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], 1)


def update_forces(mass, inertia, vel, gravity, h, applied=None):
    """UpdateRigidBodies force part: ChBody::UpdateForces + ComputeGyro + VariablesFbLoadForces
    (src/chrono/physics/ChBody.cpp:552-574,423-426,247-257): hf = h * (m g + F, T - w x (J w))."""
    w = vel[:, 3:6]
    gyro = cross(w, inertia * w)
    F = np.asarray(gravity)[None, :] * mass[:, None]
    T = np.zeros_like(F)
    if applied is not None:
        F = F + applied[:, :3]
        T = T + applied[:, 3:]
    return np.concatenate([h * F, h * (T - gyro)], 1)


def accumulate_contact_forces(nb, ct_bid, ct_force, ct_torque):
    """ProcessContacts steps 2-3 (ChIterativeSolverMulticoreSMC.cpp:696-736): stable sort by body, sum per body in
    contact order.  Returns dense per-body (force, torque) and the contact count per body."""
    F, T, cnt = np.zeros((nb, 3)), np.zeros((nb, 3)), np.zeros(nb, dtype=np.int64)
    order = np.argsort(ct_bid, kind="stable")
    first = np.ones(nb, dtype=bool)
    for k in order:  # sequential: defines the summation order
        b = ct_bid[k]
        if first[b]:
            F[b], T[b] = ct_force[k], ct_torque[k]
            first[b] = False
        else:
            F[b] = F[b] + ct_force[k]
            T[b] = T[b] + ct_torque[k]
        cnt[b] += 1
    return F, T, cnt


def accumulate_fast(nb, ct_bid, ct_force, ct_torque):
    """Same sums without a python loop (np.add.at adds in index order per target): used for large cases where
    agreement is checked to a tolerance rather than to the bit."""
    F, T = np.zeros((nb, 3)), np.zeros((nb, 3))
    order = np.argsort(ct_bid, kind="stable")
    np.add.at(F, ct_bid[order], ct_force[order])
    np.add.at(T, ct_bid[order], ct_torque[order])
    cnt = np.bincount(ct_bid, minlength=nb)
    return F, T, cnt


def velocity_update(vel, hf, mass, inv_inertia, active):
    """ComputeInvMassMatrix + ComputeImpulses without bilaterals (ChIterativeSolverMulticore.cpp:61-147,
    ChIterativeSolverMulticoreSMC.cpp:854-868): v = v + M^-1 hf on active bodies; inactive bodies have no rows."""
    minv = np.concatenate([np.repeat((1.0 / mass)[:, None], 3, 1), inv_inertia], 1)
    out = vel + minv * hf
    out[~np.asarray(active, dtype=bool)] = vel[~np.asarray(active, dtype=bool)]
    return out


def rotmat(q):
    """ChMatrix33::SetFromQuaternion (src/chrono/core/ChMatrix33.h:253-274), q = (e0,e1,e2,e3)."""
    e0, e1, e2, e3 = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.empty((q.shape[0], 3, 3))
    R[:, 0, 0] = (e0 * e0 + e1 * e1) * 2 - 1
    R[:, 0, 1] = (e1 * e2 - e0 * e3) * 2
    R[:, 0, 2] = (e1 * e3 + e0 * e2) * 2
    R[:, 1, 0] = (e1 * e2 + e0 * e3) * 2
    R[:, 1, 1] = (e0 * e0 + e2 * e2) * 2 - 1
    R[:, 1, 2] = (e2 * e3 - e0 * e1) * 2
    R[:, 2, 0] = (e1 * e3 - e0 * e2) * 2
    R[:, 2, 1] = (e2 * e3 + e0 * e1) * 2
    R[:, 2, 2] = (e0 * e0 + e3 * e3) * 2 - 1
    return R


def integrate(pos, rot, vel, h, active):
    """ChBody::VariablesQbIncrementPosition (src/chrono/physics/ChBody.cpp:288-310) on active bodies."""
    act = np.asarray(active, dtype=bool)
    p = pos + vel[:, :3] * h
    R = rotmat(rot)
    w = vel[:, 3:6]
    wa = np.stack([R[:, i, 0] * w[:, 0] + R[:, i, 1] * w[:, 1] + R[:, i, 2] * w[:, 2] for i in range(3)], 1)
    ln = np.sqrt(wa[:, 0] * wa[:, 0] + wa[:, 1] * wa[:, 1] + wa[:, 2] * wa[:, 2])
    angle = ln * h
    tiny = ln < np.finfo(np.float64).tiny
    with np.errstate(divide="ignore", invalid="ignore"):
        s = 1 / ln
        axis = wa * s[:, None]
    axis[tiny] = [1.0, 0.0, 0.0]
    half = angle / 2
    sh = np.sin(half)
    d = np.stack([np.cos(half), axis[:, 0] * sh, axis[:, 1] * sh, axis[:, 2] * sh], 1)
    a, b = d, rot  # ChQuaternion::Cross(a, b) (ChQuaternion.h:891-903)
    q = np.stack([
        a[:, 0] * b[:, 0] - a[:, 1] * b[:, 1] - a[:, 2] * b[:, 2] - a[:, 3] * b[:, 3],
        a[:, 0] * b[:, 1] + a[:, 1] * b[:, 0] - a[:, 3] * b[:, 2] + a[:, 2] * b[:, 3],
        a[:, 0] * b[:, 2] + a[:, 2] * b[:, 0] + a[:, 3] * b[:, 1] - a[:, 1] * b[:, 3],
        a[:, 0] * b[:, 3] + a[:, 3] * b[:, 0] - a[:, 2] * b[:, 1] + a[:, 1] * b[:, 2]], 1)
    p[~act] = pos[~act]
    q[~act] = rot[~act]
    return p, q


def smc_step(scene, ref, pos, rot, vel, fast=True, history=None):
    """One full SMC step on the CPU: reference collision + reference force function + the glue above.
    Returns (pos, rot, vel, info)."""
    st = scene["settings"]
    h = st["step_size"]
    nb = scene["mass"].size
    active = scene["active"].astype(bool)
    inv_inertia = 1.0 / scene["inertia"]
    hf = update_forces(scene["mass"], scene["inertia"], vel, st["gravity"], h)
    ref.set_bodies(pos, rot, scene["active"], scene["collide"])
    ref.run(7)
    c = ref.contacts()
    nc = c["depth"].size
    F = T = np.zeros((nb, 3))
    if nc:
        sa = (c["shape_ids"] >> 32).astype(np.int64)
        sb = (c["shape_ids"] & 0xFFFFFFFF).astype(np.int64)
        mats = composite_smc(scene["materials"], scene["shape_mat"][sa], scene["shape_mat"][sb])
        bid, cf, ct = ref.smc_contact_forces(c, mats, st, scene["mass"], pos, rot, vel, history=history)
        acc = accumulate_fast if fast else accumulate_contact_forces
        F, T, cnt = acc(nb, bid, cf, ct)
        touched = cnt > 0
        hf[touched, :3] += (h * F)[touched]
        hf[touched, 3:] += (h * T)[touched]
    elif history is not None:
        history["neigh"][:, 0] = -1  # no contact touched any slot
    v2 = velocity_update(vel, hf, scene["mass"], inv_inertia, active)
    p2, q2 = integrate(pos, rot, v2, h, active)
    return p2, q2, v2, dict(contacts=c, hf=hf, body_force=F, body_torque=T)


def contact_plane(normal):
    """ChMatrix33::SetFromAxisX(normal, VECT_Y) (src/chrono/core/ChMatrix33.h:389-395 over
    ChVector3::GetDirectionAxesAsX, ChVector3.h:899-921): columns X (the normal), Y, Z."""
    tiny = np.finfo(float).tiny

    def normalize(v):
        n = np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
        if n < tiny:
            return np.array([1.0, 0.0, 0.0]), False
        return v * (1.0 / n), True

    X, ok = normalize(np.asarray(normal, float))
    sugg = np.array([0.0, 1.0, 0.0])
    Z, ok = normalize(np.cross(X, sugg))
    idx = 0
    while not ok:
        sugg[idx] += 1.0
        Z, ok = normalize(np.cross(X, sugg))
        idx += 1
    Y = np.cross(Z, X)
    return np.stack([X, Y, Z], axis=1)


def report_contacts(method, contacts, pos, rot, h, ct_force=None, ct_torque=None, gamma=None, rows=3):
    """ChContactContainerMulticore::ReportAllContacts (MC/collision/ChContactContainerMulticore.cpp:65-149): the
    arguments of OnReportContact per contact.  `gamma` in the reference's layout (normals | u v pairs | spinning)."""
    n = contacts["depth"].size
    planes, F, T = np.empty((n, 3, 3)), np.zeros((n, 3)), np.zeros((n, 3))
    for i in range(n):
        P = contact_plane(contacts["normal"][i])
        planes[i] = P
        if method == "nsc":
            F[i, 0] = gamma[i] / h
            if rows >= 3:
                F[i, 1], F[i, 2] = gamma[n + 2 * i] / h, gamma[n + 2 * i + 1] / h
            if rows >= 6:
                T[i] = gamma[3 * n + 3 * i:3 * n + 3 * i + 3] / h
        else:
            b = contacts["body_ids"][i, 1]
            R = rotmat(np.asarray(rot[b], float)[None])[0]
            f_abs, t_loc = ct_force[2 * i + 1], ct_torque[2 * i + 1]
            f_loc = R.T @ f_abs
            pB_loc = R.T @ (contacts["pt_b"][i] - pos[b])
            t_abs = R @ (t_loc - np.cross(pB_loc, f_loc))
            F[i], T[i] = P.T @ f_abs, P.T @ t_abs
    return dict(pA=contacts["pt_a"], pB=contacts["pt_b"], plane=planes, distance=contacts["depth"],
                eff_radius=contacts["eff_radius"], force=F, torque=T, body_ids=contacts["body_ids"])


def minres(N, mb, max_iter, tol):
    """ChSolverMulticoreMinRes::Solve (MC/solver/ChSolverMulticoreMINRES.cpp:19-112) on an explicit matrix, x0 = 0.
    Returns (x, iterations, residual)."""
    n = mb.size
    x = np.zeros(n)
    v_hat = mb - N @ x
    beta = np.linalg.norm(v_hat)
    c, eta, norm_rMR, norm_r0, c_old, s_old, s = 1.0, beta, beta, beta, 1.0, 0.0, 0.0
    v, w, w_old = np.zeros(n), np.zeros(n), np.zeros(n)
    if beta == 0 or norm_rMR / norm_r0 < tol:
        return x, 0, (0.0 if beta == 0 else 1.0)
    residual, it = 1.0, 0
    for it in range(max_iter):
        v_old = v
        v = 1.0 / beta * v_hat
        Av = N @ v
        alpha = v @ Av
        v_hat = Av - alpha * v
        v_hat = v_hat - beta * v_old
        beta_old = beta
        beta = np.linalg.norm(v_hat)
        c_oold, c_old, s_oold, s_old = c_old, c, s_old, s
        r1_hat = c_old * alpha - c_oold * s_old * beta_old
        r1 = 1 / np.sqrt(r1_hat * r1_hat + beta * beta)
        r2 = s_old * alpha + c_oold * c_old * beta_old
        r3 = s_oold * beta_old
        c, s = r1_hat * r1, beta * r1
        w_oold, w_old = w_old, w
        w = (v - r3 * w_oold - r2 * w_old) * r1
        x = x + c * eta * w
        norm_rMR = norm_rMR * abs(s)
        eta = -s * eta
        residual = norm_rMR / norm_r0
        if residual < tol:
            break
    else:
        it = max_iter
    return x, it, residual


def bilateral_solve(ids, cq_a, cq_b, rhs, vel, hf, mass, inv_inertia, active, max_iter=100, tol=1e-4):
    """ChIterativeSolverMulticoreSMC::RunTimeStep :808-845 for BODY_BODY rows: M_invk = v + M^-1 hf (active bodies),
    R = -b - D^T M_invk (ComputeR :772-783), N = D^T M^-1 D (ChSchurProductBilateral), MINRES (PerformStabilization),
    v = M_invk + M^-1 D gamma (ComputeImpulses :854-868).  Returns (v, gamma, iterations, residual)."""
    nb, n = mass.size, rhs.size
    act = np.asarray(active, dtype=bool)
    minv = np.concatenate([np.repeat(1.0 / mass[:, None], 3, 1), inv_inertia.reshape(nb, 3)], 1) * act[:, None]
    M_invk = vel + minv * hf
    D = np.zeros((n, 6 * nb))
    for i in range(n):
        D[i, 6 * ids[i, 0]:6 * ids[i, 0] + 6] += cq_a[i]
        D[i, 6 * ids[i, 1]:6 * ids[i, 1] + 6] += cq_b[i]
    Minv = minv.reshape(-1)
    R = -rhs - D @ M_invk.reshape(-1)
    N = (D * Minv) @ D.T
    gamma, it, res = minres(N, R, max_iter, tol)
    v = M_invk.reshape(-1) + Minv * (D.T @ gamma)
    return v.reshape(nb, 6), gamma, it, res
