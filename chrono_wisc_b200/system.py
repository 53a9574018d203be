"""Host-side mirror of the reference's system interface for the granular-contact hot path.

``ChSystemMulticoreSMC`` / ``ChSystemMulticoreNSC`` keep the names and meaning of the reference classes
(src/chrono_multicore/physics/ChSystemMulticore.h:52-267): ``GetSettings()``, ``SetGravitationalAcceleration``,
``DoStepDynamics``, ``GetNumContacts``, ``GetBodyContactForce``, ``GetTimerCollisionBroad`` ... but bodies and
shapes are added in bulk as arrays (the C++ shim in INTEGRATION.md does the per-``ChBody`` gather).  All state
lives on the B200; every method is a thin call into the C ABI (include/chrono_b200.h).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import cb200_counts, cb200_settings, check

# ChCollisionShape::Type values (src/chrono/collision/ChCollisionShape.h:31-53)
SPHERE, ELLIPSOID, BOX, CYLINDER, CYLSHELL, CONVEXHULL = 0, 1, 2, 3, 4, 5
TRIANGLE, CAPSULE, CONE, ROUNDEDBOX, ROUNDEDCYL = 10, 12, 13, 14, 15

# ChBroadphase::GridType, ChNarrowphase::Algorithm
FIXED_RESOLUTION, FIXED_BIN_SIZE, FIXED_DENSITY = 0, 1, 2
MPR, PRIMS, HYBRID = 0, 1, 2
# ChSystemSMC enums
HOOKE, HERTZ, PLAIN_COULOMB, FLORES = 0, 1, 2, 3
ADHESION_CONSTANT, ADHESION_DMT, ADHESION_PERKO = 0, 1, 2
DISPL_NONE, DISPL_ONE_STEP, DISPL_MULTI_STEP = 0, 1, 2
# SolverMode / SolverType
MODE_NORMAL, MODE_SLIDING, MODE_SPINNING = 0, 1, 2
SOLVER_APGD, SOLVER_APGDREF = 0, 1


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# composite material records handed to an AddContactCallback (ChContactMaterialCompositeSMC / NSC members)
COMPOSITE_FIELDS_SMC = ("mu_eff", "muRoll_eff", "muSpin_eff", "E_eff", "G_eff", "adhesion_eff", "adhesionMultDMT_eff",
                        "adhesionSPerko_eff", "cr_eff", "kn", "kt", "gn", "gt")
COMPOSITE_FIELDS_NSC = ("sliding_friction", "rolling_friction", "spinning_friction", "cohesion", "compliance",
                        "complianceT", "complianceRoll", "complianceSpin")


class ChSystemMulticore:
    """Common part (reference: ChSystemMulticore, src/chrono_multicore/physics/ChSystemMulticore.{h,cpp})."""

    system_type = 0

    def __init__(self, device=0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        check(self._lib.cb200_create(int(device), C.byref(self._h)))
        self.device = int(device)
        self.settings = cb200_settings()
        self._lib.cb200_default_settings(C.byref(self.settings))
        self.settings.system_type = self.system_type
        self.num_bodies = 0
        self.num_shapes = 0
        self.ch_time = 0.0
        self._narrow_callback = None
        self._add_contact_callback = None

    def __del__(self):
        try:
            if getattr(self, "_h", None) and self._h.value:
                self._lib.cb200_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ---- configuration -------------------------------------------------------------------------------------
    def GetSettings(self):
        """``sys.GetSettings()`` (ChSystemMulticore.cpp:684-686); pushed to the device at every step."""
        return self.settings

    def SetGravitationalAcceleration(self, g):
        self.settings.gravity[:] = [float(g[0]), float(g[1]), float(g[2])]

    def SetStep(self, h):
        self.settings.step_size = float(h)

    def _push_settings(self):
        check(self._lib.cb200_set_settings(self._h, C.byref(self.settings)))

    # ---- model building ------------------------------------------------------------------------------------
    def AddBodies(self, mass, inertia_xx, pos, rot=None, vel=None, active=None, collide=None, inv_inertia_xx=None):
        """Bulk ``AddBody`` (ChSystemMulticore.cpp:205-226). Returns the index of the first new body."""
        mass = _f64(mass).ravel()
        n = mass.size
        inertia_xx = _f64(inertia_xx, (n, 3))
        inv = None if inv_inertia_xx is None else _f64(inv_inertia_xx, (n, 3))
        act = None if active is None else np.ascontiguousarray(active, dtype=np.int8)
        col = None if collide is None else np.ascontiguousarray(collide, dtype=np.int8)
        first = C.c_int(0)
        check(self._lib.cb200_append_bodies(self._h, n, _ptr(mass), _ptr(inertia_xx), _ptr(inv), _ptr(act), _ptr(col),
                                            C.byref(first)))
        self.num_bodies += n
        self.SetBodyState(first.value, pos, rot, vel, n)
        return first.value

    def SetInertiaFull(self, first, inertia, inv_inertia=None):
        """``ChBody::SetInertia`` with products of inertia (``use_full_inertia_tensor``, MC/ChSettings.h:107): n x 3 x 3
        symmetric matrices in the body frame; the inverse defaults to numpy's (symmetrised)."""
        J = _f64(inertia, (-1, 3, 3))
        if inv_inertia is None:
            K = np.linalg.inv(J)
            K = np.ascontiguousarray(0.5 * (K + K.transpose(0, 2, 1)))
        else:
            K = _f64(inv_inertia, (-1, 3, 3))
        check(self._lib.cb200_set_inertia_full(self._h, int(first), J.shape[0], _ptr(J), _ptr(K)))

    def SetBodyState(self, first, pos=None, rot=None, vel=None, n=None):
        """Gather of UpdateRigidBodies (ChSystemMulticore.cpp:329-377): host pose/velocity -> device."""
        arrs = [None if a is None else _f64(a) for a in (pos, rot, vel)]
        if n is None:
            n = next(a.size // k for a, k in zip(arrs, (3, 4, 6)) if a is not None)
        check(self._lib.cb200_upload_state(self._h, int(first), int(n), _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2])))

    def SetActiveBoundingBox(self, aabb_min=None, aabb_max=None):
        """``ChCollisionSystemMulticore::SetActiveBoundingBox`` (ChCollisionSystemMulticore.cpp:72-77); None: off."""
        if aabb_min is None:
            check(self._lib.cb200_set_active_aabb(self._h, 0, None, None))
        else:
            lo, hi = _f64(aabb_min), _f64(aabb_max)
            check(self._lib.cb200_set_active_aabb(self._h, 1, _ptr(lo), _ptr(hi)))

    def SetSpeedLimits(self, first, max_speed, max_wvel):
        """``ChBody::SetLimitSpeed(true)`` + ``SetMaxLinVel`` / ``SetMaxAngVel`` (ChBody.cpp:319-329) per body."""
        if max_speed is None:
            check(self._lib.cb200_set_speed_limits(self._h, 0, 0, None, None))
            return
        a, b = _f64(max_speed), _f64(max_wvel)
        check(self._lib.cb200_set_speed_limits(self._h, int(first), int(a.size), _ptr(a), _ptr(b)))

    def GetBodyFlags(self, first=0, n=None):
        n = self.num_bodies - first if n is None else n
        act, col = np.zeros(n, dtype=np.int8), np.zeros(n, dtype=np.int8)
        check(self._lib.cb200_download_body_flags(self._h, int(first), int(n), _ptr(act), _ptr(col)))
        return act.astype(bool), col.astype(bool)

    def SetAppliedForces(self, first, force_torque):
        ft = None if force_torque is None else _f64(force_torque)
        n = self.num_bodies - first if ft is None else ft.size // 6
        check(self._lib.cb200_upload_applied_forces(self._h, int(first), int(n), _ptr(ft)))

    def AddShapes(self, type, body_id, dims, pos_local=None, rot_local=None, family=None, material_id=None):
        """Bulk ``ChCollisionSystemMulticore::Add`` (ChCollisionSystemMulticore.cpp:86-187)."""
        type = np.ascontiguousarray(type, dtype=np.int32).ravel()
        n = type.size
        body_id = np.ascontiguousarray(body_id, dtype=np.uint32).ravel()
        dims = _f64(dims, (n, 4))
        pl = np.zeros((n, 3)) if pos_local is None else _f64(pos_local, (n, 3))
        if rot_local is None:
            rl = np.zeros((n, 4))
            rl[:, 0] = 1
        else:
            rl = _f64(rot_local, (n, 4))
        if family is None:
            fam = np.empty((n, 2), dtype=np.int16)
            fam[:, 0] = 1  # family group bit 0, mask all (ChCollisionModel defaults)
            fam[:, 1] = 0x7FFF
        else:
            fam = np.ascontiguousarray(family, dtype=np.int16).reshape(n, 2)
        mat = np.zeros(n, dtype=np.int32) if material_id is None else np.ascontiguousarray(material_id, dtype=np.int32)
        check(self._lib.cb200_append_shapes(self._h, n, _ptr(type), _ptr(body_id), _ptr(fam), _ptr(pl), _ptr(rl),
                                            _ptr(dims), _ptr(mat)))
        self.num_shapes += n

    def AddShapeVertices(self, xyz):
        """Append points to the vertex pool of triangles (body frame) and convex hulls (shape frame) --
        shape_container::triangle_rigid / convex_rigid (CMC/ChCollisionData.h:35-59).  Returns the first index."""
        xyz = _f64(xyz, (-1, 3))
        first = C.c_int(0)
        check(self._lib.cb200_append_shape_vertices(self._h, xyz.shape[0], _ptr(xyz), C.byref(first)))
        return first.value

    def UpdateShapeVertices(self, first, xyz):
        """Overwrite part of the vertex pool in place (the terrain node's per-exchange write into
        shape_data.triangle_rigid, ChVehicleCosimTerrainNodeGranularOMP.cpp:858-911)."""
        xyz = _f64(xyz, (-1, 3))
        check(self._lib.cb200_update_shape_vertices(self._h, int(first), xyz.shape[0], _ptr(xyz)))

    def SetMaterials(self, table):
        table = np.ascontiguousarray(table, dtype=np.float32)
        if self.system_type == 1:
            table = table.reshape(-1, 14)
            check(self._lib.cb200_set_materials_smc(self._h, table.shape[0], _ptr(table)))
        else:
            table = table.reshape(-1, 11)
            check(self._lib.cb200_set_materials_nsc(self._h, table.shape[0], _ptr(table)))

    # ---- stepping ------------------------------------------------------------------------------------------
    def DoStepDynamics(self, step=None):
        """``ChSystem::DoStepDynamics`` -> AdvanceDynamics (ChSystemMulticore.cpp:73-200)."""
        if step is not None:
            self.settings.step_size = float(step)
        self.DoSteps(1)

    def StepHost(self, pos, rot, vel):
        """One DoStepDynamics on host-resident state (in place): the per-step gather / scatter of the reference
        (ChSystemMulticore.cpp:329-377, :134-153) as one pipelined call.  Arrays must be C-contiguous float64,
        ideally pinned (``_lib.load().cb200_host_alloc``)."""
        for a, w in ((pos, 3), (rot, 4), (vel, 6)):
            if a.dtype != np.float64 or not a.flags.c_contiguous or a.size != w * self.num_bodies:
                raise ValueError("StepHost needs C-contiguous float64 arrays of nb x (3, 4, 6)")
        self._push_settings()
        check(self._lib.cb200_step_host(self._h, _ptr(pos), _ptr(rot), _ptr(vel)))
        self.ch_time += self.settings.step_size

    def DoSteps(self, n):
        self._push_settings()
        if self._narrow_callback is not None or self._add_contact_callback is not None:
            for _ in range(int(n)):  # the callbacks sit between the collision stage and the solver: staged steps
                self.Update()
                self.RunCollision(7)
                self.RunContactCallbacks()
                self.RunSolver()
                self.Integrate()
            return
        check(self._lib.cb200_do_steps(self._h, int(n)))
        self.ch_time += n * self.settings.step_size

    # ---- bilateral constraints --------------------------------------------------------------------
    def SetBilaterals(self, body_ids, cq_a, cq_b, rhs, max_iteration_bilateral=100, tol_speed=1e-4):
        """BODY_BODY rows of ``ChConstraintBilateral`` (Build_D / Build_b): ``cq_a . v_a + cq_b . v_b + rhs = 0`` per
        row; ``None`` removes all rows.  Defaults of ``max_iteration_bilateral`` / ``tol_speed``: MC/ChSettings.h."""
        if body_ids is None:
            check(self._lib.cb200_set_bilaterals(self._h, 0, None, None, None, None, 0, 0.0))
            return
        self._push_settings()
        ids = np.ascontiguousarray(body_ids, dtype=np.int32).reshape(-1, 2)
        a, b, r = _f64(cq_a, (-1, 6)), _f64(cq_b, (-1, 6)), _f64(rhs).ravel()
        check(self._lib.cb200_set_bilaterals(self._h, ids.shape[0], _ptr(ids), _ptr(a), _ptr(b), _ptr(r),
                                             int(max_iteration_bilateral), float(tol_speed)))

    def GetBilateralGamma(self, n):
        g, it, res = np.zeros(n), C.c_int(), C.c_double()
        check(self._lib.cb200_download_bilateral_gamma(self._h, _ptr(g), C.byref(it), C.byref(res)))
        return g, it.value, res.value

    # ---- contact callbacks -----------------------------------------------------------------------
    def RegisterNarrowphaseCallback(self, callback):
        """``ChCollisionSystem::RegisterNarrowphaseCallback``: ``callback(cinfo) -> bool``; False drops the contact
        (ChCollisionSystemMulticore.cpp:361-368).  ``cinfo``: dict(bodyA, bodyB, shapeA, shapeB, vN, vpA, vpB, distance,
        eff_radius).  None unregisters."""
        self._narrow_callback = callback

    def RegisterAddContactCallback(self, callback):
        """``ChContactContainer::RegisterAddContactCallback``: ``callback(cinfo, material)`` may modify the composite
        material record of the contact in place (a numpy row; fields in ``COMPOSITE_FIELDS_SMC / _NSC``).  As in the
        reference (ChContactContainerMulticoreSMC.cpp:147-165) ``cinfo['vpA'] / ['vpB']`` passed to THIS callback are
        the contact points plus the body positions."""
        self._add_contact_callback = callback

    def RunContactCallbacks(self):
        if self._narrow_callback is None and self._add_contact_callback is None:
            return
        c = self.GetContacts()
        n = c["depth"].size
        if n == 0:
            return
        sa, sb = (c["shape_ids"] >> 32).astype(np.int64), (c["shape_ids"] & 0xFFFFFFFF).astype(np.int64)

        def cinfo(i, offset_a=0.0, offset_b=0.0):
            return dict(bodyA=int(c["body_ids"][i, 0]), bodyB=int(c["body_ids"][i, 1]), shapeA=int(sa[i]),
                        shapeB=int(sb[i]), vN=c["normal"][i], vpA=c["pt_a"][i] + offset_a, vpB=c["pt_b"][i] + offset_b,
                        distance=float(c["depth"][i]), eff_radius=float(c["eff_radius"][i]))

        keep = np.ones(n, dtype=np.uint8)
        if self._narrow_callback is not None:
            for i in range(n):
                keep[i] = 1 if self._narrow_callback(cinfo(i)) else 0
            if not keep.all():
                check(self._lib.cb200_filter_contacts(self._h, n, _ptr(keep)))
        if self._add_contact_callback is not None:
            kept = np.flatnonzero(keep)
            stride = C.c_int()
            check(self._lib.cb200_download_contact_materials(self._h, 0, None, C.byref(stride)))
            mats = np.empty((kept.size, stride.value))
            check(self._lib.cb200_download_contact_materials(self._h, kept.size, _ptr(mats), C.byref(stride)))
            pos, _, _ = self.GetBodyState()
            for j, i in enumerate(kept):
                ci = cinfo(i, pos[c["body_ids"][i, 0]], pos[c["body_ids"][i, 1]])
                self._add_contact_callback(ci, mats[j])
            check(self._lib.cb200_upload_contact_materials(self._h, kept.size, _ptr(mats)))

    # the reference's own seams, for stage-level parity tests
    def Update(self):
        self._push_settings()
        check(self._lib.cb200_update_forces(self._h))

    def RunCollision(self, stages=7):
        self._push_settings()
        check(self._lib.cb200_run_collision(self._h, int(stages)))

    def RunSolver(self):
        self._push_settings()
        if self.system_type == 1:
            check(self._lib.cb200_run_solver_smc(self._h))
        else:
            check(self._lib.cb200_run_solver_nsc(self._h))

    def Integrate(self):
        self._push_settings()
        check(self._lib.cb200_integrate(self._h))

    def Synchronize(self):
        check(self._lib.cb200_synchronize(self._h))

    # ---- read-backs ----------------------------------------------------------------------------------------
    def GetCounts(self):
        c = cb200_counts()
        check(self._lib.cb200_get_counts(self._h, C.byref(c)))
        return c

    def GetNumContacts(self):
        return int(self.GetCounts().num_rigid_contacts)

    def GetBodyState(self, first=0, n=None):
        n = self.num_bodies - first if n is None else n
        pos, rot, vel = np.empty((n, 3)), np.empty((n, 4)), np.empty((n, 6))
        check(self._lib.cb200_download_state(self._h, int(first), int(n), _ptr(pos), _ptr(rot), _ptr(vel)))
        return pos, rot, vel

    def GetAABB(self):
        mn, mx = np.empty((self.num_shapes, 3)), np.empty((self.num_shapes, 3))
        check(self._lib.cb200_download_aabb(self._h, _ptr(mn), _ptr(mx)))
        return mn, mx

    def GetBins(self):
        c = self.GetCounts()
        ni, na = c.num_bin_aabb_intersections, c.num_active_bins
        bn, ba = np.empty(ni, np.uint32), np.empty(ni, np.uint32)
        act, st = np.empty(na, np.uint32), np.empty(na + 1, np.uint32)
        check(self._lib.cb200_download_bins(self._h, _ptr(bn), _ptr(ba), _ptr(act), _ptr(st)))
        return bn, ba, act, st

    def GetPairs(self):
        n = self.GetCounts().num_possible_collisions
        p = np.empty(n, np.int64)
        check(self._lib.cb200_download_pairs(self._h, _ptr(p)))
        return p

    def GetContacts(self):
        n = self.GetCounts().num_rigid_contacts
        out = dict(normal=np.empty((n, 3)), pt_a=np.empty((n, 3)), pt_b=np.empty((n, 3)), depth=np.empty(n),
                   eff_radius=np.empty(n), body_ids=np.empty((n, 2), np.int32), shape_ids=np.empty(n, np.int64))
        check(self._lib.cb200_download_contacts(self._h, _ptr(out["normal"]), _ptr(out["pt_a"]), _ptr(out["pt_b"]),
                                                _ptr(out["depth"]), _ptr(out["eff_radius"]), _ptr(out["body_ids"]),
                                                _ptr(out["shape_ids"])))
        return out

    def GetContactReport(self):
        """The batch behind ``ReportAllContacts``: one row per contact with the arguments of
        ``ReportContactCallback::OnReportContact`` (ChContactContainerMulticore.cpp:65-149)."""
        n = self.GetCounts().num_rigid_contacts
        out = dict(pA=np.empty((n, 3)), pB=np.empty((n, 3)), plane=np.empty((n, 3, 3)), distance=np.empty(n),
                   eff_radius=np.empty(n), force=np.empty((n, 3)), torque=np.empty((n, 3)),
                   body_ids=np.empty((n, 2), np.int32))
        cnt = C.c_int()
        check(self._lib.cb200_report_contacts(self._h, n, _ptr(out["pA"]), _ptr(out["pB"]), _ptr(out["plane"]),
                                              _ptr(out["distance"]), _ptr(out["eff_radius"]), _ptr(out["force"]),
                                              _ptr(out["torque"]), _ptr(out["body_ids"]), C.byref(cnt)))
        return out

    def ReportAllContacts(self, callback):
        """``ChContactContainer::ReportAllContacts``: ``callback(pA, pB, plane, distance, eff_radius, force, torque,
        bodyA, bodyB)`` once per contact, in list order; a False return stops the walk (as ``proceed`` does)."""
        r = self.GetContactReport()
        for i in range(r["distance"].size):
            if callback(r["pA"][i], r["pB"][i], r["plane"][i], r["distance"][i], r["eff_radius"][i], r["force"][i],
                        r["torque"][i], int(r["body_ids"][i, 0]), int(r["body_ids"][i, 1])) is False:
                break

    def GetContactForces(self):
        """host_data.ct_force / ct_torque, two rows per contact (body 1, body 2)."""
        n = self.GetCounts().num_rigid_contacts
        f, t = np.empty((2 * n, 3)), np.empty((2 * n, 3))
        check(self._lib.cb200_download_contact_forces(self._h, _ptr(f), _ptr(t)))
        return f, t

    def GetBodyContactForces(self):
        f, t = np.empty((self.num_bodies, 3)), np.empty((self.num_bodies, 3))
        check(self._lib.cb200_download_body_contact_forces(self._h, _ptr(f), _ptr(t)))
        return f, t

    def GetBodyContactForce(self, body):
        """``ChSystemMulticoreSMC::GetBodyContactForce`` (ChSystemMulticoreSMC.cpp:74-81)."""
        return self.GetBodyContactForces()[0][body]

    def GetBodyContactTorque(self, body):
        return self.GetBodyContactForces()[1][body]

    def GetHf(self):
        hf = np.empty((self.num_bodies, 6))
        check(self._lib.cb200_download_hf(self._h, _ptr(hf)))
        return hf

    # ---- NSC read-backs (host_data.gamma / R_full / D_T, measures.solver) ------------------------------------
    def _nsc_rows(self):
        return {MODE_NORMAL: 1, MODE_SLIDING: 3, MODE_SPINNING: 6}[int(self.settings.solver_mode)]

    def GetGamma(self):
        n = self._nsc_rows() * self.GetCounts().num_rigid_contacts
        g = np.zeros(n)
        check(self._lib.cb200_download_gamma(self._h, _ptr(g), int(n)))
        return g

    def GetRhs(self):
        g = np.zeros(self._nsc_rows() * self.GetCounts().num_rigid_contacts)
        check(self._lib.cb200_download_nsc_rhs(self._h, _ptr(g)))
        return g

    def GetJacobianRows(self):
        n = self._nsc_rows() * self.GetCounts().num_rigid_contacts
        rows, bids = np.zeros((n, 12)), np.zeros((n, 2), np.int32)
        check(self._lib.cb200_download_nsc_jacobian(self._h, _ptr(rows), _ptr(bids)))
        return rows, bids

    def GetSolverHistory(self):
        """(residual, objective) per APGD iteration of the last solve (measures.solver.maxd_hist)."""
        n = C.c_int(0)
        check(self._lib.cb200_download_nsc_history(self._h, None, None, 0, C.byref(n)))
        r, o = np.zeros(n.value), np.zeros(n.value)
        check(self._lib.cb200_download_nsc_history(self._h, _ptr(r), _ptr(o), n.value, C.byref(n)))
        return r, o

    def SchurProduct(self, x, local_mode=None):
        """``ChSchurProduct::operator()`` on the system of the last NSC solve."""
        x = _f64(x)
        out = np.zeros_like(x)
        lm = int(self.settings.solver_mode) if local_mode is None else int(local_mode)
        check(self._lib.cb200_nsc_schur_product(self._h, lm, _ptr(x), _ptr(out)))
        return out

    def Project(self, gamma, local_mode=None):
        """``ChProjectConstraints::operator()``."""
        g = _f64(gamma).copy()
        lm = int(self.settings.solver_mode) if local_mode is None else int(local_mode)
        check(self._lib.cb200_nsc_project(self._h, lm, _ptr(g)))
        return g

    def GetTimer(self, name):
        s = C.c_double(0)
        check(self._lib.cb200_get_timer(self._h, name.encode(), C.byref(s)))
        return s.value

    # reference timer accessors (ChSystemMulticore.cpp:656-682)
    def GetTimerStep(self):
        return self.GetTimer("step")

    def GetTimerAdvance(self):
        return self.GetTimer("advance")

    def GetTimerUpdate(self):
        return self.GetTimer("update")

    def GetTimerCollision(self):
        return self.GetTimer("collision")

    def GetTimerCollisionBroad(self):
        return self.GetTimer("collision_broad")

    def GetTimerCollisionNarrow(self):
        return self.GetTimer("collision_narrow")

    def KernelLaunches(self):
        return int(self._lib.cb200_kernel_launches(self._h))

    def DeviceArray(self, name):
        """(device pointer, element count) of a resident array, for zero-copy tooling."""
        p, n = C.c_void_p(), C.c_int64()
        check(self._lib.cb200_device_array(self._h, name.encode(), C.byref(p), C.byref(n)))
        return p.value, n.value


class ChSystemMulticoreSMC(ChSystemMulticore):
    """Penalty (SMC) system, reference: src/chrono_multicore/physics/ChSystemMulticoreSMC.cpp."""

    system_type = 1


class ChSystemMulticoreNSC(ChSystemMulticore):
    """Complementarity (NSC / DVI) system, reference: src/chrono_multicore/physics/ChSystemMulticoreNSC.cpp."""

    system_type = 0


def load_scene(sys_, scene):
    """Push a scene dict produced by ``chrono_wisc_b200.scenes`` into a system."""
    st = sys_.GetSettings()
    for k, v in scene["settings"].items():
        cur = getattr(st, k)
        if hasattr(cur, "__len__"):
            cur[:] = list(v)
        else:
            setattr(st, k, v)
    sys_.AddBodies(scene["mass"], scene["inertia"], scene["pos"], scene["rot"], scene["vel"], scene["active"],
                   scene["collide"])
    if scene.get("shape_verts") is not None and len(scene["shape_verts"]):
        assert sys_.AddShapeVertices(scene["shape_verts"]) == 0
    sys_.AddShapes(scene["shape_type"], scene["shape_body"], scene["shape_dims"], scene["shape_pos"],
                   scene["shape_rot"], scene["shape_fam"], scene["shape_mat"])
    sys_.SetMaterials(scene["materials"])
    return sys_
