"""ctypes access to the two CPU checkers -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

  * ``Ref``    : oracle/_ref/libchrono_ref.so, the reference's own collision translation units + extracted
                 SMC force / cone functions (built from /root/reference by oracle/Makefile).
  * ``Oracle`` : oracle/liboracle.so, the CPU restatement (oracle/oracle.cpp) of the whole hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
Nothing under chrono_wisc_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libchrono_ref.so")
ORACLE_LIB = os.path.join(HERE, "liboracle.so")
P = C.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(P)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def build(ref_root="/root/reference"):
    """Compile liboracle.so always, and _ref/libchrono_ref.so when the reference tree is present."""
    subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, capture_output=True)
    if os.path.isdir(ref_root):
        subprocess.run(["make", "-C", HERE, "-j8", "ref", f"REF_ROOT={ref_root}"], check=True, capture_output=True)


def have_ref():
    return os.path.exists(REF_LIB)


class OrcSettings(C.Structure):
    """The checkers' own mirror of ``cb200_settings`` (include/chrono_b200.h -- the boundary contract, a header; the
    oracle never loads the product library).  tests/test_abi.py checks it field by field against the product's mirror."""

    _fields_ = [
        ("collision_envelope", C.c_double), ("broadphase_grid", C.c_int32), ("bins_per_axis", C.c_int32 * 3),
        ("bin_size", C.c_double * 3), ("grid_density", C.c_double), ("narrowphase_algorithm", C.c_int32),
        ("system_type", C.c_int32), ("step_size", C.c_double), ("gravity", C.c_double * 3),
        ("contact_force_model", C.c_int32), ("adhesion_force_model", C.c_int32), ("tangential_displ_mode", C.c_int32),
        ("use_material_properties", C.c_int32), ("characteristic_vel", C.c_double), ("min_slip_vel", C.c_double),
        ("min_roll_vel", C.c_double), ("min_spin_vel", C.c_double), ("solver_type", C.c_int32),
        ("solver_mode", C.c_int32), ("max_iteration_normal", C.c_int32), ("max_iteration_sliding", C.c_int32),
        ("max_iteration_spinning", C.c_int32), ("cache_step_length", C.c_int32), ("alpha", C.c_double),
        ("contact_recovery_speed", C.c_double), ("tolerance", C.c_double), ("deterministic", C.c_int32),
        ("test_objective", C.c_int32), ("tolerance_objective", C.c_double), ("reserved", C.c_int32 * 4),
    ]


class Ref:
    """The reference collision pipeline (GenerateAABB + ChBroadphase + ChNarrowphase) and helpers."""

    def __init__(self):
        if not have_ref():
            raise RuntimeError("oracle/_ref/libchrono_ref.so not built (needs /root/reference at build time)")
        L = C.CDLL(REF_LIB)
        L.ref_create.restype = P
        for name in ("ref_destroy", "ref_set_bodies", "ref_set_shapes", "ref_set_collision_settings",
                     "ref_run_collision", "ref_get_counts", "ref_get_grid", "ref_get_aabb", "ref_get_bins",
                     "ref_get_pairs", "ref_get_contacts", "ref_smc_contact_forces", "ref_orthogonalize",
                     "ref_compute_jacobian", "ref_rotate", "ref_rotate_t", "ref_abs_rotate", "ref_quat_mult",
                     "ref_set_num_threads", "ref_smc_setup", "ref_smc_step", "ref_smc_get_state",
                     "ref_smc_get_timers"):
            getattr(L, name).restype = None
        L.ref_prims_pair.restype = C.c_int
        L.ref_mpr_pair.restype = C.c_int
        L.ref_cone_generalized.restype = C.c_int
        L.ref_get_max_threads.restype = C.c_int
        self.L = L
        self.h = P(L.ref_create())
        self.ns = 0

    def __del__(self):
        try:
            self.L.ref_destroy(self.h)
        except Exception:
            pass

    def set_num_threads(self, n):
        self.L.ref_set_num_threads(C.c_int(int(n)))

    def max_threads(self):
        return int(self.L.ref_get_max_threads())

    def set_scene(self, scene):
        self.set_shapes(scene)
        self.set_bodies(scene["pos"], scene["rot"], scene["active"], scene["collide"])
        st = scene["settings"]
        env = 0.0 if st.get("system_type", 0) == 1 else st.get("collision_envelope", 0.0)
        self.set_settings(env, st.get("broadphase_grid", 0), st.get("bins_per_axis", (10, 10, 10)),
                          st.get("bin_size", (1, 1, 1)), st.get("grid_density", 5.0),
                          st.get("narrowphase_algorithm", 2))

    def set_bodies(self, pos, rot, active, collide):
        pos, rot = _f64(pos), _f64(rot)
        act = np.ascontiguousarray(active, dtype=np.int8)
        col = np.ascontiguousarray(collide, dtype=np.int8)
        self.L.ref_set_bodies(self.h, C.c_int(act.size), _p(pos), _p(rot), _p(act), _p(col))

    def set_shapes(self, scene):
        typ = np.ascontiguousarray(scene["shape_type"], dtype=np.int32)
        body = np.ascontiguousarray(scene["shape_body"], dtype=np.uint32)
        fam = np.ascontiguousarray(scene["shape_fam"], dtype=np.int16)
        oba, obr, dims = _f64(scene["shape_pos"]), _f64(scene["shape_rot"]), _f64(scene["shape_dims"])
        self.ns = typ.size
        # vertex pool (scene["shape_verts"], dims = (first vertex, count)) -> the driver's conventions: triangles
        # index a list of 9-double triangles, hulls give (count, offset) into convex_rigid (the whole pool)
        verts = _f64(scene["shape_verts"]) if scene.get("shape_verts") is not None else np.zeros((0, 3))
        tri = np.zeros((0, 9))
        if verts.size:
            dims = dims.reshape(-1, 4).copy()
            tsel = np.flatnonzero(typ == 10)
            tri = np.ascontiguousarray(
                np.stack([verts.reshape(-1, 3)[int(dims[i, 0]):int(dims[i, 0]) + 3].ravel() for i in tsel])
                if tsel.size else np.zeros((0, 9)))
            dims[tsel, 0] = np.arange(tsel.size)
            hsel = np.flatnonzero(typ == 5)
            dims[hsel, 0], dims[hsel, 1] = dims[hsel, 1].copy(), dims[hsel, 0].copy()
            dims = np.ascontiguousarray(dims)
        self._keep = (verts, tri, dims)
        self.L.ref_set_shapes(self.h, C.c_int(typ.size), _p(typ), _p(body), _p(fam), _p(oba), _p(obr), _p(dims),
                              _p(tri) if tri.size else None, C.c_int(verts.size // 3),
                              _p(verts) if verts.size else None)

    def set_settings(self, envelope, grid_type, res, bin_size, density, algorithm):
        res = np.ascontiguousarray(res, dtype=np.int32)
        bs = _f64(bin_size)
        self.L.ref_set_collision_settings(self.h, C.c_double(envelope), C.c_int(grid_type), _p(res), _p(bs),
                                          C.c_double(density), C.c_int(algorithm))

    def run(self, stages=7):
        self.L.ref_run_collision(self.h, C.c_int(stages))

    # ---- whole SMC step: reference collision + reference force function + restated glue, all in C++ / OpenMP ----
    SMC_PHASES = ("update", "collision_broad", "collision_narrow", "materials", "forces", "accumulate", "velocity",
                  "integrate")

    def smc_setup(self, scene):
        """set_scene + everything the step needs beside the collision data (ref_smc_setup)."""
        self.set_scene(scene)
        st = scene["settings"]
        g = lambda k, d: st.get(k, d)
        iset = np.array([g("contact_force_model", 1), g("adhesion_force_model", 0), g("tangential_displ_mode", 1),
                         g("use_material_properties", 1)], dtype=np.int32)
        if iset[2] == 2:
            raise ValueError("the whole-step reference driver keeps no MultiStep history")
        grav = g("gravity", (0, 0, -9.81))
        dset = np.array([g("characteristic_vel", 1.0), g("min_slip_vel", 1e-4), g("min_roll_vel", 1e-4),
                         g("min_spin_vel", 1e-4), g("step_size", 1e-3), grav[0], grav[1], grav[2]], dtype=np.float64)
        mats = np.ascontiguousarray(scene["materials"], dtype=np.float32)
        smat = np.ascontiguousarray(scene["shape_mat"], dtype=np.int32)
        self.nb = scene["mass"].size
        self.L.ref_smc_setup(self.h, C.c_int(self.nb), _p(_f64(scene["mass"])), _p(_f64(scene["inertia"])),
                             _p(_f64(scene["vel"])), C.c_int(mats.shape[0]), _p(mats), C.c_int(smat.size), _p(smat),
                             _p(iset), _p(dset))

    def smc_step(self, n=1):
        self.L.ref_smc_step(self.h, C.c_int(int(n)))

    def smc_state(self):
        pos, rot, vel = np.empty((self.nb, 3)), np.empty((self.nb, 4)), np.empty((self.nb, 6))
        self.L.ref_smc_get_state(self.h, _p(pos), _p(rot), _p(vel))
        return pos, rot, vel

    def smc_timers(self):
        """seconds per step and phase since smc_setup"""
        t, n = np.zeros(8), C.c_long(0)
        self.L.ref_smc_get_timers(self.h, _p(t), C.byref(n))
        k = max(n.value, 1)
        return {name: float(v) / k for name, v in zip(self.SMC_PHASES, t)}

    def counts(self):
        c = np.zeros(8, dtype=np.uint32)
        self.L.ref_get_counts(self.h, _p(c))
        return dict(N_i=int(c[0]), N_a=int(c[1]), N_p=int(c[2]), N_c=int(c[3]), bins=tuple(int(x) for x in c[4:7]),
                    num_bins=int(c[7]))

    def grid(self):
        g = np.zeros(18)
        self.L.ref_get_grid(self.h, _p(g))
        g = g.reshape(6, 3)
        return dict(min_bounding_point=g[0], max_bounding_point=g[1], bin_size=g[2], inv_bin_size=g[3],
                    rigid_min=g[4], rigid_max=g[5])

    def aabb(self):
        mn, mx = np.empty((self.ns, 3)), np.empty((self.ns, 3))
        self.L.ref_get_aabb(self.h, _p(mn), _p(mx))
        return mn, mx

    def bins(self):
        c = self.counts()
        bn, ba = np.empty(c["N_i"], np.uint32), np.empty(c["N_i"], np.uint32)
        act, st = np.empty(c["N_a"], np.uint32), np.empty(c["N_a"] + 1, np.uint32)
        self.L.ref_get_bins(self.h, _p(bn), _p(ba), _p(act), _p(st))
        return bn, ba, act, st

    def pairs(self):
        p = np.empty(self.counts()["N_p"], np.int64)
        self.L.ref_get_pairs(self.h, _p(p))
        return p

    def contacts(self):
        n = self.counts()["N_c"]
        out = dict(normal=np.empty((n, 3)), pt_a=np.empty((n, 3)), pt_b=np.empty((n, 3)), depth=np.empty(n),
                   eff_radius=np.empty(n), body_ids=np.empty((n, 2), np.int32), shape_ids=np.empty(n, np.int64))
        self.L.ref_get_contacts(self.h, _p(out["normal"]), _p(out["pt_a"]), _p(out["pt_b"]), _p(out["depth"]),
                                _p(out["eff_radius"]), _p(out["body_ids"]), _p(out["shape_ids"]))
        return out

    # ---- single-pair narrowphase (golden vectors) ----
    def prims_pair(self, typeA, posA, rotA, dimsA, typeB, posB, rotB, dimsB, separation):
        a = [_f64(x) for x in (posA, rotA, np.resize(np.asarray(dimsA, float), 4) if np.size(dimsA) == 4 else np.pad(np.asarray(dimsA, float), (0, 4 - np.size(dimsA))))]
        b = [_f64(x) for x in (posB, rotB, np.resize(np.asarray(dimsB, float), 4) if np.size(dimsB) == 4 else np.pad(np.asarray(dimsB, float), (0, 4 - np.size(dimsB))))]
        n, p1, p2, d, e = np.zeros((8, 3)), np.zeros((8, 3)), np.zeros((8, 3)), np.zeros(8), np.zeros(8)
        nc = self.L.ref_prims_pair(C.c_int(typeA), _p(a[0]), _p(a[1]), _p(a[2]), C.c_int(typeB), _p(b[0]), _p(b[1]),
                                   _p(b[2]), C.c_double(separation), _p(n), _p(p1), _p(p2), _p(d), _p(e))
        k = max(nc, 0)
        return nc, dict(normal=n[:k], pt_a=p1[:k], pt_b=p2[:k], depth=d[:k], eff_radius=e[:k])

    def mpr_pair(self, typeA, posA, rotA, dimsA, typeB, posB, rotB, dimsB, envelope):
        pad = lambda v: np.pad(np.asarray(v, float).ravel(), (0, 4 - np.size(v)))
        a = [_f64(posA), _f64(rotA), _f64(pad(dimsA))]
        b = [_f64(posB), _f64(rotB), _f64(pad(dimsB))]
        n, p1, p2, d = np.zeros(3), np.zeros(3), np.zeros(3), np.zeros(1)
        hit = self.L.ref_mpr_pair(C.c_int(typeA), _p(a[0]), _p(a[1]), _p(a[2]), C.c_int(typeB), _p(b[0]), _p(b[1]),
                                  _p(b[2]), C.c_double(envelope), _p(n), _p(p1), _p(p2), _p(d))
        return hit, dict(normal=n, pt_a=p1, pt_b=p2, depth=d[0])

    # ---- SMC per-contact forces through the reference's function_CalcContactForces ----
    @staticmethod
    def new_shear_history(nb, max_shear=20):
        """host_container::shear_neigh / shear_disp / contact_relvel_init / contact_duration as
        ChSystemMulticoreSMC::AddMaterialSurfaceData initialises them (MC/physics/ChSystemMulticoreSMC.cpp:39-50)."""
        return dict(neigh=np.full((max_shear * nb, 3), -1, np.int32), disp=np.zeros((max_shear * nb, 3)),
                    relvel=np.zeros(max_shear * nb), duration=np.zeros(max_shear * nb))

    def smc_contact_forces(self, contacts, mats, settings, mass, pos, rot, vel, history=None):
        """history (MultiStep): updated in place, including the release of untouched slots
        (ChIterativeSolverMulticoreSMC::ProcessContacts, ...SMC.cpp:687-695)."""
        nc = contacts["depth"].size
        nb = np.asarray(mass).size
        bids = np.ascontiguousarray(contacts["body_ids"], dtype=np.int32)
        fric, mod, adh = _f64(mats["fric"]), _f64(mats["modulus"]), _f64(mats["adhesion"])
        cr, smc = _f64(mats["cr"]), _f64(mats["smc"])
        ct_bid = np.zeros(2 * nc, np.int32)
        ct_f, ct_t = np.zeros((2 * nc, 3)), np.zeros((2 * nc, 3))
        g = lambda k, d: settings.get(k, d)
        sp = touch = None
        hp = [None] * 5
        if history is not None:
            sid = np.asarray(contacts["shape_ids"]).astype(np.int64)
            sp = np.ascontiguousarray(np.stack([sid >> 32, sid & 0xFFFFFFFF], 1), dtype=np.int32)
            touch = np.zeros(history["relvel"].size, np.int8)
            hp = [_p(history["neigh"]), _p(touch), _p(history["disp"]), _p(history["relvel"]), _p(history["duration"])]
        self.L.ref_smc_contact_forces(
            C.c_int(nc), _p(bids), _p(sp) if sp is not None else None, C.c_int(g("contact_force_model", 1)), C.c_int(g("adhesion_force_model", 0)),
            C.c_int(g("tangential_displ_mode", 1)), C.c_int(g("use_material_properties", 1)),
            C.c_double(g("characteristic_vel", 1.0)), C.c_double(g("min_slip_vel", 1e-4)),
            C.c_double(g("min_roll_vel", 1e-4)), C.c_double(g("min_spin_vel", 1e-4)), C.c_double(g("step_size", 1e-3)),
            C.c_int(nb), _p(_f64(mass)), _p(_f64(pos)), _p(_f64(rot)), _p(_f64(vel)), _p(fric), _p(mod), _p(adh),
            _p(cr), _p(smc), _p(_f64(contacts["pt_a"])), _p(_f64(contacts["pt_b"])), _p(_f64(contacts["normal"])),
            _p(_f64(contacts["depth"])), _p(_f64(contacts["eff_radius"])), hp[0], hp[1], hp[2], hp[3], hp[4],
            _p(ct_bid), _p(ct_f), _p(ct_t))
        if history is not None:
            history["neigh"][touch == 0, 0] = -1
        return ct_bid, ct_f, ct_t

    def rotate(self, v, q):
        o = np.zeros(3)
        self.L.ref_rotate(_p(_f64(v)), _p(_f64(q)), _p(o))
        return o

    def abs_rotate(self, q, v):
        o = np.zeros(3)
        self.L.ref_abs_rotate(_p(_f64(q)), _p(_f64(v)), _p(o))
        return o

    def orthogonalize(self, U):
        V, W = np.zeros(3), np.zeros(3)
        self.L.ref_orthogonalize(_p(_f64(U)), _p(V), _p(W))
        return V, W

    def compute_jacobian(self, quat, U, V, W, point):
        T1, T2, T3 = np.zeros(3), np.zeros(3), np.zeros(3)
        self.L.ref_compute_jacobian(_p(_f64(quat)), _p(_f64(U)), _p(_f64(V)), _p(_f64(W)), _p(_f64(point)), _p(T1),
                                    _p(T2), _p(T3))
        return T1, T2, T3

    def cone_generalized(self, gn, gu, gv, mu):
        a, b, c = C.c_double(gn), C.c_double(gu), C.c_double(gv)
        r = self.L.ref_cone_generalized(C.byref(a), C.byref(b), C.byref(c), C.c_double(mu))
        return r, (a.value, b.value, c.value)


class Oracle:
    """The CPU restatement (oracle/oracle.cpp) of the whole SMC hot path."""

    def __init__(self, scene=None, threads=None):
        if not os.path.exists(ORACLE_LIB):
            build()
        L = C.CDLL(ORACLE_LIB)
        L.orc_create.restype = P
        L.orc_max_threads.restype = C.c_int
        for name in ("orc_destroy", "orc_default_settings", "orc_set_threads", "orc_set_settings", "orc_set_bodies", "orc_set_shapes",
                     "orc_set_materials_smc", "orc_set_state", "orc_get_state", "orc_collide", "orc_smc_step",
                     "orc_get_counts", "orc_get_grid", "orc_get_aabb", "orc_get_pairs", "orc_get_contacts",
                     "orc_get_contact_forces", "orc_get_body_forces", "orc_get_hf", "orc_get_timers",
                     "orc_set_materials_nsc", "orc_nsc_solve", "orc_nsc_step", "orc_nsc_info", "orc_nsc_get_vectors",
                     "orc_nsc_get_rows", "orc_nsc_schur", "orc_nsc_project"):
            getattr(L, name).restype = None
        L.orc_nsc_hist.restype = C.c_int
        self.L = L
        self.h = P(L.orc_create())
        self.nb = self.ns = 0
        if threads is not None:
            self.set_threads(threads)
        if scene is not None:
            self.set_scene(scene)

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    def set_threads(self, n):
        self.L.orc_set_threads(C.c_int(int(n)))

    def max_threads(self):
        return int(self.L.orc_max_threads())

    def set_scene(self, scene):
        st = OrcSettings()
        self.L.orc_default_settings(C.byref(st))
        for k, v in scene["settings"].items():
            cur = getattr(st, k)
            if hasattr(cur, "__len__"):
                cur[:] = list(v)
            else:
                setattr(st, k, v)
        self.settings = st
        self.L.orc_set_settings(self.h, C.byref(st))
        self.nb, self.ns = scene["mass"].size, scene["shape_type"].size
        act = np.ascontiguousarray(scene["active"], dtype=np.int8)
        col = np.ascontiguousarray(scene["collide"], dtype=np.int8)
        self.L.orc_set_bodies(self.h, C.c_int(self.nb), _p(_f64(scene["mass"])), _p(_f64(scene["inertia"])), _p(act),
                              _p(col))
        self.L.orc_set_shapes(self.h, C.c_int(self.ns), _p(np.ascontiguousarray(scene["shape_type"], np.int32)),
                              _p(np.ascontiguousarray(scene["shape_body"], np.uint32)),
                              _p(np.ascontiguousarray(scene["shape_fam"], np.int16)), _p(_f64(scene["shape_pos"])),
                              _p(_f64(scene["shape_rot"])), _p(_f64(scene["shape_dims"])),
                              _p(np.ascontiguousarray(scene["shape_mat"], np.int32)))
        m = np.ascontiguousarray(scene["materials"], dtype=np.float32)
        if m.shape[-1] == 14:
            self.L.orc_set_materials_smc(self.h, C.c_int(m.shape[0]), _p(m))
        else:
            self.L.orc_set_materials_nsc(self.h, C.c_int(m.shape[0]), _p(m))
        self.set_state(scene["pos"], scene["rot"], scene["vel"])

    def set_state(self, pos, rot, vel):
        self.L.orc_set_state(self.h, _p(_f64(pos)), _p(_f64(rot)), _p(_f64(vel)))

    def state(self):
        pos, rot, vel = np.empty((self.nb, 3)), np.empty((self.nb, 4)), np.empty((self.nb, 6))
        self.L.orc_get_state(self.h, _p(pos), _p(rot), _p(vel))
        return pos, rot, vel

    def collide(self):
        self.L.orc_collide(self.h)

    def step(self, n=1):
        if self.settings.system_type == 1:
            self.L.orc_smc_step(self.h, C.c_int(int(n)))
        else:
            self.L.orc_nsc_step(self.h, C.c_int(int(n)))

    # ---- NSC (oracle_nsc.inc) ----
    def nsc_rows(self):
        return {0: 1, 1: 3, 2: 6}[int(self.settings.solver_mode)]

    def nsc_solve(self):
        """Forces + NSC solve on the current contact list (call collide() first); updates the velocities."""
        self.L.orc_nsc_solve(self.h)

    def nsc_info(self):
        i = np.zeros(4)
        self.L.orc_nsc_info(self.h, _p(i))
        return dict(iterations=int(i[0]), residual=i[1], objective=i[2], L=i[3])

    def nsc_hist(self):
        n = self.L.orc_nsc_hist(self.h, None, None, C.c_int(0))
        r, o = np.zeros(n), np.zeros(n)
        self.L.orc_nsc_hist(self.h, _p(r), _p(o), C.c_int(n))
        return r, o

    def nsc_vectors(self):
        n = self.nsc_rows() * self.counts()["N_c"]
        out = dict(gamma=np.zeros(n), R=np.zeros(n), b=np.zeros(n), E=np.zeros(n))
        self.L.orc_nsc_get_vectors(self.h, _p(out["gamma"]), _p(out["R"]), _p(out["b"]), _p(out["E"]))
        return out

    def nsc_jacobian_rows(self):
        """D_T rows in the reference's row order: 12 values (block of the lower body id first) + the two body ids."""
        n = self.nsc_rows() * self.counts()["N_c"]
        v, b = np.zeros((n, 12)), np.zeros((n, 2), np.int32)
        self.L.orc_nsc_get_rows(self.h, _p(v), _p(b))
        return v, b

    def nsc_schur(self, x, local_mode=None):
        x = _f64(x)
        out = np.zeros_like(x)
        lm = int(self.settings.solver_mode) if local_mode is None else int(local_mode)
        self.L.orc_nsc_schur(self.h, C.c_int(lm), _p(x), _p(out))
        return out

    def nsc_project(self, gamma, local_mode=None):
        g = _f64(gamma).copy()
        lm = int(self.settings.solver_mode) if local_mode is None else int(local_mode)
        self.L.orc_nsc_project(self.h, C.c_int(lm), _p(g))
        return g

    def counts(self):
        c = np.zeros(8, dtype=np.uint32)
        self.L.orc_get_counts(self.h, _p(c))
        return dict(N_i=int(c[0]), N_a=int(c[1]), N_p=int(c[2]), N_c=int(c[3]), bins=tuple(int(x) for x in c[4:7]),
                    num_bins=int(c[7]))

    def grid(self):
        g = np.zeros(12)
        self.L.orc_get_grid(self.h, _p(g))
        g = g.reshape(4, 3)
        return dict(min_bounding_point=g[0], max_bounding_point=g[1], bin_size=g[2], inv_bin_size=g[3])

    def aabb(self):
        mn, mx = np.empty((self.ns, 3)), np.empty((self.ns, 3))
        self.L.orc_get_aabb(self.h, _p(mn), _p(mx))
        return mn, mx

    def pairs(self):
        p = np.empty(self.counts()["N_p"], np.int64)
        self.L.orc_get_pairs(self.h, _p(p))
        return p

    def contacts(self):
        n = self.counts()["N_c"]
        out = dict(normal=np.empty((n, 3)), pt_a=np.empty((n, 3)), pt_b=np.empty((n, 3)), depth=np.empty(n),
                   eff_radius=np.empty(n), body_ids=np.empty((n, 2), np.int32), shape_ids=np.empty(n, np.int64))
        self.L.orc_get_contacts(self.h, _p(out["normal"]), _p(out["pt_a"]), _p(out["pt_b"]), _p(out["depth"]),
                                _p(out["eff_radius"]), _p(out["body_ids"]), _p(out["shape_ids"]))
        return out

    def contact_forces(self):
        n = self.counts()["N_c"]
        f, t = np.zeros((2 * n, 3)), np.zeros((2 * n, 3))
        self.L.orc_get_contact_forces(self.h, _p(f), _p(t))
        return f, t

    def body_forces(self):
        f, t = np.zeros((self.nb, 3)), np.zeros((self.nb, 3))
        self.L.orc_get_body_forces(self.h, _p(f), _p(t))
        return f, t

    def hf(self):
        hf = np.empty((self.nb, 6))
        self.L.orc_get_hf(self.h, _p(hf))
        return hf

    def timers(self):
        t = np.zeros(6)
        self.L.orc_get_timers(self.h, _p(t))
        return dict(zip(("update", "collision_broad", "collision_narrow", "forces", "accumulate", "integrate"), t))


def canonical_pairs(p):
    """Canonical order for pair lists: ascending packed key."""
    return np.sort(np.asarray(p, dtype=np.int64))


def canonical_contacts(c):
    """Canonical order for contact lists: by (shape pair key, then point coordinates for multi-contact pairs)."""
    order = np.lexsort((c["pt_a"][:, 2], c["pt_a"][:, 1], c["pt_a"][:, 0], c["shape_ids"]))
    return {k: v[order] for k, v in c.items()}
