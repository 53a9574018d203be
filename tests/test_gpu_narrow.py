"""Device narrowphase of non-sphere pairs -- multi-contact analytic tests (capsule-capsule, box-capsule, box-box)
and Minkowski Portal Refinement -- against the reference's own narrowphase (oracle/_ref: ChNarrowphasePRIMS.cpp,
ChNarrowphaseMPR.cpp compiled from /root/reference) on identical states.

Bar: the contact list (shape ids, body ids, normal, points, depth, effective radius) is BIT-EXACT and in the
reference's order for the three algorithms ChNarrowphase::Algorithm::{MPR, PRIMS, HYBRID}.
"""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from chrono_wisc_b200.system import (BOX, CAPSULE, CONE, CONVEXHULL, CYLINDER, CYLSHELL, ELLIPSOID, ROUNDEDBOX,
                                     ROUNDEDCYL, SPHERE, TRIANGLE)

pytestmark = pytest.mark.gpu

MPR, PRIMS, HYBRID = 0, 1, 2
KEYS = ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius")

DIMS = {
    SPHERE: [0.1, 0, 0, 0],
    BOX: [0.1, 0.08, 0.06, 0],
    CAPSULE: [0.06, 0.06, 0, 0],
    CYLINDER: [0.07, 0.07, 0.09, 0],
    CONE: [0.08, 0.08, 0.1, 0],
    ELLIPSOID: [0.1, 0.07, 0.05, 0],
    ROUNDEDBOX: [0.08, 0.07, 0.05, 0.02],
    ROUNDEDCYL: [0.06, 0.06, 0.08, 0.02],
    CYLSHELL: [0.06, 0.08, 0.08, 0],  # PRIMS reads (radius, hlen) = (x, y), MPR (radius, hheight) = (x, z)
}


def cluttered(kinds, n_side, spacing, seed, algorithm, envelope, aligned=False, grid_density=5.0):
    """A jittered lattice of randomly oriented bodies cycling through `kinds`, spacing below the shape size so that
    neighbours interpenetrate; a fixed floor box below."""
    b = scenes.SceneBuilder()
    rng = np.random.default_rng(seed)
    floor = b.add_bodies(1.0, (1, 1, 1), [[0, 0, -0.2]], active=0)
    b.add_shapes(BOX, [floor], [[n_side * spacing, n_side * spacing, 0.15, 0]], pos=[[0.5 * n_side * spacing] * 2 + [0]])
    i = 0
    for ix in range(n_side):
        for iy in range(n_side):
            for iz in range(n_side):
                p = np.array([ix, iy, iz]) * spacing + rng.uniform(-0.15, 0.15, 3) * spacing
                if aligned == 2:  # identity, quarter turn about z, quarter turn about x (axis along -y)
                    q = [np.array([1.0, 0, 0, 0]), np.array([np.sqrt(0.5), 0, 0, np.sqrt(0.5)]),
                         np.array([np.sqrt(0.5), np.sqrt(0.5), 0, 0])][(i // 2) % 3]
                elif aligned:
                    q = np.array([1.0, 0, 0, 0]) if (i % 3) else np.array([np.sqrt(0.5), 0, 0, np.sqrt(0.5)])
                else:
                    q = rng.normal(size=4)
                    q /= np.linalg.norm(q)
                bid = b.add_bodies(1.0, (0.01, 0.01, 0.01), [p], rot=[q])
                t = kinds[i % len(kinds)]
                if t == TRIANGLE:  # three body-frame vertices around the body origin
                    v = rng.uniform(-0.12, 0.12, (3, 3))
                    b.add_triangle(bid, v[0], v[1], v[2])
                elif t == CONVEXHULL:  # 4..12 shape-frame points, offset shape frame
                    pts = rng.uniform(-0.09, 0.09, (int(rng.integers(4, 13)), 3))
                    ql = rng.normal(size=4)
                    b.add_convex_hull(bid, pts, pos=[rng.uniform(-0.02, 0.02, 3)], rot=[ql / np.linalg.norm(ql)])
                else:
                    b.add_shapes(t, [bid], [DIMS[t]])
                i += 1
    # the SMC system runs with a zero envelope (ChSystemMulticoreSMC.cpp:27); a non-zero one is an NSC scene
    nsc = envelope > 0
    settings = dict(system_type=0 if nsc else 1, step_size=1e-3, gravity=(0, 0, -9.81), broadphase_grid=2,
                    grid_density=grid_density, narrowphase_algorithm=algorithm, collision_envelope=envelope)
    return b.finish([scenes.nsc_material() if nsc else scenes.smc_material()], settings)


def compare(scene, ref, min_contacts):
    cls = cw.ChSystemMulticoreSMC if scene["settings"]["system_type"] == 1 else cw.ChSystemMulticoreNSC
    sys = cw.load_scene(cls(), scene)
    sys.RunCollision(7)
    ref.set_scene(scene)
    ref.run(7)
    k, rk = sys.GetContacts(), ref.contacts()
    assert rk["depth"].size >= min_contacts, rk["depth"].size
    assert sys.GetCounts().num_rigid_contacts == rk["depth"].size
    for key in KEYS:
        assert np.array_equal(k[key], rk[key]), key
    return scene, rk


@pytest.mark.parametrize("envelope", [0.0, 0.01])
@pytest.mark.parametrize("algorithm", [MPR, PRIMS, HYBRID])
def test_box_capsule_sphere_mix(ref, algorithm, envelope):
    """config E's shape families (demo_MCORE_mixerSMC.cpp:118-168): spheres, boxes, capsules."""
    scene = cluttered([SPHERE, BOX, CAPSULE], 7, 0.15, 3, algorithm, envelope)
    scene, rk = compare(scene, ref, 300)
    if algorithm != MPR:
        types = scene["shape_type"]
        seen = set(zip(types[rk["shape_ids"] >> 32], types[rk["shape_ids"] & 0xFFFFFFFF]))
        for pair in [(BOX, BOX), (CAPSULE, CAPSULE)]:
            assert pair in seen, seen
        assert (BOX, CAPSULE) in seen or (CAPSULE, BOX) in seen
        # multi-contact pairs were produced
        _, counts = np.unique(rk["shape_ids"], return_counts=True)
        assert counts.max() >= 2


@pytest.mark.parametrize("algorithm", [MPR, HYBRID])
def test_all_convex_types_mpr(ref, algorithm):
    """Everything without an analytic test falls through to MPR in HYBRID mode (ChNarrowphase.cpp:303-320)."""
    kinds = [SPHERE, BOX, CAPSULE, CYLINDER, CONE, ELLIPSOID, ROUNDEDBOX, ROUNDEDCYL]
    scene = cluttered(kinds, 8, 0.14, 5, algorithm, 0.005)
    compare(scene, ref, 500)


def test_face_face_boxes_many_contacts(ref):
    """Axis-aligned / quarter-turned boxes: face-face and face-edge branches of box_box (up to 8 contacts a pair),
    parallel capsules (two contacts)."""
    scene = cluttered([BOX, BOX, CAPSULE], 6, 0.13, 9, PRIMS, 0.01, aligned=True)
    scene, rk = compare(scene, ref, 300)
    _, counts = np.unique(rk["shape_ids"], return_counts=True)
    assert counts.max() >= 4, counts.max()


@pytest.mark.parametrize("algorithm", [PRIMS, HYBRID])
@pytest.mark.parametrize("aligned", [2, False])
def test_box_cylshell(ref, algorithm, aligned):
    """box_cylshell (ChNarrowphasePRIMS.cpp:719-915): analytic only when the shell's axis is parallel or perpendicular
    to a box face (up to 8 contacts); any other configuration makes PRIMSCollision return false, i.e. MPR in HYBRID
    mode and no contact in PRIMS mode."""
    scene = cluttered([BOX, CYLSHELL], 6, 0.13, 11, algorithm, 0.01, aligned=aligned)
    scene, rk = compare(scene, ref, 50 if (aligned or algorithm == HYBRID) else 0)
    if aligned:
        types = scene["shape_type"]
        seen = set(zip(types[rk["shape_ids"] >> 32], types[rk["shape_ids"] & 0xFFFFFFFF]))
        assert (BOX, CYLSHELL) in seen or (CYLSHELL, BOX) in seen, seen


@pytest.mark.parametrize("envelope", [0.0, 0.01])
@pytest.mark.parametrize("algorithm", [MPR, PRIMS, HYBRID])
def test_triangles_and_hulls(ref, algorithm, envelope):
    """Triangles (triangle_sphere, ChNarrowphasePRIMS.cpp:379-434; GetSupportPoint_Triangle / GetCenter_Triangle under
    MPR) and convex hulls (GetSupportPoint_Convex, GetCenter_Convex) among spheres, boxes and capsules.  A
    triangle-box pair reserves 6 slots but PRIMSCollision reports it unsupported (:1590-1603): MPR decides it in
    HYBRID mode, PRIMS mode gives no contact."""
    scene = cluttered([SPHERE, TRIANGLE, BOX, CONVEXHULL, SPHERE, CAPSULE, TRIANGLE], 7, 0.14, 21, algorithm, envelope)
    scene, rk = compare(scene, ref, 150)
    types = scene["shape_type"]
    seen = set(zip(types[rk["shape_ids"] >> 32], types[rk["shape_ids"] & 0xFFFFFFFF]))
    assert (SPHERE, TRIANGLE) in seen or (TRIANGLE, SPHERE) in seen, seen
    if algorithm != PRIMS:
        assert any(CONVEXHULL in p for p in seen), seen
        assert any(TRIANGLE in p and SPHERE not in p for p in seen), seen


def test_triangle_mesh_floor_step(ref):
    """A bed of spheres dropped on a triangulated floor (the tire-proxy / mesh-terrain use of triangle_sphere): contact
    list vs the reference, then one SMC step vs the reference force function on that list (1e-10 relative)."""
    from oracle import np_oracle

    b = scenes.SceneBuilder()
    rng = np.random.default_rng(5)
    floor = b.add_bodies(1.0, (1, 1, 1), [[0, 0, 0]], active=0)
    n, s = 8, 0.05
    z = rng.uniform(-0.004, 0.004, (n + 1, n + 1))
    for i in range(n):
        for j in range(n):
            p = lambda a, c: [a * s, c * s, z[a, c]]
            b.add_triangle(floor, p(i, j), p(i + 1, j), p(i + 1, j + 1))
            b.add_triangle(floor, p(i, j), p(i + 1, j + 1), p(i, j + 1))
    r = 0.02
    pos = [[(i + 0.5) * 0.041 + 0.02, (j + 0.5) * 0.041 + 0.02, r - 0.003 + 0.002 * rng.random()] for i in range(8)
           for j in range(8)]
    ids = b.add_bodies(0.01, (1e-5, 1e-5, 1e-5), pos) + np.arange(len(pos))
    b.add_shapes(SPHERE, ids, [[r, 0, 0, 0]])
    settings = dict(system_type=1, step_size=1e-4, gravity=(0, 0, -9.81), broadphase_grid=2, grid_density=5.0,
                    narrowphase_algorithm=HYBRID, collision_envelope=0.0)
    scene = b.finish([scenes.smc_material()], settings)
    scene, rk = compare(scene, ref, 60)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoStepDynamics()
    _, _, vel = sys.GetBodyState()
    ref.set_scene(scene)
    p, q, v, info = np_oracle.smc_step(scene, ref, scene["pos"], scene["rot"], scene["vel"])
    assert info["contacts"]["depth"].size == sys.GetNumContacts()
    err = np.abs(vel - v).max() / np.abs(v).max()
    assert err < 1e-10, err


def test_spheres_only_under_mpr(ref):
    """Algorithm::MPR sends sphere-sphere and sphere-box pairs through MPR as well (ChNarrowphase.cpp:232-261)."""
    scene = scenes.settling_smc(10, 10, 4, spacing_factor=1.98, jitter_factor=0.03)
    scene["pos"][1:, 2] -= 0.0005
    scene["settings"]["narrowphase_algorithm"] = MPR
    compare(scene, ref, 100)


def test_mixed_step_matches_oracle_forces(ref):
    """One SMC step of the mixed scene: contact forces from the reference force function on the reference contact
    list vs the CUDA step (1e-10 relative)."""
    from oracle import np_oracle

    scene = cluttered([SPHERE, BOX, CAPSULE], 6, 0.16, 4, HYBRID, 0.0)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoStepDynamics()
    _, _, vel = sys.GetBodyState()
    ref.set_scene(scene)
    p, q, v, info = np_oracle.smc_step(scene, ref, scene["pos"], scene["rot"], scene["vel"])
    assert info["contacts"]["depth"].size == sys.GetNumContacts()
    err = np.abs(vel - v).max() / np.abs(v).max()
    assert err < 1e-10, err


def test_multistep_history_with_multi_contact_pairs(ref):
    """MultiStep history when one shape pair has several contacts (box-box, capsule pairs): they share one history
    slot and the reference visits them in list order -- the CUDA kernel walks such a run in one thread."""
    from oracle import np_oracle

    scene = cluttered([SPHERE, BOX, CAPSULE, BOX], 5, 0.17, 8, HYBRID, 0.0, aligned=True)
    scene["settings"].update(tangential_displ_mode=2, step_size=1e-4)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.5, restitution=0.3)])
    rng = np.random.default_rng(2)
    scene["vel"][1:, :3] = rng.normal(scale=0.05, size=(scene["mass"].size - 1, 3))
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    ref.set_scene(scene)
    hist = ref.new_shear_history(scene["mass"].size)
    p, q, v = scene["pos"].copy(), scene["rot"].copy(), scene["vel"].copy()
    multi = 0
    for step in range(8):
        sys.SetBodyState(0, p, q, v)
        sys.DoStepDynamics()
        _, _, gv = sys.GetBodyState()
        p, q, v, info = np_oracle.smc_step(scene, ref, p, q, v, history=hist)
        _, counts = np.unique(info["contacts"]["shape_ids"], return_counts=True)
        multi = max(multi, counts.max())
        err = np.abs(gv - v).max() / np.abs(v).max()
        assert err < 1e-10, (step, err)
    assert multi >= 2


@pytest.mark.parametrize("algorithm", [MPR, HYBRID])
def test_complex_pairs_in_long_bins(ref, algorithm):
    """Bins with more than 64 entries keep neither pair masks nor complex-pair records on the device: pairs without a
    sphere are then evaluated in place by the count and store passes.  Same list as the reference, bit for bit."""
    scene = cluttered([SPHERE, BOX, CAPSULE, BOX], 7, 0.15, 13, algorithm, 0.01, grid_density=0.2)
    scene, rk = compare(scene, ref, 300)
    _, _, _, start = ref.bins()
    assert np.diff(start.astype(np.int64)).max() > 64
    _, counts = np.unique(rk["shape_ids"], return_counts=True)
    assert algorithm == MPR or counts.max() >= 2


def test_bodies_and_shapes_added_between_steps():
    """ChSystemMulticore::AddBody after the system has stepped (MC/physics/ChSystemMulticore.cpp:205-226): a system
    that receives its bodies and shapes in two batches, with a collision pass in between, ends up with the same
    contact list and takes the same step as one that got them all at once."""
    scene = cluttered([SPHERE, BOX, CAPSULE, TRIANGLE, CONVEXHULL, SPHERE], 6, 0.15, 17, HYBRID, 0.0)
    whole = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    nb = scene["mass"].size
    nb1 = nb // 2
    sel1, sel2 = scene["shape_body"] < nb1, scene["shape_body"] >= nb1
    part = cw.ChSystemMulticoreSMC()
    st = part.GetSettings()
    for k, v in scene["settings"].items():
        cur = getattr(st, k)
        if hasattr(cur, "__len__"):
            cur[:] = list(v)
        else:
            setattr(st, k, v)
    part.SetMaterials(scene["materials"])
    part.AddShapeVertices(scene["shape_verts"])
    for lo, hi, sel in ((0, nb1, sel1), (nb1, nb, sel2)):
        part.AddBodies(scene["mass"][lo:hi], scene["inertia"][lo:hi], scene["pos"][lo:hi], scene["rot"][lo:hi],
                       scene["vel"][lo:hi], scene["active"][lo:hi], scene["collide"][lo:hi])
        part.AddShapes(scene["shape_type"][sel], scene["shape_body"][sel], scene["shape_dims"][sel],
                       scene["shape_pos"][sel], scene["shape_rot"][sel], scene["shape_fam"][sel], scene["shape_mat"][sel])
        if lo == 0:  # the system steps with the first batch only, which is then put back where it started
            part.DoStepDynamics()
            part.SetBodyState(0, scene["pos"][:nb1], scene["rot"][:nb1], scene["vel"][:nb1])
    whole.RunCollision(7)
    part.SetBodyState(0, scene["pos"], scene["rot"], scene["vel"])
    part.RunCollision(7)
    a, b = whole.GetContacts(), part.GetContacts()
    assert a["depth"].size > 200
    for key in KEYS:
        assert np.array_equal(a[key], b[key]), key
    whole.DoStepDynamics()
    part.DoStepDynamics()
    for x, y in zip(whole.GetBodyState(), part.GetBodyState()):
        assert np.array_equal(x, y)
