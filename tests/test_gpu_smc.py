"""CUDA SMC force / accumulation / integration vs the reference's own force function and the CPU restatement of
the glue around it, on identical states.

Tolerances (BASELINE.json north_star): contact lists bit-exact; per-contact forces, per-body sums and velocities
within 1e-10 relative (libm pow/log/sin/cos differ from glibc in the last ulp); observables over many steps within
1e-6 relative for short runs (trajectories are chaotic over long ones).
"""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import np_oracle as no

pytestmark = pytest.mark.gpu

RTOL = 1e-10


def bed(nx, ny, nz, seed=1, **kw):
    s = scenes.settling_smc(nx, ny, nz, spacing_factor=1.98, jitter_factor=0.03, seed=seed, **kw)
    s["pos"][1:, 2] -= 0.025 * 0.02
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.05, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    q = rng.normal(size=(n - 1, 4))
    s["rot"][1:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    return s


def relerr(a, b):
    scale = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / scale


def reference_contact_forces(scene, ref, c):
    sa, sb = (c["shape_ids"] >> 32), (c["shape_ids"] & 0xFFFFFFFF)
    mats = no.composite_smc(scene["materials"], scene["shape_mat"][sa], scene["shape_mat"][sb])
    return ref.smc_contact_forces(c, mats, scene["settings"], scene["mass"], scene["pos"], scene["rot"], scene["vel"])


@pytest.mark.parametrize("model", [cw.system.HERTZ, cw.system.HOOKE, cw.system.FLORES, cw.system.PLAIN_COULOMB])
@pytest.mark.parametrize("use_mat", [1, 0])
def test_contact_forces_match_reference_function(model, use_mat, ref):
    scene = bed(12, 12, 5)
    scene["settings"].update(contact_force_model=model, use_material_properties=use_mat)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.4, restitution=0.4, rolling=0.05,
                                                       spinning=0.03, adhesion=0.01, kn=1e5, kt=4e4, gn=30, gt=10)])
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    ref.set_scene(scene)
    ref.run(7)
    c = ref.contacts()
    assert c["depth"].size > 500
    bid, f, t = reference_contact_forces(scene, ref, c)
    gf, gt = sys.GetContactForces()
    assert relerr(gf, f) < RTOL
    assert relerr(gt, t) < RTOL


@pytest.mark.parametrize("adhesion", [cw.system.ADHESION_CONSTANT, cw.system.ADHESION_DMT, cw.system.ADHESION_PERKO])
def test_adhesion_models(adhesion, ref):
    scene = bed(8, 8, 4, seed=3)
    scene["settings"].update(adhesion_force_model=adhesion)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, adhesion=0.02, dmt=0.5, perko=3.0)])
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    ref.set_scene(scene)
    ref.run(7)
    bid, f, t = reference_contact_forces(scene, ref, ref.contacts())
    gf, gt = sys.GetContactForces()
    assert relerr(gf, f) < RTOL and relerr(gt, t) < RTOL


def test_two_materials_composite(ref):
    scene = bed(10, 10, 4, seed=4)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.4, restitution=0.4),
                                   scenes.smc_material(young=5e5, poisson=0.25, friction=0.2, restitution=0.7)])
    scene["shape_mat"] = (np.arange(scene["shape_type"].size) % 2).astype(np.int32)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    ref.set_scene(scene)
    ref.run(7)
    bid, f, t = reference_contact_forces(scene, ref, ref.contacts())
    gf, gt = sys.GetContactForces()
    assert relerr(gf, f) < RTOL and relerr(gt, t) < RTOL


def test_one_step_against_cpu(ref):
    """One whole step (update, collision, forces, accumulation, velocity update, integration)."""
    scene = bed(14, 14, 5, seed=7)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoStepDynamics()
    ref.set_scene(scene)
    p, q, v, info = no.smc_step(scene, ref, scene["pos"], scene["rot"], scene["vel"], fast=False)
    gp, gq, gv = sys.GetBodyState()
    # per-body contact sums incl. the container body (summed by the masked block reduction)
    bf, bt = sys.GetBodyContactForces()
    assert relerr(bf, info["body_force"]) < RTOL
    assert relerr(bt, info["body_torque"]) < RTOL
    assert np.abs(bf[0]).max() > 0, "container must feel the bed"
    assert relerr(sys.GetHf(), info["hf"]) < RTOL
    assert relerr(gv, v) < RTOL
    assert relerr(gp, p) < 1e-12
    assert relerr(gq, q) < 1e-13
    nb_touched = int((np.bincount(info["contacts"]["body_ids"].ravel(), minlength=scene["mass"].size) > 0).sum())
    assert sys.GetCounts().ct_body_count == nb_touched


def test_free_fall_matches_closed_form():
    """utest_MCORE_gravity (src/tests/unit_tests/multicore/utest_MCORE_gravity.cpp:27-49): semi-implicit Euler free
    fall, 1000 steps of 1e-5 s."""
    b = scenes.SceneBuilder()
    b.add_bodies(1.0, (1, 1, 1), [[0, 0, 0]], active=0)
    b.add_shapes(cw.system.BOX, [0], [[1, 1, 0.1]], pos=[[0, 0, -5]])
    bid = b.add_bodies(1.0, (0.4, 0.4, 0.4), [[0, 0, 0]])
    b.add_shapes(cw.system.SPHERE, [bid], [[1.0]])
    h, n = 1e-5, 1000
    scene = b.finish([scenes.smc_material()], dict(system_type=1, step_size=h, gravity=(0, -9.80665, 0)))
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoSteps(n)
    p, q, v = sys.GetBodyState()
    t = n * h
    assert abs(v[1, 1] - (-9.80665 * t)) < 1e-9
    # symplectic Euler: y_n = -g h^2 n(n+1)/2
    assert abs(p[1, 1] - (-9.80665 * h * h * n * (n + 1) / 2)) < 1e-12
    assert np.array_equal(p[0], [0, 0, 0])


def test_short_run_tracks_cpu(ref):
    """25 steps in lock-step with the CPU path: contact counts equal every step, state within 1e-8."""
    scene = bed(8, 8, 4, seed=9)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    ref.set_scene(scene)
    p, q, v = scene["pos"].copy(), scene["rot"].copy(), scene["vel"].copy()
    for step in range(25):
        sys.DoStepDynamics()
        p, q, v, info = no.smc_step(scene, ref, p, q, v)
        assert sys.GetNumContacts() == info["contacts"]["depth"].size, step
    gp, gq, gv = sys.GetBodyState()
    assert relerr(gp, p) < 1e-8 and relerr(gv, v) < 1e-6


def test_weight_on_container_after_settling():
    """utest_MCORE_contact_forces-style acceptance (src/tests/unit_tests/multicore/utest_MCORE_contact_forces.cpp:
    154-177): once a small bed is at rest, the container's contact force equals the total weight (rel 1e-3)."""
    scene = scenes.settling_smc(4, 4, 2, spacing_factor=2.02, young=2e6, cr=0.1)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoSteps(8000)
    f, t = sys.GetBodyContactForces()
    weight = scene["mass"][1:].sum() * 9.81
    p, q, v = sys.GetBodyState()
    print("settling: max |v| =", np.abs(v[1:, :3]).max(), " F_z/W =", -f[0, 2] / weight)
    assert np.abs(v[1:, :3]).max() < 5e-2
    assert abs(-f[0, 2] - weight) / weight < 2e-2
    # Newton's third law over the whole contact set
    assert np.abs(f.sum(0)).max() < 1e-9 * weight + 1e-12


def test_determinism():
    scene = bed(16, 16, 6, seed=2)
    outs = []
    for _ in range(2):
        sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
        sys.DoSteps(20)
        outs.append(sys.GetBodyState())
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


def test_step_host_equals_upload_step_download():
    """cb200_step_host (host state in, one step, host state out, velocity upload overlapped with the collision stages)
    returns exactly what upload_state -> do_steps(1) -> download_state returns."""
    scene = scenes.settling_smc(24, 24, 6, spacing_factor=1.98, jitter_factor=0.03)
    a = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    b = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    a.DoSteps(5)
    b.DoSteps(5)
    p, q, v = (np.ascontiguousarray(x) for x in a.GetBodyState())
    for _ in range(3):
        a.StepHost(p, q, v)
        b.DoSteps(1)
    pb, qb, vb = b.GetBodyState()
    assert np.array_equal(p, pb) and np.array_equal(q, qb) and np.array_equal(v, vb)
    assert a.GetNumContacts() == b.GetNumContacts() > 0


@pytest.mark.parametrize("model", [cw.system.HERTZ, cw.system.HOOKE, cw.system.FLORES])
def test_multistep_tangential_history(model, ref):
    """ChSystemSMC::TangentialDisplacementModel::MultiStep (...SMC.cpp:188-250, :460-467, :687-695): per-body contact
    history (accumulated shear displacement, initial approach velocity, contact duration).  The CUDA system and the
    reference force function step the same state for 30 steps, each keeping its own history; velocities agree to
    1e-10 after every step (the state is re-synchronised each step, the histories never are)."""
    scene = bed(9, 9, 4, seed=7)
    scene["settings"].update(contact_force_model=model, tangential_displ_mode=2, step_size=2e-4)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.5, restitution=0.3, rolling=0.04,
                                                       spinning=0.02)])
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    ref.set_scene(scene)
    nb = scene["mass"].size
    hist = ref.new_shear_history(nb)
    p, q, v = scene["pos"].copy(), scene["rot"].copy(), scene["vel"].copy()
    worst = 0.0
    for step in range(30):
        sys.SetBodyState(0, p, q, v)
        sys.DoStepDynamics()
        gp, gq, gv = sys.GetBodyState()
        p, q, v, info = no.smc_step(scene, ref, p, q, v, history=hist)
        assert info["contacts"]["depth"].size == sys.GetNumContacts()
        worst = max(worst, relerr(gv, v), relerr(gp, p))
        assert worst < RTOL, (step, worst)
    # persistent contacts exist and carry accumulated displacement
    live = hist["neigh"][:, 0] >= 0
    assert live.sum() > 200 and (hist["duration"][live] > 0).any() and np.abs(hist["disp"][live]).max() > 0
    # and the history matters: the same state stepped with OneStep gives different velocities
    st1 = dict(scene, settings=dict(scene["settings"], tangential_displ_mode=1))
    ref.set_scene(st1)
    _, _, v1, _ = no.smc_step(st1, ref, gp, gq, gv)
    _, _, v2, _ = no.smc_step(scene, ref, gp, gq, gv, history=hist)
    assert relerr(v1, v2) > 1e-6


def test_multistep_graph_replay_equals_single_steps():
    """DoSteps replays a CUDA graph of the step (history memset, force kernel, slot sweep included): same bits as
    stepping one call at a time; slot assignment by atomics does not leak into the forces."""
    scene = bed(9, 9, 4, seed=11)
    scene["settings"].update(tangential_displ_mode=2, step_size=2e-4)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.5, restitution=0.3)])
    a = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    b = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    a.DoSteps(24)
    for _ in range(24):
        b.DoStepDynamics()
    for x, y in zip(a.GetBodyState(), b.GetBodyState()):
        assert np.array_equal(x, y)


def test_contacts_after_graph_replay_are_current(monkeypatch):
    """A contact download between graph-replayed steps returns the records of the LAST step's key list (the records are
    produced on demand from the collision stage's shape records, which the next step's preparation must not touch),
    and stepping through the graph equals stepping without it bit for bit."""
    def run(no_graph):
        if no_graph:
            monkeypatch.setenv("CB200_NO_GRAPH", "1")
        else:
            monkeypatch.delenv("CB200_NO_GRAPH", raising=False)
        scene = scenes.settling_smc(14, 14, 5, spacing_factor=1.98, jitter_factor=0.03)
        s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
        out = []
        for n in (12, 1, 9):
            s.DoSteps(n)
            k = s.GetContacts()
            out.append({key: np.array(v, copy=True) for key, v in k.items()})
            out.append(s.GetBodyState())
        return out

    a, b = run(False), run(True)
    for x, y in zip(a, b):
        if isinstance(x, dict):
            assert x["depth"].size > 0
            for key in x:
                assert np.array_equal(x[key], y[key]), key
        else:
            for u, v in zip(x, y):
                assert np.array_equal(u, v)
