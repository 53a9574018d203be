"""Contact callbacks: NarrowphaseCallback drops contacts, AddContactCallback rewrites the composite
material per contact; the forces that follow are checked against the reference's own force function fed with the same
per-contact materials."""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import np_oracle as no

pytestmark = pytest.mark.gpu


def bed(seed=9):
    s = scenes.settling_smc(10, 10, 4, spacing_factor=1.98, jitter_factor=0.03, seed=seed)
    s["pos"][1:, 2] -= 0.025 * 0.02
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.05, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    s["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.6, restitution=0.4)])
    return s


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_add_contact_callback_changes_friction_like_demo_callback(ref):
    """demo_MCORE_callbackSMC.cpp:82-93: friction 0.3 / 0.8 by the sign of a coordinate of cinfo.vpA."""
    scene = bed()
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    seen = []

    def on_add(cinfo, mat):
        seen.append(cinfo["vpA"].copy())
        mat[0] = 0.3 if cinfo["vpA"][0] > 0 else 0.8  # mu_eff

    sys.RegisterAddContactCallback(on_add)
    sys.Update()
    sys.RunCollision(7)
    sys.RunContactCallbacks()
    sys.RunSolver()
    c = sys.GetContacts()
    n = c["depth"].size
    assert n > 300 and len(seen) == n
    # the reference adds the body position to the contact point it shows this callback (…SMC.cpp:149-150)
    assert np.array_equal(np.array(seen), c["pt_a"] + scene["pos"][c["body_ids"][:, 0]])
    sa, sb = (c["shape_ids"] >> 32), (c["shape_ids"] & 0xFFFFFFFF)
    mats = no.composite_smc(scene["materials"], scene["shape_mat"][sa], scene["shape_mat"][sb])
    mu = np.where(np.array(seen)[:, 0] > 0, 0.3, 0.8)
    assert 0 < (mu == 0.3).sum() < n
    mats = override_friction(mats, mu)
    bid, f, t = ref.smc_contact_forces(c, mats, scene["settings"], scene["mass"], scene["pos"], scene["rot"],
                                       scene["vel"])
    gf, gt = sys.GetContactForces()
    assert relerr(gf, f) < 1e-10 and relerr(gt, t) < 1e-10
    # without the callback the forces differ
    sys2 = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys2.Update(); sys2.RunCollision(7); sys2.RunSolver()
    gf2, _ = sys2.GetContactForces()
    assert relerr(gf2, f) > 1e-4
    # and the override does not outlive its contact list
    import ctypes as C

    from chrono_wisc_b200._lib import check

    def composites():
        stride, m = C.c_int(), np.empty((sys.GetCounts().num_rigid_contacts, 13))
        check(sys._lib.cb200_download_contact_materials(sys._h, m.shape[0], m.ctypes.data_as(C.c_void_p), C.byref(stride)))
        assert stride.value == 13
        return m

    assert np.array_equal(composites()[:, 0], mu)
    sys.RegisterAddContactCallback(None)
    sys.RunCollision(7)
    assert np.all(composites()[:, 0] == np.float32(0.6))


def override_friction(mats, mu):
    """composite_smc returns the per-contact arrays of host_data (fric_rigid_rigid = mu, muRoll, muSpin); friction is
    the field the demo's callback changes."""
    out = {k: v.copy() for k, v in mats.items()}
    out["fric"][:, 0] = mu
    return out


def test_narrowphase_callback_drops_contacts(ref):
    scene = bed(seed=10)
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.Update()
    sys.RunCollision(7)
    full = sys.GetContacts()
    drop = lambda ci: not (ci["bodyA"] == 0 or ci["bodyB"] == 0 or ci["bodyA"] % 3 == 0)  # noqa: E731
    sys.RegisterNarrowphaseCallback(drop)
    sys.RunContactCallbacks()
    sys.RunSolver()
    c = sys.GetContacts()
    keep = np.array([drop(dict(bodyA=a, bodyB=b)) for a, b in full["body_ids"]])
    assert 0 < keep.sum() < keep.size
    for k in ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius"):
        assert np.array_equal(c[k], full[k][keep]), k
    sa, sb = (c["shape_ids"] >> 32), (c["shape_ids"] & 0xFFFFFFFF)
    mats = no.composite_smc(scene["materials"], scene["shape_mat"][sa], scene["shape_mat"][sb])
    bid, f, t = ref.smc_contact_forces(c, mats, scene["settings"], scene["mass"], scene["pos"], scene["rot"],
                                       scene["vel"])
    gf, gt = sys.GetContactForces()
    assert relerr(gf, f) < 1e-10
    # bodies that only had dropped contacts feel no contact force: the floor (body 0) among them
    bf, _ = sys.GetBodyContactForces()
    assert np.all(bf[0] == 0)


def test_callbacks_inside_do_steps_and_nsc():
    scene = scenes.balls_nsc(6, 6, 3, iterations=60)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    calls = [0]

    def on_add(cinfo, mat):
        calls[0] += 1
        mat[0] = 0.0  # frictionless

    sys.RegisterAddContactCallback(on_add)
    sys.DoSteps(60)
    sys.Synchronize()
    assert calls[0] > 0
    r = sys.GetContactReport()
    loaded = np.abs(r["force"][:, 0]) > 1e-9
    assert loaded.any()
    assert np.abs(r["force"][:, 1:]).max() <= 1e-12 * np.abs(r["force"][:, 0]).max()  # no tangential multipliers
