"""Settings the reference applies around the hot path: the active bounding box of the collision system and the
per-body speed clamp.  Both are restated here in numpy from the reference lines they follow; the shape boxes the
active-box test works on come from the reference's own AABB generator (oracle/_ref)."""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

pytestmark = pytest.mark.gpu


def overlapping_bodies(mn, mx, shape_body, nb, lo, hi):
    """GetOverlappingAABB (ChCollisionSystemMulticore.cpp:547-562): inclusive box test per shape, OR per body."""
    hit = np.all((lo <= mx) & (mn <= hi), axis=1)
    out = np.zeros(nb, dtype=bool)
    np.logical_or.at(out, shape_body[hit], True)
    return out


def test_active_box_deactivates_bodies_outside(ref):
    scene = scenes.settling_smc(10, 10, 4, spacing_factor=2.2, jitter_factor=0.05, seed=4)
    nb = scene["mass"].size
    lo, hi = np.array([-0.05, -0.04, 0.0]), np.array([0.06, 0.05, 0.05])
    ref.set_scene(scene)
    ref.run(1)
    mn, mx = ref.aabb()
    keep = overlapping_bodies(mn, mx, scene["shape_body"], nb, lo, hi)
    act0, col0 = scene["active"].astype(bool), scene["collide"].astype(bool)
    expect = np.where(act0 & col0, keep, act0)  # :297-303
    assert 0 < (expect & act0).sum() < act0.sum()

    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.SetActiveBoundingBox(lo, hi)
    sys.Update()
    sys.RunCollision(7)
    act, col = sys.GetBodyFlags()
    assert np.array_equal(act, expect) and np.array_equal(col, col0)

    # the narrowphase sees the new flags in the same pass: no contact between two inactive bodies
    c = sys.GetContacts()
    a, b = c["body_ids"][:, 0], c["body_ids"][:, 1]
    assert np.all(act[a] | act[b])

    # deactivated bodies stay where they are, and stay deactivated once the box is switched off again
    p0, _, _ = sys.GetBodyState()
    sys.DoSteps(5)
    sys.SetActiveBoundingBox(None)
    sys.DoSteps(5)
    sys.Synchronize()
    p1, _, v1 = sys.GetBodyState()
    frozen = act0 & ~expect
    assert np.array_equal(p1[frozen], p0[frozen])
    moving = expect & act0
    assert np.all(np.any(p1[moving] != p0[moving], axis=1))
    act2, _ = sys.GetBodyFlags()
    assert np.array_equal(act2, expect)


def clamp_reference(q, v, max_speed, max_wvel):
    """ChBody::ClampSpeed (ChBody.cpp:319-329) on the quaternion rate the reference stores: rot_dt = 0.5 q (0, w)."""
    v = v.copy()
    for i in range(v.shape[0]):
        if max_speed[i] <= 0:
            continue
        w = v[i, 3:]
        e0, e1, e2, e3 = q[i]
        # quaternion product q * (0, w) / 2
        rdt = 0.5 * np.array([-e1 * w[0] - e2 * w[1] - e3 * w[2], e0 * w[0] + e2 * w[2] - e3 * w[1],
                              e0 * w[1] + e3 * w[0] - e1 * w[2], e0 * w[2] + e1 * w[1] - e2 * w[0]])
        wn = 2.0 * np.linalg.norm(rdt)
        if wn > max_wvel[i]:
            v[i, 3:] *= max_wvel[i] / wn
        sp = np.linalg.norm(v[i, :3])
        if sp > max_speed[i]:
            v[i, :3] *= max_speed[i] / sp
    return v


@pytest.mark.parametrize("system", ["smc", "nsc"])
def test_speed_clamp_matches_chbody(system):
    rng = np.random.default_rng(11)
    n = 64
    pos = np.zeros((n, 3))
    pos[:, 0] = np.arange(n) * 1.0  # far apart: no contacts, the step is force-free apart from gravity
    pos[:, 2] = 5.0
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    vel = np.concatenate([rng.normal(scale=0.6, size=(n, 3)), rng.normal(scale=6.0, size=(n, 3))], axis=1)
    cls = cw.ChSystemMulticoreSMC if system == "smc" else cw.ChSystemMulticoreNSC
    sys = cls()
    h = 1e-3
    sys.SetStep(h)
    sys.SetGravitationalAcceleration([0, 0, -9.81])
    sys.AddBodies(np.full(n, 2.0), np.full((n, 3), 0.01), pos, q, vel)
    sys.AddShapes(np.zeros(n, dtype=np.int32), np.arange(n), np.tile([0.05, 0, 0, 0], (n, 1)))
    if system == "smc":
        sys.SetMaterials(np.array([scenes.smc_material()]))
    else:
        sys.SetMaterials(np.array([scenes.nsc_material()]))
    max_speed = np.where(np.arange(n) % 3 == 0, -1.0, 0.5)  # every third body: SetLimitSpeed(false)
    max_wvel = np.full(n, 2.0 * np.pi)
    sys.SetSpeedLimits(0, max_speed, max_wvel)
    sys.DoSteps(1)
    sys.Synchronize()
    p1, q1, v1 = sys.GetBodyState()
    free = vel.copy()
    free[:, 2] -= 9.81 * h
    expect = clamp_reference(q, free, max_speed, max_wvel)
    assert np.any(np.abs(expect - free) > 1e-3)
    assert np.allclose(v1, expect, rtol=1e-13, atol=1e-15)
    # ChSystemMulticore.cpp:145-146: the pose moves with the solver's velocity, then the stored velocity is clamped
    assert np.allclose(p1, pos + h * free[:, :3], rtol=1e-13, atol=1e-15)
    # the graph-replayed steps keep clamping
    sys.DoSteps(4)
    sys.Synchronize()
    _, _, v5 = sys.GetBodyState()
    lim = max_speed > 0
    assert np.all(np.linalg.norm(v5[lim, :3], axis=1) <= 0.5 * (1 + 1e-12))
    assert np.all(np.linalg.norm(v5[lim, 3:], axis=1) <= 2 * np.pi * (1 + 1e-12))
    sys.SetSpeedLimits(0, None, None)
    sys.DoSteps(200)
    sys.Synchronize()
    _, _, v6 = sys.GetBodyState()
    assert np.all(np.linalg.norm(v6[:, :3], axis=1) > 0.5)
