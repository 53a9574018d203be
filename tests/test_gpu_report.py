"""Contact reporting: the batch of OnReportContact arguments computed on the device against the numpy
restatement of ChContactContainerMulticore::ReportAllContacts (oracle/np_oracle.py), SMC and NSC."""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import np_oracle as no

pytestmark = pytest.mark.gpu


def close(a, b, tol=1e-12, scale=None):
    return np.abs(a - b).max() <= tol * max(np.abs(b).max() if scale is None else scale, 1e-300)


def test_contact_plane_fallbacks_cpu_restatement():
    for n in ([0, 1, 0], [0, -1, 0], [1, 0, 0], [0.3, 0.9, -0.2]):
        P = no.contact_plane(np.array(n, float) / np.linalg.norm(n))
        assert np.allclose(P.T @ P, np.eye(3), atol=1e-14) and np.linalg.det(P) > 0.999


def smc_system():
    s = scenes.settling_smc(10, 10, 4, spacing_factor=1.98, jitter_factor=0.03, seed=5)
    s["pos"][1:, 2] -= 0.025 * 0.02
    rng = np.random.default_rng(5)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.05, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    q = rng.normal(size=(n - 1, 4))
    s["rot"][1:] = q / np.linalg.norm(q, axis=1, keepdims=True)
    s["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.4, restitution=0.4)])
    return s


def test_report_smc_matches_reference_formulas():
    scene = smc_system()
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    c = sys.GetContacts()
    assert c["depth"].size > 300
    f, t = sys.GetContactForces()
    pos, rot, _ = sys.GetBodyState()
    want = no.report_contacts("smc", c, pos, rot, scene["settings"]["step_size"], ct_force=f.reshape(-1, 3),
                              ct_torque=t.reshape(-1, 3))
    got = sys.GetContactReport()
    assert np.array_equal(got["body_ids"], c["body_ids"])
    assert np.array_equal(got["pA"], c["pt_a"]) and np.array_equal(got["pB"], c["pt_b"])
    assert np.array_equal(got["distance"], c["depth"]) and np.array_equal(got["eff_radius"], c["eff_radius"])
    assert close(got["plane"], want["plane"], 1e-14)
    # a sphere contact has no torque about its own contact point: what is left is the rounding of r x F - r x F
    assert close(got["force"], want["force"]) and close(got["torque"], want["torque"], 1e-12, scale=np.abs(t).max())
    # X axis of the plane is the contact normal; the normal force component carries (almost) all of a sphere contact
    assert close(got["plane"][:, :, 0], c["normal"], 1e-15)

    seen = []

    def cb(pA, pB, plane, dist, erad, force, torque, bA, bB):
        seen.append((bA, bB, dist))
        return len(seen) < 7  # proceed == false stops the walk

    sys.ReportAllContacts(cb)
    assert len(seen) == 7 and seen[3][0] == c["body_ids"][3, 0] and seen[3][2] == c["depth"][3]


def test_report_after_whole_steps_uses_the_steps_contacts():
    scene = smc_system()
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.DoSteps(6)
    sys.Synchronize()
    c = sys.GetContacts()
    f, t = sys.GetContactForces()
    pos, rot, _ = sys.GetBodyState()
    want = no.report_contacts("smc", c, pos, rot, 1e-3, ct_force=f.reshape(-1, 3), ct_torque=t.reshape(-1, 3))
    got = sys.GetContactReport()
    assert close(got["force"], want["force"]) and close(got["torque"], want["torque"], 1e-12, scale=np.abs(t).max())


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_report_nsc_matches_reference_formulas(mode):
    scene = scenes.balls_nsc(8, 8, 3, iterations=60)
    its = [0, 0, 0]
    its[mode] = 60
    scene["settings"].update(solver_mode=mode, max_iteration_normal=its[0], max_iteration_sliding=its[1],
                             max_iteration_spinning=its[2])
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
    sys.DoSteps(80)  # until the balls rest on each other
    sys.Synchronize()
    c = sys.GetContacts()
    assert c["depth"].size > 50
    gamma = sys.GetGamma()
    rows = (1, 3, 6)[mode]
    pos, rot, _ = sys.GetBodyState()
    want = no.report_contacts("nsc", c, pos, rot, scene["settings"]["step_size"], gamma=gamma, rows=rows)
    got = sys.GetContactReport()
    assert close(got["plane"], want["plane"], 1e-14)
    assert np.array_equal(got["force"], want["force"]) and np.array_equal(got["torque"], want["torque"])
    assert np.abs(got["force"][:, 0]).max() > 0
