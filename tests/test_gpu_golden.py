"""The reference's own known-answer vectors through the CUDA narrowphase (two-body scenes through the C ABI).

* tests/golden/prims_ref_tests.json  -- utest_COLL_narrow_prims.cpp, 1e-10 (Algorithm::PRIMS)
* tests/golden/mpr_ref_tests.json    -- utest_COLL_narrow_mpr.cpp, 5e-6   (Algorithm::MPR)

Every case is also run through the reference's own pipeline (oracle/_ref) on the same scene, and the CUDA contact list
must equal it bit for bit -- including the hand-picked edge cases (separation 0 / 0.1, exactly touching shapes).
"""
import json
import os

import numpy as np
import pytest

import chrono_wisc_b200 as cw
from test_oracle_cpu import GOLD, single_pair_scene

pytestmark = pytest.mark.gpu
KEYS = ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius")


def both(scene, ref):
    s = cw.load_scene(cw.ChSystemMulticoreNSC(), {**scene, "materials": np.array([cw.scenes.nsc_material()])})
    s.RunCollision(7)
    ref.set_scene(scene)
    ref.run(7)
    k, rk = s.GetContacts(), ref.contacts()
    assert s.GetCounts().num_possible_collisions == ref.counts()["N_p"]
    for key in KEYS:
        assert np.array_equal(k[key], rk[key]), key
    return k, ref.counts()["N_p"]


def test_prims_unit_test_vectors_through_cuda(ref):
    cases = json.load(open(os.path.join(GOLD, "prims_ref_tests.json")))
    checked = 0
    for c in cases:
        A, B = c["A"], c["B"]
        scene = single_pair_scene(A["type"], A["pos"], A["rot"], A["dims"], B["type"], B["pos"], B["rot"], B["dims"],
                                  c["separation"])
        k, n_pairs = both(scene, ref)
        if n_pairs == 0 and c["nC"] > 0:
            # the pair never reaches the narrowphase of the full pipeline (the reference's capsule AABB extends along
            # local Y while its narrowphase axis is local Z): nothing to compare with the vector
            continue
        if c["separation"] == 0 and c["nC"] == 0 and c["name"].startswith("ss_touch"):
            continue  # exact touching without envelope sits on the AABB-overlap boundary of the full pipeline
        assert k["depth"].size == c["nC"], c["name"]
        for key, v in c["expect"].items():
            assert np.allclose(k[key][0], v, rtol=0, atol=1e-10), (c["name"], key)
        checked += 1
    assert checked >= 25


def test_mpr_unit_test_vectors_through_cuda(ref):
    cases = json.load(open(os.path.join(GOLD, "mpr_ref_tests.json")))
    checked = 0
    for c in cases:
        A, B = c["A"], c["B"]
        scene = single_pair_scene(A["type"], A["pos"], A["rot"], A["dims"], B["type"], B["pos"], B["rot"], B["dims"], 0.0)
        scene["settings"]["narrowphase_algorithm"] = 0  # Algorithm::MPR
        k, n_pairs = both(scene, ref)
        if c["hit"] and n_pairs == 0:
            continue  # exactly touching boxes of the two shapes: not a candidate pair of the full pipeline
        assert (k["depth"].size > 0) == c["hit"], c["name"]
        if c["expect"] and c["hit"]:
            n, p1, p2 = k["normal"][0], k["pt_a"][0], k["pt_b"][0]
            d = float(np.dot(n, p2 - p1)) if c["depth_is_dot"] else float(k["depth"][0])
            assert np.allclose(n, c["expect"]["normal"], rtol=0, atol=c["tol"]), c["name"]
            assert np.allclose(p1, c["expect"]["pt_a"], rtol=0, atol=c["tol"]), c["name"]
            assert np.allclose(p2, c["expect"]["pt_b"], rtol=0, atol=c["tol"]), c["name"]
            assert abs(d - c["expect"]["depth"]) <= c["tol"], c["name"]
        checked += 1
    assert checked >= 12


@pytest.mark.slow
def test_benchmark_bed_lists_bit_exact(ref):
    """The 1,000,000-sphere bench bed after a short run: candidate pair list and contact list (3+ M contacts, long bins,
    wall bins, 32-bit offsets at full size) equal to the reference's own collision code, in order, bit for bit."""
    from chrono_wisc_b200 import scenes

    scene = scenes.settling_smc(250, 250, 16, seed=12345, spacing_factor=2.0, jitter_factor=0.005, grid="bin_size",
                                bin_edge=2.5)
    s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    s.DoSteps(60)
    pos, rot, vel = s.GetBodyState()
    s.RunCollision(7)
    state = dict(scene, pos=pos, rot=rot, vel=vel)
    ref.set_num_threads(os.cpu_count() or 1)
    ref.set_scene(state)
    ref.run(7)
    c, rc = s.GetCounts(), ref.counts()
    assert (c.num_bin_aabb_intersections, c.num_active_bins, c.num_possible_collisions, c.num_rigid_contacts) == \
           (rc["N_i"], rc["N_a"], rc["N_p"], rc["N_c"])
    assert rc["N_c"] > 2_000_000
    assert np.array_equal(s.GetPairs(), ref.pairs())
    k, rk = s.GetContacts(), ref.contacts()
    for key in KEYS:
        assert np.array_equal(k[key], rk[key]), key
