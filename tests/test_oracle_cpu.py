"""Pins the CPU checkers (no GPU needed).

1. the reference's own collision code (oracle/_ref) reproduces the known-answer cases transcribed from the
   reference's unit test utest_COLL_narrow_prims.cpp (tests/golden/prims_ref_tests.json, tolerance 1e-10 as in the
   reference test) and the committed outputs in tests/golden/prims_random_pairs.npz;
2. the CPU restatement (oracle/oracle.cpp + oracle/np_oracle.py) equals the reference's own code bit for bit on
   seeded scenes: AABBs, grid, pair list, contact list, per-contact SMC forces; and the two restatements of the
   step glue agree with each other.
"""
import json
import os

import numpy as np
import pytest

from chrono_wisc_b200 import scenes
from oracle import np_oracle as no
from oracle import pyoracle as po

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SPHERE, BOX, CAPSULE = 0, 2, 12


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


def single_pair_scene(tA, pA, qA, dA, tB, pB, qB, dB, sep):
    """Two one-shape bodies, one bin: runs a pair through the full pipeline (separation = 2 * envelope)."""
    b = scenes.SceneBuilder()
    b.add_bodies(1.0, (1, 1, 1), [pA], rot=[qA])
    b.add_bodies(1.0, (1, 1, 1), [pB], rot=[qB])
    b.add_shapes(tA, [0], [dA])
    b.add_shapes(tB, [1], [dB])
    return b.finish([scenes.smc_material()],
                    dict(system_type=0, collision_envelope=sep / 2, broadphase_grid=0, bins_per_axis=(1, 1, 1),
                         narrowphase_algorithm=1, step_size=1e-3, gravity=(0, 0, 0)))


# ---- 1. golden vectors -------------------------------------------------------------------------------------------
def test_reference_code_reproduces_its_unit_test_vectors(ref):
    cases = json.load(open(os.path.join(GOLD, "prims_ref_tests.json")))
    assert len(cases) >= 30
    for c in cases:
        A, B = c["A"], c["B"]
        nC, out = ref.prims_pair(A["type"], A["pos"], A["rot"], A["dims"], B["type"], B["pos"], B["rot"], B["dims"],
                                 c["separation"])
        assert nC == c["nC"], c["name"]
        for k, v in c["expect"].items():
            got = out[k][0]
            assert np.allclose(got, v, rtol=0, atol=1e-10), (c["name"], k, got, v)


def test_oracle_restatement_reproduces_unit_test_vectors():
    cases = json.load(open(os.path.join(GOLD, "prims_ref_tests.json")))
    for c in cases:
        A, B = c["A"], c["B"]
        if c["separation"] == 0 and c["nC"] == 0 and c["name"].startswith("ss_touch"):
            continue  # exact touching without envelope sits on the AABB-overlap boundary of the full pipeline
        o = po.Oracle(single_pair_scene(A["type"], A["pos"], A["rot"], A["dims"], B["type"], B["pos"], B["rot"],
                                        B["dims"], c["separation"]))
        o.collide()
        k = o.contacts()
        if A["type"] == CAPSULE and o.counts()["N_p"] == 0:
            # the reference's capsule AABB extends along local Y while its narrowphase axis is local Z
            # (SURVEY Appendix A.1), so the full pipeline never sees this pair; the direct call above covers it
            continue
        assert k["depth"].size == c["nC"], c["name"]
        for key, v in c["expect"].items():
            assert np.allclose(k[key][0], v, rtol=0, atol=1e-10), (c["name"], key)


def test_reference_code_matches_committed_random_pairs(ref):
    g = np.load(os.path.join(GOLD, "prims_random_pairs.npz"))
    for i in range(g["nC"].size):
        nC, out = ref.prims_pair(int(g["typeA"][i]), g["posA"][i], g["rotA"][i], g["dimsA"][i], int(g["typeB"][i]),
                                 g["posB"][i], g["rotB"][i], g["dimsB"][i], float(g["sep"][i]))
        assert nC == g["nC"][i]
        if nC == 1:
            for k in ("normal", "pt_a", "pt_b"):
                assert np.array_equal(out[k][0], g[k][i]), (i, k)
            assert out["depth"][0] == g["depth"][i] and out["eff_radius"][0] == g["eff_radius"][i]


def test_oracle_restatement_matches_committed_random_pairs():
    g = np.load(os.path.join(GOLD, "prims_random_pairs.npz"))
    hits = 0
    for i in range(g["nC"].size):
        o = po.Oracle(single_pair_scene(int(g["typeA"][i]), g["posA"][i], g["rotA"][i], g["dimsA"][i],
                                        int(g["typeB"][i]), g["posB"][i], g["rotB"][i], g["dimsB"][i],
                                        float(g["sep"][i])))
        o.collide()
        k = o.contacts()
        if g["nC"][i] == 1 and o.counts()["N_p"] == 1:
            hits += 1
            assert k["depth"].size == 1, i
            for key in ("normal", "pt_a", "pt_b", "depth", "eff_radius"):
                assert np.array_equal(k[key][0], g[key][i]), (i, key)
        elif g["nC"][i] == 0:
            assert k["depth"].size == 0, i
    assert hits > 250


# ---- 2. restatement == reference code on whole scenes --------------------------------------------------------------
SCENES = {
    "bed_density": lambda: bed(14, 14, 5),
    "bed_fixed_res": lambda: bed(16, 16, 4, grid="fixed", bins=(10, 10, 1)),
    "bed_bin_size": lambda: bed(12, 12, 6, grid="bin_size", bin_edge=2.0),
    "no_contacts": lambda: scenes.settling_smc(6, 6, 3),
}


@pytest.mark.parametrize("name", list(SCENES))
def test_oracle_collision_equals_reference_code(name, ref):
    scene = SCENES[name]()
    o = po.Oracle(scene)
    o.collide()
    ref.set_scene(scene)
    ref.run(7)
    assert o.counts() == ref.counts()
    g, rg = o.grid(), ref.grid()
    for k in ("min_bounding_point", "max_bounding_point", "bin_size", "inv_bin_size"):
        assert np.array_equal(g[k], rg[k]), k
    for a, b in zip(o.aabb(), ref.aabb()):
        assert np.array_equal(a, b)
    assert np.array_equal(o.pairs(), ref.pairs())
    k, rk = o.contacts(), ref.contacts()
    for key in rk:
        assert np.array_equal(k[key], rk[key]), key


@pytest.mark.parametrize("model", [0, 1, 2, 3])
@pytest.mark.parametrize("use_mat", [0, 1])
def test_oracle_forces_equal_reference_function(model, use_mat, ref):
    scene = bed(10, 10, 4, seed=5)
    scene["settings"].update(contact_force_model=model, use_material_properties=use_mat)
    scene["materials"] = np.array([scenes.smc_material(young=2e6, friction=0.4, rolling=0.05, spinning=0.03,
                                                       adhesion=0.01, kn=1e5, kt=4e4, gn=30, gt=10)])
    o = po.Oracle(scene)
    o.step(1)
    ref.set_scene(scene)
    ref.run(7)
    c = ref.contacts()
    sa, sb = c["shape_ids"] >> 32, c["shape_ids"] & 0xFFFFFFFF
    mats = no.composite_smc(scene["materials"], scene["shape_mat"][sa], scene["shape_mat"][sb])
    bid, f, t = ref.smc_contact_forces(c, mats, scene["settings"], scene["mass"], scene["pos"], scene["rot"],
                                       scene["vel"])
    of, ot = o.contact_forces()
    assert np.allclose(of, f, rtol=1e-14, atol=0) and np.allclose(ot, t, rtol=1e-13, atol=1e-300)


def test_oracle_step_equals_numpy_glue(ref):
    scene = bed(10, 10, 4, seed=6)
    o = po.Oracle(scene)
    o.step(1)
    ref.set_scene(scene)
    p, q, v, info = no.smc_step(scene, ref, scene["pos"], scene["rot"], scene["vel"], fast=False)
    op, oq, ov = o.state()
    bf, bt = o.body_forces()
    assert np.allclose(bf, info["body_force"], rtol=1e-14, atol=1e-300)
    assert np.allclose(o.hf(), info["hf"], rtol=1e-14, atol=1e-300)
    assert np.allclose(ov, v, rtol=1e-14, atol=1e-300)
    assert np.allclose(op, p, rtol=1e-15, atol=0) and np.allclose(oq, q, rtol=1e-15, atol=1e-18)


def test_composite_material_float_arithmetic():
    """ChContactMaterialCompositeSMC in float: E* of two equal materials is E / (2 (1 - nu^2)) evaluated in float."""
    t = np.array([scenes.smc_material(young=2e6, poisson=0.3), scenes.smc_material(young=5e5, poisson=0.25)])
    m = no.composite_smc(t, [0, 0, 1], [0, 1, 1])
    f = np.float32
    inv = (f(1) - f(0.3) * f(0.3)) / f(2e6) + (f(1) - f(0.3) * f(0.3)) / f(2e6)
    assert m["modulus"][0, 0] == float(f(1) / inv)
    assert m["modulus"].dtype == np.float64 and m["fric"][1, 0] == float(f(0.6))


def test_oracle_omp_threads_do_not_change_results():
    scene = bed(12, 12, 4, seed=8)
    outs = []
    for th in (1, 4):
        o = po.Oracle(scene, threads=th)
        o.step(3)
        outs.append(o.state())
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


# ---- world_size-2 host logic is covered in tests/test_slab_gloo.py -------------------------------------------------


def test_reference_code_reproduces_its_mpr_unit_test_vectors(ref):
    """tests/golden/mpr_ref_tests.json: transcribed from utest_COLL_narrow_mpr.cpp (tolerance 5e-6 as there)."""
    cases = json.load(open(os.path.join(GOLD, "mpr_ref_tests.json")))
    assert len(cases) >= 17
    for c in cases:
        A, B = c["A"], c["B"]
        hit, out = ref.mpr_pair(A["type"], A["pos"], A["rot"], A["dims"], B["type"], B["pos"], B["rot"], B["dims"],
                                c["envelope"])
        assert bool(hit) == c["hit"], c["name"]
        if c["expect"]:
            d = float(np.dot(out["normal"], out["pt_b"] - out["pt_a"])) if c["depth_is_dot"] else out["depth"]
            for k in ("normal", "pt_a", "pt_b"):
                assert np.allclose(out[k], c["expect"][k], rtol=0, atol=c["tol"]), (c["name"], k)
            assert abs(d - c["expect"]["depth"]) <= c["tol"], c["name"]
