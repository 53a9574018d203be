"""CUDA collision pipeline vs the reference's own collision code (oracle/_ref) on identical states.

Bar (BASELINE.json north_star): AABBs, grid, broadphase pair list and contact list are BIT-EXACT -- integer lists
equal after canonical sort (they are in fact emitted in the reference's order, which is asserted too), FP64 contact
geometry equal to the last bit.
"""
import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


def overlapping_bed(nx, ny, nz, **kw):
    """A lattice with spacing below one diameter and strong jitter: plenty of sphere-sphere and sphere-wall
    contacts, ragged bins."""
    s = scenes.settling_smc(nx, ny, nz, spacing_factor=1.98, jitter_factor=0.03, **kw)
    s["pos"][1:, 2] -= 0.025 * 0.02
    return s


def run_both(scene, ref):
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.RunCollision(7)
    ref.set_scene(scene)
    ref.run(7)
    return sys


CASES = {
    "tiny": lambda: overlapping_bed(4, 4, 2),
    "small_density": lambda: overlapping_bed(16, 16, 6),
    "fixed_10x10x1": lambda: overlapping_bed(24, 24, 4, grid="fixed", bins=(10, 10, 1)),
    "fixed_1x1x1": lambda: overlapping_bed(6, 6, 3, grid="fixed", bins=(1, 1, 1)),
    # bins of 65..128 entries (128-bit masks left by k_long_rows, keys written by k_pair_store) and of more than 128
    # (both walks by k_long_rows, records from global memory)
    "fixed_2x2x1_long": lambda: overlapping_bed(12, 12, 3, grid="fixed", bins=(2, 2, 1)),
    "fixed_1x1x1_longer": lambda: overlapping_bed(8, 8, 3, grid="fixed", bins=(1, 1, 1)),
    "fixed_3x3x1_mixed": lambda: overlapping_bed(16, 16, 3, grid="fixed", bins=(3, 3, 1)),
    "mid_density": lambda: overlapping_bed(48, 48, 10),
    "no_contacts": lambda: scenes.settling_smc(8, 8, 3),
}


@pytest.mark.parametrize("name", list(CASES))
def test_collision_bit_exact(name, ref):
    scene = CASES[name]()
    sys = run_both(scene, ref)
    c, rc = sys.GetCounts(), ref.counts()
    g = ref.grid()
    # grid
    assert tuple(c.bins_per_axis) == rc["bins"]
    assert np.array_equal(np.array(c.min_bounding_point), g["min_bounding_point"])
    assert np.array_equal(np.array(c.max_bounding_point), g["max_bounding_point"])
    assert np.array_equal(np.array(c.bin_size), g["bin_size"])
    # AABBs (relative to global_origin, as the reference leaves them)
    mn, mx = sys.GetAABB()
    rmn, rmx = ref.aabb()
    assert np.array_equal(mn, rmn) and np.array_equal(mx, rmx)
    # counters
    assert c.num_bin_aabb_intersections == rc["N_i"]
    assert c.num_active_bins == rc["N_a"]
    assert c.num_possible_collisions == rc["N_p"]
    assert c.num_rigid_contacts == rc["N_c"]
    # bins: (bin, shape) lists sorted by bin then shape, run-length encoding
    if rc["N_i"]:
        bn, ba, act, st = sys.GetBins()
        rbn, rba, ract, rst = ref.bins()
        assert np.array_equal(bn, rbn) and np.array_equal(ba, rba)
        assert np.array_equal(act, ract) and np.array_equal(st, rst)
    # pairs
    p, rp = sys.GetPairs(), ref.pairs()
    assert np.array_equal(po.canonical_pairs(p), po.canonical_pairs(rp))
    assert np.array_equal(p, rp), "pair emission order differs from the reference"
    # contacts
    k, rk = sys.GetContacts(), ref.contacts()
    ck, crk = po.canonical_contacts(k), po.canonical_contacts(rk)
    for key in ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius"):
        assert np.array_equal(ck[key], crk[key]), key
        assert np.array_equal(k[key], rk[key]), f"{key}: contact order differs from the reference"


def test_inactive_and_noncolliding_bodies(ref):
    scene = overlapping_bed(10, 10, 4)
    rng = np.random.default_rng(5)
    n = scene["mass"].size
    scene["active"][1:] = rng.random(n - 1) > 0.3
    scene["collide"][1:] = rng.random(n - 1) > 0.2
    sys = run_both(scene, ref)
    assert np.array_equal(sys.GetPairs(), ref.pairs())
    assert sys.GetCounts().num_rigid_contacts == ref.counts()["N_c"]


def test_family_masks(ref):
    scene = overlapping_bed(10, 10, 4)
    ns = scene["shape_type"].size
    fam = scene["shape_fam"].copy()
    # half the spheres in family 2 which does not collide with family 1 (demo_MCORE_mixerSMC-style masking)
    odd = np.arange(ns) % 2 == 1
    fam[odd, 0] = 1 << 2
    fam[odd, 1] = 0x7FFF & ~(1 << 0)
    fam[~odd, 1] = 0x7FFF & ~(1 << 2)
    scene["shape_fam"] = fam
    sys = run_both(scene, ref)
    assert np.array_equal(sys.GetPairs(), ref.pairs())
    assert sys.GetCounts().num_possible_collisions > 0


def test_rotated_bodies_and_mixed_sphere_shapes(ref):
    """Everything that is analytic against a sphere: box, capsule, cylinder, rounded box, rounded cylinder."""
    b = scenes.SceneBuilder()
    rng = np.random.default_rng(11)
    kinds = [cw.system.BOX, cw.system.CAPSULE, cw.system.CYLINDER, cw.system.ROUNDEDBOX, cw.system.ROUNDEDCYL]
    n = 60
    for i in range(n):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        p = np.array([(i % 6) * 0.5, (i // 6) * 0.5, 0.0]) + rng.uniform(-0.02, 0.02, 3)
        bid = b.add_bodies(1.0, (0.1, 0.1, 0.1), [p], rot=[q])
        t = kinds[i % len(kinds)]
        dims = {cw.system.BOX: [0.15, 0.1, 0.08, 0], cw.system.CAPSULE: [0.08, 0.08, 0, 0],
                cw.system.CYLINDER: [0.1, 0.1, 0.12, 0], cw.system.ROUNDEDBOX: [0.12, 0.1, 0.07, 0.03],
                cw.system.ROUNDEDCYL: [0.08, 0.08, 0.1, 0.02]}[t]
        lq = rng.normal(size=4)
        lq /= np.linalg.norm(lq)
        b.add_shapes(t, [bid], [dims], pos=[rng.uniform(-0.01, 0.01, 3)], rot=[lq], fam=(1 << 1, 0x7FFF & ~(1 << 1)))
        # spheres around it (family 1: collide with everything but each other is fine too)
        for k in range(6):
            d = rng.normal(size=3)
            d /= np.linalg.norm(d)
            sb = b.add_bodies(0.1, (0.01, 0.01, 0.01), [p + d * rng.uniform(0.10, 0.2)])
            b.add_shapes(cw.system.SPHERE, [sb], [[0.05]])
    scene = b.finish([scenes.smc_material()], dict(system_type=1, step_size=1e-3, gravity=(0, 0, -9.81),
                                                   broadphase_grid=2, grid_density=5.0, narrowphase_algorithm=1))
    sys = run_both(scene, ref)
    k, rk = sys.GetContacts(), ref.contacts()
    assert rk["depth"].size > 100
    types = scene["shape_type"]
    seen = set(zip(types[(rk["shape_ids"] >> 32)], types[rk["shape_ids"] & 0xFFFFFFFF]))
    assert len(seen) >= 6, seen
    for key in ("shape_ids", "body_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius"):
        assert np.array_equal(k[key], rk[key]), key


def test_envelope_nsc_style(ref):
    """With an envelope (NSC), near-contacts with positive depth are reported too (separation = 2*envelope)."""
    scene = scenes.settling_smc(12, 12, 4, spacing_factor=2.05, jitter_factor=0.02)
    scene["settings"].update(system_type=0, collision_envelope=0.002)
    sys = cw.load_scene(cw.ChSystemMulticoreNSC(), {**scene, "materials": np.array([scenes.nsc_material()])})
    sys.RunCollision(7)
    ref.set_scene(scene)
    ref.run(7)
    k, rk = sys.GetContacts(), ref.contacts()
    assert (rk["depth"] > 0).any()
    for key in ("shape_ids", "normal", "pt_a", "pt_b", "depth", "eff_radius"):
        assert np.array_equal(k[key], rk[key]), key


def test_empty_system():
    sys = cw.ChSystemMulticoreSMC()
    sys.RunCollision(7)
    assert sys.GetNumContacts() == 0


@pytest.mark.parametrize("n", [0, 1, 17, 8192, 8193, 300_000, 2_750_000, 4_300_000, 9_000_000])
def test_device_scan_against_host_scan(n):
    """The one-pass scan behind every list offset: sizes around one tile (8192 items), the bench's entry count, and
    more than 512 tiles (the look-back then needs the inclusive prefix in front of its window); repeated scans reuse
    the scratch without clearing it."""
    import ctypes as C

    from chrono_wisc_b200._lib import check

    sys = cw.ChSystemMulticoreSMC()
    bad = C.c_int64(-1)
    check(sys._lib.cb200_selftest_scan(sys._h, n, 7, 4, C.byref(bad)))
    assert bad.value == 0
