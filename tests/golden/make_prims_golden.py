"""Generates the narrowphase golden fixtures (run in the build container, where /root/reference exists).

  prims_ref_tests.json     known-answer cases transcribed from the reference's own unit test
                           src/tests/unit_tests/collision/utest_COLL_narrow_prims.cpp (sphere_sphere :996-1094,
                           box_sphere :1098-1250, capsule_sphere :1254-1336), tolerance 1e-10 (:29-33).
  prims_random_pairs.npz   outputs of the reference's PRIMSCollision (through oracle/_ref) on seeded random
                           sphere-vs-{sphere,box,capsule,cylinder,rounded box,rounded cylinder} pairs in both
                           argument orders: bit-exact targets for the oracle restatement and the CUDA kernels.

Usage: python tests/golden/make_prims_golden.py
"""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import pyoracle  # noqa: E402

SPHERE, BOX, CYLINDER, CAPSULE, RBOX, RCYL = 0, 2, 3, 12, 14, 15
Q1 = [1.0, 0.0, 0.0, 0.0]


def quat_z(a):
    return [math.cos(a / 2), 0.0, 0.0, math.sin(a / 2)]


def quat_y(a):
    return [math.cos(a / 2), 0.0, math.sin(a / 2), 0.0]


def transcribed_cases():
    cases = []

    def add(name, A, B, sep, nC, **exp):
        cases.append(dict(name=name, A=A, B=B, separation=sep, nC=nC, expect=exp))

    def sph(pos, r):
        return dict(type=SPHERE, pos=list(pos), rot=Q1, dims=[r, 0, 0, 0])

    for sep_on in (False, True):
        sep = 0.1 if sep_on else 0.0
        tag = f"sep{int(sep_on)}"
        # ---- sphere_sphere (:996-1094)
        add(f"ss_far_{tag}", sph((2, 2, 0), 1), sph((2, 0, 0), 0.5), sep, 0)
        if sep_on:
            add(f"ss_near_{tag}", sph((2, 2, 0), 1), sph((2, 0, 0), 0.95), sep, 1, normal=[0, -1, 0], depth=0.05,
                pt_a=[2, 1, 0], pt_b=[2, 0.95, 0], eff_radius=0.95 / 1.95)
            add(f"ss_touch_{tag}", sph((2, 2, 0), 1), sph((2, 0, 0), 1), sep, 1, normal=[0, -1, 0], depth=0.0,
                pt_a=[2, 1, 0], pt_b=[2, 1, 0], eff_radius=0.5)
        else:
            add(f"ss_near_{tag}", sph((2, 2, 0), 1), sph((2, 0, 0), 0.95), sep, 0)
            add(f"ss_touch_{tag}", sph((2, 2, 0), 1), sph((2, 0, 0), 1), sep, 0)
        add(f"ss_pen_{tag}", sph((1, 1, 0), 1), sph((2.5, 1, 0), 1), sep, 1, normal=[1, 0, 0], depth=-0.5,
            pt_a=[2, 1, 0], pt_b=[1.5, 1, 0], eff_radius=0.5)

        # ---- box_sphere (:1098-1250): box hdims (1,2,3) at (sqrt2,0,0) rotated 45 deg about z; sphere r = 1.5
        o = math.sqrt(0.5)
        box = dict(type=BOX, pos=[math.sqrt(2.0), 0, 0], rot=quat_z(math.pi / 4), dims=[1, 2, 3, 0])
        sr, er = 1.5, 0.1
        edge_rad = sr * er / (sr + er)
        add(f"bs_inside_{tag}", box, sph((0.5, 0.5, 1.0), sr), sep, 0)
        add(f"bs_face_far_{tag}", box, sph((3.5, 2.5, 1.0), sr), sep, 0)
        if sep_on:
            add(f"bs_face_near_{tag}", box, sph((4.55 * o, 2.55 * o, 1.0), sr), sep, 1, normal=[o, o, 0], depth=0.05,
                pt_a=[3 * o, o, 1], pt_b=[3.05 * o, 1.05 * o, 1])
        else:
            add(f"bs_face_near_{tag}", box, sph((4.55 * o, 2.55 * o, 1.0), sr), sep, 0)
        add(f"bs_face_pen_{tag}", box, sph((4 * o, 2 * o, 1.0), sr), sep, 1, normal=[o, o, 0], depth=-0.5,
            pt_a=[3 * o, o, 1], pt_b=[2.5 * o, 0.5 * o, 1], eff_radius=sr)
        add(f"bs_edge_far_{tag}", box, sph((o, 4.0, 1.0), sr), sep, 0)
        if sep_on:
            add(f"bs_edge_near_{tag}", box, sph((o, 3 * o + 1.55, 1.0), sr), sep, 1, normal=[0, 1, 0], depth=0.05,
                pt_a=[o, 3 * o, 1], pt_b=[o, 3 * o + 0.05, 1])
        else:
            add(f"bs_edge_near_{tag}", box, sph((o, 3 * o + 1.55, 1.0), sr), sep, 0)
        add(f"bs_edge_pen_{tag}", box, sph((o, 3 * o + 1.0, 1.0), sr), sep, 1, normal=[0, 1, 0], depth=-0.5,
            pt_a=[o, 3 * o, 1], pt_b=[o, 3 * o - 0.5, 1], eff_radius=edge_rad)
        add(f"bs_corner_far_{tag}", box, sph((o, 4.0, 4.0), sr), sep, 0)
        sp = (o, 4 * o, 3.0 + o)
        n = [0, o, o]
        add(f"bs_corner_pen_{tag}", box, sph(sp, sr), sep, 1, normal=n, depth=-0.5, pt_a=[o, 3 * o, 3],
            pt_b=[sp[i] - sr * n[i] for i in range(3)], eff_radius=edge_rad)

    # ---- capsule_sphere (:1254-1336): capsule r=.5 hlen=2 along x at (2,0,0); sphere r = 1; separation 0
    o = math.sqrt(0.5)
    cap = dict(type=CAPSULE, pos=[2.0, 0, 0], rot=quat_y(math.pi / 2), dims=[0.5, 2.0, 0.5, 0])
    er = 1.0 * 0.5 / 1.5
    add("cs_axis", cap, sph((3, 0, 0), 1), 0.0, 0)
    add("cs_cap_sep", cap, sph((5, 1.5, 0), 1), 0.0, 0)
    add("cs_cap_pen", cap, sph((5, 1, 0), 1), 0.0, 1, normal=[o, o, 0], depth=math.sqrt(2.0) - 1.5,
        pt_a=[4 + 0.5 * o, 0.5 * o, 0], pt_b=[5 - o, 1 - o, 0], eff_radius=er)
    add("cs_side_sep", cap, sph((2.5, 2.0, 0), 1), 0.0, 0)
    add("cs_side_pen", cap, sph((2.5, 1.25, 0), 1), 0.0, 1, normal=[0, 1, 0], depth=-0.25, pt_a=[2.5, 0.5, 0],
        pt_b=[2.5, 0.25, 0], eff_radius=er)
    return cases


def random_pairs(n=600, seed=2024):
    rng = np.random.default_rng(seed)
    ref = pyoracle.Ref()
    kinds = [SPHERE, BOX, CAPSULE, CYLINDER, RBOX, RCYL]
    rec = dict(typeA=[], typeB=[], posA=[], posB=[], rotA=[], rotB=[], dimsA=[], dimsB=[], sep=[], nC=[], normal=[],
               pt_a=[], pt_b=[], depth=[], eff_radius=[])

    def dims_for(t):
        if t == SPHERE:
            return [rng.uniform(0.05, 0.3), 0, 0, 0]
        if t == CAPSULE:
            return [rng.uniform(0.05, 0.2), rng.uniform(0.05, 0.4), 0, 0]
        d = list(rng.uniform(0.08, 0.4, 3)) + [0.0]
        if t in (RBOX, RCYL):
            d[3] = rng.uniform(0.01, 0.05)
        return d

    for i in range(n):
        other = kinds[i % len(kinds)]
        sphere_first = (i // len(kinds)) % 2 == 1
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        dims_o, dims_s = dims_for(other), dims_for(SPHERE)
        p_o = rng.uniform(-1, 1, 3)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        reach = max(dims_o[:3]) + dims_s[0]
        p_s = p_o + d * rng.uniform(0.2, 1.3) * reach
        sep = float(rng.choice([0.0, 0.02]))
        A = (other, p_o, q, dims_o)
        B = (SPHERE, p_s, np.array(Q1), dims_s)
        if sphere_first:
            A, B = B, A
        nC, out = ref.prims_pair(A[0], A[1], A[2], A[3], B[0], B[1], B[2], B[3], sep)
        rec["typeA"].append(A[0]), rec["typeB"].append(B[0])
        rec["posA"].append(A[1]), rec["posB"].append(B[1]), rec["rotA"].append(A[2]), rec["rotB"].append(B[2])
        rec["dimsA"].append(A[3]), rec["dimsB"].append(B[3]), rec["sep"].append(sep), rec["nC"].append(nC)
        for k in ("normal", "pt_a", "pt_b"):
            rec[k].append(out[k][0] if nC == 1 else np.zeros(3))
        for k in ("depth", "eff_radius"):
            rec[k].append(out[k][0] if nC == 1 else 0.0)
    return {k: np.asarray(v) for k, v in rec.items()}


if __name__ == "__main__":
    with open(os.path.join(HERE, "prims_ref_tests.json"), "w") as f:
        json.dump(transcribed_cases(), f, indent=1)
    r = random_pairs()
    np.savez_compressed(os.path.join(HERE, "prims_random_pairs.npz"), **r)
    print("cases:", len(transcribed_cases()), " random pairs:", len(r["nC"]), " hits:", int((r["nC"] == 1).sum()))
