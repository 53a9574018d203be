"""Transcribes the known-answer cases of the reference's MPR unit test into tests/golden/mpr_ref_tests.json.

Source: src/tests/unit_tests/collision/utest_COLL_narrow_mpr.cpp (TESTs sphere_sphere :112-149, ellipsoid_ellipsoid
:153-190, sphere_box :194-229, cylinder_sphere :305-418, roundedcyl_sphere :471-590; tolerance `precision` = 5e-6, :41).
Every case calls ChNarrowphase::MPRCollision(shapeA, shapeB, envelope = 0, n, p1, p2, d); where the reference test
recomputes the depth as Dot(n, p2 - p1) before asserting, the case carries "depth_is_dot".  Cases the reference keeps
commented out, or asserts nothing about (ASSERT_TRUE(res || !res)), are not transcribed.
Shape type codes follow src/chrono/collision/ChCollisionShape.h:31-53.
"""
import json
import math
import os

SPHERE, ELLIPSOID, BOX, CYLINDER, ROUNDEDCYL = 0, 1, 2, 3, 15
I = [1.0, 0.0, 0.0, 0.0]
s = math.sin(math.pi / 4.0)
QY = [math.cos(math.pi / 4.0), 0.0, math.sin(math.pi / 4.0), 0.0]  # QuatFromAngleY(CH_PI_2)


def shape(t, pos, rot, dims):
    return {"type": t, "pos": list(map(float, pos)), "rot": list(rot), "dims": list(map(float, dims)) + [0.0] * (4 - len(dims))}


def case(name, A, B, hit, normal=None, pt_a=None, pt_b=None, depth=None, depth_is_dot=False):
    e = {}
    if normal is not None:
        e = {"normal": list(map(float, normal)), "pt_a": list(map(float, pt_a)), "pt_b": list(map(float, pt_b)), "depth": float(depth)}
    return {"name": name, "A": A, "B": B, "envelope": 0.0, "hit": hit, "expect": e, "depth_is_dot": depth_is_dot, "tol": 5e-6}


cases = []
for nm, T, d in (("sphere", SPHERE, [1, 0, 0]), ("ellipsoid", ELLIPSOID, [1, 1, 1])):
    cases.append(case(f"{nm}_{nm}_touching", shape(T, [2, 2, 0], I, d), shape(T, [2, 0, 0], I, d), True,
                      [0, -1, 0], [2, 1, 0], [2, 1, 0], 0.0))
    nc = [s, -s, 0.0]
    p1 = [1 + nc[0], 1 + nc[1], 0.0]
    p2 = [2 - nc[0], 0 - nc[1], 0.0]
    cases.append(case(f"{nm}_{nm}_oblique", shape(T, [1, 1, 0], I, d), shape(T, [2, 0, 0], I, d), True, nc, p1, p2,
                      sum(n * (b - a) for n, a, b in zip(nc, p1, p2)), True))
cases.append(case("sphere_box_penetrating", shape(SPHERE, [0, 1.5, 0], I, [1, 0, 0]), shape(BOX, [0, 0, 0], I, [1, 1, 1]), True,
                  [0, -1, 0], [0, 0.5, 0], [0, 1, 0], -0.5))
cases.append(case("sphere_box_touching", shape(SPHERE, [0.1, 2, 0], I, [1, 0, 0]), shape(BOX, [0, 0, 0], I, [1, 1, 1]), True,
                  [0, -1, 0], [0.1, 1, 0], [0.1, 1, 0], 0.0))
for nm, T, dims, extra in (("cylinder", CYLINDER, [2.0, 2.0, 1.5], 0.0), ("roundedcyl", ROUNDEDCYL, [2.0, 2.0, 1.5, 0.1], 0.1)):
    cyl = shape(T, [1.5, 0, 0], QY, dims)
    sph = lambda p: shape(SPHERE, p, I, [1.0, 0, 0])
    if T == CYLINDER:
        cases.append(case(f"{nm}_sphere_centre_inside", cyl, sph([2.5, 1.5, 0]), True))
    cases.append(case(f"{nm}_sphere_cap_separated", cyl, sph([4.5, 1.5, 0]), False))
    cases.append(case(f"{nm}_sphere_cap_penetrated", cyl, sph([3.75, 1.5, 0]), True, [1, 0, 0], [3 + extra, 1.5, 0],
                      [2.75, 1.5, 0], -0.25 - extra, True))
    cases.append(case(f"{nm}_sphere_side_separated", cyl, sph([2.5, 3.5, 0]), False))
    cases.append(case(f"{nm}_sphere_side_penetrated", cyl, sph([2.5, 2.5, 0]), True, [0, 1, 0], [2.5, 2.0 + extra, 0],
                      [2.5, 1.5, 0], -0.5 - extra, True))
    cases.append(case(f"{nm}_sphere_edge_separated", cyl, sph([4, 3, 0]), False))

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mpr_ref_tests.json")
json.dump(cases, open(out, "w"), indent=1)
print(len(cases), "cases ->", out)
