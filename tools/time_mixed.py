"""Scratch timing helper (not part of the product): BASELINE.json configs[4] flavour -- a bed of mixed spheres / boxes /
capsules (demo_MCORE_mixerSMC.cpp:118-168 shape families) settling in an open box, without the mixer blade (its motor is
a bilateral constraint).  usage: tools/time_mixed.py nx ny nz warm steps algorithm(0 MPR, 2 HYBRID)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sys, time
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from chrono_wisc_b200.system import BOX, CAPSULE, SPHERE, HERTZ, FIXED_BIN_SIZE

nx, ny, nz, warm, steps, alg = (int(a) for a in sys.argv[1:7])
radius, rho = 0.1, 1000.0
spacing = 2.4 * radius
b = scenes.SceneBuilder()
sx, sy, sz = nx * spacing + radius, ny * spacing + radius, (nz + 2) * spacing
bin_body = b.add_bodies(1.0, (1.0, 1.0, 1.0), [[0.0, 0.0, 0.0]], active=0, collide=1)
b.add_box_container(bin_body, (0, 0, sz / 2), (sx, sy, sz), 0.2, faces=(2, 2, -1))
pos = scenes.lattice_positions(nx, ny, nz, spacing, 0.02 * radius, 1.2 * radius, 7)
rng = np.random.Generator(np.random.MT19937(8))
n = pos.shape[0]
q = rng.normal(size=(n, 4))
q /= np.linalg.norm(q, axis=1, keepdims=True)
kind = np.arange(n) % 3
for k in range(3):
    sel = np.flatnonzero(kind == k)
    if k == 0:
        m = rho * 4 / 3 * np.pi * radius ** 3
        I = (0.4 * m * radius ** 2,) * 3
        first = b.add_bodies(m, I, pos[sel], rot=q[sel])
        b.add_shapes(SPHERE, first + np.arange(sel.size), [[radius]])
    elif k == 1:
        h = np.array([radius, 0.8 * radius, 0.6 * radius])
        m = rho * 8 * h.prod()
        I = m / 3 * np.array([h[1] ** 2 + h[2] ** 2, h[0] ** 2 + h[2] ** 2, h[0] ** 2 + h[1] ** 2])
        first = b.add_bodies(m, I, pos[sel], rot=q[sel])
        b.add_shapes(BOX, first + np.arange(sel.size), [h])
    else:
        r, hl = 0.6 * radius, 0.6 * radius
        m = rho * (np.pi * r * r * 2 * hl + 4 / 3 * np.pi * r ** 3)
        I = (0.5 * m * r * r,) * 3
        first = b.add_bodies(m, I, pos[sel], rot=q[sel])
        b.add_shapes(CAPSULE, first + np.arange(sel.size), [[r, hl]])
settings = dict(system_type=1, step_size=1e-3, gravity=(0.0, 0.0, -9.81), contact_force_model=HERTZ,
                narrowphase_algorithm=alg, broadphase_grid=FIXED_BIN_SIZE, bin_size=(float(sys.argv[7]) * radius if len(sys.argv) > 7 else 3 * radius,) * 3)
scene = b.finish([scenes.smc_material(young=2e5, friction=0.4, restitution=0.1)], settings)
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
t0 = time.time(); s.DoSteps(warm); s.Synchronize(); dt = time.time() - t0
c = s.GetCounts()
print(f"warm: {dt/warm*1e3:.3f} ms/step wall; N_b={n} N_i={c.num_bin_aabb_intersections} N_p={c.num_possible_collisions} N_c={c.num_rigid_contacts} bins={tuple(c.bins_per_axis)}")
import os
if os.environ.get("CB200_PROFILE"):  # ncu --profile-from-start off: one directly enqueued step
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaProfilerStart(); s.DoSteps(1); s.Synchronize(); rt.cudaProfilerStop()
t0 = time.time(); s.DoSteps(steps); s.Synchronize(); dt = time.time() - t0
c = s.GetCounts()
print(f"timed (algorithm {alg}): {dt/steps*1e3:.3f} ms/step wall, {n*steps/dt/1e6:.1f} M particle-steps/s; N_c={c.num_rigid_contacts}")
for name in ("step", "update", "collision_broad", "collision_narrow", "advance", "integrate", "k_pair_masks", "k_pair_store",
             "k_smc_force", "k_body_reduce", "k_bin_sort"):
    print(f"  {name:18s} {s.GetTimer(name)*1e3:.4f} ms")
p, _, v = s.GetBodyState()
print("max z", p[1:, 2].max(), "mean z", p[1:, 2].mean(), "max |v|", np.abs(v[1:, :3]).max())
