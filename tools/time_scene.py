"""Scratch timing helper (not part of the product): per-stage GPU timers on a settling bed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sys, time
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

nx, ny, nz = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (250, 250, 16)
warm = int(sys.argv[4]) if len(sys.argv) > 4 else 200
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 100
edge = float(sys.argv[6]) if len(sys.argv) > 6 else 0
scene = scenes.settling_smc(nx, ny, nz, grid='bin_size', bin_edge=edge) if edge else scenes.settling_smc(nx, ny, nz)
t0 = time.time()
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
print("load", time.time() - t0)
for blk in range(2):
    t0 = time.time(); s.DoSteps(warm); s.Synchronize(); dt = time.time() - t0
    c = s.GetCounts()
    print(f"block {blk}: {dt/warm*1e3:.3f} ms/step wall; N_i={c.num_bin_aabb_intersections} N_a={c.num_active_bins} N_p={c.num_possible_collisions} N_c={c.num_rigid_contacts} bins={tuple(c.bins_per_axis)}")
t0 = time.time(); s.DoSteps(steps); s.Synchronize(); dt = time.time() - t0
print(f"timed: {dt/steps*1e3:.3f} ms/step wall, {scene['mass'].size*steps/dt/1e6:.1f} M particle-steps/s")
for name in ("step", "update", "collision", "collision_broad", "collision_narrow", "advance", "integrate"):
    print(f"  {name:18s} {s.GetTimer(name)*1e3:.4f} ms")
p, q, v = s.GetBodyState()
print("max z", p[1:,2].max(), "mean z", p[1:,2].mean(), "KE/particle", 0.5*(scene['mass'][1:]*(v[1:,:3]**2).sum(1)).mean())
