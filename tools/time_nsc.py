"""Scratch timing helper (not part of the product): NSC APGD step on a ball bed (BASELINE.json configs[3]).
usage: tools/time_nsc.py nx ny nz warm steps [iterations]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sys, time
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

nx, ny, nz = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (128, 128, 16)
warm = int(sys.argv[4]) if len(sys.argv) > 4 else 20
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 10
iters = int(sys.argv[6]) if len(sys.argv) > 6 else 100
scene = scenes.balls_nsc(nx, ny, nz, iterations=iters)
s = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
t0 = time.time(); s.DoSteps(warm); s.Synchronize(); dt = time.time() - t0
c = s.GetCounts()
print(f"warm: {dt/warm*1e3:.3f} ms/step wall; N_c={c.num_rigid_contacts} iters={c.solver_iterations} res={c.solver_residual:.3e}")
l0 = s.KernelLaunches()
t0 = time.time(); s.DoSteps(steps); s.Synchronize(); dt = time.time() - t0
c = s.GetCounts()
nb = scene['mass'].size
print(f"timed: {dt/steps*1e3:.3f} ms/step wall, {nb*steps/dt/1e6:.2f} M particle-steps/s, N_c={c.num_rigid_contacts} iters={c.solver_iterations} slots={c.solver_slots} launches/step={(s.KernelLaunches()-l0)/steps:.0f}")
for name in ("step", "update", "collision", "collision_broad", "collision_narrow", "advance", "integrate"):
    print(f"  {name:18s} {s.GetTimer(name)*1e3:.4f} ms")
nc = c.num_rigid_contacts
it = max(1, c.solver_iterations)
adv = s.GetTimer("advance")
print(f"  per APGD iteration: {adv/it*1e6:.1f} us; SURVEY 8d bytes/iter = {(2*(nc*(2*144+48)+nb*96)+240*nc)/1e9:.3f} GB -> {(2*(nc*(2*144+48)+nb*96)+240*nc)/(adv/it)/1e9:.0f} GB/s")
