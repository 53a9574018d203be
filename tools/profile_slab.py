"""Two slabs of one bed in one process on one GPU (LocalGroup: device copies stand in for send / recv): the kernel list of
a slab step for `ncu --profile-from-start off`.   usage: tools/profile_slab.py nx ny nz warm steps"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from chrono_wisc_b200 import scenes, slab

nx, ny, nz, warm, steps = (int(a) for a in sys.argv[1:6])
rt = ctypes.CDLL("libcudart.so.12")
scene = scenes.settling_smc(nx, ny, nz, grid="bin_size", bin_edge=2.5)
cuts = (-1e9, 0.0, 1e9)
systems = [slab.SlabSystemSMC(slab.SlabLayout(2, r, cuts, axis=0, width=0.05), device=0).LoadSlab(scene) for r in range(2)]
grp = slab.LocalGroup(systems)
grp.Exchange()
grp.DoSteps(warm)
for s in systems:
    s.Synchronize()
rt.cudaProfilerStart()
grp.DoSteps(steps)
for s in systems:
    s.Synchronize()
rt.cudaProfilerStop()
for s in systems:
    s.Check()
    c = s.GetCounts()
    print(f"N_b={c.num_bodies} N_c={c.num_rigid_contacts} step={s.GetTimer('step') * 1e3:.3f} ms")
