"""Scratch profiling helper: warm up, then run a few steps inside a cudaProfilerStart/Stop range
(use with: ncu --profile-from-start off ...)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes, sys
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

nx, ny, nz, warm, steps = (int(a) for a in sys.argv[1:6])
edge = float(sys.argv[6]) if len(sys.argv) > 6 else 2.0
rt = ctypes.CDLL("libcudart.so.12")
sp = float(sys.argv[7]) if len(sys.argv) > 7 else 2.02
scene = scenes.settling_smc(nx, ny, nz, grid="bin_size", bin_edge=edge, spacing_factor=sp)
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
s.DoSteps(warm)
s.Synchronize()
rt.cudaProfilerStart()
s.DoSteps(steps)
s.Synchronize()
rt.cudaProfilerStop()
c = s.GetCounts()
print(f"N_b={scene['mass'].size} N_i={c.num_bin_aabb_intersections} N_a={c.num_active_bins} N_p={c.num_possible_collisions} N_c={c.num_rigid_contacts} bins={tuple(c.bins_per_axis)} step={s.GetTimer('step')*1e3:.3f} ms")
