"""Scratch profiling helper: the mixed bed of bench.py --config mixed, warm up, then steps inside a
cudaProfilerStart/Stop range (use with ncu --profile-from-start off)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
nx, ny, nz, warm, steps = (int(a) for a in sys.argv[1:6])
rt = ctypes.CDLL("libcudart.so.12")
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scenes.mixed_bed(nx, ny, nz))
s.DoSteps(warm); s.Synchronize()
rt.cudaProfilerStart()
s.DoSteps(steps); s.Synchronize()
rt.cudaProfilerStop()
c = s.GetCounts()
print(f"N_i={c.num_bin_aabb_intersections} N_p={c.num_possible_collisions} N_c={c.num_rigid_contacts} step={s.GetTimer('step')*1e3:.3f} ms")
