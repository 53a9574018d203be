"""Does a StepHost (host-state step) leave the device-resident path slower?  usage: tools/dbg_stephost.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

scene = scenes.settling_smc(250, 250, 16, grid="bin_size", bin_edge=2.5)
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
s.DoSteps(500); s.Synchronize()
s.DoSteps(100); s.Synchronize()
print("before   ", round(s.GetTimer("do_steps") * 1e3, 4), s.GetNumContacts())
p, q, v = s.GetBodyState()
p, q, v = np.ascontiguousarray(p), np.ascontiguousarray(q), np.ascontiguousarray(v)
for _ in range(20):
    s.StepHost(p, q, v)
s.Synchronize()
for k in range(4):
    s.DoSteps(100); s.Synchronize()
    print("after", k, "  ", round(s.GetTimer("do_steps") * 1e3, 4), s.GetNumContacts(), {n: round(s.GetTimer(n) * 1e3, 4) for n in ("k_pair_masks", "k_smc_force", "k_body_reduce", "k_integrate")})
