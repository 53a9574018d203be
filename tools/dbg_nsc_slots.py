import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
scene = scenes.balls_nsc(32, 32, 8, iterations=100)
s = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
for k in range(6):
    l0 = s.KernelLaunches()
    s.DoSteps(1); s.Synchronize()
    c = s.GetCounts()
    res, obj = s.GetSolverHistory()
    print(k, "launches", s.KernelLaunches() - l0, "iters", c.solver_iterations, "hist", len(res), "res", c.solver_residual, "obj", c.solver_objective, "first/last", res[:2], res[-2:])
