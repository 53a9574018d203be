import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from oracle import pyoracle
scene = scenes.settling_pd(half=(0.5, 0.5), layers=8)
nb = scene["mass"].size
m, I, r = scene["mass"][1:], scene["inertia"][1:], 0.02
def obs(p, v):
    ke = float((0.5*m*(v[1:, :3]**2).sum(1) + 0.5*(I*v[1:, 3:]**2).sum(1)).sum())
    z = p[1:, 2]
    h = float(np.percentile(z, 99) + r)
    phi = (nb-1)*4/3*np.pi*r**3/(1.0*1.0*h)
    return ke, float(z.mean()), h, phi
g = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
ref = pyoracle.Ref(); ref.set_num_threads(os.cpu_count()); ref.smc_setup(scene)
t0 = time.time()
for k in range(10):
    g.DoSteps(500); ref.smc_step(500)
    pg, qg, vg = g.GetBodyState(); pr, qr, vr = ref.smc_state()
    print(500*(k+1), ["%.4e" % x for x in obs(pg, vg)], ["%.4e" % x for x in obs(pr, vr)], "%.1fs" % (time.time()-t0))
