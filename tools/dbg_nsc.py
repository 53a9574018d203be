import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes, slab
from test_gpu_slab import nsc_bed
scene = nsc_bed(iters=500)
one = cw.load_scene(cw.ChSystemMulticoreNSC(), scene)
cuts=(-10.0, 0.05, 10.0)
systems = [slab.SlabSystemNSC(slab.SlabLayout(2, r, cuts, axis=0, width=0.35), device=0).LoadSlab(scene) for r in range(2)]
grp = slab.LocalGroup(systems); grp.Exchange()
for step in range(6):
    one.DoSteps(1); c = one.GetCounts(); print('one', c.num_rigid_contacts, c.solver_iterations, c.solver_residual, c.solver_objective)
    p1,q1,v1 = one.GetBodyState()
    grp.DoSteps(1)
    tot = 0
    for s in systems:
        c = s.GetCounts(); k = s.GetContacts()
        ids = np.empty(s.num_bodies, np.int64); fl = np.empty(s.num_bodies, np.uint8)
        import ctypes as C
        s._lib.cb200_download_body_ids(s._h, 0, s.num_bodies, ids.ctypes.data_as(C.c_void_p), fl.ctypes.data_as(C.c_void_p))
        bi = k['body_ids']; ref = np.where(bi[:,0] >= s.nb_static, bi[:,0], bi[:,1]); own = (fl[ref] & 1).astype(bool)
        tot += own.sum()
        print('  slab', c.num_rigid_contacts, 'owned contacts', own.sum(), c.solver_iterations, c.solver_residual, c.solver_objective, 'owned bodies', s.GetOwned()[0].size, 'ghosts', s.GetGhostIds().size)
    print('  owned contacts total', tot)
    ids = np.concatenate([s.GetOwned()[0] for s in systems]); vel = np.concatenate([s.GetOwned()[3] for s in systems]); pos = np.concatenate([s.GetOwned()[1] for s in systems])
    o = np.argsort(ids); d = np.abs(vel[o]-v1[1:]).max(1)
    bad = ids[o][d>1e-7]
    print('  max vel diff', d.max(), 'n bad', bad.size, 'bad x range', (p1[bad,0].min(), p1[bad,0].max()) if bad.size else None)
