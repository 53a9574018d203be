import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import ctypes as C
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import slab
from chrono_wisc_b200._lib import check
from test_gpu_slab import moving_bed
scene = moving_bed()
cuts=(-10.0,0.0,10.0)
one = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
systems=[slab.SlabSystemSMC(slab.SlabLayout(2,r,cuts,width=0.05),device=0).LoadSlab(scene) for r in range(2)]
def exchange():
    for s in systems: s.Pack()
    slab.LocalExchange.run(systems)
    for s in systems: s.Unpack()
def ids_flags(s):
    nb=s.num_bodies; ids=np.empty(nb,np.int64); fl=np.empty(nb,np.uint8)
    check(s._lib.cb200_download_body_ids(s._h,0,nb,ids.ctypes.data_as(C.c_void_p),fl.ctypes.data_as(C.c_void_p)))
    ids[:s.nb_static]=np.arange(s.nb_static)
    return ids,fl
exchange()
for step in range(4):
    pre=[ids_flags(s) for s in systems]
    one.DoSteps(1)
    for s in systems: s.StepOnly()
    for s in systems: s.Check()
    f1,t1=one.GetBodyContactForces()
    c1=one.GetContacts()
    for r,s in enumerate(systems):
        ids,fl=pre[r]
        f,t=s.GetBodyContactForces()
        own=(fl&1)==1
        own[:s.nb_static]=False
        g=ids[own]
        ef=np.abs(f[own]-f1[g]).max(1); et=np.abs(t[own]-t1[g]).max(1)
        w=np.argmax(ef+et)
        print(step,"rank",r,"force err %.2e torque err %.2e worst gid %d"%(ef.max(),et.max(),g[w]))
        if ef.max()>1e-6 or et.max()>1e-6:
            gid=g[w]
            c=s.GetContacts()
            gb=ids[c["body_ids"]]
            m=(gb==gid).any(1)
            print(" slab contacts:", sorted(map(tuple,np.sort(gb[m],1))), "depths", np.round(c["depth"][m],8))
            m1=(c1["body_ids"]==gid).any(1)
            print(" one  contacts:", sorted(map(tuple,np.sort(c1["body_ids"][m1],1))), "depths", np.round(c1["depth"][m1],8))
            fl_of={int(i):int(f_) for i,f_ in zip(ids,fl)}
            print(" flags of partners:", {int(b):fl_of.get(int(b)) for b in np.unique(gb[m])})
            u,cnt=np.unique(ids[(fl&4)==0][s.nb_static:],return_counts=True)
            print(" duplicate live gids:", u[cnt>1])
    exchange()
