"""Kinetic energy / height / contact count of the bench bed over time.  usage: tools/dbg_settle.py nx ny nz chunks"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes

nx, ny, nz, chunks = (int(a) for a in sys.argv[1:5])
sp = float(sys.argv[5]) if len(sys.argv) > 5 else 2.02
scene = scenes.settling_smc(nx, ny, nz, grid="bin_size", bin_edge=2.5, spacing_factor=sp, jitter_factor=0.005)
s = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
m, I = scene["mass"][1:], scene["inertia"][1:]
print("materials", scene["materials"], "settings", {k: scene["settings"][k] for k in ("step_size", "contact_force_model", "tangential_displ_mode")})
for c in range(chunks):
    s.DoSteps(500)
    s.Synchronize()
    p, q, v = s.GetBodyState()
    v = v[1:]
    ke_t = (0.5 * m * (v[:, :3] ** 2).sum(1)).mean()
    ke_r = (0.5 * (I * v[:, 3:] ** 2).sum(1)).mean()
    sp = np.linalg.norm(v[:, :3], axis=1)
    print(f"step {500*(c+1):5d} KE_t {ke_t:.3e} KE_r {ke_r:.3e} |v| med {np.median(sp):.3e} max {sp.max():.3e} z [{p[1:,2].min():.4f}, {p[1:,2].max():.4f}] "
          f"x [{p[1:,0].min():.3f}, {p[1:,0].max():.3f}] N_c {s.GetNumContacts()} step_ms {s.GetTimer('do_steps')*1e3:.3f}")
    if os.environ.get("KTIMERS"):
        print("   ", {n: round(s.GetTimer(n) * 1e3, 4) for n in ("k_bin_count", "k_bin_fill", "k_pair_masks", "k_pair_queue", "k_pair_store", "k_smc_force", "k_inc_fill", "k_body_reduce", "k_integrate", "update", "collision_broad", "collision_narrow", "advance", "integrate", "step")}, "N_i", s.GetCounts().num_bin_aabb_intersections, "N_p", s.GetCounts().num_possible_collisions)
