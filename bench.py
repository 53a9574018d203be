#!/usr/bin/env python
"""bench.py -- particle-steps/s of the Chrono::Multicore SMC settling hot path on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W                 # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W  # the reference's CPU path on the host cores

One "step" = one DoStepDynamics of the whole bed: body forces, AABBs, grid broadphase, narrowphase, SMC contact
forces, per-body accumulation, velocity update, position integration -- all device-resident.

Workload at N=1: BASELINE.json configs[1], 1,000,000 monodisperse spheres (250 x 250 x 16 jittered lattice, r = 0.02,
Hertz SMC, FP64) settling in a 5-wall box; the bed is advanced `--settle` steps before warm-up so that the timed
window sees a contact-rich bed (N_c is printed in `config`).  At N>1 every rank owns one slab of that size
(weak scaling, configs[2] = 8M over 8 GPUs).

Output: ONE JSON line on rank 0 (see DESIGN.md "Measurement" for every field).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "particle-steps/s"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def make_scene(args, rank=0, world=1):
    from chrono_wisc_b200 import scenes

    if args.config == "mixed":
        return scenes.mixed_bed(args.mixed_nx, args.mixed_ny, args.mixed_nz)
    # near-touching jittered lattice: the bed forms its contact network within the settle window
    if args.bin_edge > 0:  # FIXED_BIN_SIZE, cubic bins of --bin-edge sphere diameters
        return scenes.settling_smc(args.nx, args.ny, args.nz, seed=12345 + rank, spacing_factor=args.spacing,
                                   jitter_factor=0.005, grid="bin_size", bin_edge=args.bin_edge)
    return scenes.settling_smc(args.nx, args.ny, args.nz, seed=12345 + rank, spacing_factor=args.spacing,
                               jitter_factor=0.005, grid="density")


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.stop_flag = threading.Event()

    def run(self):
        # NVML in-process (no nvidia-smi fork per sample: a fork every 200 ms perturbs the launch thread)
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown")
                    else nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown",
                                                   getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0)),
                    "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown",
                                                   getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0)),
                    "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap",
                                            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0))}
            while not self.stop_flag.is_set():
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for n, b in bits.items():
                    if b and (r & b):
                        self.reasons.add(n)
                self.stop_flag.wait(0.05)
            return
        except Exception:
            pass
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                f = [x.strip() for x in out.split(",")]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(n)
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def result(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def algorithmic_bytes(nb, c):
    """Bytes_SMC = N_b*250 + N_i*16 + N_p*16 + N_c*192 + N_b*96 (materials from a table)."""
    return (nb * 250 + c.num_bin_aabb_intersections * 16 + c.num_possible_collisions * 16 +
            c.num_rigid_contacts * 192 + nb * 96)


def kernel_algorithmic_bytes(nb, ns, c):
    """Per-kernel algorithmic bytes: what the kernel must move given its inputs and outputs, unpadded, every logical
    array once (a body or shape record gathered by many contacts counts once per body).  DESIGN.md
    section 4 has the same table."""
    ni, nc = c.num_bin_aabb_intersections, c.num_rigid_contacts
    return {
        "k_bin_count": ns * (32 + 16 + 4),                    # shape world record + integer record in, bins touched out
        "k_bin_fill": ns * (32 + 16 + 4) + ni * 8,            # same in, (shape, bin) per entry out
        "k_pair_masks": ni * 8 + ns * 48 + ni * (4 + 8 + 4),  # entries in, shape records once, sorted id + mask + count out
        "k_pair_queue": 0,                                    # wall-bin pairs: latency-bound, no streaming term of its own
        "k_pair_store": ni * (8 + 4 + 4 + 4) + ns * 16 + nc * (8 + 8 + 4),  # mask, offset, id, bin in; key, body ids, materials out
        "k_smc_force": nc * (8 + 8 + 4) + ns * (32 + 16) + nb * (32 + 48 + 8) + nc * 96,  # keys/ids in, shape + body state once, record out
        "k_inc_fill": nc * 8 + nb * 4 + nc * 8,               # body ids + offsets in, two incidence codes out
        "k_body_reduce": nc * (96 + 8) + nb * (4 + 48 + 8 + 24) + nb * 96,  # records + codes in, body state in, sums + hf out
        "k_integrate": nb * (48 + 48 + 24 + 32 + 8 + 24 + 1) + nb * 104 + nb * (4 + 8 + 32),  # + the next step's sphere records
        "k_contact_gen": nc * (8 + 96),
    }


KERNEL_TIMERS = ("k_bin_count", "k_bin_fill", "k_pair_masks", "k_pair_queue", "k_pair_store", "k_smc_force", "k_inc_fill",
                 "k_body_reduce", "k_integrate", "k_contact_gen")


def load_traffic():
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of this round's build, from the committed
    ncu --set full capture (profiles/r2_traffic.json; tools/ncu_traffic.py makes it)."""
    for name in ("r2_traffic.json",):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return json.load(f).get("dram_bytes_per_launch", {}), name
        except Exception:
            pass
    return {}, None


def kernel_rooflines(sys_, nb, ns, c, peak):
    """Per-kernel roofline entries: live CUDA-event durations (events around the launch, on the library stream, in the
    directly enqueued steps of the timed region) against each kernel's algorithmic bytes.  `traffic` = measured DRAM
    bytes per launch (ncu); algorithmic bytes above the measured traffic would claim bandwidth DRAM never delivered, so
    `achieved` then uses the traffic and the entry says so."""
    alg = kernel_algorithmic_bytes(nb, ns, c)
    traffic, src = load_traffic()
    out = {}
    for k in KERNEL_TIMERS:
        try:
            ms = sys_.GetTimer(k) * 1e3
        except Exception:
            continue
        if ms <= 0:
            continue
        t = traffic.get(k)
        used = alg[k]
        capped = t is not None and used > t
        if capped:
            used = t
        a = used / (ms * 1e-3) / 1e9
        out[k] = {"ms": ms, "algorithmic_bytes": alg[k], "achieved": a, "frac": a / peak, "traffic": t,
                  "capped_to_traffic": capped}
    return out, src


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def ref_cpu_steps(scene, state, n_settle, n_warm, n_steps, seconds):
    """The reference's CPU path on the host cores ("_ref+glue": oracle/_ref = the reference's own collision translation
    units + its extracted force function under the reference's OpenMP loop + Thrust-OMP sort/reduce, restated glue for
    the Eigen/ChBody parts; oracle/ref_driver.cpp ref_smc_step).  Bounded: the step counts shrink to the time budget."""
    from oracle import pyoracle

    ref = pyoracle.Ref()
    cores = host_threads()
    ref.set_num_threads(cores)  # (torchrun exports OMP_NUM_THREADS=1: the arm sets its own team size)
    sc = dict(scene)
    if state is not None:
        sc["pos"], sc["rot"], sc["vel"] = state
    ref.smc_setup(sc)
    t0 = time.time()
    ref.smc_step(1)
    t1 = max(time.time() - t0, 1e-6)
    settle = int(max(0, min(n_settle, 0.35 * seconds / t1 - 1)))
    warm = int(max(1, min(n_warm, 0.15 * seconds / t1)))
    steps = int(max(2, min(n_steps, 0.5 * seconds / t1)))
    ref.smc_step(settle)
    ref.smc_step(warm)
    ref.smc_setup(dict(sc, **dict(zip(("pos", "rot", "vel"), ref.smc_state()))))  # (resets the phase timers)
    t0 = time.time()
    ref.smc_step(steps)
    dt = time.time() - t0
    nb = scene["mass"].size
    phases = {k: v * 1e3 for k, v in ref.smc_timers().items()}
    cn = ref.counts()
    return {"value": nb * steps / dt, "unit": METRIC, "cores": cores, "kind": "_ref+glue", "ms_per_step": dt / steps * 1e3,
            "phase_ms": phases, "steps": steps, "settle_steps": 1 + settle + warm, "N_c": cn["N_c"], "N_p": cn["N_p"],
            "N_i": cn["N_i"]}


def cpu_baseline_sample(scene, state, seconds_budget=25.0):
    r = ref_cpu_steps(scene, state, 0, 1, 20, seconds_budget)
    r["sample"] = (f"{r['steps']} steps from the same {scene['mass'].size}-body bed state as the timed window "
                   f"(reference collision TUs + extracted force function + Thrust-OMP accumulation, OpenMP {r['cores']} "
                   f"threads), {r['ms_per_step']:.0f} ms/step")
    return r


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.config == "nsc":
        return run_reference_nsc(args)
    scene = make_scene(args)
    nb = scene["mass"].size
    r = ref_cpu_steps(scene, None, args.settle, args.warmup, args.steps, args.ref_seconds)
    sample = (f"{r['steps']} timed steps after {r['settle_steps']} settle + warm-up steps of the SAME {workload_name(args)} bed "
              f"(step counts bounded by --ref-seconds {args.ref_seconds:.0f}; N_c = {r['N_c']} in the window), "
              f"reference collision TUs + extracted function_CalcContactForces + Thrust-OMP sort/reduce + restated glue, "
              f"OpenMP {r['cores']} threads")
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": METRIC, "n_gpus": args.gpus,
            "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": dict(workload_config(args, None, nb), N_i=r["N_i"], N_p=r["N_p"], N_c=r["N_c"]),
            "cpu_baseline": {"value": r["value"], "unit": METRIC, "cores": r["cores"], "kind": "_ref+glue", "sample": sample,
                             "phase_ms": r["phase_ms"]},
            "e2e": {"value": r["value"], "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_reference_nsc(args):
    """configs[3] on the host cores.  The NSC linear algebra of the reference is Eigen (absent here), so this
    arm is the CPU restatement (oracle/liboracle.so, kind "port") over the same scene."""
    from chrono_wisc_b200 import scenes
    from oracle import pyoracle

    scene = scenes.balls_nsc(args.nsc_nx, args.nsc_ny, args.nsc_nz, iterations=args.nsc_iters)
    nb = scene["mass"].size
    orc = pyoracle.Oracle(scene)
    cores = host_threads()
    orc.set_threads(cores)
    t0 = time.time()
    orc.step(1)
    t1 = max(time.time() - t0, 1e-6)
    steps = int(max(1, min(args.steps, 0.6 * args.ref_seconds / t1)))
    t0 = time.time()
    orc.step(steps)
    dt = time.time() - t0
    value = nb * steps / dt
    sample = (f"{steps} steps of the same {nb - 1}-sphere NSC bed, APGD {args.nsc_iters} iterations/step, CPU restatement "
              f"(Eigen absent: the reference's own NSC code cannot be built), OpenMP {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus, "steps": steps,
            "warmup": 1, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": nsc_config(args, nb, None),
            "cpu_baseline": {"value": value, "unit": METRIC, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def workload_name(args):
    if args.config == "mixed":
        return f"{args.mixed_nx}x{args.mixed_ny}x{args.mixed_nz} mixed sphere/box/capsule"
    return f"{args.nx}x{args.ny}x{args.nz} = {args.nx * args.ny * args.nz}-sphere"


def workload_config(args, counts, nb):
    if args.config == "mixed":
        wl = (f"{args.mixed_nx}x{args.mixed_ny}x{args.mixed_nz} = {nb - 1} bodies, one third spheres r=0.1 / boxes "
              "(0.1,0.08,0.06) / capsules r=0.06, Hertz SMC, E=2e5, mu=0.4, cr=0.1, settling in a 5-wall box, HYBRID "
              "narrowphase (PRIMS + MPR), FP64, h=1e-3, FIXED_BIN_SIZE 3 radii (BASELINE.json configs[4] without the "
              "motor-driven blade)")
    else:
        wl = (f"{args.nx}x{args.ny}x{args.nz} = {args.nx * args.ny * args.nz} monodisperse spheres r=0.02 on a jittered "
              f"lattice (spacing {args.spacing} r, jitter 0.005 r, seed 12345), Hertz SMC settling in a 5-wall box, FP64, "
              "h=1e-3, " + (f"FIXED_BIN_SIZE {args.bin_edge} diameters" if args.bin_edge > 0 else
                            "FIXED_DENSITY 5 shapes/bin") + " (BASELINE.json configs[1])")
    cfg = {"workload": wl, "bodies_per_gpu": nb, "settle_steps": args.settle,
           "l2": "inputs larger than L2 (body+contact state >> 126 MB); no flush needed"}
    if counts is not None:
        cfg.update({"N_i": counts.num_bin_aabb_intersections, "N_p": counts.num_possible_collisions,
                    "N_c": counts.num_rigid_contacts, "bins": list(counts.bins_per_axis)})
    return cfg


def nsc_config(args, nb, counts):
    cfg = {"workload": f"{args.nsc_nx}x{args.nsc_ny}x{args.nsc_nz} = {nb - 1} spheres r=0.15, m=1, mu=0.4, envelope 0.01, "
                       f"NSC DVI, SLIDING, alpha=0, recovery speed 1e4, APGD exactly {args.nsc_iters} iterations/step, "
                       "h=1e-2, FIXED_DENSITY 5 shapes/bin (BASELINE.json configs[3], demo_MCORE_ballsNSC-style)",
           "bodies_per_gpu": nb, "settle_steps": args.nsc_settle,
           "l2": "the iteration's working set (contact frames + vectors) is larger than L2 at this size"}
    if counts is not None:
        cfg.update({"N_c": counts.num_rigid_contacts, "iterations": counts.solver_iterations,
                    "residual": counts.solver_residual})
    return cfg


def kinetic_energy_per_particle(scene, vel):
    m, I = scene["mass"][1:], scene["inertia"][1:]
    v = vel[1:]
    return float((0.5 * m * (v[:, :3] ** 2).sum(1) + 0.5 * (I * v[:, 3:] ** 2).sum(1)).mean())


def run_b200(args):
    import torch
    import torch.distributed as dist

    import chrono_wisc_b200 as cw
    from chrono_wisc_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the chrono_wisc_b200 hot path has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lib = _lib.load()
    if args.config == "nsc":
        return run_b200_nsc(args, rank, world, local, barrier)
    if world > 1:
        if args.config != "smc":
            raise SystemExit("bench.py: --config mixed is a single-GPU workload")
        return run_b200_slabs(args, rank, world, local, barrier)
    scene = make_scene(args, rank, world)
    nb, ns = scene["mass"].size, scene["shape_type"].size
    sys_ = cw.load_scene(cw.ChSystemMulticoreSMC(local), scene)

    # ---- window 1: the falling / colliding transient, from step 0 --------------------------------
    sys_.DoSteps(args.warmup)
    sys_.Synchronize()
    n_tr = max(1, min(args.steps, args.settle))
    sys_.DoSteps(n_tr)
    sys_.Synchronize()
    tr_ms = sys_.GetTimer("do_steps") * 1e3
    tr_counts = sys_.GetCounts()
    sys_.DoSteps(max(0, args.settle - n_tr))
    sys_.DoSteps(args.warmup)
    sys_.Synchronize()

    # ---- window 2, the headline: K steps of the contact-rich bed, CUDA events on the library's own stream --------
    sampler = ClockSampler(local)
    sampler.start()
    l0 = sys_.KernelLaunches()
    barrier()
    t0 = time.perf_counter()
    sys_.DoSteps(args.steps)
    sys_.Synchronize()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.result()
    launches = sys_.KernelLaunches() - l0
    dev_ms = sys_.GetTimer("do_steps") * 1e3  # CUDA events bracketing all K steps on the library stream, per step
    counts = sys_.GetCounts()
    stage_ms = {n: sys_.GetTimer(n) * 1e3 for n in ("update", "collision_broad", "collision_narrow", "advance",
                                                   "integrate")}
    value = nb * args.steps / (dev_ms * 1e-3 * args.steps)
    # per-kernel timers of THIS window (read before the e2e and late windows enqueue further timed steps)
    kr, traffic_src = kernel_rooflines(sys_, nb, ns, counts, measured_peak_gbs()[0])

    # ---- e2e: same steps through the public API with HOST state: upload pose+velocity, step, download ----------
    n_e2e = max(1, min(args.steps, args.e2e_steps))
    pin = []

    def pinned(shape):
        n = int(np.prod(shape)) * 8
        p = C.c_void_p()
        _lib.check(lib.cb200_host_alloc(n, C.byref(p)))
        pin.append(p)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(int(np.prod(shape)),)).reshape(shape)

    hp, hq, hv = pinned((nb, 3)), pinned((nb, 4)), pinned((nb, 6))
    p0, q0, v0 = sys_.GetBodyState()
    hp[:], hq[:], hv[:] = p0, q0, v0
    barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        sys_.StepHost(hp, hq, hv)  # cb200_step_host: H2D pose + velocity, one step, D2H pose + velocity
    barrier()
    e2e_wall = time.perf_counter() - t0
    e2e_value = nb * n_e2e / e2e_wall
    for p in pin:
        lib.cb200_host_free(p)

    peak, peak_kind = measured_peak_gbs()
    bytes_step = algorithmic_bytes(nb, counts)
    achieved = bytes_step / (dev_ms * 1e-3) / 1e9
    ke = kinetic_energy_per_particle(scene, v0)

    # ---- window 3: the settled bed, kinetic energy per particle below 1e-9 J -- or as close as
    # --settle-max further steps bring it (untimed, a look at the energy every 500), then K' timed steps.  With the
    # OneStep tangential model at h = 1e-3 the criterion is out of reach for the reference's scheme as for this one:
    # the translational energy falls to ~1e-5 J, the rotational one stays at ~4.6e-4 J (explicit friction at its Coulomb
    # limit flips the spin of a loaded sphere every step).  What the window does show is the
    # COMPACTED bed: a quarter more contacts and thousands of bins of more than 64 entries.
    settled = None
    if args.settle_max > 0:
        ke_s, more = ke, 0
        while ke_s >= 1e-9 and more < args.settle_max:
            sys_.DoSteps(500)
            more += 500
            sys_.Synchronize()
            ke_s = kinetic_energy_per_particle(scene, sys_.GetBodyState()[2])
        n_s = max(1, min(args.steps, 100))
        sys_.DoSteps(args.warmup)
        sys_.Synchronize()
        sys_.DoSteps(n_s)
        sys_.Synchronize()
        s_ms = sys_.GetTimer("do_steps") * 1e3
        settled = {"extra_settle_steps": more, "steps": n_s, "ms_per_step": s_ms, "value": nb / (s_ms * 1e-3),
                   "N_c": sys_.GetCounts().num_rigid_contacts, "ke_per_particle_J": ke_s,
                   "settled_by_SURVEY_8d_criterion": bool(ke_s < 1e-9)}
    line = {
        "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms, "wall_ms_per_step": wall / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, counts, nb),
        "windows": {
            "transient": {"steps": f"{args.warmup}..{args.warmup + n_tr}", "ms_per_step": tr_ms,
                          "value": nb / (tr_ms * 1e-3), "N_c": tr_counts.num_rigid_contacts},
            "timed": {"steps": f"{2 * args.warmup + args.settle}..{2 * args.warmup + args.settle + args.steps}",
                      "ms_per_step": dev_ms, "N_c": counts.num_rigid_contacts, "ke_per_particle_J": ke,
                      "settled_by_SURVEY_8d_criterion": bool(ke < 1e-9)},
            **({"late": settled} if settled else {})},
        "stage_ms": stage_ms,
        "stage_ms_note": "stage and kernel timers are CUDA-event pairs of the directly enqueued (non-graph) steps only, "
                         "one per 64-step chunk; ms_per_step covers all K steps, graph replays included",
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": nb * 13 * 8,
                "d2h_bytes_per_step": nb * 13 * 8, "steps": n_e2e},
        "gpu_launches": int(launches),
        "roofline_step": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                          "frac": achieved / peak, "peak_kind": peak_kind, "kernel": "whole step (all kernels)",
                          "algorithmic_bytes_per_step": bytes_step},
    }
    line["kernels"] = kr
    line["kernels_traffic_source"] = traffic_src
    if kr:
        dom = max(kr, key=lambda k: kr[k]["ms"])
        line["roofline"] = {"bound": "hbm", "achieved": kr[dom]["achieved"], "peak": peak, "unit": "GB/s",
                            "frac": kr[dom]["frac"], "traffic": kr[dom]["traffic"], "peak_kind": peak_kind,
                            "kernel": dom, "ms_per_launch": kr[dom]["ms"],
                            "algorithmic_bytes_per_launch": kr[dom]["algorithmic_bytes"]}
    else:
        line["roofline"] = dict(line["roofline_step"], traffic=None)
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_sample(scene, (p0, q0, v0), args.cpu_seconds)
    print(json.dumps(line))


def run_b200_nsc(args, rank, world, local, barrier):
    """BASELINE.json configs[3]: NSC DVI ball bed, APGD with a fixed iteration count per step."""
    import chrono_wisc_b200 as cw
    from chrono_wisc_b200 import _lib, scenes

    if world > 1:
        return run_b200_nsc_slabs(args, rank, world, local, barrier)
    lib = _lib.load()
    scene = scenes.balls_nsc(args.nsc_nx, args.nsc_ny, args.nsc_nz, iterations=args.nsc_iters)
    nb = scene["mass"].size
    sys_ = cw.load_scene(cw.ChSystemMulticoreNSC(local), scene)
    sys_.DoSteps(args.nsc_settle)
    sys_.DoSteps(min(args.warmup, 5))
    sys_.Synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    steps = max(1, min(args.steps, args.nsc_steps))
    l0 = sys_.KernelLaunches()
    barrier()
    t0 = time.perf_counter()
    sys_.DoSteps(steps)
    sys_.Synchronize()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.result()
    launches = sys_.KernelLaunches() - l0
    dev_ms = sys_.GetTimer("do_steps") * 1e3
    counts = sys_.GetCounts()
    adv_ms = sys_.GetTimer("advance") * 1e3
    nc, it = counts.num_rigid_contacts, max(1, counts.solver_iterations)
    # e2e: host pose + velocity in, one step, out
    pin = []

    def pinned(shape):
        n = int(np.prod(shape)) * 8
        p = C.c_void_p()
        _lib.check(lib.cb200_host_alloc(n, C.byref(p)))
        pin.append(p)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(int(np.prod(shape)),)).reshape(shape)

    hp, hq, hv = pinned((nb, 3)), pinned((nb, 4)), pinned((nb, 6))
    p0, q0, v0 = sys_.GetBodyState()
    hp[:], hq[:], hv[:] = p0, q0, v0
    n_e2e = max(1, min(steps, 5))
    barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        sys_.StepHost(hp, hq, hv)
    barrier()
    e2e_value = nb * n_e2e / (time.perf_counter() - t0)
    for p in pin:
        lib.cb200_host_free(p)
    peak, peak_kind = measured_peak_gbs()
    # per APGD iteration 2 Schur products + the fused vector algebra; per step the setup on top
    bytes_it = 2 * (nc * (2 * 144 + 48) + nb * 96) + 240 * nc
    bytes_step = it * bytes_it + nc * (96 + 144 + 72)
    ach_it = bytes_it / (adv_ms * 1e-3 / it) / 1e9
    line = {"metric": METRIC, "value": nb * steps / (dev_ms * 1e-3 * steps), "unit": METRIC, "n_gpus": 1, "steps": steps,
            "warmup": min(args.warmup, 5), "ms_per_step": dev_ms, "wall_ms_per_step": wall / steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": nsc_config(args, nb, counts), "clocks": clocks,
            "stage_ms": {n: sys_.GetTimer(n) * 1e3 for n in ("update", "collision_broad", "collision_narrow", "advance",
                                                           "integrate")},
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": nb * 13 * 8, "d2h_bytes_per_step": nb * 13 * 8,
                    "steps": n_e2e},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": ach_it, "peak": peak, "unit": "GB/s", "frac": ach_it / peak,
                         "traffic": None, "peak_kind": peak_kind,
                         "kernel": "one APGD iteration (k_nsc_scatter/gather x2 + fused vector kernels)",
                         "us_per_iteration": adv_ms * 1e3 / it, "algorithmic_bytes_per_iteration": bytes_it},
            "roofline_step": {"bound": "hbm", "achieved": bytes_step / (dev_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": bytes_step / (dev_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_step": bytes_step}}
    if not args.no_cpu:
        from oracle import pyoracle

        orc = pyoracle.Oracle(dict(scene, pos=p0, rot=q0, vel=v0))
        cores = host_threads()
        orc.set_threads(cores)
        t0 = time.time()
        orc.step(1)
        t1 = time.time() - t0
        n = int(max(1, min(5, args.cpu_seconds / max(t1, 1e-6) - 1)))
        t0 = time.time()
        orc.step(n)
        dt = time.time() - t0
        line["cpu_baseline"] = {"value": nb * n / dt, "unit": METRIC, "cores": cores, "kind": "port",
                                "sample": f"{n} steps of the same bed state, CPU restatement of the NSC path (Eigen absent), "
                                          f"OpenMP {cores} threads, {dt / n * 1e3:.0f} ms/step"}
    print(json.dumps(line))


def run_b200_nsc_slabs(args, rank, world, local, barrier):
    """configs[3] at N > 1: the SAME 262,144-sphere NSC bed cut into N slabs along x (strong scaling).  Every Schur
    product exchanges the body-space vector of the ghost bodies with the neighbours, every APGD slot ends in one NCCL
    all-reduce of the 16 scalars (csrc/comm.cu, nsc.cu)."""
    import torch
    import torch.distributed as dist

    from chrono_wisc_b200 import scenes, slab

    scene = scenes.balls_nsc(args.nsc_nx, args.nsc_ny, args.nsc_nz, iterations=args.nsc_iters)
    nb = scene["mass"].size
    spacing = 2.05 * 0.15
    x_of = lambda i: (i - (args.nsc_nx - 1) / 2) * spacing
    cuts = [x_of(round(g * args.nsc_nx / world)) - spacing / 2 for g in range(world + 1)]
    cuts[0], cuts[-1] = -1e3, 1e3
    layout = slab.SlabLayout(world, rank, cuts, axis=0, width=0.35)
    sys_ = slab.SlabSystemNSC(layout, device=local, exchange=slab.LibraryExchange(layout), own_stream=True)
    sys_.LoadSlab(scene)
    sys_.ExchangeGhosts()
    sys_.DoSteps(args.nsc_settle)
    sys_.DoSteps(min(args.warmup, 5))
    sys_.Check()
    sys_.ResetTimers()
    steps = max(1, min(args.steps, args.nsc_steps))
    sampler = ClockSampler(local)
    sampler.start()
    l0 = sys_.KernelLaunches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.perf_counter()
    e0.record(sys_.stream)
    sys_.DoSteps(steps)
    e1.record(sys_.stream)
    torch.cuda.synchronize()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.result()
    sys_.Check()
    launches = sys_.KernelLaunches() - l0
    dev_ms = e0.elapsed_time(e1) / steps
    counts = sys_.GetCounts()
    adv_ms = sys_.GetTimer("advance") * 1e3
    t = torch.tensor([dev_ms, wall, adv_ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall, adv_ms = (float(x) for x in t)
    own = torch.tensor([sys_.GetOwned()[0].size, counts.num_rigid_contacts], dtype=torch.float64, device="cuda")
    dist.all_reduce(own)
    assert int(own[0].item()) == nb - 1, "slabs lost or duplicated bodies"
    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        it = max(1, counts.solver_iterations)
        nc_rank, nb_rank = counts.num_rigid_contacts, sys_.GetOwned()[0].size + sys_.GetGhostIds().size
        bytes_it = 2 * (nc_rank * (2 * 144 + 48) + nb_rank * 96) + 240 * nc_rank
        ach = bytes_it / (adv_ms * 1e-3 / it) / 1e9
        cfg = nsc_config(args, nb, counts)
        cfg.update({"parallelism": f"slab{world}", "N_c_rank0": nc_rank, "N_c_sum_over_ranks": int(own[1].item()),
                    "ghost_width": 0.35, "per_slot": "2 halo exchanges of 48 B per ghost body + 1 all-reduce of 16 doubles"})
        line = {"metric": METRIC, "value": (nb - 1) * steps / (dev_ms * 1e-3 * steps), "unit": METRIC, "n_gpus": world,
                "steps": steps, "warmup": min(args.warmup, 5), "ms_per_step": dev_ms, "wall_ms_per_step": wall / steps * 1e3,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": (nb - 1) * steps / wall, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                        "note": "device-resident slabs; wall clock of the same steps incl. the per-step host sync"},
                "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                             "peak_kind": peak_kind, "kernel": "one APGD iteration of rank 0's slab, halo + all-reduce included",
                             "us_per_iteration": adv_ms * 1e3 / it, "algorithmic_bytes_per_iteration": bytes_it}}
        print(json.dumps(line))
    dist.destroy_process_group()


def run_b200_slabs(args, rank, world, local, barrier):
    """N > 1: BASELINE.json configs[2] -- one bed of world * nx x ny x nz spheres cut into slabs along x, one slab per
    GPU, ghost layers exchanged with NCCL send/recv after every step (weak scaling: 1M spheres per GPU)."""
    import torch
    import torch.distributed as dist

    from chrono_wisc_b200 import _lib, scenes, slab

    lib = _lib.load()
    scene, gids, cuts, width = scenes.settling_smc_slab(rank, world, args.nx, args.ny, args.nz, seed=12345,
                                                        bin_edge=args.bin_edge if args.bin_edge > 0 else 2.5,
                                                        spacing_factor=args.spacing, jitter_factor=0.005)
    spacing = args.spacing * 0.02
    # records per exchange buffer: the band holds ny * nz * width / spacing bodies (4,900 at the default sizes, fewer
    # than 5,500 once the bed has compacted); the buffers travel at capacity (the counts live on the device), so the
    # head-room is paid on the link at every step: twice the expected count, not the five-fold of round 1
    ghost_cap = int(2.0 * args.ny * args.nz * (width / spacing)) + 1024
    layout = slab.SlabLayout(world, rank, cuts, axis=0, width=width)
    n_owned0 = gids.size
    # the data plane (pack, NCCL send/recv, unpack) is enqueued by the library itself (csrc/comm.cu)
    sys_ = slab.SlabSystemSMC(layout, device=local, exchange=slab.LibraryExchange(layout), ghost_cap=ghost_cap,
                              own_stream=True,
                              owned_cap=int(n_owned0 * 1.1) + 4096)
    nb_static = 1
    ns_static = int((scene["shape_body"] < nb_static).sum())
    sys_.LoadPart(scene, gids, nb_static, ns_static)
    nb_total = world * n_owned0 + 1
    sys_.ExchangeGhosts()  # initial ghost layers
    sys_.DoSteps(args.settle)
    sys_.DoSteps(args.warmup)
    sys_.Check()
    sys_.ResetTimers()

    sampler = ClockSampler(local)
    sampler.start()
    l0 = sys_.KernelLaunches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0 = time.perf_counter()
    e0.record(sys_.stream)
    sys_.DoSteps(args.steps)
    e1.record(sys_.stream)
    torch.cuda.synchronize()
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.result()
    sys_.Check()
    launches = sys_.KernelLaunches() - l0
    dev_ms = e0.elapsed_time(e1) / args.steps  # step + pack + NCCL exchange + unpack, on the one stream they share
    counts = sys_.GetCounts()
    stage_ms = {n: sys_.GetTimer(n) * 1e3 for n in ("update", "collision_broad", "collision_narrow", "advance",
                                                   "integrate")}
    t = torch.tensor([dev_ms, wall], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall = float(t[0].item()), float(t[1].item())
    value = (nb_total - 1) * args.steps / (dev_ms * 1e-3 * args.steps)
    owned = torch.tensor([sys_.GetOwned()[0].size, sys_.GetGhostIds().size], dtype=torch.float64, device="cuda")
    tot = owned.clone()
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    assert int(tot[0].item()) == nb_total - 1, "slabs lost or duplicated bodies"

    # e2e: host pose + velocity of every slot in, one step with exchange, host state out (pinned buffers)
    nbs = sys_.num_bodies
    n_e2e = max(1, min(args.steps, args.e2e_steps))
    pin = []

    def pinned(shape):
        n = int(np.prod(shape)) * 8
        p = C.c_void_p()
        _lib.check(lib.cb200_host_alloc(n, C.byref(p)))
        pin.append(p)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(int(np.prod(shape)),)).reshape(shape)

    hp, hq, hv = pinned((nbs, 3)), pinned((nbs, 4)), pinned((nbs, 6))
    p0, q0, v0 = sys_.GetBodyState()
    hp[:], hq[:], hv[:] = p0, q0, v0
    barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        sys_.SetBodyState(0, hp, hq, hv, nbs)
        sys_.DoSteps(1)
        _lib.check(lib.cb200_download_state(sys_._h, 0, nbs, hp.ctypes.data_as(C.c_void_p),
                                            hq.ctypes.data_as(C.c_void_p), hv.ctypes.data_as(C.c_void_p)))
    barrier()
    e2e_wall = time.perf_counter() - t0
    t = torch.tensor([e2e_wall], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = (nb_total - 1) * n_e2e / float(t.item())
    for p in pin:
        lib.cb200_host_free(p)

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        bytes_step = algorithmic_bytes(n_owned0 + int(owned[1].item()), counts)
        achieved = bytes_step / (dev_ms * 1e-3) / 1e9
        cfg = workload_config(args, counts, n_owned0)
        cfg["workload"] = (f"{world * args.nx}x{args.ny}x{args.nz} = {world * n_owned0} monodisperse spheres r=0.02, Hertz SMC "
                           f"settling in a 5-wall box, FP64, h=1e-3, FIXED_BIN_SIZE {args.bin_edge} diameters, {world} slabs "
                           "along x with NCCL ghost-layer exchange every step (BASELINE.json configs[2] at 1M spheres/GPU)")
        cfg.update({"parallelism": f"slab{world}", "ghost_width": width, "ghost_bodies_rank0": int(owned[1].item()),
                    "exchange_bytes_per_step_per_neighbour": int(8 * (2 + 33 * ghost_cap))})
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms, "wall_ms_per_step": wall / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg, "stage_ms": stage_ms, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": nbs * 13 * 8 * world,
                    "d2h_bytes_per_step": nbs * 13 * 8 * world, "steps": n_e2e},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_kind": peak_kind, "kernel": "whole step of rank 0 (all kernels + exchange)",
                         "algorithmic_bytes_per_step": bytes_step},
        }
        print(json.dumps(line))
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="smc", choices=["smc", "nsc", "mixed"],
                    help="smc: BASELINE.json configs[1]/[2] (the metric's config); nsc: configs[3]; mixed: configs[4]")
    ap.add_argument("--nx", type=int, default=250)
    ap.add_argument("--ny", type=int, default=250)
    ap.add_argument("--nz", type=int, default=16)
    ap.add_argument("--spacing", type=float, default=2.0, help="lattice spacing in sphere radii")
    ap.add_argument("--bin-edge", type=float, default=2.5,
                    help="0: FIXED_DENSITY 5 shapes/bin; >0: FIXED_BIN_SIZE with this edge in sphere diameters")
    ap.add_argument("--settle", type=int, default=300, help="untimed steps before warm-up (bed forms its contacts)")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--settle-max", type=int, default=3000,
                    help="single-GPU smc: extra untimed steps (at most) towards the settled-bed window; 0 skips it")
    ap.add_argument("--cpu-seconds", type=float, default=25.0)
    ap.add_argument("--ref-seconds", type=float, default=150.0, help="--impl reference: wall-clock budget of the arm")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--nsc-nx", type=int, default=128)
    ap.add_argument("--nsc-ny", type=int, default=128)
    ap.add_argument("--nsc-nz", type=int, default=16)
    ap.add_argument("--nsc-iters", type=int, default=100)
    ap.add_argument("--nsc-settle", type=int, default=20)
    ap.add_argument("--nsc-steps", type=int, default=20, help="cap on the timed NSC steps (25 ms each)")
    ap.add_argument("--mixed-nx", type=int, default=180)
    ap.add_argument("--mixed-ny", type=int, default=180)
    ap.add_argument("--mixed-nz", type=int, default=16)
    args = ap.parse_args()
    if args.config == "mixed" and args.settle == 300:
        args.settle = 2500  # the mixed bed needs longer to come to rest on the floor
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
