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


KERNEL_TIMERS = ("k_pair_masks", "k_smc_force", "k_body_reduce", "k_contact_gen", "k_pair_store", "k_bin_sort")


def kernel_rooflines(sys_, nb, c, peak):
    """Per-kernel roofline entries for the heaviest kernels: live CUDA-event durations (events around the launch, on
    the library stream, in the directly enqueued steps of the timed region) against each kernel's own share of the
    SURVEY.md 8(d) algorithmic bytes.  `traffic` = DRAM bytes per launch from the committed
    ncu --set full capture (profiles/r1_traffic.json), when the capture holds that kernel."""
    ni, np_, nc = c.num_bin_aabb_intersections, c.num_possible_collisions, c.num_rigid_contacts
    alg = {
        "k_pair_masks": ni * 16 + np_ * 16,          # entries read; the pair list it decides but never materialises
        # pair key + two shape poses (centre, radius) + two body states in; body/material ids, force, torques out
        # (the contact geometry is evaluated in the kernel, no contact record is read in the step)
        "k_smc_force": nc * (8 + 2 * 32 + 2 * 112 + 16 + 72),
        "k_body_reduce": nb * 96 + nc * 2 * 56,      # per-body accumulators + (index, force, torque) per incidence
        "k_contact_gen": nc * (8 + 96),              # key in, contact record out
        "k_pair_store": ni * 16 + nc * 8,            # masks + offsets in, keys out
        "k_bin_sort": ni * 16,                       # (bin, shape) key/value moved once
    }
    traffic = {}
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch", {})
    except Exception:
        pass
    out = {}
    for k in KERNEL_TIMERS:
        try:
            ms = sys_.GetTimer(k) * 1e3
        except Exception:
            continue
        if ms <= 0:
            continue
        a = alg[k] / (ms * 1e-3) / 1e9
        out[k] = {"ms": ms, "algorithmic_bytes": alg[k], "achieved": a, "frac": a / peak, "traffic": traffic.get(k)}
    return out


def cpu_baseline_sample(scene, seconds_budget=20.0):
    """Oracle ("port": oracle/liboracle.so, OpenMP over the host cores) on a bounded sample: the same bed state,
    as many whole steps as fit the budget (at least 1)."""
    from oracle import pyoracle

    orc = pyoracle.Oracle(scene)
    cores = orc.max_threads()
    orc.set_threads(cores)
    nb = scene["mass"].size
    t0 = time.time()
    orc.step(1)
    t1 = time.time() - t0
    n = int(max(1, min(50, seconds_budget / max(t1, 1e-6) - 1)))
    t0 = time.time()
    orc.step(n)
    dt = time.time() - t0
    return {"value": nb * n / dt, "unit": METRIC, "cores": cores, "kind": "port",
            "sample": f"{n} steps of the same {nb}-body bed state (oracle/liboracle.so, OpenMP {cores} threads), "
                      f"{dt / n * 1e3:.1f} ms/step", "ms_per_step": dt / n * 1e3}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyoracle

    scene = make_scene(args)
    # the reference arm times a bounded sample: a bed of --ref-nx x --ref-ny x nz spheres (same materials, spacing,
    # grid rule and step) so that W + K steps end within minutes on the host cores
    small = dict(vars(args))
    sargs = argparse.Namespace(**small)
    sargs.nx, sargs.ny = args.ref_nx, args.ref_ny
    scene = make_scene(sargs)
    nb = scene["mass"].size
    orc = pyoracle.Oracle(scene)
    cores = orc.max_threads()
    orc.set_threads(cores)
    orc.step(min(args.settle, args.ref_settle))
    orc.step(args.warmup)
    t0 = time.time()
    orc.step(args.steps)
    dt = time.time() - t0
    value = nb * args.steps / dt
    sample = (f"{args.steps} steps of a {sargs.nx}x{sargs.ny}x{sargs.nz} = {nb - 1}-sphere bed (same materials/grid rule "
              f"as the {args.nx}x{args.ny}x{args.nz} workload), oracle/liboracle.so OpenMP {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, None, nb),
            "cpu_baseline": {"value": value, "unit": METRIC, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def workload_config(args, counts, nb):
    cfg = {"workload": f"{args.nx}x{args.ny}x{args.nz} = {args.nx * args.ny * args.nz} monodisperse spheres r=0.02, "
                       "Hertz SMC settling in a 5-wall box, FP64, h=1e-3, " +
                       (f"FIXED_BIN_SIZE {args.bin_edge} diameters" if args.bin_edge > 0 else
                        "FIXED_DENSITY 5 shapes/bin") + " (BASELINE.json configs[1])",
           "bodies_per_gpu": nb, "settle_steps": args.settle,
           "l2": "inputs larger than L2 (body+contact state >> 126 MB); no flush needed"}
    if counts is not None:
        cfg.update({"N_i": counts.num_bin_aabb_intersections, "N_p": counts.num_possible_collisions,
                    "N_c": counts.num_rigid_contacts, "bins": list(counts.bins_per_axis)})
    return cfg


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
    if world > 1:
        return run_b200_slabs(args, rank, world, local, barrier)
    scene = make_scene(args, rank, world)
    nb = scene["mass"].size
    sys_ = cw.load_scene(cw.ChSystemMulticoreSMC(local), scene)
    sys_.DoSteps(args.settle)
    sys_.DoSteps(args.warmup)
    sys_.Synchronize()

    # ---- device-resident timed region: K steps, CUDA events on the library's own stream ("step" timer) ----------
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
    # the job's time is the slowest rank's device time over the K steps (CUDA events, host sync gaps included)
    t = torch.tensor([dev_ms, wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall = float(t[0].item()), float(t[1].item())
    value = nb * world * args.steps / (dev_ms * 1e-3 * args.steps)

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
    t = torch.tensor([e2e_wall], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = nb * world * n_e2e / float(t.item())
    for p in pin:
        lib.cb200_host_free(p)

    if rank == 0:
        peak, peak_kind = measured_peak_gbs()
        bytes_step = algorithmic_bytes(nb, counts)
        achieved = bytes_step / (dev_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms, "wall_ms_per_step": wall / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, counts, nb),
            "stage_ms": stage_ms, "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": METRIC, "h2d_bytes_per_step": nb * 13 * 8,
                    "d2h_bytes_per_step": nb * 13 * 8, "steps": n_e2e},
            "gpu_launches": int(launches),
            "roofline_step": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                              "frac": achieved / peak, "peak_kind": peak_kind, "kernel": "whole step (all kernels)",
                              "algorithmic_bytes_per_step": bytes_step},
        }
        kr = kernel_rooflines(sys_, nb, counts, peak)
        line["kernels"] = kr
        if kr:
            dom = max(kr, key=lambda k: kr[k]["ms"])
            line["roofline"] = {"bound": "hbm", "achieved": kr[dom]["achieved"], "peak": peak, "unit": "GB/s",
                                "frac": kr[dom]["frac"], "traffic": kr[dom]["traffic"], "peak_kind": peak_kind,
                                "kernel": dom, "ms_per_launch": kr[dom]["ms"],
                                "algorithmic_bytes_per_launch": kr[dom]["algorithmic_bytes"]}
        else:
            line["roofline"] = dict(line["roofline_step"], traffic=None)
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline_sample(dict(scene, pos=p0, rot=q0, vel=v0), args.cpu_seconds)
        print(json.dumps(line))
    if world > 1:
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
    ghost_cap = int(2.5 * args.ny * args.nz * (width / spacing + 1)) + 1024
    layout = slab.SlabLayout(world, rank, cuts, axis=0, width=width)
    n_owned0 = gids.size
    sys_ = slab.SlabSystemSMC(layout, device=local, exchange=slab.DistExchange(layout), ghost_cap=ghost_cap,
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
    ap.add_argument("--nx", type=int, default=250)
    ap.add_argument("--ny", type=int, default=250)
    ap.add_argument("--nz", type=int, default=16)
    ap.add_argument("--spacing", type=float, default=2.0, help="lattice spacing in sphere radii")
    ap.add_argument("--bin-edge", type=float, default=2.5,
                    help="0: FIXED_DENSITY 5 shapes/bin; >0: FIXED_BIN_SIZE with this edge in sphere diameters")
    ap.add_argument("--settle", type=int, default=300, help="untimed steps before warm-up (bed forms its contacts)")
    ap.add_argument("--e2e-steps", type=int, default=20)
    ap.add_argument("--cpu-seconds", type=float, default=20.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--ref-nx", type=int, default=100)
    ap.add_argument("--ref-ny", type=int, default=100)
    ap.add_argument("--ref-settle", type=int, default=50)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
