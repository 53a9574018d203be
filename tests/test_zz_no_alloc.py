"""The allocation-free MINRES entry point (cb200_minres_noalloc: scratch in a static device array of the library)
against the existing path (cb200_set_bilaterals + the solver stage + cb200_download_bilateral_gamma), through the C ABI,
on two calls in a row.  Collected last on purpose (file name).

The existing path builds N = D^T M^-1 D and R = -b - D^T M_invk on the device; the new entry point is handed the same
system assembled in numpy (oracle/np_oracle.py, which the bilateral tests already pin the device against) in caller-owned
device buffers.  Tolerance: 1e-9 relative, the bar of tests/test_gpu_bilateral.py."""
import ctypes as C

import numpy as np
import pytest

import chrono_wisc_b200 as cw
from chrono_wisc_b200 import scenes
from chrono_wisc_b200._lib import Cb200Error, check

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def bed(seed=33):
    s = scenes.settling_smc(8, 8, 3, spacing_factor=1.98, jitter_factor=0.03, seed=seed)
    s["pos"][1:, 2] -= 0.025 * 0.02
    rng = np.random.default_rng(seed)
    n = s["mass"].size
    s["vel"][1:, :3] = rng.normal(scale=0.05, size=(n - 1, 3))
    s["vel"][1:, 3:] = rng.normal(scale=2.0, size=(n - 1, 3))
    return s


def rows_for(scene, nrows, seed):
    nb = scene["mass"].size
    rng = np.random.default_rng(seed)
    ids = np.stack([rng.integers(0, nb, nrows), rng.integers(0, nb, nrows)], 1).astype(np.int32)
    same = ids[:, 0] == ids[:, 1]
    ids[same, 1] = (ids[same, 1] + 1) % nb
    cq_a, cq_b = rng.normal(size=(nrows, 6)), rng.normal(size=(nrows, 6))
    cq_a[:, 3:] *= 0.02
    cq_b[:, 3:] *= 0.02
    return ids, cq_a, cq_b, rng.normal(scale=0.1, size=nrows)


def assemble(scene, hf, ids, cq_a, cq_b, rhs):
    """N and R of ChIterativeSolverMulticoreSMC::RunTimeStep for BODY_BODY rows (np_oracle.bilateral_solve's algebra)."""
    nb, n = scene["mass"].size, rhs.size
    act = scene["active"].astype(bool)
    minv = np.concatenate([np.repeat(1.0 / scene["mass"][:, None], 3, 1), 1.0 / scene["inertia"].reshape(nb, 3)], 1)
    minv = minv * act[:, None]
    M_invk = scene["vel"] + minv * hf
    D = np.zeros((n, 6 * nb))
    for i in range(n):
        D[i, 6 * ids[i, 0]:6 * ids[i, 0] + 6] += cq_a[i]
        D[i, 6 * ids[i, 1]:6 * ids[i, 1] + 6] += cq_b[i]
    R = -rhs - D @ M_invk.reshape(-1)
    N = (D * minv.reshape(-1)) @ D.T
    return np.ascontiguousarray(N), R


def existing_path(scene, rows, iters, tol):
    sys = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    sys.SetBilaterals(*rows, max_iteration_bilateral=iters, tol_speed=tol)
    sys.Update()
    sys.RunCollision(7)
    sys.RunSolver()
    return sys.GetBilateralGamma(rows[3].size)


def test_noalloc_minres_matches_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    scene = bed()
    nb = scene["mass"].size
    free = cw.load_scene(cw.ChSystemMulticoreSMC(), scene)
    free.Update()
    free.RunCollision(7)
    free.RunSolver()
    hf = free.GetHf().reshape(nb, 6)

    cases = [(40, 200, 1e-10, 1), (6, 100, 1e-12, 2), (40, 200, 1e-10, 1)]  # the third repeats the first: stale scratch
    want = [existing_path(scene, rows_for(scene, n, seed), it, tol) for n, it, tol, seed in cases]

    sys = cw.ChSystemMulticoreSMC()  # an empty system: it only lends its device and stream
    lib = sys._lib
    assert lib.cb200_minres_noalloc_max_rows() == 2048
    dev = torch.device("cuda", 0)
    bufs = []
    for n, it, tol, seed in cases:
        N, R = assemble(scene, hf, *rows_for(scene, n, seed))
        bufs.append((torch.from_numpy(N).to(dev), torch.from_numpy(R).to(dev),
                     torch.full((n,), float("nan"), dtype=torch.float64, device=dev),
                     torch.zeros(2, dtype=torch.float64, device=dev)))
    torch.cuda.synchronize()
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    for (n, it, tol, seed), (Nd, Rd, xd, info) in zip(cases, bufs):  # back to back on the system's stream, no sync
        check(lib.cb200_minres_noalloc(sys._h, n, p(Nd), p(Rd), p(xd), it, tol, p(info)))
    sys.Synchronize()
    for (n, it, tol, seed), (Nd, Rd, xd, info), (g, g_it, g_res) in zip(cases, bufs, want):
        x, inf = xd.cpu().numpy(), info.cpu().numpy()
        assert np.all(np.isfinite(x))
        assert np.abs(x - g).max() <= RTOL * np.abs(g).max()
        assert int(inf[0]) == g_it
        assert inf[1] < max(tol, 1e-16) or int(inf[0]) == it
    # the same problem gave the same answer before and after another call used the scratch
    assert np.array_equal(bufs[0][2].cpu().numpy(), bufs[2][2].cpu().numpy())

    # beyond the fixed scratch: refused with a clear error, nothing launched
    l0 = sys.KernelLaunches()
    with pytest.raises(Cb200Error, match="exceed the fixed scratch"):
        check(lib.cb200_minres_noalloc(sys._h, 2049, p(bufs[0][0]), p(bufs[0][1]), p(bufs[0][2]), 10, 1e-6, p(bufs[0][3])))
    assert sys.KernelLaunches() == l0
    # n = 0 is a no-op that still reports
    check(lib.cb200_minres_noalloc(sys._h, 0, None, None, None, 10, 1e-6, p(bufs[1][3])))
    sys.Synchronize()
    assert bufs[1][3].cpu().numpy()[0] == 0
