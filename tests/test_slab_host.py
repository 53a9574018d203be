"""Host side of the slab decomposition without a GPU: layout arithmetic, scene splitting, and the neighbour exchange
over torch.distributed (gloo, world_size 2 and 3 on CPU tensors) driven by a numpy emulation of the device pack /
unpack passes (tests-only restatement of csrc/slab.cu's rules)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_layout_and_ownership():
    from chrono_wisc_b200.slab import SlabLayout

    L = SlabLayout(3, 1, (-5.0, -1.0, 2.0, 5.0), axis=0, width=0.1)
    assert (L.left, L.right, L.lo, L.hi) == (0, 2, -1.0, 2.0)
    x = np.array([-7.0, -1.0, -1.0001, 1.9999, 2.0, 9.0])
    assert list(L.owner_of(x)) == [0, 1, 0, 1, 2, 2]  # outer slabs extend to infinity, planes belong to the right
    assert list(L.in_ghost_band_of(x, 0)) == [True, True, True, False, False, False]
    with pytest.raises(ValueError):
        SlabLayout(2, 0, (0.0, 0.1, 1.0), width=0.1)
    with pytest.raises(ValueError):
        SlabLayout(2, 2, (0.0, 1.0, 2.0))


def test_split_scene_partitions_the_bed():
    from chrono_wisc_b200 import scenes
    from chrono_wisc_b200.slab import SlabLayout, split_scene

    scene = scenes.settling_smc(12, 4, 2)
    seen = []
    for r in range(3):
        part, gids, nbs, nss = split_scene(scene, SlabLayout(3, r, (-9.0, -0.1, 0.1, 9.0), width=0.05))
        assert (nbs, nss) == (1, 5)
        assert part["mass"].size == 1 + gids.size and part["shape_type"].size == 5 + gids.size
        assert np.array_equal(part["pos"][1:], scene["pos"][gids])
        assert np.array_equal(part["shape_body"][5:], np.arange(1, 1 + gids.size))
        seen.append(gids)
    assert np.array_equal(np.sort(np.concatenate(seen)), np.arange(1, scene["mass"].size))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from chrono_wisc_b200.slab import HDR, REC, DistExchange, SlabLayout

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cuts = np.linspace(-1.5, 1.5, world + 1)
        L = SlabLayout(world, rank, cuts, width=0.1)
        ex = DistExchange(L)
        rng = np.random.default_rng(5)  # same on all ranks: every rank knows the global truth
        n = 400
        x = rng.uniform(-1.4, 1.4, n)
        v = rng.normal(scale=0.6, size=n)
        owned = {int(i) for i in np.nonzero(L.owns(x))[0]}
        cap = 128
        bufs = [torch.zeros(HDR + REC * cap, dtype=torch.float64) for _ in range(4)]
        sl, sr, rl, rr = bufs
        ok = True
        for step in range(40):
            x = x + 0.02 * v
            v = np.where(np.abs(x) > 1.4, -v, v)
            # pack (csrc/slab.cu k_slab_flags/k_slab_pack): records in slot (= id) order
            for buf, nb_rank, sel in ((sl, L.left, lambda c: c < L.lo + L.width), (sr, L.right, lambda c: c >= L.hi - L.width)):
                ids = [i for i in sorted(owned) if nb_rank is not None and sel(x[i])]
                buf[0] = len(ids)
                for k, i in enumerate(ids):
                    buf[HDR + REC * k] = float(i)
                    buf[HDR + REC * k + 1] = x[i]
            stay = {i for i in owned if (L.left is None or x[i] >= L.lo) and (L.right is None or x[i] < L.hi)}
            ghosts = owned - stay  # a body handed over lingers one step as a ghost (k_slab_pack)
            owned = stay
            ex(sl, sr, rl, rr)
            for buf in (rl, rr):
                for k in range(int(buf[0])):
                    i, xi = int(buf[HDR + REC * k]), float(buf[HDR + REC * k + 1])
                    ok &= xi == x[i]
                    if (L.left is None or xi >= L.lo) and (L.right is None or xi < L.hi):
                        owned.add(i)
                    else:
                        ghosts.add(i)
            want_owned = {int(i) for i in np.nonzero(L.owns(x))[0]}
            want_ghosts = {int(i) for i in range(n) if i not in want_owned and L.lo - L.width <= x[i] < L.hi + L.width}
            ok &= owned == want_owned and ghosts == want_ghosts
        q.put((rank, bool(ok), len(owned)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_exchange_over_gloo(world):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
    assert sum(n for _, _, n in res) == 400


def _nsc_worker(rank, world, port, q):
    """The algebra of the decomposed NSC solve (csrc/comm.cu + nsc.cu) restated in numpy over gloo: contacts in the
    overlap are duplicated, the body-space vector w = M^-1 D x of a ghost body comes from its owner (halo BROADCAST, no
    sum), dot products count every contact on the rank that owns its dynamic body with the smaller global id, and one
    all-reduce gives every rank the global scalars.  Checked against the undecomposed product N x = D^T M^-1 D x."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from chrono_wisc_b200.slab import SlabLayout

    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(17)  # same on all ranks: every rank can form the global truth
        n, width = 120, 0.12
        xs = np.sort(rng.uniform(-1.4, 1.4, n))
        L = SlabLayout(world, rank, np.linspace(-1.5, 1.5, world + 1), width=width)
        pairs = [(i, j) for i in range(n) for j in range(i + 1, n) if xs[j] - xs[i] <= width]  # "contacts"
        m = len(pairs)
        J = rng.normal(size=(m, 2, 6))          # the two 1x6 blocks of a contact's row of D^T
        Minv = rng.uniform(0.5, 2.0, size=(n, 6))
        x = rng.normal(size=m)                  # multipliers, replicated wherever the contact is visible
        # ---- global truth
        w = np.zeros((n, 6))
        for c, (i, j) in enumerate(pairs):
            w[i] += J[c, 0] * x[c]
            w[j] += J[c, 1] * x[c]
        w *= Minv
        Nx = np.array([J[c, 0] @ w[i] + J[c, 1] @ w[j] for c, (i, j) in enumerate(pairs)])
        # ---- this rank: owned bodies + ghost band, the contacts among them
        owned = L.owns(xs)
        visible = owned | ((xs >= L.lo - width) & (xs < L.hi + width))
        mine = [c for c, (i, j) in enumerate(pairs) if visible[i] and visible[j]]
        wl = np.zeros((n, 6))
        for c in mine:
            i, j = pairs[c]
            wl[i] += J[c, 0] * x[c]
            wl[j] += J[c, 1] * x[c]
        wl *= Minv
        assert np.allclose(wl[owned], w[owned], rtol=0, atol=1e-13)  # every contact of an owned body is visible
        # halo: owners send the rows of their bodies inside the neighbour's band, ghosts take them
        for nb, band in ((L.left, lambda c: c < L.lo + width), (L.right, lambda c: c >= L.hi - width)):
            if nb is None:
                continue
            ids = np.nonzero(owned & band(xs))[0]
            send = torch.from_numpy(np.concatenate([[float(ids.size)], ids.astype(float), wl[ids].ravel()]))
            recv = torch.zeros(1 + 7 * n, dtype=torch.float64)
            pad = torch.zeros(1 + 7 * n, dtype=torch.float64)
            pad[:send.numel()] = send
            ops = [dist.P2POp(dist.isend, pad, nb), dist.P2POp(dist.irecv, recv, nb)]
            for r in dist.batch_isend_irecv(ops):
                r.wait()
            k = int(recv[0])
            gid = recv[1:1 + k].numpy().astype(int)
            wl[gid] = recv[1 + k:1 + 7 * k].numpy().reshape(k, 6)
        Nx_l = np.array([J[c, 0] @ wl[pairs[c][0]] + J[c, 1] @ wl[pairs[c][1]] for c in mine])
        ok = np.allclose(Nx_l, Nx[mine], rtol=0, atol=1e-12)
        # dot product with every contact counted once: owner of the body with the smaller global id
        own_c = np.array([owned[min(pairs[c])] for c in mine])
        part = torch.tensor([float((x[mine] * Nx_l)[own_c].sum()), float(own_c.sum())], dtype=torch.float64)
        dist.all_reduce(part)
        ok &= abs(part[0].item() - float(x @ Nx)) <= 1e-10 * abs(float(x @ Nx)) and int(part[1].item()) == m
        q.put((rank, bool(ok), int(own_c.sum())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_nsc_halo_and_allreduce_algebra_over_gloo(world):
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nsc_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok, _ in res), res
