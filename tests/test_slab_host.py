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
