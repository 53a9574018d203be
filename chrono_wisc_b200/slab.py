"""Slab decomposition of one granular system across the GPUs of a box (BASELINE.json configs[2]).

One process per GPU.  Rank g owns the bodies whose centre lies in ``[cuts[g], cuts[g+1])`` along ``axis``; static
geometry (inactive bodies: container walls) is replicated.  Every step each rank

  1. advances its slab (``cb200_enqueue_steps``: forces, collision, contact solve, integration on owned + ghost bodies;
     ghost bodies are inactive, so contacts with them push the owned body only -- no force exchange),
  2. packs the bodies within ``width`` of a cut plane (``cb200_slab_pack_dev``, device buffers),
  3. swaps the buffers with its two neighbours (``torch.distributed`` send/recv: NCCL over NVLink on the GPU box,
     gloo in the CPU tests of the exchange layer),
  4. rebuilds its ghost layer and adopts bodies that crossed into its interval (``cb200_slab_unpack_dev``).

Everything is enqueued on one CUDA stream (torch's current stream is handed to the library), so a step has no host
synchronisation; ``Check()`` synchronises and raises if a capacity overflowed.

The reference has no counterpart (Chrono::Multicore is single-node OpenMP); the interface mirrors
``ChSystemMulticoreSMC`` so a driver switches by constructing ``SlabSystemSMC`` instead.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import check
from .system import ChSystemMulticoreNSC, ChSystemMulticoreSMC, load_scene

REC = 33   # doubles per record (csrc/slab.cu)
HDR = 2


class SlabLayout:
    """Cut planes and neighbours of a 1-D decomposition."""

    def __init__(self, world, rank, cuts, axis=0, width=0.1):
        cuts = [float(c) for c in cuts]
        if len(cuts) != world + 1 or any(b <= a for a, b in zip(cuts, cuts[1:])):
            raise ValueError("cuts must be world+1 increasing plane positions")
        if not 0 <= rank < world:
            raise ValueError("rank out of range")
        if min(b - a for a, b in zip(cuts, cuts[1:])) < 2 * width:
            raise ValueError("a slab is thinner than two ghost widths")
        self.world, self.rank, self.axis, self.width, self.cuts = world, rank, int(axis), float(width), cuts
        self.lo, self.hi = cuts[rank], cuts[rank + 1]
        self.left = rank - 1 if rank > 0 else None
        self.right = rank + 1 if rank < world - 1 else None

    def owner_of(self, coord):
        """Rank owning a centre coordinate (the outermost slabs extend to infinity)."""
        inner = np.asarray(self.cuts[1:-1])
        return np.searchsorted(inner, coord, side="right")

    def owns(self, coord):
        return self.owner_of(coord) == self.rank

    def in_ghost_band_of(self, coord, neighbour):
        """Bodies of this rank that the given neighbour (rank-1 or rank+1) needs as ghosts."""
        c = np.asarray(coord)
        if neighbour == self.rank - 1:
            return c < self.lo + self.width
        if neighbour == self.rank + 1:
            return c >= self.hi - self.width
        raise ValueError("not a neighbour")


def split_scene(scene, layout, gids=None):
    """The part of a global scene rank ``layout.rank`` starts with: all static (inactive) bodies with their shapes,
    then the active bodies it owns, one shape each.  Returns (scene, gids of the dynamic bodies, nb_static, ns_static)."""
    active = scene["active"].astype(bool)
    nb_static = int((~active).sum())
    if active[:nb_static].any():
        raise ValueError("static (inactive) bodies must come first")
    sb = scene["shape_body"]
    static_shape = sb < nb_static
    ns_static = int(static_shape.sum())
    if static_shape[ns_static:].any() or not np.array_equal(sb[ns_static:], np.arange(nb_static, scene["mass"].size)):
        raise ValueError("dynamic bodies need exactly one shape each, in body order, after the static shapes")
    if gids is None:
        gids = np.arange(scene["mass"].size, dtype=np.int64)
    mine = np.zeros(scene["mass"].size, bool)
    mine[nb_static:] = layout.owns(scene["pos"][nb_static:, layout.axis])
    keep_b = mine.copy()
    keep_b[:nb_static] = True
    keep_s = np.concatenate([np.ones(ns_static, bool), mine[nb_static:]])
    out = dict(scene)
    for k in ("mass", "inertia", "pos", "rot", "vel", "active", "collide"):
        out[k] = scene[k][keep_b]
    for k in ("shape_type", "shape_dims", "shape_pos", "shape_rot", "shape_fam", "shape_mat"):
        out[k] = scene[k][keep_s]
    n_dyn = int(mine.sum())
    out["shape_body"] = np.concatenate([sb[:ns_static], np.arange(nb_static, nb_static + n_dyn)]).astype(np.uint32)
    return out, gids[mine], nb_static, ns_static


def nccl_library_path():
    """The NCCL shared library the process already uses (torch's bundled copy), for cb200_comm_init_nccl."""
    import glob
    import os

    import torch

    base = os.path.dirname(os.path.dirname(torch.__file__))
    hits = sorted(glob.glob(os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so*")))
    return hits[0] if hits else "libnccl.so.2"


class LibraryExchange:
    """The data plane in the library (csrc/comm.cu): ghost exchange, NSC halo and all-reduce are NCCL calls the
    library itself enqueues on the system's stream.  torch.distributed is used once, to hand rank 0's NCCL id round."""

    def __init__(self, layout):
        self.layout = layout

    def attach(self, system):
        import torch
        import torch.distributed as dist

        lib, L = system._lib, self.layout
        path = nccl_library_path().encode()
        dev = torch.device("cuda", system.device)
        ident = torch.zeros(128, dtype=torch.uint8, device=dev)
        if L.rank == 0:
            buf = (C.c_char * 128)()
            check(lib.cb200_comm_unique_id(path, C.cast(buf, C.c_void_p)))
            ident = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev)
        if L.world > 1:
            dist.broadcast(ident, 0)
        raw = bytes(ident.cpu().tolist())
        check(lib.cb200_comm_init_nccl(system._h, path, L.rank, L.world, C.cast(C.c_char_p(raw), C.c_void_p)))


class LocalGroup:
    """All slabs of a decomposed system in ONE process, on one stream (tests, single-GPU emulation): the library steps
    them in lockstep and the transfers are device copies (cb200_comm_init_local)."""

    def __init__(self, systems):
        self.systems = list(systems)
        s0 = self.systems[0]
        for s in self.systems[1:]:
            check(s._lib.cb200_set_stream(s._h, C.c_void_p(s0.stream.cuda_stream), 0))
            s.stream = s0.stream
        self._arr = (C.c_void_p * len(self.systems))(*[s._h for s in self.systems])
        check(s0._lib.cb200_comm_init_local(self._arr, len(self.systems)))

    def Exchange(self):
        check(self.systems[0]._lib.cb200_comm_exchange(self._arr, len(self.systems)))

    def DoSteps(self, n):
        for s in self.systems:
            s._push_settings()
        check(self.systems[0]._lib.cb200_comm_steps(self._arr, len(self.systems), int(n)))
        for s in self.systems:
            s.ch_time += n * s.settings.step_size


class DistExchange:
    """Neighbour exchange over ``torch.distributed`` (NCCL on GPUs, gloo on CPU tensors)."""

    def __init__(self, layout, group=None):
        import torch.distributed as dist

        self.dist, self.layout, self.group = dist, layout, group

    def __call__(self, send_left, send_right, recv_left, recv_right):
        dist, L = self.dist, self.layout
        ops = []
        if L.left is not None:
            ops.append(dist.P2POp(dist.isend, send_left, L.left, self.group))
            ops.append(dist.P2POp(dist.irecv, recv_left, L.left, self.group))
        if L.right is not None:
            ops.append(dist.P2POp(dist.isend, send_right, L.right, self.group))
            ops.append(dist.P2POp(dist.irecv, recv_right, L.right, self.group))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()  # NCCL: orders the current CUDA stream after the transfer, does not block the host


class LocalExchange:
    """All ranks live in one process (tests, single-GPU emulation): buffers are copied between the systems directly.
    Call ``run(systems)`` once per step instead of the per-rank callable."""

    @staticmethod
    def run(systems):
        for a, b in zip(systems, systems[1:]):
            b.recv_left.copy_(a.send_right)
            a.recv_right.copy_(b.send_left)


class _SlabMixin:
    """Stepping with ghost exchange on top of a ChSystemMulticore{SMC,NSC}."""

    def _slab_init(self, layout, exchange=None, owned_cap=None, ghost_cap=None, own_stream=False):
        self.layout = layout
        # own_stream: step, pack/unpack and the NCCL transfers run on a dedicated (capturable) CUDA stream, which lets
        # the library replay its step as a CUDA graph; otherwise everything is ordered on torch's current stream
        self._own_stream, self.stream = bool(own_stream), None
        self.exchange = exchange
        self._owned_cap, self._ghost_cap = owned_cap, ghost_cap
        self.nb_static = self.ns_static = 0
        self.send_left = self.send_right = self.recv_left = self.recv_right = None

    def LoadSlab(self, global_scene, gids=None):
        """Load this rank's part of a global scene and set up the slab (buffers on the system's device)."""
        import torch

        L = self.layout
        part, my_gids, nb_static, ns_static = split_scene(global_scene, L, gids)
        if self._ghost_cap is None:
            # must be the same on every rank (it sizes the exchange buffers): four times the mean population of a
            # band of the ghost width in the global scene, plus head room
            x = global_scene["pos"][nb_static:, L.axis]
            span = max(float(x.max() - x.min()), 2 * L.width)
            self._ghost_cap = int(4 * x.size * L.width / span) + 4096
        return self.LoadPart(part, my_gids, nb_static, ns_static)

    def LoadPart(self, part, my_gids, nb_static, ns_static):
        """Load a scene that already holds only this rank's share: ``nb_static`` static bodies (``ns_static`` shapes)
        followed by the owned bodies, one shape each; ``my_gids`` are their global ids.  ``ghost_cap`` must have been
        given to the constructor (the same value on every rank)."""
        import torch

        L = self.layout
        if self._ghost_cap is None:
            raise ValueError("ghost_cap must be set (identical on all ranks) when loading a per-rank scene")
        self.nb_static, self.ns_static = int(nb_static), int(ns_static)
        load_scene(self, part)
        my_gids = np.ascontiguousarray(my_gids, dtype=np.int64)
        n_dyn = my_gids.size
        owned_cap = self._owned_cap or int(n_dyn * 1.25) + 1024
        ids = np.full(self.num_bodies, -1, np.int64)
        ids[self.nb_static:] = my_gids
        check(self._lib.cb200_upload_body_ids(self._h, 0, self.num_bodies, ids.ctypes.data_as(C.c_void_p)))
        self._push_settings()
        check(self._lib.cb200_slab_configure(self._h, self.nb_static, self.ns_static, L.axis, L.lo, L.hi, L.width,
                                             int(L.left is not None), int(L.right is not None), owned_cap,
                                             self._ghost_cap))
        self._owned_cap = owned_cap
        self.num_bodies = self.nb_static + owned_cap + 2 * self._ghost_cap
        self.num_shapes = self.ns_static + owned_cap + 2 * self._ghost_cap
        n = int(self._lib.cb200_slab_buffer_doubles(self._h))
        dev = torch.device("cuda", self.device)
        with torch.cuda.device(dev):
            self.send_left, self.send_right, self.recv_left, self.recv_right = (
                torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(4))
            # everything the library enqueues goes on one torch stream: ordered with the NCCL transfers
            self.stream = torch.cuda.Stream(dev) if self._own_stream else torch.cuda.current_stream(dev)
            if self._own_stream:
                self.stream.wait_stream(torch.cuda.current_stream(dev))
            check(self._lib.cb200_set_stream(self._h, C.c_void_p(self.stream.cuda_stream), 0))
        if isinstance(self.exchange, LibraryExchange):
            self.exchange.attach(self)
            self._one = (C.c_void_p * 1)(self._h)
        return self

    def _on_stream(self):
        import contextlib

        import torch

        return torch.cuda.stream(self.stream) if self._own_stream else contextlib.nullcontext()

    def _ptr(self, t, side):
        return C.c_void_p(t.data_ptr()) if side is not None else None

    def Pack(self):
        L = self.layout
        check(self._lib.cb200_slab_pack_dev(self._h, self._ptr(self.send_left, L.left), self._ptr(self.send_right, L.right)))

    def Unpack(self):
        L = self.layout
        check(self._lib.cb200_slab_unpack_dev(self._h, self._ptr(self.recv_left, L.left), self._ptr(self.recv_right, L.right)))

    def ExchangeGhosts(self):
        if isinstance(self.exchange, LibraryExchange):
            check(self._lib.cb200_comm_exchange(self._one, 1))
            return
        with self._on_stream():
            self.Pack()
            if self.exchange is not None:
                self.exchange(self.send_left, self.send_right, self.recv_left, self.recv_right)
            self.Unpack()

    def DoSteps(self, n):
        """n steps, each followed by the ghost exchange; asynchronous (call Check() or Synchronize())."""
        self._push_settings()
        if isinstance(self.exchange, LibraryExchange):
            check(self._lib.cb200_comm_steps(self._one, 1, int(n)))  # step + exchange enqueued by the library
        else:
            with self._on_stream():
                for _ in range(int(n)):
                    check(self._lib.cb200_enqueue_steps(self._h, 1))
                    self.ExchangeGhosts()
        self.ch_time += n * self.settings.step_size

    def StepOnly(self):
        self._push_settings()
        check(self._lib.cb200_enqueue_steps(self._h, 1))

    def Check(self):
        check(self._lib.cb200_check(self._h))

    def ResetTimers(self):
        check(self._lib.cb200_reset_timers(self._h))

    def GetOwned(self):
        """(gids, pos, rot, vel) of the bodies this rank owns, sorted by global id."""
        nb = self.num_bodies
        ids, fl = np.empty(nb, np.int64), np.empty(nb, np.uint8)
        check(self._lib.cb200_download_body_ids(self._h, 0, nb, ids.ctypes.data_as(C.c_void_p),
                                                fl.ctypes.data_as(C.c_void_p)))
        pos, rot, vel = self.GetBodyState()
        own = (fl & 1).astype(bool) & ~(fl & 4).astype(bool)
        own[:self.nb_static] = False
        order = np.argsort(ids[own], kind="stable")
        return ids[own][order], pos[own][order], rot[own][order], vel[own][order]

    def GetGhostIds(self):
        nb = self.num_bodies
        ids, fl = np.empty(nb, np.int64), np.empty(nb, np.uint8)
        check(self._lib.cb200_download_body_ids(self._h, 0, nb, ids.ctypes.data_as(C.c_void_p),
                                                fl.ctypes.data_as(C.c_void_p)))
        g = ((fl & 7) == 2)
        g[:self.nb_static] = False
        return np.sort(ids[g])


class SlabSystemSMC(_SlabMixin, ChSystemMulticoreSMC):
    def __init__(self, layout, device=0, exchange=None, owned_cap=None, ghost_cap=None, own_stream=False):
        ChSystemMulticoreSMC.__init__(self, device)
        self._slab_init(layout, exchange, owned_cap, ghost_cap, own_stream)


class SlabSystemNSC(_SlabMixin, ChSystemMulticoreNSC):
    """One slab of an NSC (DVI) system.  The APGD solve runs over all slabs at once: the exchange must be the
    library's (``LibraryExchange`` over NCCL, or a ``LocalGroup`` when all slabs live in one process)."""

    def __init__(self, layout, device=0, exchange=None, owned_cap=None, ghost_cap=None, own_stream=False):
        ChSystemMulticoreNSC.__init__(self, device)
        self._slab_init(layout, exchange, owned_cap, ghost_cap, own_stream)
