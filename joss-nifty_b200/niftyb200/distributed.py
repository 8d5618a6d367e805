"""Slab-decomposed correlated field across the GPUs of one box (one process per GPU).

The reference has no domain decomposition (its MPI layer shards *samples*,
``/root/reference/src/utilities.py:295-390``, ``src/minimization/kl_energies.py:136-141``);
this module adds what ``BASELINE.json:north_star`` asks for: a 3-D grid split across ranks, ONE
all-to-all (NCCL over NVLink) per harmonic transform and an all-reduce of a few fp64 scalars
per CG inner product.

Decomposition.  The genuine Hartley transform needs, for every row ``(k0, k1, :)``, its mirror
row ``(-k0, -k1, :)`` (see ``csrc/nb2_passes.cuh``).  Each rank therefore owns a
*mirror-symmetric* set of planes of the split axis: rank ``r`` of ``P`` owns
``k in [r h, (r+1) h)`` and their mirrors ``N - k`` (``h = N / 2P``; rank 0 also holds the
self-mirrored planes ``0`` and ``N/2``).  Two layouts are used:

* layout X (harmonic side, e.g. ``xi``): axis 0 split, axes 1, 2 complete -- local ``(N0/P, N1, N2)``
* layout Y (position side, e.g. the signal and the likelihood weight): axis 1 split -- local ``(N0, N1/P, N2)``

A transform X -> Y is: strided pass over axis 1 (local, storing straight into the send blocks) ->
all-to-all -> strided pass over axis 0 (local, loading straight from the receive blocks) -> row pass (mirror in k0 by index negation, in k1 by reflection inside the local
symmetric part); Y -> X is the same with the roles of the axes exchanged.  Hartley is symmetric,
so X -> Y serves the forward apply and Y -> X its adjoint.
"""

import ctypes
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from .transforms import Plan, PointOp, RawPointer, device as _device


class SymmetricSplit:
    """Mirror-symmetric ownership of an axis of length ``N`` by ``P`` ranks."""

    def __init__(self, N, P):
        if N % (2 * P) != 0 or N // (2 * P) < 1:
            raise ValueError(f"axis length {N} must be a multiple of 2 * {P}")
        self.N, self.P, self.h = N, P, N // (2 * P)

    def owned(self, r):
        """Global indices owned by rank ``r`` in local order; local mirror of ``l`` is
        ``(2h - l - shift) mod 2h`` with ``shift = 0`` on rank 0 and ``1`` elsewhere."""
        N, h = self.N, self.h
        if r == 0:
            return np.concatenate((np.arange(h), [N // 2], np.arange(N - h + 1, N))).astype(np.int64)
        return np.concatenate((np.arange(r * h, (r + 1) * h), np.arange(N - (r + 1) * h + 1, N - r * h + 1))).astype(np.int64)

    def shift(self, r):
        return 0 if r == 0 else 1


class Comm:
    """The two collectives of the path on a ``torch.distributed`` group."""

    def __init__(self, group=None):
        self.group = group
        self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)

    def allreduce_(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def allreduce_max_(self, t):
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t

    def broadcast_flag(self, flag):
        """Rank 0's boolean, on every rank (stopping decisions that depend on a local clock)."""
        t = torch.tensor([1.0 if flag else 0.0], dtype=torch.float32, device=_device())
        dist.broadcast(t, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0, group=self.group)
        return bool(t.item() != 0.0)

    def all_to_all(self, recv, send, async_op=False):
        """Returns the work handle when ``async_op`` (the caller ``wait()``s before reading)."""
        return dist.all_to_all_single(recv, send, group=self.group, async_op=async_op)


class PeerExchange:
    """Receive buffers of every rank mapped into every process of the box (CUDA IPC), so that the
    strided pass before the exchange stores straight into the peers' HBM over NVLink: the
    all-to-all is part of the pass (``nb2_strided_pass_remap`` with row offsets that point into
    the peers' buffers).  Two buffers alternate between consecutive transforms; one all-reduce of
    a single scalar per transform is the barrier that tells a rank its blocks have arrived."""

    NBUF = 2

    def __init__(self, comm, nbytes):
        self.lib, self.comm, self.nbytes = _lib.get(), comm, int(nbytes)
        self.own, self.peers, self.turn = [], [], 0
        handles = []
        for _ in range(self.NBUF):
            p = ctypes.c_void_p()
            self.lib.check(self.lib.nb2_device_alloc(self.nbytes, ctypes.byref(p)), "nb2_device_alloc")
            h = ctypes.create_string_buffer(64)
            self.lib.check(self.lib.nb2_ipc_export(p, h), "nb2_ipc_export")
            self.own.append(p.value)
            handles.append(h.raw)
        everyone = [None] * comm.size
        dist.all_gather_object(everyone, handles, group=comm.group)
        for b in range(self.NBUF):
            row = []
            for q in range(comm.size):
                if q == comm.rank:
                    row.append(self.own[b])
                    continue
                p = ctypes.c_void_p()
                self.lib.check(self.lib.nb2_ipc_open(ctypes.create_string_buffer(everyone[q][b], 64), ctypes.byref(p)),
                               "nb2_ipc_open")
                row.append(p.value)
            self.peers.append(row)
        self.flag = torch.zeros(1, dtype=torch.float32, device=_device())

    def rel(self, b, q, elem):
        """Offset of rank q's buffer b relative to this rank's buffer b, in elements of `elem` bytes."""
        d = self.peers[b][q] - self.own[b]
        assert d % elem == 0
        return d // elem

    def mine(self, b):
        return RawPointer(self.own[b])

    def arrive(self):
        """Stream-ordered barrier: returns (on the stream) once every rank's pass has completed."""
        self.comm.allreduce_(self.flag)

    def close(self):
        for b, row in enumerate(self.peers):
            for q, p in enumerate(row):
                if q != self.comm.rank:
                    self.lib.nb2_ipc_close(p)
        for p in self.own:
            self.lib.nb2_device_free(p)
        self.peers, self.own = [], []


class SlabHartley:
    """Distributed genuine 3-D Hartley transform between layouts X and Y."""

    def __init__(self, shape, dtype, comm, chunks=None):
        if len(shape) != 3:
            raise ValueError("slab decomposition is implemented for 3-D grids")
        self.shape, self.dtype, self.comm = tuple(int(s) for s in shape), dtype, comm
        N0, N1, N2 = self.shape
        P, r = comm.size, comm.rank
        self.s0, self.s1 = SymmetricSplit(N0, P), SymmetricSplit(N1, P)
        dev = _device()
        self.own0 = [torch.from_numpy(self.s0.owned(q)).to(dev) for q in range(P)]
        self.own1 = [torch.from_numpy(self.s1.owned(q)).to(dev) for q in range(P)]
        self.n0, self.n1 = 2 * self.s0.h, 2 * self.s1.h
        self.shape_x, self.shape_y = (self.n0, N1, N2), (N0, self.n1, N2)
        self.plan_x, self.plan_y = Plan(self.shape_x, None, dtype), Plan(self.shape_y, None, dtype)
        if not (self.plan_x.is_fast and self.plan_y.is_fast):
            raise ValueError("slab decomposition needs power-of-two extents")
        self.plan_x.set_mirror(self.s0.shift(r), 0)
        self.plan_y.set_mirror(0, self.s1.shift(r))
        # The exchange is pipelined in `C` chunks of complex columns (last axis): the strided pass
        # before it stores chunk c straight into its send blocks, the all-to-all of chunk c runs on
        # NCCL's stream while chunk c+1 is being transformed, and the strided pass after it loads
        # straight from the receive blocks (nb2_strided_pass_remap) -- no pack / unpack sweeps.
        M = N2 // 2
        # peer-memory exchange (GPUs of one box): the pass before the exchange stores into the peers
        use_p2p = os.environ.get("NIFTYB200_SLAB_P2P", "1") != "0" and P > 1 and dev.type == "cuda"
        if use_p2p:
            chunks = 1
        if chunks is None:
            chunks = int(os.environ.get("NIFTYB200_SLAB_CHUNKS", "0"))
        if not chunks:
            local_bytes = self.n0 * N1 * N2 * torch.empty((), dtype=dtype).element_size()
            chunks = 4 if (local_bytes >= (64 << 20) and M >= 256) else 1
        if chunks < 1 or M % chunks or (M // chunks) & (M // chunks - 1):
            raise ValueError(f"chunks={chunks} must split the {M} complex columns into powers of two")
        self.chunks, self.mc, self.m = chunks, M // chunks, M
        blk = (chunks, P, self.n0, self.n1, 2 * self.mc)
        self.send = None if use_p2p else torch.empty(blk, dtype=dtype, device=dev)
        self.recv = None if use_p2p else torch.empty(blk, dtype=dtype, device=dev)
        # row-offset tables (complex elements) into one chunk's blocks [peer][l0][l1][m];
        # global plane k0 = own0[q][l0], k1 = own1[q][l1]
        t0 = np.empty(N0, dtype=np.int64)
        t1 = np.empty(N1, dtype=np.int64)
        for q in range(P):
            t0[self.s0.owned(q)] = (q * self.n0 + np.arange(self.n0)) * self.n1 * self.mc
            t1[self.s1.owned(q)] = (q * self.n0 * self.n1 + np.arange(self.n1)) * self.mc
        self.rowoff0, self.rowoff1 = torch.from_numpy(t0).to(dev), torch.from_numpy(t1).to(dev)
        self.ostride1 = self.n1 * self.mc

        self.peer = None
        if use_p2p:
            esz = torch.empty((), dtype=dtype).element_size()
            self.peer = PeerExchange(comm, P * self.n0 * self.n1 * N2 * esz)
            # The sending pass writes every row into its owner's buffer IN THE OWNER'S NATURAL LAYOUT
            # (nb2_strided_pass_scatter), so the receiving pass is a plain pass (TMA kernel, no table):
            #   X -> Y: row k1 of local plane l0 -> rank q = owner(k1), element (k0, l1, m) of its (N0, n1, M)
            #   Y -> X: row k0 of local column block l1 -> rank q = owner(k0), element (l0, k1, m) of its (n0, N1, M)
            self.p2p_x, self.p2p_y = [], []       # per buffer: out-row tables of the pass before the exchange
            for b in range(PeerExchange.NBUF):
                tx = np.empty(N1, dtype=np.int64)
                ty = np.empty(N0, dtype=np.int64)
                for q in range(P):
                    rel = self.peer.rel(b, q, 2 * esz)
                    tx[self.s1.owned(q)] = rel + np.arange(self.n1) * M
                    ty[self.s0.owned(q)] = rel + np.arange(self.n0) * N1 * M
                self.p2p_x.append(torch.from_numpy(tx).to(dev))
                self.p2p_y.append(torch.from_numpy(ty).to(dev))
            # origin of local plane l0 (global k0) in a Y-layout buffer / of local column block l1 (global k1)
            # in an X-layout buffer
            self.p2p_x_outer = torch.from_numpy(self.s0.owned(r) * self.n1 * M).to(dev)
            self.p2p_y_col = torch.from_numpy(self.s1.owned(r) * M).to(dev)

    def _exchange(self, before, after):
        """before(c) fills send[c]; after(c) consumes recv[c]; the all-to-alls overlap both."""
        works = []
        for c in range(self.chunks):
            before(c)
            works.append(self.comm.all_to_all(self.recv[c], self.send[c], async_op=self.chunks > 1))
        for c in range(self.chunks):
            if works[c] is not None:
                works[c].wait()
            after(c)

    def x_to_y(self, x, out, convention=None, pro=None, epi=None, product_last=False):
        """Hartley transform of ``x`` (layout X) into ``out`` (layout Y).  ``product_last``: (axes 0, 1) x
        (axis 2) product of spaces (``Plan.hartley(product_last=True)``): only the row pass differs."""
        mc, pl = self.mc, product_last
        if self.peer is not None:
            b = self.peer.turn
            self.peer.turn = (b + 1) % PeerExchange.NBUF
            mine = self.peer.mine(b)
            # axis 1 (local) with the stores going to the owners of the k1 planes, over NVLink
            self.plan_x.run_strided_scatter(1, x, mine, pro, out_rowoff=self.p2p_x[b], out_outer_off=self.p2p_x_outer)
            self.peer.arrive()
            self.plan_y.run_pass(0, mine, out)
            return self.plan_y.run_pass(2, out, out, convention, epi=epi, product_last=pl)
        self._exchange(
            lambda c: self.plan_x.run_strided_remap(1, x, self.send[c], pro, out_rowoff=self.rowoff1,
                                                    out_ostride=self.ostride1, col0=c * mc, ncols=mc),
            lambda c: self.plan_y.run_strided_remap(0, self.recv[c], out, in_rowoff=self.rowoff0, col0=c * mc,
                                                    ncols=self.n1 * mc, run=mc, run_stride=self.m))
        return self.plan_y.run_pass(2, out, out, convention, epi=epi, product_last=pl)   # rows + Hartley combine

    def y_to_x(self, y, out, convention=None, pro=None, epi=None, product_last=False):
        """Hartley transform of ``y`` (layout Y) into ``out`` (layout X)."""
        mc, pl = self.mc, product_last
        if self.peer is not None:
            b = self.peer.turn
            self.peer.turn = (b + 1) % PeerExchange.NBUF
            mine = self.peer.mine(b)
            self.plan_y.run_strided_scatter(0, y, mine, pro, out_rowoff=self.p2p_y[b], out_col_off=self.p2p_y_col)
            self.peer.arrive()
            self.plan_x.run_pass(1, mine, out)
            return self.plan_x.run_pass(2, out, out, convention, epi=epi, product_last=pl)
        self._exchange(
            lambda c: self.plan_y.run_strided_remap(0, y, self.send[c], pro, out_rowoff=self.rowoff0, col0=c * mc,
                                                    ncols=self.n1 * mc, run=mc, run_stride=self.m),
            lambda c: self.plan_x.run_strided_remap(1, self.recv[c], out, in_rowoff=self.rowoff1,
                                                    in_ostride=self.ostride1, col0=c * mc, ncols=mc))
        return self.plan_x.run_pass(2, out, out, convention, epi=epi, product_last=pl)

    def close(self):
        """Unmap the peers' receive buffers and free this rank's (collective: every rank calls it, after a
        barrier, so that no peer is still storing into a buffer that goes away)."""
        if self.peer is not None:
            torch.cuda.synchronize()
            self.comm.allreduce_(self.peer.flag)        # barrier on the communicator the exchange uses
            torch.cuda.synchronize()
            self.peer.close()
            self.peer = None

    # ---- host-side helpers: natural global array <-> local parts ----------------------------
    def scatter_x(self, a):
        return torch.as_tensor(np.ascontiguousarray(np.asarray(a)[self.s0.owned(self.comm.rank)])).to(_device())

    def scatter_y(self, a):
        return torch.as_tensor(np.ascontiguousarray(np.asarray(a)[:, self.s1.owned(self.comm.rank)])).to(_device())

    def gather_x(self, t):
        parts = [torch.empty_like(t) for _ in range(self.comm.size)]
        dist.all_gather(parts, t.contiguous(), group=self.comm.group)
        out = np.empty(self.shape, dtype=t.cpu().numpy().dtype)
        for q, p in enumerate(parts):
            out[self.s0.owned(q)] = p.cpu().numpy()
        return out

    def gather_y(self, t):
        parts = [torch.empty_like(t) for _ in range(self.comm.size)]
        dist.all_gather(parts, t.contiguous(), group=self.comm.group)
        out = np.empty(self.shape, dtype=t.cpu().numpy().dtype)
        for q, p in enumerate(parts):
            out[:, self.s1.owned(q)] = p.cpu().numpy()
        return out


class SlabTables:
    """Streamed amplitude field ``azm * amp[pindex]`` of the local X part."""

    def __init__(self, field, amps, azm):
        self.amp_mode, self.amp, self.azm, self.amps, self.azm_value = 2, field, 1.0, amps, azm


class SlabCorrelatedField:
    """The hot tail of a 3-D correlated field on a slab-decomposed grid: one isotropic 3-D space, or the
    product of a 2-D space with a 1-D space (``cf.joint``: the space x frequency fields of
    correlated_field.py:663-688; the joint transform differs from the single-space one in the row pass only).

    Wraps a :class:`niftyb200.correlated_field.CorrelatedField` (which supplies the O(n_bins)
    amplitude chain and the geometry) and offers the same ``apply_xi`` / ``adjoint_xi`` /
    ``streaming_tables`` interface, so :class:`niftyb200.metric.CFMetric` and the CG solvers
    work unchanged; latent vectors live in layout X, position-space fields in layout Y."""

    def __init__(self, cf, group=None):
        self.product = getattr(cf, "joint", None) is not None and len(cf.shape) == 3 and len(cf.grids[0].shape) == 2
        if not self.product and (not cf.single or len(cf.shape) != 3 or not cf.grids[0].isotropic):
            raise ValueError("slab decomposition: one isotropic 3-D sub-space, or a power-of-two 2-D x 1-D product")
        self.cf, self.comm = cf, Comm(group)
        self.dtype, self.device, self.dvol, self.offset_mean = cf.dtype, cf.device, cf.dvol, cf.offset_mean
        self.ht = SlabHartley(cf.shape, cf.dtype, self.comm)
        self.shape = cf.shape
        self.latent_shape, self.position_shape = self.ht.shape_x, self.ht.shape_y

    def close(self):
        """Release the peer-memory mappings of the exchange (collective)."""
        self.ht.close()

    def tables(self, p):
        if self.product:       # only the O(n_bins) amplitude chain: the full outer product is never built
            from types import SimpleNamespace
            amps, azm, _ = self.cf._amplitudes(p)
            return self.streaming_tables(SimpleNamespace(amps=[self.cf._as_dev(a) for a in amps], azm=float(azm)))
        return self.streaming_tables(self.cf.tables(p))

    def streaming_tables(self, tab):
        """Local part of ``azm * amp[pindex]`` from the ksq table (global coordinates)."""
        if isinstance(tab, SlabTables):
            return tab
        if self.product:
            # outer product of the two expanded spectra (correlated_field.py:681-688), own planes of axis 0
            g0, g1 = self.cf.grids
            e0 = g0.expand(tab.amps[0], torch.float64)[self.ht.own0[self.comm.rank]]     # (n0, N1)
            e1 = g1.expand(tab.amps[1], torch.float64) * float(getattr(tab, "azm_value", tab.azm))
            field = torch.empty(self.latent_shape, dtype=self.dtype, device=self.device)
            for i in range(self.ht.n0):
                field[i] = (e0[i][:, None] * e1[None, :]).to(self.dtype)
            return SlabTables(field, tab.amps, float(getattr(tab, "azm_value", tab.azm)))
        grid = self.cf.grids[0]
        table = grid.ksq_table(tab.amps[0], torch.float64)
        N0, N1, N2 = self.shape

        def sq(idx, n):
            k = torch.minimum(idx, n - idx)
            return k * k
        dev = self.device
        f0 = sq(self.ht.own0[self.comm.rank], N0)
        f12 = sq(torch.arange(N1, device=dev), N1)[:, None] + sq(torch.arange(N2, device=dev), N2)[None, :]
        field = torch.empty(self.latent_shape, dtype=self.dtype, device=dev)
        for i in range(self.ht.n0):                      # plane by plane: bounded temporaries
            field[i] = (tab.azm * table[f12 + f0[i]]).to(self.dtype)
        return SlabTables(field, tab.amps, tab.azm)

    # The first pass of x_to_y is a strided pass over natural local rows and CAN carry d <- r + beta d
    # (apply_xi(pre=...) works and is parity-tested), but MEASURED at 2 GPUs the sending pass, already bound
    # by its NVLink stores, loses more than the saved axpby sweep gains (82.9 vs 90.4 CG it/s,
    # profiles/r1_bench_f32_n2_fused_vs_unfused.log) -> off unless NIFTYB200_SLAB_FUSED_AXPBY=1
    fuses_direction_update = os.environ.get("NIFTYB200_SLAB_FUSED_AXPBY", "0") == "1"

    def apply_xi(self, xi, tab, out=None, convention=None, offset=None, mul=None, pre=None):
        off = self.offset_mean if offset is None else offset
        out = torch.empty(self.position_shape, dtype=self.dtype, device=self.device) if out is None else out
        pro = PointOp(2, tab.amp, 1.0)
        if pre is not None:
            pro.pre_add, pro.pre_out = pre[0], xi
            if isinstance(pre[1], torch.Tensor):       # beta lives on the device (device-resident CG)
                pro.pre_scale_dev = pre[1]
            else:
                pro.pre_scale = float(pre[1])
        return self.ht.x_to_y(xi, out, convention, pro=pro, product_last=self.product,
                              epi=PointOp(scale=self.dvol, add_scalar=off, mul=mul))

    def adjoint_xi(self, ybar, tab, out=None, convention=None, add=None, dot=None, dot_out=None):
        out = torch.empty(self.latent_shape, dtype=self.dtype, device=self.device) if out is None else out
        res = self.ht.y_to_x(ybar, out, convention, product_last=self.product,
                             epi=PointOp(2, tab.amp, self.dvol, add=add, dot=dot, dot_out=dot_out))
        if dot is not None:
            self.comm.allreduce_(dot_out)                 # CG curvature: a single fp64 scalar
        return res
