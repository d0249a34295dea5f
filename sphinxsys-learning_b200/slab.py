"""Slab decomposition of the dam-break path over several GPUs (the reference is single-address-space).

1-D slabs along x.  x is the slowest index of the linear cell index (meshes/base_mesh.hxx:68-78), so the
cell pages [page_lo, page_hi) a rank owns are one contiguous run of its cell-sorted particles, and the
two pages next to it (one cut-off radius each) are its halo.  Every rank uses the GLOBAL mesh, so cell
indices, neighbour sets and sums are those of the single-GPU run.

Per acoustic sub-step:  1st-half initialization -> halo records rec_a (new p) from the owners -> 1st-half
interaction+update -> halo records rec_b (new v) -> 2nd half -> min all-reduce of dt.
Per advection step:     one exchange that carries migrating particles AND the fresh halo (each rank
sends its particles lying in the two pages either side of a cut), then sort + relations + halo index maps.

The ranks' kernels are libsphb200's (host.py), including the halo pack/unpack (sphb200_rows_gather/scatter), the
re-decomposition (sphb200_slab_select/pack/rebuild: one cell-list update per advection step) and the sub-step loop's
bookkeeping (dt, relaxation time and the exit test stay on the device; the ranks all-reduce the bit pattern of
max(c0+|v|) on the device, the host is not in the sub-step loop).  torch is the plumbing only: views of device
memory, buffers, torch.distributed (NCCL) send/recv and all-reduce.
`LocalComm` runs all slabs inside one process on one GPU (tests, single-GPU boxes); `DistComm` is one
slab per process.
"""
import math
import os
import sys

import numpy as np

from . import cases

try:
    import torch
    import torch.distributed as dist
except Exception:  # pragma: no cover - torch is part of the image
    torch = None
    dist = None


class _Raw:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (int(ptr), False), "shape": tuple(shape), "typestr": typestr, "version": 2}


def dev_tensor(ptr, shape, np_dtype, device):
    """torch view of raw device memory owned by libsphb200"""
    if int(np.prod(shape)) == 0:
        return torch.empty(tuple(shape), dtype=torch.from_numpy(np.zeros(0, np_dtype)).dtype, device=device)
    return torch.as_tensor(_Raw(ptr, shape, np.dtype(np_dtype).str), device=device)


# ------------------------------------------------------------------------------------------------
# page cuts (pure host logic, tested on CPU)
# ------------------------------------------------------------------------------------------------
def page_cuts(fluid_per_page, n_slabs):
    """cut pages so every slab owns ~the same number of fluid particles.
    returns n_slabs+1 page indices, first 0, last len(fluid_per_page)"""
    c = np.concatenate([[0], np.cumsum(np.asarray(fluid_per_page, dtype=np.int64))])
    total = int(c[-1])
    cuts = [0]
    for r in range(1, n_slabs):
        target = total * r / n_slabs
        p = int(np.searchsorted(c, target, side="left"))
        # pick the nearer page boundary, keep cuts strictly increasing
        if p > 0 and abs(c[p - 1] - target) <= abs(c[min(p, len(c) - 1)] - target):
            p -= 1
        cuts.append(max(p, cuts[-1] + 1))
    cuts.append(len(fluid_per_page))
    return cuts


def recut(fluid_per_page, old_cuts):
    """new page cuts for the current fluid distribution, every interior cut moved by at most half the width of the
    narrower slab next to it minus one page: then each page a slab needs (its own and one halo page either side) is
    owned, before the move, by the slab itself or a direct neighbour, and one neighbour exchange re-decomposes the run"""
    n = len(old_cuts) - 1
    want = page_cuts(fluid_per_page, n)
    new = list(old_cuts)
    for r in range(1, n):
        room = max(min(old_cuts[r] - old_cuts[r - 1], old_cuts[r + 1] - old_cuts[r]) // 2 - 1, 0)
        new[r] = min(max(want[r], old_cuts[r] - room), old_cuts[r] + room)
    return new


def imbalance(fluid_per_page, cuts):
    """max over slabs of owned fluid / mean - 1"""
    c = np.concatenate([[0], np.cumsum(np.asarray(fluid_per_page, dtype=np.int64))])
    owned = np.array([c[b] - c[a] for a, b in zip(cuts, cuts[1:])], dtype=np.float64)
    return float(owned.max() / max(owned.mean(), 1.0) - 1.0)


def match_by_gid(send_gid, recv_gid):
    """orders that make sender rows and receiver slots line up: both sides sort by global id"""
    return np.argsort(send_gid, kind="stable"), np.argsort(recv_gid, kind="stable")


# ------------------------------------------------------------------------------------------------
# communicators: exchange(list over local parts of (to_left, to_right)) -> list of (from_left, from_right)
# ------------------------------------------------------------------------------------------------
class LocalComm:
    """all slabs in this process (one GPU): neighbours are list neighbours, the 'transfer' is a clone"""

    def __init__(self, n_slabs):
        self.n_slabs, self.rank0, self.world = n_slabs, 0, 1

    def local_slabs(self):
        return list(range(self.n_slabs))

    def exchange(self, sends):
        out = []
        for k in range(self.n_slabs):
            from_left = sends[k - 1][1].clone() if k > 0 else None
            from_right = sends[k + 1][0].clone() if k + 1 < self.n_slabs else None
            out.append((from_left, from_right))
        return out

    exchange_fixed = None  # same as exchange (sizes are implicit in-process)

    def allreduce(self, values, op):
        v = min(values) if op == "min" else (max(values) if op == "max" else sum(values))
        return v

    def allreduce_sum_array(self, arrays):
        """arrays: one host int64 array per local slab; returns their element-wise sum over all slabs"""
        return np.sum(np.stack(arrays), axis=0)

    def peer_bases(self, systems, bases):
        """(left, right) neighbour's landing allocation as seen from each local slab: same process, the pointer itself"""
        n = len(bases)
        return [(bases[k - 1] if k > 0 else None, bases[k + 1] if k + 1 < n else None) for k in range(n)]

    def allreduce_max_device(self, slots):
        """slots: one int64 device tensor per local slab (bit pattern of a non-negative double): all get the maximum"""
        if len(slots) > 1:
            m = torch.stack(slots).max(dim=0).values
            for t in slots:
                t.copy_(m)

    def barrier(self):
        pass


class DistComm:
    """one slab per process over torch.distributed (NCCL): rank r owns slab r"""

    def __init__(self, device):
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.n_slabs, self.device = self.world, device

    def local_slabs(self):
        return [self.rank]

    def exchange(self, sends):
        to_left, to_right = sends[0]
        left = self.rank - 1 if self.rank > 0 else None
        right = self.rank + 1 if self.rank + 1 < self.world else None
        # row counts first (the payload shapes differ every advection step)
        cnt_out = torch.tensor([to_left.shape[0] if left is not None else 0, to_right.shape[0] if right is not None else 0],
                               dtype=torch.int64, device=self.device)
        cnt_in = torch.zeros(2, dtype=torch.int64, device=self.device)
        ops = []
        if left is not None:
            ops += [dist.P2POp(dist.isend, cnt_out[0:1], left), dist.P2POp(dist.irecv, cnt_in[0:1], left)]
        if right is not None:
            ops += [dist.P2POp(dist.isend, cnt_out[1:2], right), dist.P2POp(dist.irecv, cnt_in[1:2], right)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        n_left, n_right = (int(x) for x in cnt_in.tolist())
        from_left = from_right = None
        ops = []
        if left is not None:
            from_left = torch.empty((n_left,) + tuple(to_left.shape[1:]), dtype=to_left.dtype, device=self.device)
            ops += [dist.P2POp(dist.isend, to_left.contiguous(), left), dist.P2POp(dist.irecv, from_left, left)]
        if right is not None:
            from_right = torch.empty((n_right,) + tuple(to_right.shape[1:]), dtype=to_right.dtype, device=self.device)
            ops += [dist.P2POp(dist.isend, to_right.contiguous(), right), dist.P2POp(dist.irecv, from_right, right)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return [(from_left, from_right)]

    def exchange_fixed(self, sends, recv_rows):
        """payload whose row counts both sides already know (the per-sub-step halo records)"""
        to_left, to_right = sends[0]
        n_left, n_right = recv_rows[0]
        left = self.rank - 1 if self.rank > 0 else None
        right = self.rank + 1 if self.rank + 1 < self.world else None
        from_left = from_right = None
        ops = []
        if left is not None:
            from_left = torch.empty((n_left,) + tuple(to_left.shape[1:]), dtype=to_left.dtype, device=self.device)
            ops += [dist.P2POp(dist.isend, to_left, left), dist.P2POp(dist.irecv, from_left, left)]
        if right is not None:
            from_right = torch.empty((n_right,) + tuple(to_right.shape[1:]), dtype=to_right.dtype, device=self.device)
            ops += [dist.P2POp(dist.isend, to_right, right), dist.P2POp(dist.irecv, from_right, right)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return [(from_left, from_right)]

    def allreduce(self, values, op):
        t = torch.tensor([values[0]], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op={"min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX, "sum": dist.ReduceOp.SUM}[op])
        return float(t.item())

    def allreduce_sum_array(self, arrays):
        a = np.asarray(arrays[0])
        t = torch.as_tensor(a.astype(np.float64 if a.dtype.kind == "f" else np.int64), device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.cpu().numpy()

    def peer_bases(self, systems, bases):
        """every rank exports its landing allocation (cudaIpcGetMemHandle), the handles go round once, and each rank maps
        its two neighbours' allocations into its own address space (cudaIpcOpenMemHandle, peer access over NVLink)"""
        mine = torch.frombuffer(bytearray(systems[0].peer_export(bases[0])), dtype=torch.uint8).to(self.device)
        everyone = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(everyone, mine)
        def opened(r):
            return systems[0].peer_open(bytes(everyone[r].cpu().numpy().tobytes())) if 0 <= r < self.world else None
        return [(opened(self.rank - 1), opened(self.rank + 1))]

    def allreduce_max_device(self, slots):
        dist.all_reduce(slots[0], op=dist.ReduceOp.MAX)  # stays on the device: no host synchronisation

    def barrier(self):
        dist.barrier()


def _exchange_fixed(comm, sends, recv_rows):
    if getattr(comm, "exchange_fixed", None):
        return comm.exchange_fixed(sends, recv_rows)
    return comm.exchange(sends)


# ------------------------------------------------------------------------------------------------
# the decomposed 3-D dam break
# ------------------------------------------------------------------------------------------------
WALL_VARS = ["Position", "VolumetricMeasure", "Mass", "NormalDirection", "Velocity", "Acceleration"]
MIGRATE_VARS = ["Position", "VolumetricMeasure", "Mass", "Density", "Pressure", "DensityChangeRate", "Velocity", "Force",
                "ForcePrior", "PreviousGravityForce", "GlobalID"]


class _Part:
    """one slab: a DamBreak restricted to its pages +- one halo page, on its own context"""
    pass


class SlabDamBreak:
    """tests/3d_examples/test_3d_dambreak/dambreak.cpp:159-195 over n_slabs slabs; same loop as cases.DamBreak"""

    def __init__(self, backend, comm, dp=0.05, real_bytes=8, device=0, stream=None, ndim=3, device_loop=True, fast_reconfigure=True,
                 recut_interval=20, recut_threshold=0.05, peer_halo=None, probes=None):
        """recut_interval / recut_threshold: every recut_interval advection steps the owned fluid per page is summed over
        the slabs and, when the fullest slab holds more than (1 + threshold) x the mean, the cuts move (0 = never);
        probes: observation points (True: the reference main's six pressure probes).  Every slab carries the observer
        body; a probe is evaluated by the slab that owns the page of its position (its neighbourhood lies in that slab's
        owned + halo particles) and the values are summed over the slabs (owner-rank evaluation);
        peer_halo: the per-sub-step halo records are pushed straight into the neighbour GPU's memory by the pack kernel
        (csrc/peer.cu; False: pack -> torch.distributed send/recv -> unpack; None: on unless env SPHB200_PEER_HALO=0);
        device_loop: dt / relaxation time / exit test of the sub-step loop on the device (False: one host round trip
        per sub-step, the round-1 path); fast_reconfigure: sphb200_slab_select/pack/rebuild and one cell-list update
        per advection step (False: keep_range + two cell-list updates + torch.cat, the round-1 path)"""
        assert torch is not None
        self.comm, self.dp, self.ndim, self.real_bytes = comm, dp, ndim, real_bytes
        self.device_loop, self.fast_reconfigure, self._guess = device_loop, fast_reconfigure, 5
        self.recut_interval, self.recut_threshold, self.n_recuts = (recut_interval if fast_reconfigure else 0), recut_threshold, 0
        self.device = torch.device("cuda", device)
        self.np_real = np.float64 if real_bytes == 8 else np.float32
        self.np_int = np.int64 if real_bytes == 8 else np.int32
        self.t_int = torch.int64 if real_bytes == 8 else torch.int32
        if stream is None:
            stream = torch.cuda.current_stream(self.device).cuda_stream
        # ---- global mesh (query the library with an empty probe case) and the page cuts
        probe = cases.DamBreak(backend, ndim=ndim, dp=dp, real_bytes=real_bytes, with_observer=False, x_range=(1e30, 2e30),
                               defer_configuration=True, device=device, stream=stream)
        info = probe.water.info()
        self.mesh_lo, self.spacing, self.cells = info["mesh_lo"], info["spacing"], info["cells"]
        self.sys_lo = probe.sys.lo.copy()
        self.U_ref, self.gravity = probe.U_ref, probe.gravity
        n_lat = np.ceil((probe.sys.hi - probe.sys.lo) / dp).astype(np.int64)
        self.n_lat = n_lat
        xs = probe.sys.lo[0] + (np.arange(n_lat[0]) + 0.5) * dp
        in_fluid = (xs >= 0.0) & (xs <= 2.0)
        page_of = np.clip(np.floor((xs - self.mesh_lo[0]) / self.spacing).astype(np.int64), 0, self.cells[0] - 1)
        fluid_per_page = np.bincount(page_of[in_fluid], minlength=self.cells[0])
        probe.sys.close()
        self.cuts = page_cuts(fluid_per_page, comm.n_slabs)
        # ---- the local slabs
        self.parts = []
        for r in comm.local_slabs():
            P = _Part()
            P.rank, P.p_lo, P.p_hi = r, self.cuts[r], self.cuts[r + 1]
            P.has_left, P.has_right = r > 0, r + 1 < comm.n_slabs
            x_lo = self.mesh_lo[0] + (P.p_lo - 1) * self.spacing - dp
            x_hi = self.mesh_lo[0] + (P.p_hi + 1) * self.spacing + dp
            P.case = c = cases.DamBreak(backend, ndim=ndim, dp=dp, real_bytes=real_bytes, with_observer=probes is not None,
                                        probes=None if probes is True else probes,
                                        x_range=(x_lo, x_hi), defer_configuration=True, device=device, stream=stream)
            w, s = c.water, c.wall
            pos = w.get("Position").astype(np.float64) if w.n else np.zeros((0, ndim))
            idx = np.rint((pos - self.sys_lo) / dp - 0.5).astype(np.int64)
            gid = idx[:, 0]
            for d in range(1, ndim):
                gid = gid * int(n_lat[d]) + idx[:, d]
            w.set("GlobalID", gid.astype(self.np_int).view(self.np_real))  # bit pattern, moved verbatim by the sort
            for nm, wd in (("DensitySummation", 1), ("GravityForce", ndim)):
                w.register(nm, wd, False)
            # trim both bodies to pages [p_lo-1, p_hi+1) exactly, then own [p_lo, p_hi)
            for body in (w, s):
                body.update_cell_list()
                a, b = body.page_begin(P.p_lo - 1), body.page_begin(P.p_hi + 1)
                body.keep_range(a, b)
            w.set_owned_pages(P.p_lo, P.p_hi)
            s.set_owned_pages(P.p_lo, P.p_hi)   # the wall is only ever a target; its owned pages serve the re-cut
            w.update_cell_list()
            s.update_cell_list()
            if c.observer is not None:
                c.observer.update_cell_list()
            c.update_configuration(observer=c.observer is not None)
            c.sys.gravity_force(w, c.gravity)
            self.parts.append(P)
        if peer_halo is None:
            peer_halo = os.environ.get("SPHB200_PEER_HALO", "1") != "0"
        self.peer_halo, self._epoch = bool(peer_halo), [0, 0]
        if self.peer_halo:
            # the mapping of a neighbour's memory can be refused (no peer access between two GPUs, IPC disabled in a
            # container): every rank must then take the send/recv transport - agree on it before the first exchange
            ok, why = 1.0, ""
            try:
                self._peer_setup()
            except Exception as e:  # noqa: BLE001 - any failure of the set-up means "no peer path here"
                ok, why = 0.0, repr(e)
            if comm.allreduce([ok], "min") < 1.0:
                if why:
                    print(f"[sphb200 slab] halo push into peer memory unavailable ({why}); using send/recv", file=sys.stderr, flush=True)
                self.peer_halo = False
        self._halo_maps()
        self.time, self.dt, self.n_advection, self.n_acoustic = 0.0, 0.0, 0, 0

    # ---- views ----------------------------------------------------------------------------------
    def _var(self, P, name):
        w = P.case.water
        width = w.width(name)
        return dev_tensor(w.device_ptr(name), (w.n, width), self.np_real, self.device)

    def _gid(self, P):
        return self._var(P, "GlobalID").view(self.t_int).reshape(-1)

    def _halo_maps(self):
        """index lists (sorted by global id on both sides) for the per-sub-step halo record exchange"""
        for P in self.parts:
            w = P.case.water
            ob, oe = w.owned_range()
            P.ob, P.oe = ob, oe
            gid = self._gid(P)
            e1 = w.page_begin(P.p_lo + 1) if P.has_left else ob       # owned particles of my first page
            b1 = w.page_begin(P.p_hi - 1) if P.has_right else oe      # owned particles of my last page
            e1, b1 = min(max(e1, ob), oe), min(max(b1, ob), oe)
            P.send_left = (ob + torch.argsort(gid[ob:e1], stable=True)) if P.has_left else None
            P.send_right = (b1 + torch.argsort(gid[b1:oe], stable=True)) if P.has_right else None
            P.recv_left = torch.argsort(gid[0:ob], stable=True) if P.has_left else None
            P.recv_right = (oe + torch.argsort(gid[oe:w.n], stable=True)) if P.has_right else None
            P.rec = [dev_tensor(w.record_ptr(k), (w.n, 8), self.np_real, self.device) for k in (0, 1)]
            P.rec_ptr = [w.record_ptr(k) for k in (0, 1)]
            t_real = P.rec[0].dtype
            P.send_buf = [torch.empty((ix.numel() if ix is not None else 0, 8), dtype=t_real, device=self.device)
                          for ix in (P.send_left, P.send_right)]
            P.recv_buf = [torch.empty((ix.numel() if ix is not None else 0, 8), dtype=t_real, device=self.device)
                          for ix in (P.recv_left, P.recv_right)]
            P.slot = dev_tensor(P.case.sys.reduce_slot_ptr(0), (1,), np.int64, self.device)
            if self.peer_halo:
                for ix in (P.send_left, P.send_right, P.recv_left, P.recv_right):
                    if ix is not None and ix.numel() > self._peer_rows:
                        raise RuntimeError(f"halo page of {ix.numel()} rows exceeds the landing buffer ({self._peer_rows})")

    # ---- halo records over peer memory -------------------------------------------------------------
    def _peer_setup(self):
        """one exportable allocation per slab: 4 flag words + 4 landing buffers ({from left, from right} x {rec_a, rec_b}),
        each large enough for a completely filled cell page (a page cannot hold more than its volume / dp^3 particles, so
        the buffers never have to grow and the sender needs no answer from the receiver)"""
        per_page = int(np.prod(self.n_lat[1:])) * (int(math.ceil(self.spacing / self.dp)) + 1)
        self._peer_rows = int(1.1 * per_page) + 1024
        self._row_bytes = 8 * self.real_bytes
        self._peer_flags = 256
        nbytes = self._peer_flags + 4 * self._peer_rows * self._row_bytes
        for P in self.parts:
            P.peer_base = P.case.sys.peer_alloc(nbytes)
        got = self.comm.peer_bases([P.case.sys for P in self.parts], [P.peer_base for P in self.parts])
        for P, (left, right) in zip(self.parts, got):
            P.peer_left, P.peer_right = (left if P.has_left else None), (right if P.has_right else None)

    def _peer_slot(self, base, side, which):
        """(landing buffer, flag word) of the area `side` (0: rows that came from the left, 1: from the right) of record `which`"""
        k = side * 2 + which
        return base + self._peer_flags + k * self._peer_rows * self._row_bytes, base + 8 * k

    def _halo_records_peer(self, which):
        self._epoch[which] += 1
        ep = self._epoch[which]
        for P in self.parts:            # every push is enqueued before any wait (one stream serves all slabs of a LocalComm)
            S = P.case.sys
            for side, ix, peer in ((0, P.send_left, P.peer_left), (1, P.send_right, P.peer_right)):
                if peer is not None:
                    landing, flag = self._peer_slot(peer, 1 - side, which)   # what I send left arrives "from the right"
                    n = ix.numel()
                    S.rows_push(P.rec_ptr[which], self._row_bytes, ix.data_ptr() if n else 0, n, landing, flag, ep)
        for P in self.parts:
            S = P.case.sys
            for side, ix, there in ((0, P.recv_left, P.has_left), (1, P.recv_right, P.has_right)):
                if there:
                    landing, flag = self._peer_slot(P.peer_base, side, which)
                    n = ix.numel()
                    S.rows_wait_scatter(P.rec_ptr[which], self._row_bytes, ix.data_ptr() if n else 0, n, landing, flag, ep)

    def _halo_records(self, which):
        """owner's records of the rows in the neighbour's halo: pack (sphb200_rows_gather) -> send/recv -> unpack (rows_scatter)"""
        if self.peer_halo:
            return self._halo_records_peer(which)
        sends, rows = [], []
        row_bytes = 8 * self.real_bytes
        for P in self.parts:
            S = P.case.sys
            for side, ix in ((0, P.send_left), (1, P.send_right)):
                if ix is not None and ix.numel():
                    S.rows_gather(P.rec_ptr[which], row_bytes, ix.data_ptr(), ix.numel(), P.send_buf[side].data_ptr())
            sends.append((P.send_buf[0] if P.has_left else None, P.send_buf[1] if P.has_right else None))
            rows.append((P.recv_left.numel() if P.has_left else 0, P.recv_right.numel() if P.has_right else 0))
        recvs = _exchange_fixed(self.comm, sends, rows)
        for P, (fl, fr) in zip(self.parts, recvs):
            S = P.case.sys
            for ix, t in ((P.recv_left, fl), (P.recv_right, fr)):
                if ix is not None and ix.numel():
                    S.rows_scatter(P.rec_ptr[which], row_bytes, ix.data_ptr(), ix.numel(), t.data_ptr())

    # ---- the loop ---------------------------------------------------------------------------------
    def _min_over_slabs(self, fn):
        vals = []
        for P in self.parts:
            vals.append(fn(P) if P.oe > P.ob else math.inf)
        return self.comm.allreduce(vals, "min")

    def acoustic_substep(self, dt):
        for P in self.parts:
            P.case.sys.integration_1st_half_init(P.case.inner, dt)
        self._halo_records(0)
        for P in self.parts:
            P.case.sys.integration_1st_half_interact(P.case.inner, P.case.wall_contact, True, dt)
        self._halo_records(1)
        for P in self.parts:
            P.case.sys.integration_2nd_half(P.case.inner, P.case.wall_contact, True, dt)
        self.n_acoustic += 1

    def _substep_device(self):
        """one sub-step with dt < 0 = the device-resident dt; the ranks' maxima of c0+|v| are all-reduced on the device"""
        for P in self.parts:
            if P.oe <= P.ob:
                P.slot.zero_()
            P.case.sys.integration_1st_half_init(P.case.inner, -1.0)
        self._halo_records(0)
        for P in self.parts:
            P.case.sys.integration_1st_half_interact(P.case.inner, P.case.wall_contact, True, -1.0)
        self._halo_records(1)
        for P in self.parts:
            P.case.sys.integration_2nd_half(P.case.inner, P.case.wall_contact, True, -1.0)
        self.comm.allreduce_max_device([P.slot for P in self.parts])
        for P in self.parts:
            P.case.sys.substep_ctl(P.case.water, 0.6, 1, 1)

    def advection_step(self, reconfigure=True):
        """reconfigure=False leaves out the closing migration / re-decomposition (a caller that reloads the slabs from the
        host before the next step configures them then anyway)"""
        Dt = self._min_over_slabs(lambda P: P.case.sys.advection_dt(P.case.water, self.U_ref))
        for P in self.parts:
            P.case.sys.density_summation(P.case.inner, P.case.wall_contact, free_surface=True)
        relaxation, k = 0.0, 0
        if self.device_loop:
            for P in self.parts:
                P.case.sys.relaxation_begin(Dt, self.dt)
            batch = self._guess
            while True:
                for _ in range(batch):
                    self._substep_device()
                st = self.parts[0].case.sys.relaxation_state()   # the one host synchronisation of the loop
                if st["done"]:
                    break
                batch = 1
            k, self.dt, relaxation = st["substeps"], st["dt"], st["relaxation_time"]
            if self.peer_halo and any(P.case.sys.peer_timeouts() for P in self.parts):
                raise RuntimeError("a halo push from a neighbour slab did not arrive within the time-out")
            self._guess = k
            self.time += relaxation
            self.n_acoustic += k
        while relaxation < Dt:  # dambreak.cpp:168-177: step with the previous dt, then reduce
            self.acoustic_substep(self.dt)
            self.dt = self._min_over_slabs(lambda P: P.case.sys.acoustic_dt(P.case.water))
            relaxation += self.dt
            self.time += self.dt
            k += 1
        self.n_advection += 1
        if reconfigure:
            self.reconfigure()
        return Dt, k

    def fluid_per_page(self):
        """owned fluid particles per cell page, summed over the slabs (every rank gets the same array)"""
        local = []
        for P in self.parts:
            off = P.case.water.page_offsets()
            cnt = np.diff(off)
            cnt[:P.p_lo] = 0
            cnt[P.p_hi:] = 0
            local.append(cnt)
        return self.comm.allreduce_sum_array(local)

    def plan_recut(self, force=False):
        """the new cuts, or None when the slabs are balanced (needs the cell lists of the current arrangement: call it
        before the particles move, i.e. right after a reconfigure; the cuts take effect in the next one)"""
        per_page = self.fluid_per_page()
        if not force and imbalance(per_page, self.cuts) <= self.recut_threshold:
            return None
        new = recut(per_page, self.cuts)
        return None if new == list(self.cuts) else new

    def _move(self, body_of, names, t_real):
        """select -> pack -> neighbour exchange -> rebuild for one body of every local slab"""
        sends = []
        for P in self.parts:
            b = body_of(P)
            _, n_l, n_r = b.slab_select(P.has_left, P.has_right)
            width = sum(b.width(nm) for nm in names)
            bl = torch.empty((n_l, width), dtype=t_real, device=self.device)
            br = torch.empty((n_r, width), dtype=t_real, device=self.device)
            if n_l:
                b.slab_pack(1, names, bl.data_ptr())
            if n_r:
                b.slab_pack(2, names, br.data_ptr())
            sends.append((bl if P.has_left else None, br if P.has_right else None))
        recvs = self.comm.exchange(sends)
        for P, (fl, fr) in zip(self.parts, recvs):
            b = body_of(P)
            nl = fl.shape[0] if fl is not None else 0
            nr = fr.shape[0] if fr is not None else 0
            b.slab_rebuild(names, fl.data_ptr() if nl else 0, nl, fr.data_ptr() if nr else 0, nr)

    def reconfigure(self):
        """migration + fresh halo in one exchange (and, when a re-cut is pending, whole pages of fluid and wall with
        it), then ONE sort, relations and halo maps"""
        if not self.fast_reconfigure:
            return self.reconfigure_two_sorts()
        t_real = torch.float64 if self.real_bytes == 8 else torch.float32
        new_cuts, self._pending_cuts = getattr(self, "_pending_cuts", None), None
        if new_cuts is not None:
            self.cuts = list(new_cuts)
            for P in self.parts:
                P.p_lo, P.p_hi = self.cuts[P.rank], self.cuts[P.rank + 1]
                P.case.water.slab_recut(P.p_lo, P.p_hi)
                P.case.wall.slab_recut(P.p_lo, P.p_hi)
            self._move(lambda P: P.case.wall, WALL_VARS, t_real)
            for P in self.parts:
                P.case.wall.update_cell_list()
            self.n_recuts += 1
        self._move(lambda P: P.case.water, MIGRATE_VARS, t_real)
        for P in self.parts:
            P.case.water.update_cell_list()
            P.case.update_configuration(observer=P.case.observer is not None)
        self._halo_maps()
        if self.recut_interval and self.n_advection % self.recut_interval == 0:
            self._pending_cuts = self.plan_recut()

    def reconfigure_two_sorts(self):
        """the round-1 path (kept for A/B timing and as a cross-check of the fast one)"""
        sends = []
        for P in self.parts:
            w = P.case.water
            w.keep_range(P.ob, P.oe)          # drop the old halo copies
            w.update_cell_list()              # owned particles by their new cells
            a = w.page_begin(P.p_lo + 1) if P.has_left else 0      # pages <= p_lo: migrants to / halo of the left slab
            b = w.page_begin(P.p_hi - 1) if P.has_right else w.n   # pages >= p_hi-1: same for the right slab
            cols = [self._var(P, nm) for nm in MIGRATE_VARS]
            sends.append((torch.cat([c[0:a] for c in cols], dim=1) if P.has_left else None,
                          torch.cat([c[b:w.n] for c in cols], dim=1) if P.has_right else None))
        recvs = self.comm.exchange(sends)
        for P, (fl, fr) in zip(self.parts, recvs):
            w = P.case.water
            n_old = w.n
            extra = [t for t in (fl, fr) if t is not None and t.shape[0] > 0]
            n_new = n_old + sum(t.shape[0] for t in extra)
            w.resize(n_new)
            if extra:
                block = torch.cat(extra, dim=0)
                col = 0
                for nm in MIGRATE_VARS:
                    v = self._var(P, nm)
                    v[n_old:n_new] = block[:, col:col + v.shape[1]]
                    col += v.shape[1]
            w.update_cell_list()
            P.case.update_configuration(observer=P.case.observer is not None)
        self._halo_maps()

    def close(self):
        """release the peer mappings and landing buffers (every rank: the mappings first, then - after a barrier - the
        allocations they point into) and the slabs' contexts"""
        if self.peer_halo:
            torch.cuda.synchronize()
            inside_one_process = isinstance(self.comm, LocalComm)
            for P in self.parts:
                for ptr in (P.peer_left, P.peer_right):
                    if ptr is not None and not inside_one_process:
                        P.case.sys.peer_close(ptr)
            self.comm.barrier()
            for P in self.parts:
                P.case.sys.peer_free(P.peer_base)
            self.peer_halo = False
        for P in self.parts:
            P.case.sys.close()
        self.parts = []

    # ---- observables / inspection -------------------------------------------------------------------
    def n_fluid(self):
        return int(self.comm.allreduce([float(sum(P.oe - P.ob for P in self.parts))], "sum"))

    def mechanical_energy(self):
        vals = [P.case.sys.total_mechanical_energy(P.case.water, self.gravity) if P.oe > P.ob else 0.0 for P in self.parts]
        return self.comm.allreduce([sum(vals)], "sum")

    def probe_pressure(self):
        """ObservedQuantityRecording<Real>("Pressure") of the observation points (general_interpolation.h:63-83): each probe
        from the slab that owns the page of its position, summed over the slabs"""
        vals = []
        for P in self.parts:
            c = P.case
            pos = np.asarray(c.probes, dtype=np.float64)
            page = np.clip(np.floor((pos[:, 0] - self.mesh_lo[0]) / self.spacing).astype(np.int64), 0, self.cells[0] - 1)
            mine = (page >= P.p_lo) & (page < P.p_hi)
            v = np.asarray(c.probe_pressure(), dtype=np.float64)
            vals.append(np.where(mine, v, 0.0))
        return self.comm.allreduce_sum_array(vals)

    def owned(self, name):
        """(global ids, values) of the locally owned particles, on the host"""
        ids, vals = [], []
        for P in self.parts:
            torch.cuda.synchronize()
            ids.append(self._gid(P)[P.ob:P.oe].cpu().numpy().astype(np.int64))
            vals.append(self._var(P, name)[P.ob:P.oe].cpu().numpy())
        return np.concatenate(ids), np.concatenate(vals, axis=0)

    def launches(self):
        return sum(P.case.sys.launches for P in self.parts)
