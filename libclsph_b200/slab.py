"""Slab-decomposed stepping: one process per GPU, `torch.distributed` for the plumbing.

The reference is single-device (libclsph/sph_simulation.cpp:273-275); this is the layer that
spreads its step over the GPUs of one box.  Rank r owns the particles with x in
[cuts[r], cuts[r+1]).  Per step (include/sphb.h, "slab engines"):

    pack -> send/recv one fixed-size message with each slab neighbour (NCCL over NVLink)
         -> unpack -> forces -> { advect(dt) -> all-reduce MAX of 8 floats -> dt' } -> commit

`SlabDriver` holds only the protocol; the per-rank work is behind a small local-engine
interface, implemented by `CudaSlab` (the sphb C ABI) here and by a numpy stand-in in
tests/ (world_size-2 gloo tests of the protocol on CPU).
"""
import ctypes as C

import time

import numpy as np

RECORD_BYTES = 48
HEADER_BYTES = 16


def message_bytes(exchange_capacity):
    return HEADER_BYTES + RECORD_BYTES * int(exchange_capacity)


def needs_retry(dt, computed):
    """`dt - computed > 0.00001` as the reference evaluates it: a float subtraction compared
    with a double literal (libclsph/sph_simulation.cpp:260)."""
    return float(np.float32(dt) - np.float32(computed)) > 0.00001


class CudaSlab:
    """Local-engine interface over the sphb C ABI (one slab engine on one GPU)."""

    def __init__(self, params, terms, scene, device, capacity, exchange_capacity):
        import torch
        from .engine import Engine
        self.torch = torch
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.e = Engine(params, terms, device, slab_capacity=capacity, exchange_capacity=exchange_capacity)
        self.e.set_scene(scene["bbs"], scene["transforms"], scene["rvertices"], scene["total_gridpoints"])
        self.params = np.ascontiguousarray(params).copy()
        self.exchange_capacity = int(exchange_capacity)
        self.capacity = int(capacity)
        self.stream = torch.cuda.ExternalStream(self.e.stream(), device=self.device)

    # -- buffers the driver communicates with ------------------------------------------------
    def new_message(self):
        return self.torch.zeros(message_bytes(self.exchange_capacity), dtype=self.torch.uint8, device=self.device)

    def new_red8(self):
        return self.torch.zeros(8, dtype=self.torch.float32, device=self.device)

    def red8_to_host(self, red8):
        return red8.cpu().numpy()  # synchronises the current (= the engine's) stream

    def stream_context(self):
        return self.torch.cuda.stream(self.stream)

    # -- the engine ---------------------------------------------------------------------------
    def _chk(self, rc):
        self.e._check(rc)

    def configure(self, xlo, xhi, has_left, has_right):
        self._chk(self.e.L.sphb_slab_configure(self.e.h, xlo, xhi, int(has_left), int(has_right)))

    def upload(self, particles, ids):
        from .engine import PARTICLE, _ptr
        particles = np.ascontiguousarray(particles, PARTICLE)
        ids = np.ascontiguousarray(ids, np.uint32)
        assert len(particles) == len(ids)
        self._chk(self.e.L.sphb_slab_upload(self.e.h, _ptr(particles), _ptr(ids), len(ids)))

    def bbox(self, red8):
        self._chk(self.e.L.sphb_slab_bbox(self.e.h, red8.data_ptr()))

    def pack(self, send_left, send_right):
        self._chk(self.e.L.sphb_slab_pack(self.e.h, send_left.data_ptr() if send_left is not None else None,
                                          send_right.data_ptr() if send_right is not None else None))

    def unpack(self, recv_left, recv_right):
        no, ng = C.c_uint32(), C.c_uint32()
        self._chk(self.e.L.sphb_slab_unpack(self.e.h, recv_left.data_ptr() if recv_left is not None else None,
                                            recv_right.data_ptr() if recv_right is not None else None,
                                            C.byref(no), C.byref(ng)))
        return no.value, ng.value

    def forces(self, gmin, gmax):
        gmin = np.ascontiguousarray(gmin, np.float32)
        gmax = np.ascontiguousarray(gmax, np.float32)
        self._chk(self.e.L.sphb_slab_forces(self.e.h, gmin.ctypes.data, gmax.ctypes.data))

    def advect(self, dt, red8):
        self._chk(self.e.L.sphb_slab_advect(self.e.h, dt, red8.data_ptr()))

    def commit(self):
        self._chk(self.e.L.sphb_slab_commit(self.e.h))

    def compute_dt(self, max_v2, max_a2):
        return float(self.e.L.sphb_compute_dt(self.params.ctypes.data, float(max_v2), float(max_a2)))

    def download(self, out=None, ids=None):
        """(owned particles in local order, their ids).  `out` / `ids`: caller-owned arrays to fill
        (page-locked ones make the copy run at PCIe speed); views of their first n entries return."""
        from .engine import PARTICLE, _ptr
        n = C.c_uint32()
        self._chk(self.e.L.sphb_slab_download(self.e.h, None, None, 0, C.byref(n)))
        if out is None:
            out = np.zeros(n.value, PARTICLE)
            ids = np.zeros(n.value, np.uint32)
        assert len(out) >= n.value and len(ids) >= n.value
        if n.value:
            self._chk(self.e.L.sphb_slab_download(self.e.h, _ptr(out), _ptr(ids), len(out), C.byref(n)))
        return out[: n.value], ids[: n.value]

    # -- the native driver (csrc/slab_nccl.cu): exchange and all-reduce inside the library ------
    def comm_init(self, rank, world, dist=None, group=None):
        """Collective.  Rank 0 creates the NCCL id; `dist` (torch.distributed) only carries its
        128 bytes to the other ranks."""
        ident = np.zeros(128, np.uint8)
        if rank == 0:
            self._chk(self.e.L.sphb_nccl_unique_id(ident.ctypes.data))
        if world > 1:
            box = [ident.tobytes()]
            dist.broadcast_object_list(box, src=0, group=group)
            ident = np.frombuffer(box[0], np.uint8).copy()
        self._chk(self.e.L.sphb_slab_comm_init(self.e.h, ident.ctypes.data, int(rank), int(world)))
        self.native = True

    def native_start(self):
        self._chk(self.e.L.sphb_slab_start(self.e.h))

    def native_step(self, dt):
        """(computed dt, attempts, owned, ghosts) — one C call: pack, exchange, unpack, forces,
        advection attempts with their all-reduce, commit."""
        out, att, no, ng = C.c_float(), C.c_int(), C.c_uint32(), C.c_uint32()
        self._chk(self.e.L.sphb_slab_step(self.e.h, dt, C.byref(out), C.byref(att), C.byref(no), C.byref(ng)))
        return out.value, att.value, no.value, ng.value

    def native_phase_ms(self, reset=False):
        out = np.zeros(3, np.float64)
        self._chk(self.e.L.sphb_slab_phase_ms(self.e.h, out.ctypes.data, int(reset)))
        return out

    SEGMENTS = ["pack", "halo_wait", "count_readback_unpack", "sort_to_forces", "first_advect",
                "reduce_allreduce", "readback_and_retries", "idle_between_steps"]

    def native_gpu_segments(self, reset=False):
        """GPU-side ms per segment of sphb_slab_step, averaged over the profiled steps"""
        import ctypes as C
        out = np.zeros(8, np.float64)
        steps = C.c_uint64()
        self._chk(self.e.L.sphb_slab_gpu_segments(self.e.h, out.ctypes.data, C.byref(steps), int(reset)))
        k = max(1, int(steps.value))
        return {name: float(v) / k for name, v in zip(self.SEGMENTS, out)}

    def p2p_active(self):
        """the halo exchange runs over peer memory (no ncclSend / ncclRecv)"""
        return bool(self.e.L.sphb_slab_p2p_active(self.e.h))

    def native_resends(self):
        return int(self.e.L.sphb_slab_resends(self.e.h))

    def native_bytes_sent(self):
        return int(self.e.L.sphb_slab_bytes_sent(self.e.h))

    def close(self):
        self.e.close()


class SlabDriver:
    """The per-step exchange protocol of one rank.  `local` implements the local-engine
    interface (CudaSlab, or a CPU stand-in); `dist` is torch.distributed (any backend)."""

    def __init__(self, local, cuts, rank, world, dist=None, group=None):
        assert len(cuts) == world + 1
        self.local, self.rank, self.world, self.dist, self.group = local, rank, world, dist, group
        self.left = rank - 1 if rank > 0 else None
        self.right = rank + 1 if rank + 1 < world else None
        self.cuts = [float(np.float32(c)) for c in cuts]
        local.configure(self.cuts[rank], self.cuts[rank + 1], self.left is not None, self.right is not None)
        # the library drives the step itself when the local engine holds a communicator
        # (CudaSlab.comm_init); otherwise the protocol below runs here, over `dist`
        self.native = bool(getattr(local, "native", False))
        self.n_owned = self.n_ghost = 0
        self.attempts = 0
        self._bytes_sent = 0
        self.profile = False
        self.phase_s = dict(exchange=0.0, forces_advect=0.0, retries_commit=0.0)
        if self.native:
            return
        self.send_l = local.new_message() if self.left is not None else None
        self.recv_l = local.new_message() if self.left is not None else None
        self.send_r = local.new_message() if self.right is not None else None
        self.recv_r = local.new_message() if self.right is not None else None
        self.red8 = local.new_red8()
        self.gmin = self.gmax = None
        # phase_s: wall-clock seconds per phase, split at the step's host synchronisation points
        # (filled when `profile` is set: pack + exchange + unpack | sort .. forces + first
        # advection attempt + all-reduce | further attempts + commit)

    @property
    def bytes_sent(self):
        return self.local.native_bytes_sent() if self.native else self._bytes_sent

    # -- collectives ---------------------------------------------------------------------------
    def _allreduce_max(self, t):
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)

    def _exchange(self):
        if self.world == 1:
            return
        d = self.dist
        ops = []
        # a fixed order (left pair, then right pair) on every rank keeps the pairs matched
        if self.left is not None:
            ops.append(d.P2POp(d.isend, self.send_l, self.left, self.group))
            ops.append(d.P2POp(d.irecv, self.recv_l, self.left, self.group))
        if self.right is not None:
            ops.append(d.P2POp(d.isend, self.send_r, self.right, self.group))
            ops.append(d.P2POp(d.irecv, self.recv_r, self.right, self.group))
        for req in d.batch_isend_irecv(ops):
            req.wait()
        self._bytes_sent += sum(int(t.numel()) for t in (self.send_l, self.send_r) if t is not None)

    def _set_bbox(self, r):
        self.gmin = np.array([-r[0], -r[1], -r[2]], np.float32)
        self.gmax = np.array([r[3], r[4], r[5]], np.float32)

    # -- life cycle ----------------------------------------------------------------------------
    def start(self, particles, ids):
        """Uploads this rank's particles and agrees on the initial global bbox."""
        if self.native:
            self.local.upload(particles, ids)
            self.local.native_start()
            return
        with self.local.stream_context():
            self.local.upload(particles, ids)
            self.local.bbox(self.red8)
            self._allreduce_max(self.red8)
            self._set_bbox(self.local.red8_to_host(self.red8))

    def step(self, dt):
        """One simulate_single_frame over all slabs; returns the computed dt (identical on
        every rank)."""
        loc = self.local
        if self.native:
            if self.profile:
                loc.native_phase_ms(reset=True)
            computed, self.attempts, self.n_owned, self.n_ghost = loc.native_step(dt)
            if self.profile:
                ph = loc.native_phase_ms()
                for k, v in zip(("exchange", "forces_advect", "retries_commit"), ph):
                    self.phase_s[k] += v * 1e-3
            return computed
        t0 = time.perf_counter() if self.profile else 0.0
        with loc.stream_context():
            loc.pack(self.send_l, self.send_r)
            self._exchange()
            self.n_owned, self.n_ghost = loc.unpack(self.recv_l, self.recv_r)  # host sync
            t1 = time.perf_counter() if self.profile else 0.0
            loc.forces(self.gmin, self.gmax)
            computed = dt
            self.attempts = 0
            t2 = t1
            while True:
                dt = computed
                loc.advect(dt, self.red8)
                self._allreduce_max(self.red8)
                r = loc.red8_to_host(self.red8)  # host sync
                if self.profile and self.attempts == 0:
                    t2 = time.perf_counter()
                computed = loc.compute_dt(r[6], r[7])
                self.attempts += 1
                if not needs_retry(dt, computed):
                    break
            loc.commit()
            self._set_bbox(r)
        if self.profile:
            t3 = time.perf_counter()
            self.phase_s["exchange"] += t1 - t0
            self.phase_s["forces_advect"] += t2 - t1
            self.phase_s["retries_commit"] += t3 - t2
        return computed


class LoopbackWorld:
    """All ranks of a slab decomposition inside ONE process (one GPU, or the CPU stand-in):
    the same per-rank calls as SlabDriver.step, run phase by phase over the ranks, with the
    messages moved by a copy and the all-reduce done by hand.  For tests, and for running a
    domain larger than one engine's capacity on a single device."""

    def __init__(self, locals_, cuts):
        self.world = len(locals_)
        self.ranks = [SlabDriver(loc, cuts, r, self.world) for r, loc in enumerate(locals_)]
        self.attempts = 0

    def _allreduce(self):
        hosts = []
        for d in self.ranks:  # the read-back must be ordered behind the rank's own (non-blocking) stream
            with d.local.stream_context():
                hosts.append(d.local.red8_to_host(d.red8))
        return np.maximum.reduce(hosts).astype(np.float32)

    def start(self, parts):
        """parts: per rank (particles, ids)"""
        for d, (p, ids) in zip(self.ranks, parts):
            with d.local.stream_context():
                d.local.upload(p, ids)
                d.local.bbox(d.red8)
        r = self._allreduce()
        for d in self.ranks:
            d._set_bbox(r)

    def step(self, dt):
        R = self.ranks
        for d in R:
            with d.local.stream_context():
                d.local.pack(d.send_l, d.send_r)
        for k, d in enumerate(R):  # my right message is my right neighbour's left inbox
            with d.local.stream_context():
                if d.right is not None:
                    _sync(d.local)
                    R[k + 1].recv_l.copy_(d.send_r)
                if d.left is not None:
                    _sync(d.local)
                    R[k - 1].recv_r.copy_(d.send_l)
        for d in R:
            _sync(d.local)
        for d in R:
            with d.local.stream_context():
                d.n_owned, d.n_ghost = d.local.unpack(d.recv_l, d.recv_r)
                d.local.forces(d.gmin, d.gmax)
        computed = dt
        self.attempts = 0
        while True:
            dt = computed
            for d in R:
                with d.local.stream_context():
                    d.local.advect(dt, d.red8)
            r = self._allreduce()
            computed = R[0].local.compute_dt(r[6], r[7])
            self.attempts += 1
            if not needs_retry(dt, computed):
                break
        for d in R:
            with d.local.stream_context():
                d.local.commit()
            d._set_bbox(r)
            d.attempts = self.attempts
        return computed


def _sync(local):
    t = getattr(local, "torch", None)
    if t is not None:
        t.cuda.synchronize()


def equal_count_cuts(x, world):
    """Slab faces for `world` ranks with (nearly) equal particle counts: faces lie midway
    between neighbouring distinct x values; the outer faces are -inf / +inf."""
    xs = np.sort(np.asarray(x, np.float32))
    cuts = [-np.inf]
    for r in range(1, world):
        k = (len(xs) * r) // world
        lo = xs[k - 1] if k > 0 else xs[0]
        # first value strictly greater than xs[k-1], so that a lattice plane is never split
        j = np.searchsorted(xs, lo, side="right")
        hi = xs[j] if j < len(xs) else lo
        cuts.append(float(np.float32((np.float64(lo) + np.float64(hi)) / 2)))
    cuts.append(np.inf)
    return cuts
