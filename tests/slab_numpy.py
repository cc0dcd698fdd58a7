"""A numpy stand-in for the slab engine's local-engine interface (libclsph_b200/slab.py), used
by the world_size > 1 gloo tests of the exchange PROTOCOL on CPU.  It speaks the same message
format as libclsph_b200/csrc/kernels_slab.cu (16-byte header {count, migrants}, 48-byte records
(x, y, z, id)(vx, vy, vz, migrant)(v_half xyz, 0), ghost = bit 31 of the id) and the same
classification rule, but its "physics" is trivial: constant acceleration, and a neighbour
COUNT within h in place of the density — enough to prove that after every exchange each rank
holds exactly the particles it needs.  TEST INFRASTRUCTURE ONLY."""
import contextlib

import numpy as np
import torch

GHOST = np.uint32(0x80000000)


class NumpySlab:
    def __init__(self, h, exchange_capacity, g=(0.0, -9.8, 0.0), dt_cap=1.0 / 600.0):
        self.h = np.float32(h)
        self.halo = np.float32(2.0) * self.h
        self.cap = int(exchange_capacity)
        self.g = np.array(g, np.float32)
        self.dt_cap = np.float32(dt_cap)
        self.pos = np.zeros((0, 3), np.float32)
        self.vel = np.zeros((0, 3), np.float32)
        self.vhf = np.zeros((0, 3), np.float32)
        self.ids = np.zeros(0, np.uint32)
        self.count = np.zeros(0, np.int64)

    # -- buffers -----------------------------------------------------------------------------
    def new_message(self):
        return torch.zeros(16 + 48 * self.cap, dtype=torch.uint8)

    def new_red8(self):
        return torch.zeros(8, dtype=torch.float32)

    def red8_to_host(self, red8):
        return red8.numpy().copy()

    def stream_context(self):
        return contextlib.nullcontext()

    # -- engine ------------------------------------------------------------------------------
    def configure(self, xlo, xhi, has_left, has_right):
        self.xlo, self.xhi = np.float32(xlo), np.float32(xhi)
        self.has_left, self.has_right = bool(has_left), bool(has_right)

    def upload(self, particles, ids):
        self.pos = particles["position"][:, :3].copy()
        self.vel = particles["velocity"][:, :3].copy()
        self.vhf = particles["intermediate_velocity"][:, :3].copy()
        self.ids = np.asarray(ids, np.uint32).copy()

    def _red(self, owned, vel=None, acc=None):
        r = np.full(8, -np.inf, np.float32)
        if owned.any():
            p = self.pos_new[owned] if hasattr(self, "pos_new") and vel is not None else self.pos[owned]
            r[0:3] = -p.min(axis=0)
            r[3:6] = p.max(axis=0)
            if vel is not None:
                r[6] = np.max(np.sum(vel[owned] * vel[owned], axis=1, dtype=np.float32))
                r[7] = np.max(np.sum(acc[owned] * acc[owned], axis=1, dtype=np.float32))
        return r

    def bbox(self, red8):
        red8.copy_(torch.from_numpy(self._red((self.ids & GHOST) == 0)))

    def _write(self, msg, sel, migrant):
        m = msg.numpy()
        n = int(sel.sum())
        assert n <= self.cap, "exchange capacity exceeded"
        m[:16].view(np.uint32)[:] = [n, int(migrant[sel].sum()), 0, 0]
        rec = m[16:16 + 48 * n].view(np.float32).reshape(n, 12)
        rec[:, 0:3] = self.pos[sel]
        rec[:, 3] = (self.ids[sel] & ~GHOST).view(np.float32)
        rec[:, 4:7] = self.vel[sel]
        rec[:, 7] = migrant[sel].astype(np.float32)
        rec[:, 8:11] = self.vhf[sel]
        rec[:, 11] = 0

    def pack(self, send_left, send_right):
        x = self.pos[:, 0]
        live = (self.ids & GHOST) == 0
        mig_r = live & self.has_right & (x >= self.xhi)
        mig_l = live & self.has_left & ~mig_r & (x < self.xlo)
        to_r = live & self.has_right & (mig_r | (~mig_l & (x >= self.xhi - self.halo)))
        to_l = live & self.has_left & (mig_l | (~mig_r & (x < self.xlo + self.halo)))
        if send_right is not None:
            self._write(send_right, to_r, mig_r)
        if send_left is not None:
            self._write(send_left, to_l, mig_l)
        gone = mig_r | mig_l
        self.ids = np.where(gone, self.ids | GHOST, self.ids & ~GHOST).astype(np.uint32)

    def _read(self, msg):
        m = msg.numpy()
        n, _mig = m[:16].view(np.uint32)[:2]
        rec = m[16:16 + 48 * int(n)].view(np.float32).reshape(int(n), 12)
        ids = rec[:, 3].copy().view(np.uint32) & ~GHOST
        ids = np.where(rec[:, 7] != 0, ids, ids | GHOST).astype(np.uint32)
        return rec[:, 0:3].copy(), rec[:, 4:7].copy(), rec[:, 8:11].copy(), ids

    def unpack(self, recv_left, recv_right):
        for msg in (recv_left, recv_right):
            if msg is None:
                continue
            p, v, w, ids = self._read(msg)
            self.pos = np.concatenate([self.pos, p])
            self.vel = np.concatenate([self.vel, v])
            self.vhf = np.concatenate([self.vhf, w])
            self.ids = np.concatenate([self.ids, ids])
        owned = (self.ids & GHOST) == 0
        return int(owned.sum()), int((~owned).sum())

    def forces(self, gmin, gmax):
        """neighbour count within h (self included) over the LOCAL particles, for owned ones"""
        from scipy.spatial import cKDTree
        self.gmin, self.gmax = np.asarray(gmin), np.asarray(gmax)
        tree = cKDTree(self.pos.astype(np.float64))
        h = float(self.h) * (1 - 1e-6)
        self.count = np.array([len(c) for c in tree.query_ball_point(self.pos.astype(np.float64), h)], np.int64)
        self.acc = np.tile(self.g, (len(self.pos), 1)).astype(np.float32)

    def advect(self, dt, red8):
        dt = np.float32(dt)
        owned = (self.ids & GHOST) == 0
        w = (self.vhf + self.acc * dt).astype(np.float32)
        self.pos_new = np.where(owned[:, None], self.pos + w * dt, self.pos).astype(np.float32)
        self.vel_new = np.where(owned[:, None], (self.vhf + w) / np.float32(2), self.vel).astype(np.float32)
        self.vhf_new = np.where(owned[:, None], w, self.vhf).astype(np.float32)
        red8.copy_(torch.from_numpy(self._red(owned, self.vel_new, self.acc)))

    def commit(self):
        owned = (self.ids & GHOST) == 0
        self.pos, self.vel, self.vhf = self.pos_new[owned], self.vel_new[owned], self.vhf_new[owned]
        self.ids = self.ids[owned]
        self.count = self.count[owned]
        del self.pos_new

    def compute_dt(self, max_v2, max_a2):
        v2, a = np.float32(max_v2), np.sqrt(np.float32(max_a2))
        v = np.sqrt(v2)
        dt = (np.sqrt(np.float32(2) * a * self.h + v2) - v) / (np.float32(2) * a)
        dt = np.float32(dt)
        if dt < 0.00001:
            dt = np.float32(0.00001)
        if dt > self.dt_cap:
            dt = self.dt_cap
        return float(dt)

    def download(self):
        return self.pos.copy(), self.ids.copy(), self.count.copy()
