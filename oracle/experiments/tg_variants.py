"""ORACLE - TEST INFRASTRUCTURE ONLY.  Fixture-vintage experiments for the 2-D Taylor-Green case.

A numpy restatement of tests/2d_examples/test_2d_taylor_green/taylor_green.cpp:155-192 with switches
for scheme variants that earlier SPHinXsys revisions used, to find which scheme the stored
WaterBody_TotalKineticEnergy / WaterBody_MaximumSpeed runs (t = 0.1: 0.20646315402108778 / 0.9026147)
come from.  With every switch at its default it reproduces oracle/liboracle.so (the current sources).
    python oracle/experiments/tg_variants.py            # table of variants at t = 0.1, 0.2
"""
import itertools
import sys

import numpy as np
from scipy.spatial import cKDTree

PI = np.pi


class TG:
    def __init__(self, dp=0.01, **opt):
        self.o = dict(riemann1=True, riemann2=False, ujump_limited=False, tvc="limited", tvc_coef=0.2, visc=True,
                      vol_update=True, acc_by_rho=False, slope=3.0, pj_sign=1.0, dt_first=True, cfl_ac=0.6, cfl_ad=0.25,
                      visc_dt=True, tvc_when="before", dens_free=False, h_ratio=1.3)
        self.o.update(opt)
        self.dp = dp
        self.h = self.o["h_ratio"] * dp
        self.rc = 2 * self.h
        n1 = int(round(1 / dp))
        ax = -4 * dp + (np.arange(n1 + 8) + 0.5) * dp
        ax = ax[(ax > 0) & (ax < 1)]
        X, Y = np.meshgrid(ax, ax, indexing="ij")
        self.x = np.stack([X.ravel(), Y.ravel()], 1)
        self.n = len(self.x)
        self.rho0, self.c0, self.mu, self.U = 1.0, 10.0, 0.01, 1.0
        self.vol = np.full(self.n, dp * dp)
        self.m = self.rho0 * self.vol
        self.rho = np.full(self.n, self.rho0)
        x, y = self.x[:, 0], self.x[:, 1]
        self.v = np.stack([-np.cos(2 * PI * x) * np.sin(2 * PI * y), np.sin(2 * PI * x) * np.cos(2 * PI * y)], 1)
        self.p = np.zeros(self.n)
        self.drho = np.zeros(self.n)
        self.F = np.zeros((self.n, 2))
        self.Fp = np.zeros((self.n, 2))
        self.Fv_prev = np.zeros((self.n, 2))
        a2 = 7.0 / (4.0 * PI * self.h ** 2)
        self.fW, self.fdW = a2, a2 / self.h
        self.W0 = a2
        # sigma0: adaptation.cpp:26-40
        depth = int(self.rc / dp) + 1
        s = 0.0
        for j in range(-depth, depth + 1):
            for i in range(-depth, depth + 1):
                d = np.hypot(i * dp, j * dp)
                if d < self.rc:
                    s += self.W(d)
        self.sigma0 = s
        self.time = 0.0
        self.nadv = 0
        self.configure()

    def W(self, r):
        q = r / self.h
        return self.fW * (1 - 0.5 * q) ** 4 * (1 + 2 * q)

    def dW(self, r):
        q = r / self.h
        return self.fdW * 0.625 * (q - 2) ** 3 * q

    def configure(self):
        xw = np.mod(self.x, 1.0)
        tree = cKDTree(xw, boxsize=1.0)
        pairs = tree.query_pairs(self.rc, output_type="ndarray")
        i = np.concatenate([pairs[:, 0], pairs[:, 1]])
        j = np.concatenate([pairs[:, 1], pairs[:, 0]])
        d = self.x[i] - self.x[j]
        d -= np.round(d)
        r = np.sqrt((d * d).sum(1))
        keep = (d * d).sum(1) < self.rc ** 2
        i, j, d, r = i[keep], j[keep], d[keep], r[keep]
        order = np.argsort(i, kind="stable")
        self.i, self.j, self.r = i[order], j[order], r[order]
        self.e = d[order] / (self.r + 2.2e-16)[:, None]
        self.Wij = self.W(self.r)
        self.dWij = self.dW(self.r)

    def psum(self, val):
        if val.ndim == 1:
            return np.bincount(self.i, val, self.n)
        return np.stack([np.bincount(self.i, val[:, k], self.n) for k in range(val.shape[1])], 1)

    def advection_dt(self):
        acc = 4 * self.h * np.linalg.norm(self.F + self.Fp, axis=1) / self.m
        mx = np.sqrt(max(np.max((self.v ** 2).sum(1)), np.max(acc)))
        ref = self.U
        if self.o["visc_dt"]:
            ref = max(self.mu / self.rho0 / self.h, ref)
        return self.o["cfl_ad"] * self.h / (max(mx, ref) + 2.2e-16)

    def acoustic_dt(self):
        return self.o["cfl_ac"] * self.h / (self.c0 + np.sqrt((self.v ** 2).sum(1)).max() + 2.2e-16)

    def density_summation(self):
        sig = self.W0 + self.psum(self.Wij)
        rs = sig * self.rho0 / self.sigma0
        self.rho = np.maximum(rs, self.rho0) if self.o["dens_free"] else rs
        if self.o["vol_update"]:
            self.vol = self.m / self.rho

    def viscous(self):
        if not self.o["visc"]:
            return
        dv = (self.v[self.i] - self.v[self.j]) / (self.r + 0.01 * self.h)[:, None]
        f = self.psum(2.0 * self.mu * dv * (self.dWij * self.vol[self.j])[:, None]) * self.vol[:, None]
        self.Fp += f - self.Fv_prev
        self.Fv_prev = f

    def tvc(self, Dt):
        mode = self.o["tvc"]
        if mode == "off":
            return
        G = -self.psum(2.0 * (self.dWij * self.vol[self.j])[:, None] * self.e)
        h2 = self.h ** 2
        if mode == "limited":
            lim = np.minimum(100.0 * h2 * (G * G).sum(1), 1.0)
            self.x += self.o["tvc_coef"] * h2 * lim[:, None] * G
        elif mode == "plain":
            self.x += self.o["tvc_coef"] * h2 * G
        elif mode == "pb":  # p_background = 7 rho0 vmax^2 ; pos += 0.5 dt^2 a_trans, a = G * pb / rho
            vmax = np.sqrt((self.v ** 2).sum(1)).max()
            pb = self.o.get("pb_coef", 7.0) * self.rho0 * vmax * vmax
            self.x += 0.5 * Dt * Dt * G * (pb / self.rho)[:, None]

    def first_half(self, dt):
        o = self.o
        self.rho += self.drho * dt * 0.5
        self.p = self.rho0 * self.c0 ** 2 * (self.rho / self.rho0 - 1)
        self.x += self.v * dt * 0.5
        dWV = self.dWij * self.vol[self.j]
        f = -self.psum(((self.p[self.i] + self.p[self.j]) * dWV)[:, None] * self.e)
        diss = 0.0
        if o["riemann1"]:
            dj = (self.p[self.i] - self.p[self.j]) / (self.rho0 * self.c0)
            if o["ujump_limited"]:
                u = -((self.v[self.i] - self.v[self.j]) * self.e).sum(1)
                lim = np.minimum(o["slope"] * np.maximum(u, 0) / self.c0, 1.0)
                dj = dj * lim * lim
            diss = self.psum(dj * dWV)
        if o["acc_by_rho"]:
            self.F += f * (self.m / self.rho)[:, None]
        else:
            self.F += f * self.vol[:, None]
        self.drho = diss * self.rho
        self.v += (self.Fp + self.F) / self.m[:, None] * dt

    def second_half(self, dt):
        o = self.o
        self.x += self.v * dt * 0.5
        dWV = self.dWij * self.vol[self.j]
        u = ((self.v[self.i] - self.v[self.j]) * self.e).sum(1)
        rate = self.psum(u * dWV)
        if o["riemann2"]:
            us = o["pj_sign"] * u
            lim = np.minimum(o["slope"] * np.maximum(us, 0) / self.c0, 1.0)
            pj = self.rho0 * self.c0 * u * lim
            self.F = self.psum((pj * dWV)[:, None] * self.e) * self.vol[:, None]
        else:
            self.F = np.zeros_like(self.F)
        self.drho += rate * self.rho
        self.rho += self.drho * dt * 0.5

    def advection_step(self):
        Dt = self.advection_dt()
        self.density_summation()
        self.viscous()
        if self.o["tvc_when"] == "before":
            self.tvc(Dt)
        relax, k = 0.0, 0
        while relax < Dt:
            dt = min(self.acoustic_dt(), Dt)
            relax += dt
            self.first_half(dt)
            self.second_half(dt)
            self.time += dt
            k += 1
        self.nadv += 1
        self.x = np.where(self.x < 0, self.x + 1, self.x)
        self.x = np.where(self.x > 1, self.x - 1, self.x)
        self.configure()
        return Dt, k

    def ke(self):
        return float((0.5 * self.m * (self.v ** 2).sum(1)).sum())

    def vmax(self):
        return float(np.sqrt((self.v ** 2).sum(1)).max())


def sample(opt, t_samples=(0.1, 0.2)):
    c = TG(**opt)
    out = []
    for _ in t_samples:
        integ = 0.0
        while integ < 0.1:
            t0 = c.time
            c.advection_step()
            integ += c.time - t0
        out.append((c.nadv, c.time, c.ke(), c.vmax()))
    return out


GOLD = {0.1: (0.20646315402108778, 0.9026147), 0.2: (0.10449404996918778, 0.89899294)}

if __name__ == "__main__":
    base = sample({})
    print("default (current sources):", base)
