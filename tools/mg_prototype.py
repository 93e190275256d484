#!/usr/bin/env python
"""numpy prototype behind the design choices of the multigrid solver -- CPU only.

    python tools/mg_prototype.py [--variant volume|transpose|plain]

Convergence factor per V(2,2) cycle on several grids for three treatments of the walls on levels that were
reached by keeping every second point (odd number of unknowns per axis: 129^2, 33^3, 256^2 after one step):
  volume     (shipped) boundary rows scaled by 1 / control-volume width, restriction = volume-weighted average
  transpose  plain mirrored coarse operator, restriction = half the transpose of the interpolation
  plain      plain mirrored coarse operator, plain full weighting (1/4, 1/2, 1/4) everywhere
Measured here: plain stalls at 0.57-0.65 per cycle on the odd grids; transpose is 0.17 in 2-D but 0.44 in 3-D;
volume is 0.23 / 0.28 -- the only one that is uniform, which is why it is the one that was built.  Even grids
(pairs of cells merged) run at 0.14-0.2 with all three.
This is NOT the oracle (oracle/oracle_mg.inc is): it is vectorised numpy with its own operation order.
"""
import argparse

import numpy as np


def mirror(x):
    for ax in range(x.ndim):
        if x.shape[ax] == 1:
            continue
        s = [slice(None)] * x.ndim
        d = [slice(None)] * x.ndim
        s[ax], d[ax] = 0, 1
        x[tuple(s)] = x[tuple(d)]
        s[ax], d[ax] = -1, -2
        x[tuple(s)] = x[tuple(d)]


class Level:
    def __init__(self, shape, h, aL, aR):
        nd = len(shape)
        self.shape, self.h, self.aL, self.aR, self.kinds = tuple(shape), tuple(h), tuple(aL), tuple(aR), None
        self.act = [n > 1 for n in shape]
        self.inner = tuple(slice(1, -1) if a else slice(None) for a in self.act)
        self.vol, self.w = [None] * nd, [None] * nd
        for ax, n in enumerate(shape):
            if n == 1:
                continue
            v = np.ones(n - 2)
            v[0], v[-1] = aL[ax] + 0.5, aR[ax] + 0.5
            self.vol[ax] = v
            sh = [1] * nd
            sh[ax] = -1
            self.w[ax] = (1.0 / v).reshape(sh)
        self.diag = sum(2.0 * self.w[ax] / h[ax] ** 2 for ax in range(nd) if self.act[ax])
        self.colour = (np.indices(shape).sum(0) % 2)[self.inner]

    def nbsum(self, x):
        out = 0.0
        for ax in range(x.ndim):
            if not self.act[ax]:
                continue
            p, m = list(self.inner), list(self.inner)
            p[ax], m[ax] = slice(2, None), slice(0, -2)
            out = out + self.w[ax] * (x[tuple(p)] + x[tuple(m)]) / self.h[ax] ** 2
        return out

    def residual(self, x, b):
        mirror(x)
        r = np.zeros_like(x)
        r[self.inner] = b[self.inner] - (self.nbsum(x) - self.diag * x[self.inner])
        return r

    def sweep(self, x, b):
        for col in (0, 1):
            mirror(x)
            xn = (self.nbsum(x) - b[self.inner]) / self.diag
            x[self.inner] = np.where(self.colour == col, xn, x[self.inner])
        mirror(x)


def restrict1(r, ax, kind, vf, vc, plain=False):
    m = r.shape[ax] - 2
    r = np.moveaxis(r, ax, 0).copy()
    sh = (-1,) + (1,) * (r.ndim - 1)
    r[1:-1] *= vf.reshape(sh)
    if kind == "cc":
        out = np.zeros((m // 2 + 2,) + r.shape[1:])
        out[1:-1] = (r[1:m:2] + r[2:m + 1:2]) / (2 * vc.reshape(sh))
    else:
        out = np.zeros(((m - 1) // 2 + 2,) + r.shape[1:])
        a, c = 0.5 * r[1:m - 1:2], 0.5 * r[3:m + 1:2]
        if not plain:
            a[0], c[-1] = r[1], r[m]
        out[1:-1] = (a + r[2:m:2] + c) / (2 * vc.reshape(sh))
    return np.moveaxis(out, 0, ax)


def prolong1(E, ax, kind, m):
    E = np.moveaxis(E, ax, 0)
    mc = E.shape[0] - 2
    out = np.zeros((m + 2,) + E.shape[1:])
    if kind == "cc":
        out[1:m:2] = 0.75 * E[1:-1] + 0.25 * E[0:-2]
        out[2:m + 1:2] = 0.75 * E[1:-1] + 0.25 * E[2:]
    else:
        out[2:m:2] = E[1:-1]
        out[1:m:2] = 0.5 * (E[0:-2] + E[1:-1])
        out[m] = 0.5 * (E[mc] + E[mc + 1])
    return np.moveaxis(out, 0, ax)


def build(shape, h, correct):  # correct: carry the wall distances (variant 'volume')
    nd = len(shape)
    L = Level(shape, h, [0.5] * nd, [0.5] * nd)
    levels = [L]
    while True:
        kinds, ns, nh, aL, aR, ok, anyc = [], [], [], [], [], True, False
        for ax, n in enumerate(L.shape):
            m = n - 2
            if n == 1:
                kinds.append(None), ns.append(1), nh.append(L.h[ax]), aL.append(.5), aR.append(.5)
                continue
            if m % 2 == 0 and m // 2 >= 2:
                kinds.append("cc"), ns.append(m // 2 + 2)
                aL.append((L.aL[ax] + .5) / 2 if correct else .5), aR.append((L.aR[ax] + .5) / 2 if correct else .5)
            elif m % 2 == 1 and (m - 1) // 2 >= 3:
                kinds.append("vc"), ns.append((m - 1) // 2 + 2)
                aL.append((L.aL[ax] + 1) / 2 if correct else .5), aR.append((L.aR[ax] + 1) / 2 if correct else .5)
            else:
                ok = False
                break
            anyc = True
            nh.append(2 * L.h[ax])
        if not ok or not anyc:
            break
        L.kinds = kinds
        L = Level(ns, nh, aL, aR)
        levels.append(L)
    return levels


PLAIN = False


def vcycle(levels, l, x, b):
    L = levels[l]
    if L.kinds is None:
        for _ in range(30):
            L.sweep(x, b)
        return
    L.sweep(x, b)
    L.sweep(x, b)
    rc = L.residual(x, b)
    for ax, k in enumerate(L.kinds):
        if k:
            rc = restrict1(rc, ax, k, L.vol[ax], levels[l + 1].vol[ax], PLAIN)
    E = np.zeros(levels[l + 1].shape)
    vcycle(levels, l + 1, E, rc)
    mirror(E)
    e = E
    for ax, k in enumerate(L.kinds):
        if k:
            e = prolong1(e, ax, k, L.shape[ax] - 2)
    x[L.inner] += e[L.inner]
    L.sweep(x, b)
    L.sweep(x, b)


def run(shape, correct, cycles=12):
    rng = np.random.default_rng(0)
    h = [1.0 / (n - 1) if n > 1 else 1.0 for n in shape]
    levels = build(shape, h, correct)
    L = levels[0]
    b = np.zeros(shape)
    b[L.inner] = rng.standard_normal(b[L.inner].shape)
    b[L.inner] -= b[L.inner].mean()
    x = np.zeros(shape)
    res = [np.abs(L.residual(x, b)).max()]
    for _ in range(cycles):
        vcycle(levels, 0, x, b)
        res.append(np.abs(L.residual(x, b)).max())
    fac = [res[i + 1] / res[i] for i in range(cycles)]
    print("%-16s levels %d coarsest %-12s  residual %.1e -> %.1e   factor per cycle (last 3): %s"
          % (shape, len(levels), levels[-1].shape, res[0], res[-1], " ".join("%.3f" % f for f in fac[-3:])))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="volume", choices=["volume", "transpose", "plain"])
    a = ap.parse_args()
    PLAIN = a.variant == "plain"
    for shape in [(1, 129, 129), (1, 130, 130), (1, 256, 256), (1, 100, 100), (33, 33, 33), (34, 34, 34), (66, 66, 66)]:
        run(shape, a.variant == "volume")
