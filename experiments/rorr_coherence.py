#!/usr/bin/env python
"""How often do the 4 units of a RORR warp agree on the closed-form (no sort) decision of a merge?
CPU design aid for the unit-order choice: cfg2 columns, groups of 4 adjacent columns (same bin, slot)
against groups of 4 adjacent slots (same column, bin).  usage: experiments/rorr_coherence.py [ngroups]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth

ng = int(sys.argv[1]) if len(sys.argv) > 1 else 3
rnd = len(sys.argv) > 2 and sys.argv[2] == "random"
tb = synth.make_tables()
cols = synth.make_columns(4096, nz=100, doubled=False, **(dict(randomize=True, seed=11) if rnd else {}))
G, nw = tb.ngauss, tb.nw
w = np.asarray(tb.weights); E = np.concatenate(([0.0], np.cumsum(w)))
wxy = (w[:, None] * w[None, :]).reshape(-1)


def bracket(xt, x):
    return np.clip(np.searchsorted(xt, x, side="right") - 1, 0, len(xt) - 2)


def closed_flags(c, bins):
    """[bin][merge][layer] -> closed-form?"""
    T, P, dens, dz = cols["T"][:, c], cols["P"][:, c], cols["densities"][:, :, c], cols["dz"][:, c]
    nz = len(T)
    log10P = np.log10(P); colsd = dens * dz[:, None]
    out = np.zeros((len(bins), len(tb.k) - 1, nz), bool)
    kbr = []
    for k in tb.k:
        PP = np.clip(log10P, k.log10P[0], k.log10P[-1]); TT = np.clip(T, k.temp[0], k.temp[-1])
        ix, iy = bracket(k.log10P, PP), bracket(k.temp, TT)
        kbr.append((ix, iy, (PP - k.log10P[ix]) / (k.log10P[ix + 1] - k.log10P[ix]), (TT - k.temp[iy]) / (k.temp[iy + 1] - k.temp[iy])))
    for bi, l in enumerate(bins):
        ks = []
        for k, (ix, iy, q1, q2) in zip(tb.k, kbr):
            F = k.log10k[:, :, :, l]
            fx1 = (1 - q1)[:, None] * F[:, ix, iy].T + q1[:, None] * F[:, ix + 1, iy].T
            fx2 = (1 - q1)[:, None] * F[:, ix, iy + 1].T + q1[:, None] * F[:, ix + 1, iy + 1].T
            ks.append(10.0 ** ((1 - q2)[:, None] * fx1 + q2[:, None] * fx2))
        a = ks[0] * colsd[:, tb.k[0].sp_ind - 1][:, None]
        for jj in range(1, len(tb.k)):
            b = ks[jj] * colsd[:, tb.k[jj].sp_ind - 1][:, None]
            new = np.empty_like(a)
            for j in range(nz):
                aj, bj = a[j], b[j]
                rowmaj = np.all(aj[:-1] + bj[-1] <= aj[1:] + bj[0]); colmaj = np.all(bj[:-1] + aj[-1] <= bj[1:] + aj[0])
                out[bi, jj - 1, j] = rowmaj or colmaj
                txy = (aj[:, None] + bj[None, :]).reshape(-1)
                inds = np.argsort(txy, kind="stable")
                e = np.concatenate(([0.0], np.cumsum(wxy[inds])))
                Fc = np.concatenate(([0.0], np.cumsum(txy[inds] * np.diff(e))))
                new[j] = np.diff(np.interp(E, e, Fc)) / np.diff(E)
            a = new
    return out


bins = list(range(0, nw, 7))
tot = dict(unit=0, n=0, colgrp_all=0, colgrp_n=0, slotgrp_all=0, slotgrp_n=0)
for c0 in np.linspace(0, 4092, ng).astype(int) // 4 * 4:
    f = np.stack([closed_flags(c0 + i, bins) for i in range(4)])  # [4][bin][merge][layer]
    tot["unit"] += f.sum(); tot["n"] += f.size
    tot["colgrp_all"] += f.all(axis=0).sum(); tot["colgrp_n"] += f[0].size
    nz = f.shape[-1] // 4 * 4
    s = f[..., :nz].reshape(4, len(bins), f.shape[2], nz // 4, 4)
    tot["slotgrp_all"] += s.all(axis=-1).sum(); tot["slotgrp_n"] += s[..., 0].size
print("closed-form merges, per unit:", tot["unit"] / tot["n"])
print("warp (4 adjacent columns) all closed:", tot["colgrp_all"] / tot["colgrp_n"])
print("warp (4 adjacent slots) all closed:", tot["slotgrp_all"] / tot["slotgrp_n"])
