#!/usr/bin/env python
"""Statistics of the RORR merges of the cfg2 workload (design aid, CPU only): how many merges take
the closed form, how many fail the 32-bit-key certificate and why, how far the sorted permutation
moves between adjacent layers.  usage: experiments/rorr_stats.py [ncol_sample]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth

ncs = int(sys.argv[1]) if len(sys.argv) > 1 else 4
tb = synth.make_tables()
cols = synth.make_columns(4096, nz=100, doubled=False)
G, nw = tb.ngauss, tb.nw
w = np.asarray(tb.weights); E = np.concatenate(([0.0], np.cumsum(w)))
wxy = (w[:, None] * w[None, :]).reshape(-1)
pick = np.linspace(0, 4095, ncs).astype(int)
FB = 24
st = dict(merges=0, closed=0, fail=0, fail_edge=0, disp=[], fail_by_jj=np.zeros(8, int), merges_by_jj=np.zeros(8, int),
          unit_fail=0, units=0, unit_fail_edge=0, nswap=[])


def bracket(xt, x):
    i = np.clip(np.searchsorted(xt, x, side="right") - 1, 0, len(xt) - 2)
    return i


for c in pick:
    T, P, dens, dz = cols["T"][:, c], cols["P"][:, c], cols["densities"][:, :, c], cols["dz"][:, c]
    nz = len(T)
    log10P = np.log10(P); colsd = dens * dz[:, None]
    kbr = []
    for k in tb.k:
        PP = np.clip(log10P, k.log10P[0], k.log10P[-1]); TT = np.clip(T, k.temp[0], k.temp[-1])
        ix, iy = bracket(k.log10P, PP), bracket(k.temp, TT)
        kbr.append((ix, iy, (PP - k.log10P[ix]) / (k.log10P[ix + 1] - k.log10P[ix]), (TT - k.temp[iy]) / (k.temp[iy + 1] - k.temp[iy])))
    for l in range(0, nw, 3):
        ks = []
        for k, (ix, iy, q1, q2) in zip(tb.k, kbr):
            F = k.log10k[:, :, :, l]
            fx1 = (1 - q1)[:, None] * F[:, ix, iy].T + q1[:, None] * F[:, ix + 1, iy].T
            fx2 = (1 - q1)[:, None] * F[:, ix, iy + 1].T + q1[:, None] * F[:, ix + 1, iy + 1].T
            ks.append(10.0 ** ((1 - q2)[:, None] * fx1 + q2[:, None] * fx2))
        a = ks[0] * colsd[:, tb.k[0].sp_ind - 1][:, None]
        ufail = np.zeros(nz, bool); ufail_e = np.zeros(nz, bool)
        for jj in range(1, len(tb.k)):
            b = ks[jj] * colsd[:, tb.k[jj].sp_ind - 1][:, None]
            new = np.empty_like(a)
            prev_perm = None
            for j in range(nz):
                aj, bj = a[j], b[j]
                txy = (aj[:, None] + bj[None, :]).reshape(-1)
                st["merges"] += 1; st["merges_by_jj"][jj] += 1
                rowmaj = np.all(aj[:-1] + bj[-1] <= aj[1:] + bj[0]); colmaj = np.all(bj[:-1] + aj[-1] <= bj[1:] + aj[0])
                inds = np.argsort(txy, kind="stable")
                if rowmaj or colmaj:
                    st["closed"] += 1
                else:
                    bits = txy.view(np.uint64)
                    base, top = bits.min(), bits.max()
                    nb = int(top - base).bit_length(); shift = max(0, nb - FB)
                    key = (((bits - base) >> np.uint64(shift)) << np.uint64(8)) | np.arange(256, dtype=np.uint64)
                    ki = np.argsort(key)
                    seq = txy[ki]
                    bad = np.nonzero(seq[1:] < seq[:-1])[0]
                    if len(bad):
                        st["fail"] += 1; st["fail_by_jj"][jj] += 1; ufail[j] = True
                        # does any edge fall inside a cluster of equal quantised value that is out of order?
                        f = (key[ki] >> np.uint64(8))
                        cw = np.concatenate(([0.0], np.cumsum(wxy[ki])))
                        hit = False
                        for p in bad:
                            lo = p
                            while lo > 0 and f[lo - 1] == f[p]: lo -= 1
                            hi = p + 1
                            while hi < 255 and f[hi + 1] == f[p]: hi += 1
                            if np.any((E[1:-1] > cw[lo]) & (E[1:-1] < cw[hi + 1])): hit = True
                        st["nswap"].append(len(bad))
                        if hit: st["fail_edge"] += 1; ufail_e[j] = True
                    if prev_perm is not None:
                        # displacement of the exact order when the previous layer's permutation is applied
                        pos_prev = np.empty(256, int); pos_prev[prev_perm] = np.arange(256)
                        st["disp"].append(np.abs(pos_prev[inds] - np.arange(256)).max())
                prev_perm = inds
                e = np.concatenate(([0.0], np.cumsum(wxy[inds])))
                Fc = np.concatenate(([0.0], np.cumsum(txy[inds] * np.diff(e))))
                new[j] = np.diff(np.interp(E, e, Fc)) / np.diff(E)
            a = new
        st["units"] += nz; st["unit_fail"] += int(ufail.sum()); st["unit_fail_edge"] += int(ufail_e.sum())

m = st["merges"]
print("merges", m, "closed", st["closed"] / m, "cert fail (of all merges)", st["fail"] / m, "fail with edge in bad cluster", st["fail_edge"] / m)
print("units", st["units"], "unit fail", st["unit_fail"] / st["units"], "unit fail (edge-aware)", st["unit_fail_edge"] / st["units"])
print("fail by gas", st["fail_by_jj"][:5] / np.maximum(1, st["merges_by_jj"][:5]))
d = np.array(st["disp"]); print("perm displacement vs previous layer: median", np.median(d), "p90", np.percentile(d, 90), "p99", np.percentile(d, 99), "max", d.max())
ns = np.array(st["nswap"]); print("descents per failing merge: median", np.median(ns), "p90", np.percentile(ns, 90), "max", ns.max())
