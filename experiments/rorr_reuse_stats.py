#!/usr/bin/env python
"""Design aid (CPU only): how often could a RORR merge skip its sort?
For cfg2 / cfg3 / cfg5 columns, per merge (layer, bin, gas):
  closed   : rows (or columns) of a_i+b_j do not interleave at all
  lvl      : per bitonic level, rows separated at that level (row orientation / best orientation)
  prevlay  : the exact stable permutation of the previous (unique) layer of the same column, bin, gas
             is still a valid sorted order here (sequence non-decreasing)
  prevcol  : same with the same layer of the previous column
  inv_*    : number of adjacent descents when that permutation is applied (how far from sorted)
usage: experiments/rorr_reuse_stats.py [cfg2|cfg3|cfg5] [ncols_sample] [bin_stride]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
ncs = int(sys.argv[2]) if len(sys.argv) > 2 else 4
bstride = int(sys.argv[3]) if len(sys.argv) > 3 else 5
tb = synth.make_tables()
kw = dict(cfg2={}, cfg3=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0),
          cfg5=dict(randomize=True, seed=11))[cfg]
NC = 4096
cols = synth.make_columns(NC, nz=100, doubled=False, **kw)
G, nw = tb.ngauss, tb.nw
w = np.asarray(tb.weights); E = np.concatenate(([0.0], np.cumsum(w)))
wxy = (w[:, None] * w[None, :]).reshape(-1)


def bracket(xt, x):
    return np.clip(np.searchsorted(xt, x, side="right") - 1, 0, len(xt) - 2)


def kvals(c, l):
    T, P, dens, dz = cols["T"][:, c], cols["P"][:, c], cols["densities"][:, :, c], cols["dz"][:, c]
    log10P = np.log10(P); colsd = dens * dz[:, None]
    out = []
    for k in tb.k:
        PP = np.clip(log10P, k.log10P[0], k.log10P[-1]); TT = np.clip(T, k.temp[0], k.temp[-1])
        ix, iy = bracket(k.log10P, PP), bracket(k.temp, TT)
        q1 = (PP - k.log10P[ix]) / (k.log10P[ix + 1] - k.log10P[ix]); q2 = (TT - k.temp[iy]) / (k.temp[iy + 1] - k.temp[iy])
        F = k.log10k[:, :, :, l]
        fx1 = (1 - q1)[:, None] * F[:, ix, iy].T + q1[:, None] * F[:, ix + 1, iy].T
        fx2 = (1 - q1)[:, None] * F[:, ix, iy + 1].T + q1[:, None] * F[:, ix + 1, iy + 1].T
        out.append(10.0 ** ((1 - q2)[:, None] * fx1 + q2[:, None] * fx2) * colsd[:, k.sp_ind - 1][:, None])
    return out  # list of (nz, G)


def rebin(txy_sorted, wsorted):
    e = np.concatenate(([0.0], np.cumsum(wsorted)))
    Fc = np.concatenate(([0.0], np.cumsum(txy_sorted * np.diff(e))))
    return np.diff(np.interp(E, e, Fc)) / np.diff(E)


def merges_of(c, l):
    """yields per gas jj the arrays a (nz,G), b (nz,G) and the exact perms (nz,256)"""
    ks = kvals(c, l)
    a = ks[0]
    res = []
    for jj in range(1, len(ks)):
        b = ks[jj]
        txy = (a[:, :, None] + b[:, None, :]).reshape(a.shape[0], -1)
        perm = np.argsort(txy, axis=1, kind="stable")
        res.append((a.copy(), b.copy(), txy, perm))
        srt = np.take_along_axis(txy, perm, 1)
        new = np.empty_like(a)
        for j in range(a.shape[0]):
            new[j] = rebin(srt[j], wxy[perm[j]])
        a = new
    return res


stat = dict(n=0, closed=0, prevlay=0, prevcol=0, prev_any=0)
lvl_row = np.zeros(4, int); lvl_best = np.zeros(4, int)
inv_lay = []; inv_col = []; inv_rowmaj = []
nsortlev = []
pick = np.linspace(1, NC - 1, ncs).astype(int)
for c in pick:
    for l in range(0, nw, bstride):
        m1 = merges_of(c, l)
        m0 = merges_of(c - 1, l)
        for (a, b, txy, perm), (_, _, _, perm0) in zip(m1, m0):
            nz = a.shape[0]
            D = (b[:, -1] - b[:, 0])[:, None]; Da = (a[:, -1] - a[:, 0])[:, None]
            sep_r = (a[:, 1:] - a[:, :-1]) >= D   # (nz,15): rows i,i+1 separated
            sep_c = (b[:, 1:] - b[:, :-1]) >= Da
            closed = sep_r.all(1) | sep_c.all(1)
            stat["n"] += nz; stat["closed"] += int(closed.sum())
            # bitonic levels: level 5 merges rows (2m,2m+1): needs sep at i=0,2,4..; level 6 at i=1,5,9,13; 7 at i=3,11; 8 at i=7
            idx = {0: np.arange(0, 15, 2), 1: np.arange(1, 15, 4), 2: np.array([3, 11]), 3: np.array([7])}
            nlev = np.zeros(nz, int)
            for lv in range(4):
                r = sep_r[:, idx[lv]].all(1); cc = sep_c[:, idx[lv]].all(1)
                lvl_row[lv] += int((r & ~closed).sum())
                nlev += (~r).astype(int)
            nsortlev.append(nlev[~closed])
            # orientation chosen per merge: the one with more skippable levels
            # previous layer permutation
            seq = np.take_along_axis(txy[1:], perm[:-1], 1)
            desc = (seq[:, 1:] < seq[:, :-1]).sum(1)
            ok_lay = np.concatenate(([False], desc == 0))
            seqc = np.take_along_axis(txy, perm0, 1)
            descc = (seqc[:, 1:] < seqc[:, :-1]).sum(1)
            ok_col = descc == 0
            nc = ~closed
            stat["prevlay"] += int((ok_lay & nc).sum()); stat["prevcol"] += int((ok_col & nc).sum())
            stat["prev_any"] += int(((ok_lay | ok_col) & nc).sum())
            inv_lay.append(desc[nc[1:]]); inv_col.append(descc[nc])
            rm = (txy[:, 1:] < txy[:, :-1]).sum(1)
            inv_rowmaj.append(rm[nc])

n = stat["n"]; ns = n - stat["closed"]
print(cfg, "merges", n, "closed", round(stat["closed"] / n, 3))
print(" of the sorted merges: prev-layer perm valid", round(stat["prevlay"] / ns, 3), " prev-column perm valid", round(stat["prevcol"] / ns, 3), " either", round(stat["prev_any"] / ns, 3))
print(" levels 5..8 skippable (row orientation) among sorted merges:", np.round(lvl_row / ns, 3))
nl = np.concatenate(nsortlev); print(" levels still needed (row orient.) histogram 0..4:", np.bincount(nl, minlength=5) / len(nl))
for name, v in (("prev layer", inv_lay), ("prev column", inv_col), ("row-major", inv_rowmaj)):
    v = np.concatenate(v)
    print(f" descents with {name} order: median {np.median(v)} p75 {np.percentile(v,75)} p90 {np.percentile(v,90)} max {v.max()}  frac<=2: {np.mean(v<=2):.3f} frac<=8: {np.mean(v<=8):.3f}")

# displacement when the previous layer's permutation seeds the order (sorted merges only)
disp = []
for c in pick[:2]:
    for l in range(0, nw, bstride):
        for (a, b, txy, perm) in merges_of(c, l):
            D = (b[:, -1] - b[:, 0])[:, None]; Da = (a[:, -1] - a[:, 0])[:, None]
            closed = ((a[:, 1:] - a[:, :-1]) >= D).all(1) | ((b[:, 1:] - b[:, :-1]) >= Da).all(1)
            for j in range(1, a.shape[0]):
                if closed[j]:
                    continue
                seq = txy[j][perm[j - 1]]
                # stable sort of the seeded sequence: displacement of each element
                o = np.argsort(seq, kind="stable")
                disp.append(np.abs(o - np.arange(256)).max())
disp = np.array(disp)
print(" max displacement when seeded with the previous layer: hist 0..8+", np.round(np.bincount(np.minimum(disp, 9), minlength=10) / len(disp), 3))
