#!/usr/bin/env python
"""Design aid (CPU only): sort-free RORR rebin by "staircase cuts" carried along a chain of units.

The rebin (futils rebin after mrgrnk, clima_radtran_types.f90:840-849) only needs the running
integral I(E) = int_0^E tau(c) dc of the sorted sequence at the G-1 interior bin edges.  For edge E
that is   I(E) = S(cut) + tau_e * (E - W(cut))   where `cut` is the set of elements sorted before
the element e that contains E.  In the matrix a_i + b_j (a, b ascending) a cut is a staircase
c_i = number of elements of row i inside the cut, and
    W(cut) = sum_i w_i Wb[c_i],  S(cut) = sum_i w_i (a_i Wb[c_i] + Sb[c_i]),
    Wb[c] = sum_{j<c} w_j,  Sb[c] = sum_{j<c} w_j b_j.
A cut is the right one iff  L = max_i a_i + b[c_i - 1]  <=  R = min_i a_i + b[c_i]  (it is a
down-set of the sorted order) and W <= E < W + w_e  (e = the argmin of R).
The cut of the previous unit of a chain (previous column of a sweep / previous layer of the column) is
the initial guess; single-element moves repair it.  This script measures how many passes over the
16 rows that takes, per warp (2 units x 15 edges, worst lane), and the error against sort + rebin.

usage: experiments/rorr_cut_proto.py [cfg2|cfg3|cfg5] [col|slot] [nchains] [chainlen] [bin_stride]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth

cfg = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
mode = sys.argv[2] if len(sys.argv) > 2 else "col"
nchains = int(sys.argv[3]) if len(sys.argv) > 3 else 2
clen = int(sys.argv[4]) if len(sys.argv) > 4 else 16
bstride = int(sys.argv[5]) if len(sys.argv) > 5 else 10
tb = synth.make_tables()
kw = dict(cfg2={}, cfg3=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0),
          cfg5=dict(randomize=True, seed=11))[cfg]
NC = 4096
cols = synth.make_columns(NC, nz=100, doubled=False, **kw)
G, nw = tb.ngauss, tb.nw
w = np.asarray(tb.weights); E = np.concatenate(([0.0], np.cumsum(w)))
wxy = (w[:, None] * w[None, :]).reshape(-1)
Wb = np.concatenate(([0.0], np.cumsum(w)))


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


def rebin_ref(txy_sorted, wsorted):
    e = np.concatenate(([0.0], np.cumsum(wsorted)))
    Fc = np.concatenate(([0.0], np.cumsum(txy_sorted * np.diff(e))))
    return np.diff(np.interp(E, e, Fc)) / np.diff(E)


def merges_of(c, l):
    """per gas jj: a (nz,G), b (nz,G), reference rebinned result (nz,G)"""
    ks = kvals(c, l)
    a = ks[0]
    res = []
    for jj in range(1, len(ks)):
        b = ks[jj]
        txy = (a[:, :, None] + b[:, None, :]).reshape(a.shape[0], -1)
        perm = np.argsort(txy, axis=1, kind="stable")
        srt = np.take_along_axis(txy, perm, 1)
        new = np.empty_like(a)
        for j in range(a.shape[0]):
            new[j] = rebin_ref(srt[j], wxy[perm[j]])
        res.append((a.copy(), b.copy(), new.copy()))
        a = new
    return res


NEG, POS = -np.inf, np.inf
TOL = 2.0 ** -48  # edges that coincide with an element boundary: W and E differ by rounding only


def cut_pass(a, b, cnt):
    """one pass over the rows.  a, b: (n,G); cnt: (n, G-1 edges, G rows) ints.
    returns L, iL, R, iR, W, S  each (n, G-1)"""
    n = a.shape[0]
    bp = np.concatenate((np.full((n, 1), NEG), b, np.full((n, 1), POS)), axis=1)  # bp[c] = b[c-1]
    Sb = np.concatenate((np.zeros((n, 1)), np.cumsum(w[None, :] * b, axis=1)), axis=1)
    ar = np.arange(n)[:, None, None]
    lo = a[:, None, :] + bp[ar, cnt]        # last element inside, per row
    hi = a[:, None, :] + bp[ar, cnt + 1]    # first element outside
    iL = lo.argmax(2); iR = hi.argmin(2)
    L = np.take_along_axis(lo, iL[..., None], 2)[..., 0]
    R = np.take_along_axis(hi, iR[..., None], 2)[..., 0]
    W = (w[None, None, :] * Wb[cnt]).sum(2)
    S = (w[None, None, :] * (a[:, None, :] * Wb[cnt] + Sb[ar, cnt])).sum(2)
    return L, iL, R, iR, W, S


def solve_cuts(a, b, cnt, maxit=400):
    """repairs cnt in place; returns number of passes per (unit, edge) and I(E_g)"""
    n = a.shape[0]
    Eg = E[1:G][None, :]
    passes = np.zeros((n, G - 1), int)
    done = np.zeros((n, G - 1), bool)
    Iout = np.zeros((n, G - 1))
    for it in range(maxit):
        L, iL, R, iR, W, S = cut_pass(a, b, cnt)
        cR = np.take_along_axis(cnt, iR[..., None], 2)[..., 0]
        we = w[iR] * w[np.minimum(cR, G - 1)]
        valid = (L <= R) & (W <= Eg + TOL) & (Eg < W + we + TOL)
        newly = valid & ~done
        Iout[newly] = (S + R * (Eg - W))[newly]
        passes[~done] += 1
        done |= valid
        if done.all():
            break
        act = ~done
        rm = act & ((L > R) | (W > Eg + TOL))      # remove the largest inside element
        ad = act & ((L > R) | ((W <= Eg + TOL) & (Eg >= W + we + TOL)))  # add the smallest outside element
        ii, ee = np.nonzero(rm)
        cnt[ii, ee, iL[ii, ee]] -= 1
        ii, ee = np.nonzero(ad)
        cnt[ii, ee, iR[ii, ee]] += 1
    return passes, Iout, done


def cold_cut(a, b):
    """exact cuts from a sort (what a chain head would get from bisection)"""
    n = a.shape[0]
    txy = (a[:, :, None] + b[:, None, :]).reshape(n, -1)
    perm = np.argsort(txy, axis=1, kind="stable")
    cw = np.cumsum(wxy[perm], axis=1)
    cnt = np.zeros((n, G - 1, G), int)
    for u in range(n):
        for e in range(G - 1):
            p = np.searchsorted(cw[u], E[e + 1], side="right")  # elements wholly before the edge
            rows = perm[u, :p] // G
            cnt[u, e] = np.bincount(rows, minlength=G)
    return cnt


bins = list(range(0, nw, bstride))
allpass = []; warppass = []; errs = []
rng = np.random.default_rng(1)
if mode == "col":
    starts = np.linspace(0, NC - clen - 1, nchains).astype(int)
    layer_pick = [0, 20, 50, 80, 99]
    for c0 in starts:
        # units of one chain step: (layer in layer_pick) x bins x merges, vectorised
        prev = None
        for c in range(c0, c0 + clen):
            A = []; B = []; N = []
            for l in bins:
                for (a, b, new) in merges_of(c, l):
                    A.append(a[layer_pick]); B.append(b[layer_pick]); N.append(new[layer_pick])
            A = np.concatenate(A); B = np.concatenate(B); N = np.concatenate(N)
            if prev is None:
                cnt = cold_cut(A, B)
                p, I, d = solve_cuts(A, B, cnt)
                assert d.all()
            else:
                cnt = prev
                p, I, d = solve_cuts(A, B, cnt)
                assert d.all()
                allpass.append(p.ravel()); warppass.append(p.reshape(-1, 2 * (G - 1)).max(1) if p.shape[0] % 2 == 0 else p.max(1))
            Ifull = np.concatenate((np.zeros((A.shape[0], 1)), I, (w[None, :] * (A + (w[None, :] * B).sum(1, keepdims=True))).sum(1, keepdims=True)), axis=1)
            new = np.diff(Ifull, axis=1) / np.diff(E)[None, :]
            errs.append(np.abs(new / N - 1.0).max())
            prev = cnt
else:
    pickc = np.linspace(1, NC - 1, nchains).astype(int)
    for c in pickc:
        per_bin = [merges_of(c, l) for l in bins]
        nz = per_bin[0][0][0].shape[0]
        prev = None
        for s in range(nz):
            A = np.array([m[0][s] for mm in per_bin for m in mm]); B = np.array([m[1][s] for mm in per_bin for m in mm])
            N = np.array([m[2][s] for mm in per_bin for m in mm])
            if prev is None:
                cnt = cold_cut(A, B)
                p, I, d = solve_cuts(A, B, cnt)
                assert d.all()
            else:
                cnt = prev
                p, I, d = solve_cuts(A, B, cnt)
                assert d.all()
                allpass.append(p.ravel()); warppass.append(p.reshape(-1, 2 * (G - 1)).max(1))
            Ifull = np.concatenate((np.zeros((A.shape[0], 1)), I, (w[None, :] * (A + (w[None, :] * B).sum(1, keepdims=True))).sum(1, keepdims=True)), axis=1)
            new = np.diff(Ifull, axis=1) / np.diff(E)[None, :]
            errs.append(np.abs(new / N - 1.0).max())
            prev = cnt

ap = np.concatenate(allpass); wp = np.concatenate(warppass)
print(cfg, mode, "chain steps x edges:", ap.size, " max rel err vs sort+rebin:", max(errs))
print(" passes per edge : mean %.3f  hist 1..8+ %s" % (ap.mean(), np.round(np.bincount(np.minimum(ap, 9), minlength=10)[1:] / ap.size, 3)))
print(" passes per warp (2 units x 15 edges, worst lane): mean %.3f p50 %d p90 %d p99 %d max %d" % (wp.mean(), np.median(wp), np.percentile(wp, 90), np.percentile(wp, 99), wp.max()))
print("   hist 1..12+", np.round(np.bincount(np.minimum(wp, 13), minlength=14)[1:] / wp.size, 3))
