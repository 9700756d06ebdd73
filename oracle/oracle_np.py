"""oracle_np.py -- second, INDEPENDENT CPU restatement of Clima's RT hot path (NumPy/SciPy).

*** TEST INFRASTRUCTURE ONLY *** (same rule as oracle.cpp: only tests/ may import it).

Written from the Fortran directly, not from oracle.cpp, with deliberately different numerical
building blocks so that a transcription error in either restatement shows up as a disagreement:
  * bracket search       : numpy.searchsorted            (oracle.cpp: hand-written bisection)
  * RORR sort            : numpy.argsort(kind="stable")  (oracle.cpp: merge-sort ranking `mrgrnk`)
  * RORR rebin           : difference of the cumulative integral, numpy.interp
                                                         (oracle.cpp: overlap loop `rebin`)
  * tridiagonal system   : scipy.linalg.solve_banded (LAPACK dgbsv, pivoting)
                                                         (oracle.cpp: Thomas `tridiag`)
  * vectorised over layers instead of the reference's scalar loops.
Agreement target between the two: <= 1e-10 relative on every flux (tests/test_oracle.py).

Reference: src/radtran/clima_radtran_types.f90:574-983 (compute_opacity, k_rorr, interpolators),
src/radtran/clima_radtran_twostream.f90:10-316, src/radtran/clima_radtran_radiate.f90:7-196,
src/radtran/clima_radtran.f90:221-342, src/clima_eqns.f90:43-80, src/clima_const.f90:9-21.
PARITY STATUS: "parity unpinned" (see oracle.cpp header / DESIGN.md).
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import solve_banded

# src/clima_const.f90
K_BOLTZ_SI = 1.380649e-23
PLANK = 6.62607004e-34
C_LIGHT = 299792458.0
PI = 3.14159265358979323846
# src/radtran/clima_radtran_types.f90:9-11
MAX_W0, MAX_GT, TAU_MIN = 0.99999, 0.999999, 1.0e-20
TINY = np.finfo(np.float64).tiny


def _bracket(xt, x):
    """linear_interpolation_module.F90:348-350, 0-based left index, vectorised over x."""
    i = np.searchsorted(xt, x, side="right") - 1
    return np.clip(i, 0, len(xt) - 2)


def _interp1(xt, f, x):
    """interp_1d_real (:233-271): linear, extrapolating outside."""
    i = _bracket(xt, x)
    q = (x - xt[i]) / (xt[i + 1] - xt[i])
    return (1.0 - q) * f[i] + q * f[i + 1]


def planck(nu, T):
    """src/clima_eqns.f90:64-73"""
    return 1.0e3 * ((2.0 * PLANK * nu ** 3) / (C_LIGHT ** 2)) * (1.0 / (np.exp((PLANK * nu) / (K_BOLTZ_SI * T)) - 1.0))


def rebin_cumulative(old_e, old_v, new_e):
    """futils `rebin` semantics (integral-conserving bin average) via the cumulative integral."""
    F = np.concatenate(([0.0], np.cumsum(old_v * np.diff(old_e))))
    Fn = np.interp(new_e, old_e, F)
    return np.diff(Fn) / np.diff(new_e)


def opacity(tb, T, P, dens, dz, pdens=None, radii=None, custom=None):
    """OpticalProperties_compute_opacity + k_rorr.  Returns tau(nz,G,nw), w0(nz,G,nw), g(nz,nw),
    tau_band(nz,nw), top-first like OpticalPropertiesResult."""
    nz, nw, G = len(T), tb.nw, tb.ngauss
    wbin = np.asarray(tb.weights, dtype=np.float64)
    wbin_e = np.concatenate(([0.0], np.cumsum(wbin)))
    wxy = (wbin[:, None] * wbin[None, :]).reshape(-1)  # index b + a*G (a = running mixture)
    log10P = np.log10(P)
    cols = dens * dz[:, None]
    tau = np.zeros((nz, G, nw)); w0 = np.zeros((nz, G, nw)); gg = np.zeros((nz, nw)); tband = np.zeros((nz, nw))
    if tb.cont_LH2O > 0:
        iw = tb.cont_LH2O - 1
        foreign_col = cols.sum(axis=1) - cols[:, iw]
        # the reference accumulates species in order skipping H2O; redo in that order
        foreign_col = np.zeros(nz)
        for i in range(cols.shape[1]):
            if i != iw:
                foreign_col = foreign_col + cols[:, i]
    # brackets that do not depend on the bin
    kbr = []
    for k in tb.k:
        PP = np.clip(log10P, k.log10P[0], k.log10P[-1])
        TT = np.clip(T, k.temp[0], k.temp[-1])
        ix, iy = _bracket(k.log10P, PP), _bracket(k.temp, TT)
        q1 = (PP - k.log10P[ix]) / (k.log10P[ix + 1] - k.log10P[ix])
        q2 = (TT - k.temp[iy]) / (k.temp[iy + 1] - k.temp[iy])
        kbr.append((ix, iy, q1, q2))
    for l in range(nw):
        # ---- k-coefficients (nz,G) per gas
        ks = []
        for k, (ix, iy, q1, q2) in zip(tb.k, kbr):
            F = k.log10k[:, :, :, l]  # (G,np,nt)
            f11, f21 = F[:, ix, iy].T, F[:, ix + 1, iy].T
            f12, f22 = F[:, ix, iy + 1].T, F[:, ix + 1, iy + 1].T
            p1, p2 = 1.0 - q1, 1.0 - q2
            fx1 = p1[:, None] * f11 + q1[:, None] * f21
            fx2 = p1[:, None] * f12 + q1[:, None] * f22
            ks.append(np.exp((p2[:, None] * fx1 + q2[:, None] * fx2) * np.log(10.0)))
        # ---- continua (ground-first here; flipped at the end)
        tausg = np.zeros(nz)
        for i, sp in enumerate(tb.ray_sp_ind):
            tausg = tausg + tb.ray_xs[l, i] * cols[:, sp - 1]
        taua = np.zeros(nz)

        def xs_val(x):
            if x.dim == 0:
                return np.full(nz, x.data[l])
            TT = np.clip(T, x.temp[0], x.temp[-1])
            return np.exp(_interp1(x.temp, x.data[:, l], TT) * np.log(10.0))
        for x in tb.cia:
            taua = taua + xs_val(x) * dens[:, x.sp_ind[0] - 1] * dens[:, x.sp_ind[1] - 1] * dz
        for x in tb.pxs:
            taua = taua + xs_val(x) * cols[:, x.sp_ind[0] - 1]
        if tb.cont_LH2O > 0:
            TT = np.clip(T, tb.cont_temp[0], tb.cont_temp[-1])
            H2O = np.exp(_interp1(tb.cont_temp, tb.cont_log10_H2O[:, l], TT) * np.log(10.0))
            fo = np.exp(_interp1(tb.cont_temp, tb.cont_log10_foreign[:, l], TT) * np.log(10.0))
            taua = taua + H2O * dens[:, iw] * cols[:, iw] + fo * dens[:, iw] * foreign_col
        if custom is None:
            tauc = np.full(nz, TINY); w0c = np.full(nz, TINY); g0c = np.full(nz, TINY)
        else:
            c_log10P, c_dtau, c_w0, c_g0 = custom
            lp = np.log10(P * 1.0e6)
            tauc = _interp1(c_log10P, c_dtau[:, l], lp) * dz
            w0c = _interp1(c_log10P, c_w0[:, l], lp)
            g0c = _interp1(c_log10P, c_g0[:, l], lp)
        tausc = w0c * tauc
        taup = np.zeros(nz); tausp = np.zeros(nz); tausp_1 = []; gtp = []
        for p in tb.part:
            r = radii[:, p.p_ind - 1]
            rp = np.clip(r, p.radii[0], p.radii[-1])
            if np.any(rp != r):
                raise RuntimeError("Opacity computation failed in one or more wavelength bins.")
            w0p = _interp1(p.radii, p.w0[:, l], rp)
            qext = _interp1(p.radii, p.qext[:, l], rp)
            gtp.append(_interp1(p.radii, p.gt[:, l], rp))
            t1 = qext * PI * r ** 2.0 * pdens[:, p.p_ind - 1] * dz
            taup = taup + t1
            tausp_1.append(w0p * t1)
            tausp = tausp + tausp_1[-1]
        den = np.maximum(TAU_MIN, tausp + tausg + tausc)
        gt = np.zeros(nz)
        for a, b in zip(gtp, tausp_1):
            gt = gt + a * b / den
        gt = np.minimum(gt + g0c * tausc / den, MAX_GT)
        # ---- RORR
        tau_k = ks[0] * cols[:, tb.k[0].sp_ind - 1][:, None]
        for jj in range(1, len(tb.k)):
            kb = ks[jj] * cols[:, tb.k[jj].sp_ind - 1][:, None]
            txy = (tau_k[:, :, None] + kb[:, None, :]).reshape(nz, G * G)
            new = np.empty((nz, G))
            for j in range(nz):
                inds = np.argsort(txy[j], kind="stable")
                e = np.concatenate(([0.0], np.cumsum(wxy[inds])))
                new[j] = rebin_cumulative(e, txy[j][inds], wbin_e)
            tau_k = new
        tot = ((tausg + taua + taup)[:, None] + tau_k) + tauc[:, None]
        sc = (tausg + tausp + tausc)[:, None]
        w = np.where(tot <= TAU_MIN, 0.0, np.minimum(MAX_W0, sc / np.where(tot <= TAU_MIN, 1.0, tot)))
        tau[:, :, l] = tot[::-1]
        w0[:, :, l] = w[::-1]
        gg[:, l] = gt[::-1]
        tband[:, l] = (tot * wbin[None, :]).sum(axis=1)[::-1]
    return tau, w0, gg, tband


def _solve(e1, e2, e3, e4, cp0, cpb, cm0, cmb, Rsfc, Ssfc):
    """Builds the 2nz tridiagonal system (twostream.f90:93-117 / :251-275) and solves it with a
    banded LAPACK solve."""
    nz = len(e1)
    n = 2 * nz
    A = np.zeros(n); B = np.zeros(n); D = np.zeros(n); E = np.zeros(n)
    B[0], D[0], E[0] = e1[0], -e2[0], -cm0[0]
    i = np.arange(nz - 1)
    lo = 2 * i + 2  # Fortran odd rows l = 2i+1 (i=1..nz-1) -> 0-based 2,4,...
    A[lo] = e2[i] * e3[i] - e4[i] * e1[i]
    B[lo] = e1[i] * e1[i + 1] - e3[i] * e3[i + 1]
    D[lo] = e3[i] * e4[i + 1] - e1[i] * e2[i + 1]
    E[lo] = e3[i] * (cp0[i + 1] - cpb[i]) + e1[i] * (cmb[i] - cm0[i + 1])
    le = 2 * i + 1  # Fortran even rows l = 2i -> 0-based 1,3,...
    A[le] = e2[i + 1] * e1[i] - e3[i] * e4[i + 1]
    B[le] = e2[i] * e2[i + 1] - e4[i] * e4[i + 1]
    D[le] = e1[i + 1] * e4[i + 1] - e2[i + 1] * e3[i + 1]
    E[le] = e2[i + 1] * (cp0[i + 1] - cpb[i]) - e4[i + 1] * (cm0[i + 1] - cmb[i])
    A[n - 1] = e1[-1] - Rsfc * e3[-1]
    B[n - 1] = e2[-1] - Rsfc * e4[-1]
    E[n - 1] = Ssfc - cpb[-1] + Rsfc * cmb[-1]
    ab = np.zeros((3, n))
    ab[0, 1:] = D[:-1]
    ab[1] = B
    ab[2, :-1] = A[1:]
    Y = solve_banded((1, 1), ab, E)
    return Y[0::2], Y[1::2]


def two_stream_ir(tau, w0, gt, emissivity, hard, tau_min, B):
    nz = len(tau)
    Rsfc = 1.0 - emissivity if hard else 0.0
    g1 = 2.0 - w0 * (1.0 + gt)
    g2 = w0 * (1.0 - gt)
    lam = np.sqrt(g1 ** 2 - g2 ** 2)
    cg = g2 / (g1 + lam)
    ex = np.exp(-lam * tau)
    e1, e2, e3, e4 = 1.0 + cg * ex, 1.0 - cg * ex, cg + ex, cg - ex
    thin = tau <= tau_min
    b0 = np.where(thin, 0.5 * (B[:-1] + B[1:]), B[:-1])
    b1 = np.where(thin, 0.0, (B[1:] - B[:-1]) / np.where(thin, 1.0, tau))
    r = 1.0 / (g1 + g2)
    cp0 = PI * (b0 + b1 * r); cpb = PI * (b0 + b1 * (tau + r))
    cm0 = PI * (b0 - b1 * r); cmb = PI * (b0 + b1 * (tau - r))
    if hard:
        Ssfc = emissivity * PI * B[nz]
    else:
        b1b = 0.0 if tau[-1] <= tau_min else (B[nz] - B[nz - 1]) / tau[-1]
        Ssfc = PI * (B[nz] + 0.5 * b1b)
    y1, y2 = _solve(e1, e2, e3, e4, cp0, cpb, cm0, cmb, Rsfc, Ssfc)
    fup = np.empty(nz + 1); fdn = np.empty(nz + 1)
    fup[0] = y1[0] * e3[0] - y2[0] * e4[0] + cp0[0]
    fdn[0] = 0.0
    fup[1:] = y1 * e1 + y2 * e2 + cpb
    fdn[1:] = y1 * e3 + y2 * e4 + cmb
    return fup, fdn


def two_stream_solar(tau_in, w0_in, gt_in, u0, Rsfc):
    nz = len(tau_in)
    s3 = np.sqrt(3.0)
    f = 1.0 - w0_in * gt_in * gt_in
    tau = tau_in * f
    w0 = w0_in * (1.0 - gt_in * gt_in) / f
    gt = gt_in / (1.0 + gt_in)
    g1 = s3 * (2.0 - w0 * (1.0 + gt)) / 2.0
    g2 = s3 * w0 * (1.0 - gt) / 2.0
    g3 = (1.0 - s3 * gt * u0) / 2.0
    g4 = 1.0 - g3
    lam = np.sqrt(g1 ** 2 - g2 ** 2)
    cg = g2 / (g1 + lam)
    ex = np.exp(-lam * tau)
    e1, e2, e3, e4 = 1.0 + cg * ex, 1.0 - cg * ex, cg + ex, cg - ex
    tauc = np.concatenate(([0.0], np.cumsum(tau)))
    facp = w0 * ((g1 - 1.0 / u0) * g3 + g4 * g2)
    facm = w0 * ((g1 + 1.0 / u0) * g4 + g2 * g3)
    et0 = np.exp(-tauc[:-1] / u0)
    etb = et0 * np.exp(-tau / u0)
    den = lam ** 2 - 1.0 / u0 ** 2
    direct = np.concatenate(([u0], u0 * etb))
    cp0, cpb, cm0, cmb = et0 * facp / den, etb * facp / den, et0 * facm / den, etb * facm / den
    Ssfc = Rsfc * direct[nz]
    y1, y2 = _solve(e1, e2, e3, e4, cp0, cpb, cm0, cmb, Rsfc, Ssfc)
    amean = np.empty(nz + 1); fup = np.empty(nz + 1); fdn = np.empty(nz + 1)
    amean[0] = s3 * (y1[0] * e3[0] - y2[0] * e4[0] + cp0[0]) + direct[0] / u0
    amean[1:] = s3 * (y1 * (e1 + e3) + y2 * (e2 + e4) + cpb + cmb) + direct[1:] / u0
    fup[0] = y1[0] * e3[0] - y2[0] * e4[0] + cp0[0]
    fdn[0] = direct[0]
    fup[1:] = y1 * e1 + y2 * e2 + cpb
    fdn[1:] = y1 * e3 + y2 * e4 + cmb + direct[1:]
    return amean, fup, fdn


def radiate(tb, pr, T_surface, T, P, dens, dz, pdens=None, radii=None, custom=None):
    """Radtran_radiate + TOA_fluxes for ONE column (ground-first inputs).  Returns the same keys as
    clima_b200._abi.alloc_out(spectral=True) without the trailing column axis."""
    T = np.asarray(T, dtype=np.float64); nz = len(T)
    tau, w0, g, tband = opacity(tb, T, np.asarray(P), np.asarray(dens), np.asarray(dz), pdens, radii, custom)
    wbin = np.asarray(tb.weights)
    freq = C_LIGHT / (np.asarray(tb.wavl) * 1.0e-9)
    out = {}
    # ---- IR channel
    s, e = tb.ir_ind
    L = nz + 1
    fup_a = np.zeros((L, e - s + 1)); fdn_a = np.zeros((L, e - s + 1)); tb_ir = np.zeros((nz, e - s + 1))
    for ll, l in enumerate(range(s - 1, e)):
        nu = 0.5 * (freq[l] + freq[l + 1])
        B = np.concatenate((planck(nu, T[::-1]), [planck(nu, T_surface)]))
        up = np.zeros(L); dn = np.zeros(L)
        for i in range(tb.ngauss):
            fu, fd = two_stream_ir(tau[:, i, l], w0[:, i, l], g[:, l], pr.surface_emissivity[ll],
                                   pr.has_hard_surface, pr.ir_tau_min, B)
            up += fu * wbin[i]; dn += fd * wbin[i]
        fup_a[:, ll] = up[::-1]; fdn_a[:, ll] = dn[::-1]; tb_ir[:, ll] = tband[::-1, l]
    dfr = freq[s - 1:e] - freq[s:e + 1]
    out.update(fup_a_ir=fup_a, fdn_a_ir=fdn_a, tau_band_ir=tb_ir, fup_n_ir=fup_a @ dfr, fdn_n_ir=fdn_a @ dfr)
    # ---- solar channel
    s, e = tb.sol_ind
    fup_a = np.zeros((L, e - s + 1)); fdn_a = np.zeros((L, e - s + 1)); am_a = np.zeros((L, e - s + 1))
    tb_sol = np.zeros((nz, e - s + 1))
    for ll, l in enumerate(range(s - 1, e)):
        up = np.zeros(L); dn = np.zeros(L); am = np.zeros(L)
        for u0, wz in zip(pr.zenith_u, pr.zenith_weights):
            for i in range(tb.ngauss):
                a, fu, fd = two_stream_solar(tau[:, i, l], w0[:, i, l], g[:, l], u0, pr.surface_albedo[ll])
                up += fu * wbin[i] * wz; dn += fd * wbin[i] * wz; am += a * wbin[i] * wz
        sc = tb.photons_sol[ll] * pr.photon_scale_factor * pr.diurnal_fac
        nu = 0.5 * (freq[l] + freq[l + 1])
        lam = 1.0e9 * C_LIGHT / nu
        fup_a[:, ll] = up[::-1] * sc; fdn_a[:, ll] = dn[::-1] * sc
        am_a[:, ll] = am[::-1] * sc * (nu / lam) * (lam / (PLANK * C_LIGHT * 1.0e16)) * (tb.wavl[l + 1] - tb.wavl[l])
        tb_sol[:, ll] = tband[::-1, l]
    dfr = freq[s - 1:e] - freq[s:e + 1]
    out.update(fup_a_sol=fup_a, fdn_a_sol=fdn_a, amean_sol=am_a, tau_band_sol=tb_sol,
               fup_n_sol=fup_a @ dfr, fdn_n_sol=fdn_a @ dfr)
    out["f_total"] = (out["fdn_n_sol"] - out["fup_n_sol"]) + (out["fdn_n_ir"] - out["fup_n_ir"])
    out["ISR"] = out["fdn_n_sol"][nz] - out["fup_n_sol"][nz]
    out["OLR"] = -(out["fdn_n_ir"][nz] - out["fup_n_ir"][nz])
    out["opacity"] = (tau, w0, g, tband)
    return out
