#!/usr/bin/env python
"""Design aid (CPU): the frozen-opacity Jacobian as a DELTA solve with a shared factorisation.

All columns of a finite-difference Jacobian (src/adiabat/clima_adiabat_solve.f90:798-820) share (tau, w0, g),
hence the 2nz x 2nz tridiagonal matrix of two_stream_ir (clima_radtran_twostream.f90:251-275); the right-hand side is
linear in the Planck values, and a perturbed column differs from the base in 2-4 levels.  With the top-down
factors c'_r and the bottom-up factors a''_r of the SAME matrix,
    x_r = kappa_r (d'_r - c'_r d''_(r+1)),   kappa_r = 1 / (1 - c'_r a''_(r+1)),
where d' is the top-down and d'' the bottom-up elimination of the right-hand side.  For a right-hand side that is
non-zero only in rows r0..r1:  d'_r = 0 above the window and a pure chain d'_r = phi_r d'_(r-1) below it, d''_r = 0
below and d''_r = psi_r d''_(r+1) above.  So the solution everywhere follows from a small window solve and two
chains of one multiply per row -- no per-row storage per column.
This script checks the algebra against the full banded solve on random layers."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import oracle_np as onp

PI = onp.PI
rng = np.random.default_rng(5)
nz = 102
tau = 10.0 ** rng.uniform(-7, 2.5, nz); tau[::7] = 1e-8   # some layers under the thin-layer guard
w0 = rng.uniform(0.0, 0.9999, nz); gt = rng.uniform(0.0, 0.9, nz)
emis, hard, tmin = 0.9, True, 1e-6
Bb = np.sort(rng.uniform(1.0, 50.0, nz + 1))


def coeffs(tau, w0, gt):
    g1 = 2.0 - w0 * (1.0 + gt); g2 = w0 * (1.0 - gt)
    lam = np.sqrt(g1 ** 2 - g2 ** 2); cg = g2 / (g1 + lam); ex = np.exp(-lam * tau)
    return 1.0 + cg * ex, 1.0 - cg * ex, cg + ex, cg - ex, 1.0 / (g1 + g2)


def sources(B, tau, r):
    thin = tau <= tmin
    b0 = np.where(thin, 0.5 * (B[:-1] + B[1:]), B[:-1])
    b1 = np.where(thin, 0.0, (B[1:] - B[:-1]) / np.where(thin, 1.0, tau))
    return PI * (b0 + b1 * r), PI * (b0 + b1 * (tau + r)), PI * (b0 - b1 * r), PI * (b0 + b1 * (tau - r))


def matrix(e1, e2, e3, e4, Rsfc):
    n = 2 * nz
    A = np.zeros(n); Bd = np.zeros(n); D = np.zeros(n)
    Bd[0], D[0] = e1[0], -e2[0]
    i = np.arange(nz - 1)
    A[2 * i + 2] = e2[i] * e3[i] - e4[i] * e1[i]; Bd[2 * i + 2] = e1[i] * e1[i + 1] - e3[i] * e3[i + 1]; D[2 * i + 2] = e3[i] * e4[i + 1] - e1[i] * e2[i + 1]
    A[2 * i + 1] = e2[i + 1] * e1[i] - e3[i] * e4[i + 1]; Bd[2 * i + 1] = e2[i] * e2[i + 1] - e4[i] * e4[i + 1]; D[2 * i + 1] = e1[i + 1] * e4[i + 1] - e2[i + 1] * e3[i + 1]
    A[n - 1] = e1[-1] - Rsfc * e3[-1]; Bd[n - 1] = e2[-1] - Rsfc * e4[-1]
    return A, Bd, D


def rhs(e1, e2, e3, e4, cp0, cpb, cm0, cmb, Rsfc, Ssfc):
    n = 2 * nz
    E = np.zeros(n)
    E[0] = -cm0[0]
    i = np.arange(nz - 1)
    E[2 * i + 2] = e3[i] * (cp0[i + 1] - cpb[i]) + e1[i] * (cmb[i] - cm0[i + 1])
    E[2 * i + 1] = e2[i + 1] * (cp0[i + 1] - cpb[i]) - e4[i + 1] * (cm0[i + 1] - cmb[i])
    E[n - 1] = Ssfc - cpb[-1] + Rsfc * cmb[-1]
    return E


e1, e2, e3, e4, r12 = coeffs(tau, w0, gt)
Rsfc = 1.0 - emis
A, Bd, D = matrix(e1, e2, e3, e4, Rsfc)
n = 2 * nz
# shared factorisations
cp = np.zeros(n); den = np.zeros(n)
den[0] = Bd[0]; cp[0] = D[0] / den[0]
for r in range(1, n):
    den[r] = Bd[r] - A[r] * cp[r - 1]; cp[r] = D[r] / den[r]
app = np.zeros(n + 1); den2 = np.zeros(n)
den2[n - 1] = Bd[n - 1]; app[n - 1] = A[n - 1] / den2[n - 1]
for r in range(n - 2, -1, -1):
    den2[r] = Bd[r] - D[r] * app[r + 1]; app[r] = A[r] / den2[r]
kappa = 1.0 / (1.0 - cp * app[1:])
phi = -A / den; psi = -D / den2


def full_net(B):
    s = sources(B, tau, r12)
    Ssfc = emis * PI * B[nz]
    y1, y2 = onp._solve(e1, e2, e3, e4, *s, Rsfc, Ssfc)
    fup = y1 * e1 + y2 * e2 + s[1]; fdn = y1 * e3 + y2 * e4 + s[3]
    return fdn - fup   # at the bottom of every layer


base = full_net(Bb)
worst = 0.0
for trial in range(40):
    dB = np.zeros(nz + 1)
    j0 = int(rng.integers(0, nz + 1)); w = int(rng.choice([1, 2, 4]))
    dB[j0:min(nz + 1, j0 + w)] = rng.uniform(0.01, 0.5, min(nz + 1, j0 + w) - j0)
    ref = full_net(Bb + dB) - base
    # delta right-hand side (sources are linear in B)
    ds = sources(dB, tau, r12)
    dE = rhs(e1, e2, e3, e4, *ds, Rsfc, emis * PI * dB[nz])
    nzr = np.nonzero(dE)[0]
    r0, r1 = nzr.min(), nzr.max()
    dprime = np.zeros(n); dsec = np.zeros(n + 1)
    for r in range(r0, n):       # window forward, then the pure chain
        dprime[r] = ((dE[r] - A[r] * (dprime[r - 1] if r > 0 else 0.0)) / den[r]) if r <= r1 else phi[r] * dprime[r - 1]
    for r in range(r1, -1, -1):  # window backward, then the pure chain
        dsec[r] = ((dE[r] - D[r] * dsec[r + 1]) / den2[r]) if r >= r0 else psi[r] * dsec[r + 1]
    x = kappa * (dprime - cp * dsec[1:])
    got = x[0::2] * (e3 - e1) + x[1::2] * (e4 - e2) + (ds[3] - ds[1])
    err = np.max(np.abs(got - ref)) / np.max(np.abs(base))
    worst = max(worst, err)
print("window rows at most", "delta solve vs difference of two full solves: max error / flux scale =", worst)
