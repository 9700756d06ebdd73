"""How reproducible is `make_profile` + `compute_altitude` at the reference's default tolerances?

The oracle (oracle/adiabat_np.py) is run on the 12 columns of tests/test_adiabat.py with T_surf perturbed by
relative 0, +-1e-13, 3e-13, 1e-12 and compared with its converged solution (rtol 1e-13).  Measured (max relative
error of z over the layers, one row per column):

     eps = 0      1e-13     -1e-13    3e-13     1e-12
  0  1.28e-08  5.63e-07  3.29e-09  1.96e-08  4.15e-08
  1  1.07e-08  7.98e-08  4.70e-09  1.16e-06  3.58e-07
  4  3.02e-08  7.60e-08  1.03e-06  6.17e-07  2.50e-07
  7  9.63e-07  1.93e-07  3.96e-07  9.60e-07  8.14e-07
 10  4.67e-07  2.34e-06  2.14e-08  5.89e-07  4.57e-07     (other columns alike)

A perturbation in the 13th digit moves the altitude error between 1e-9 and 2e-6: which step of the second
integration (compute_altitude) straddles a kink of the piecewise-linear T(log10 P) is decided by rounding noise,
because z starts at 0 with atol = 1e-12 cm and the first error estimates are noise.  The device build sits in the
same band (6e-6 on z, 6e-5 on dz in the worst of these columns; T agrees to 3e-8 like the oracle's own error).
At BASELINE's sweep size the tail is longer: of 4096 columns (sweep(4096, seed=31), nz = 50) 1.7 % differ between the
two device kernels by more than 1e-4 in dz, up to 8e-4, always at the layers around the tropopause (the sharpest kink of
T(log10 P)); for the worst of them (column 3672) the ORACLE's dz error against its converged solution ranges from 1.8e-6
to 8.0e-4 over 60 perturbations of T_surf by multiples of 1e-13 (median 8.5e-6; second part of this script).  The
reference's compute_altitude is therefore accurate to about 1e-3 in dz near the tropopause at its default tolerances,
and tests/test_adiabat.py bounds z and dz by 2e-5 / 2e-3 there (T: 1e-6).

    python experiments/adiabat_noise.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import adiabat_np as A  # noqa: E402
import test_adiabat as TA  # noqa: E402

sp = TA.species()
T, P = TA.sweep(12)
T[1], P[:, 1] = 280.0, TA.P_I_CO2
for c in range(12):
    conv = A.adiabat_column(T[c], P[:, c], sp, 24, TA.EARTH_MASS, TA.EARTH_RADIUS, T_trop=150.0, rtol=1e-13)
    errs = []
    for eps in (0.0, 1e-13, -1e-13, 3e-13, 1e-12):
        a = A.adiabat_column(T[c] * (1 + eps), P[:, c], sp, 24, TA.EARTH_MASS, TA.EARTH_RADIUS, T_trop=150.0)
        errs.append(float(np.max(np.abs(a["z"] / conv["z"] - 1))))
    print(c, " ".join("%.2e" % e for e in errs))

T, P = TA.sweep(4096, seed=31)
c = 3672
conv = A.adiabat_column(T[c], P[:, c], sp, 50, TA.EARTH_MASS, TA.EARTH_RADIUS, 1.0, 180.0, None, 1e-13, 1e-12)
w = []
for k in range(60):
    o = A.adiabat_column(T[c] * (1 + (k - 30) * 1e-13), P[:, c], sp, 50, TA.EARTH_MASS, TA.EARTH_RADIUS, 1.0, 180.0)
    w.append(float(np.max(np.abs(o["dz"] / conv["dz"] - 1))))
print("column 3672, dz error over 60 perturbations: min %.2e median %.2e max %.2e" % (min(w), float(np.median(w)), max(w)))
