"""CPU oracle for row N4: `AdiabatClimate.make_profile`, restated in numpy.

TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in clima_b200/).

PARITY: PINNED FOR make_profile, UNPINNED FOR compute_altitude.  The reference cannot be compiled in this image
(Fortran + fypp + CPM-fetched dependencies, DESIGN.md "Oracle") and its own test of this path (tests/test_adiabat.f90)
holds no numbers -- but its tutorial notebook (examples/Tutorial_AdiabatClimate.ipynb, cells 6, 10, 11, 14) stores what
the real code printed for `c.make_profile(280 K, [260, 10, 1, 1e-10 ...] bar)`, P_top = 1e-2, T_trop = 150:
P_surf = 11.010013835789936 bar (reproduced to all 17 digits), N_surface(H2O) = 14696.15911658 mol/cm^2 (all 13
digits) and N_atmos of the seven gases to 9 digits (reproduced to 2e-9 at rtol 1e-12 and to 1e-7 at the default rtol,
with dz taken from make_profile's own z(P), as the notebook's version of the reference did).  N_atmos integrates T(P),
z(P), the mixing ratios, the CO2 saturation event and the tropopause over the whole column.  Fixture and generator:
tests/golden/tutorial_adiabat_outputs.json, tests/golden/make_tutorial_outputs.py; test:
tests/test_adiabat.py::test_oracle_against_the_outputs_the_reference_printed.  `compute_altitude` (added to the reference
after that notebook was run; it moves N_atmos by 3e-4) has no published numbers: it is checked against closed forms only.
Further checks of the restatement: closed forms (dry adiabat with constant cp, saturated mixing ratios, isothermal
hydrostatic altitude) and an independent scipy integration of the same equations with its own step control and event
location (1e-9).

What is restated, with the reference lines each function follows:
  * Dop853          -- the integrator is a third-party dependency that is NOT in /root/reference:
                       jacobwilliams/dop853 1.2.0 (src/dependencies/CMakeLists.txt:62-73), a modern-Fortran
                       edition of Hairer & Wanner's DOP853 (Dormand-Prince 8(5,3), dense output of order 7).
                       Restated from the published algorithm (Hairer, Norsett, Wanner, "Solving ODEs I",
                       II.5 and the dop853.f driver: step-size rule with safe=0.9, fac1=0.333, fac2=6,
                       beta=0, hinit, error = 5th-order estimate damped by the 3rd-order one); the
                       Butcher / dense-output coefficient tables are read from scipy's copy of the same
                       published tables.  Its stiffness test (every 1000 accepted steps) is left out:
                       these integrations take < 200 steps.
  * zeroin          -- futils `brent_class%find_zero` (futils 0.1.14, not in the tree either): Brent's
                       zeroin as published (Brent 1973, ch. 4; Forsythe, Malcolm, Moler 1977).
  * make_profile    -- src/adiabat/clima_adiabat_general.f90:89-272 (+ integrate :274-353, solout_dop
                       :355-449, find_root :451-481, root_fcn :483-513, update_f_i_dry :529-546,
                       mixing_ratios :548-574, right_hand_side :576-656, altitude_vs_P :659-669)
  * SaturationData  -- src/clima_saturationdata.f90:56-166
  * heat capacity   -- src/clima_eqns.f90:82-133;  gravity -- :201-211
  * compute_altitude-- src/adiabat/clima_adiabat_altitude.f90:16-243 (reference_pressure <= 0 branch),
                       solve_ivp_dop853 src/clima_useful.f90:97-236, the 1-D linear interpolant
                       src/dependencies/linear_interpolation_module.F90:236-262
  * adiabat_column  -- AdiabatClimate_make_profile, src/adiabat/clima_adiabat.f90:401-472
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
from scipy.integrate._ivp import dop853_coefficients as _co

RGAS = 8.31446261815324e7      # erg/(mol K)   clima_const.f90:9
K_BOLTZ = 1.380649e-16         # clima_const.f90:10
G_GRAV = 6.67430e-11           # clima_const.f90:12
N_AVO = 6.02214076e23          # clima_const.f90:15
DRY, CONDENSING = 0, 1


# --------------------------------------------------------------------------------------------------
# DOP853 (Hairer & Wanner), the driver loop of dop853.f / jacobwilliams' dop853_module
# --------------------------------------------------------------------------------------------------
class Dop853:
    SAFE, FAC1, FAC2, BETA, NMAX = 0.9, 0.333, 6.0, 0.0, 100000
    UROUND = 2.220446049250313e-16

    def __init__(self, fcn: Callable, n: int, solout: Optional[Callable] = None):
        self.fcn, self.n, self.solout = fcn, n, solout
        self.A, self.C, self.B = _co.A, _co.C, _co.B
        self.E3, self.E5, self.D = _co.E3[:12], _co.E5[:12], _co.D
        self.K = np.zeros((16, n))
        self.F = np.zeros((7, n))
        self.xold = self.hout = 0.0
        self.yold = np.zeros(n)
        self.nfcn = self.nstep = self.naccpt = self.nrejct = 0

    def _hinit(self, x, y, posneg, f0, hmax, atol, rtol):
        sk = atol + rtol * np.abs(y)
        dnf = float(np.sum((f0 / sk) ** 2))
        dny = float(np.sum((y / sk) ** 2))
        h = 1.0e-6 if (dnf <= 1.0e-10 or dny <= 1.0e-10) else math.sqrt(dny / dnf) * 0.01
        h = math.copysign(min(h, hmax), posneg)
        f1 = self.fcn(x + h, y + h * f0)
        der2 = math.sqrt(float(np.sum(((f1 - f0) / sk) ** 2))) / h
        der12 = max(abs(der2), math.sqrt(dnf))
        h1 = max(1.0e-6, abs(h) * 1.0e-3) if der12 <= 1.0e-15 else (0.01 / der12) ** (1.0 / 8.0)
        return math.copysign(min(100.0 * abs(h), h1, hmax), posneg)

    def contd8(self, i: int, xq: float) -> float:
        """Dense output of component i inside the last accepted step [xold, xold + hout]."""
        s = (xq - self.xold) / self.hout
        s1 = 1.0 - s
        F = self.F[:, i]
        conpar = F[3] + s * (F[4] + s1 * (F[5] + s * F[6]))
        return self.yold[i] + s * (F[0] + s1 * (F[1] + s * (F[2] + s1 * conpar)))

    def integrate(self, x: float, y: np.ndarray, xend: float, rtol: float, atol: float):
        """-> (x, y, idid): idid 1 reached xend, 2 stopped by solout, -2 nmax, -3 step too small."""
        n, A, C, K = self.n, self.A, self.C, self.K
        y = np.array(y, dtype=np.float64)
        self.nfcn = self.nstep = self.naccpt = self.nrejct = 0
        facold, expo1 = 1.0e-4, 1.0 / 8.0 - self.BETA * 0.2
        facc1, facc2 = 1.0 / self.FAC1, 1.0 / self.FAC2
        posneg = math.copysign(1.0, xend - x)
        hmax = abs(xend - x)
        last = reject = False
        K[0] = self.fcn(x, y)
        h = self._hinit(x, y, posneg, K[0], hmax, atol, rtol)
        self.nfcn += 2
        self.xold = x
        if self.solout is not None:
            self.hout = 1.0
            if self.solout(self.naccpt + 1, x, x, y) < 0:
                return x, y, 2
        while True:
            if self.nstep > self.NMAX:
                return x, y, -2
            if 0.1 * abs(h) <= abs(x) * self.UROUND:
                return x, y, -3
            if (x + 1.01 * h - xend) * posneg > 0.0:
                h = xend - x
                last = True
            self.nstep += 1
            for s in range(1, 12):
                K[s] = self.fcn(x + C[s] * h, y + h * (A[s, :s] @ K[:s]))
            self.nfcn += 11
            xph = x + h
            k4 = self.B @ K[:12]
            ynew = y + h * k4
            sk = atol + rtol * np.maximum(np.abs(y), np.abs(ynew))
            err = float(np.sum(((self.E5 @ K[:12]) / sk) ** 2))
            err2 = float(np.sum(((self.E3 @ K[:12]) / sk) ** 2))
            deno = err + 0.01 * err2
            if deno <= 0.0:
                deno = 1.0
            err = abs(h) * err * math.sqrt(1.0 / (n * deno))
            fac11 = err ** expo1
            fac = fac11 / facold ** self.BETA
            fac = max(facc2, min(facc1, fac / self.SAFE))
            hnew = h / fac
            if err <= 1.0:
                facold = max(err, 1.0e-4)
                self.naccpt += 1
                K[12] = self.fcn(xph, ynew)
                self.nfcn += 1
                # dense output: three more stages, then the seven polynomial coefficients
                for s in range(13, 16):
                    K[s] = self.fcn(x + C[s] * h, y + h * (A[s, :s] @ K[:s]))
                self.nfcn += 3
                dy = ynew - y
                self.F[0] = dy
                self.F[1] = h * K[0] - dy
                self.F[2] = 2.0 * dy - h * (K[12] + K[0])
                self.F[3:] = h * (self.D @ K)
                self.yold[:] = y
                self.hout = h
                K[0] = K[12]
                y = ynew
                self.xold = x
                x = xph
                if self.solout is not None and self.solout(self.naccpt + 1, self.xold, x, y) < 0:
                    return x, y, 2
                if last:
                    return x, y, 1
                if abs(hnew) > hmax:
                    hnew = posneg * hmax
                if reject:
                    hnew = posneg * min(abs(hnew), abs(h))
                reject = False
            else:
                hnew = h / min(facc1, fac11 / self.SAFE)
                reject = True
                if self.naccpt >= 1:
                    self.nrejct += 1
                last = False
            h = hnew


def zeroin(f: Callable[[float], float], ax: float, bx: float, tol: float):
    """Brent's zeroin on [ax, bx] -> (root, f(root), info); info = -1 when f(ax), f(bx) have one sign."""
    eps = 2.220446049250313e-16
    a, b = ax, bx
    fa, fb = f(a), f(b)
    if fa * fb > 0.0:
        return b, fb, -1
    c, fc = a, fa
    d = e = b - a
    while True:
        if (fb > 0.0 and fc > 0.0) or (fb < 0.0 and fc < 0.0):
            c, fc = a, fa
            d = e = b - a
        if abs(fc) < abs(fb):
            a, b, c = b, c, b
            fa, fb, fc = fb, fc, fb
        tol1 = 2.0 * eps * abs(b) + 0.5 * tol
        xm = 0.5 * (c - b)
        if abs(xm) <= tol1 or fb == 0.0:
            return b, fb, 0
        if abs(e) >= tol1 and abs(fa) > abs(fb):
            s = fb / fa
            if a == c:
                p = 2.0 * xm * s
                q = 1.0 - s
            else:
                q = fa / fc
                r = fb / fc
                p = s * (2.0 * xm * q * (q - r) - (b - a) * (r - 1.0))
                q = (q - 1.0) * (r - 1.0) * (s - 1.0)
            if p > 0.0:
                q = -q
            else:
                p = -p
            if 2.0 * p < min(3.0 * xm * q - abs(tol1 * q), abs(e * q)):
                e, d = d, p / q
            else:
                d = e = xm
        else:
            d = e = xm
        a, fa = b, fb
        if abs(d) > tol1:
            b = b + d
        else:
            b = b + math.copysign(tol1, xm)
        fb = f(b)


# --------------------------------------------------------------------------------------------------
# species data: thermodynamics and saturation
# --------------------------------------------------------------------------------------------------
def gravity(radius: float, mass: float, z: float) -> float:
    """clima_eqns.f90:201-211 [cm/s^2]; radius, z in cm, mass in g."""
    grav = G_GRAV * (mass / 1.0e3) / ((radius + z) / 1.0e2) ** 2.0
    return grav * 1.0e2


def heat_capacity(thermo: Dict, T: float) -> Optional[float]:
    """heat_capacity_eval (clima_eqns.f90:105-133): J/(mol K), None when T is outside every range."""
    temps, data = thermo["temps"], thermo["data"]
    for k in range(len(temps) - 1):
        if temps[k] <= T < temps[k + 1]:
            c = data[k]
            if thermo["model"] == "Shomate":
                TT = T / 1000.0
                return c[0] + c[1] * TT + c[2] * TT ** 2 + c[3] * TT ** 3 + c[4] / TT ** 2
            Rgas_J = 8.31446261815324
            return Rgas_J * (c[0] / T ** 2 + c[1] / T + c[2] + c[3] * T + c[4] * T ** 2 + c[5] * T ** 3
                             + c[6] * T ** 4)
    return None


def _integral(A: float, B: float, T: float) -> float:
    return -A / T + B * math.log(T)


def sat_pressure(s: Dict, T: float) -> float:
    """SaturationData%sat_pressure (clima_saturationdata.f90:97-155) [dynes/cm^2]."""
    if T >= s["T_critical"]:
        tmp = (_integral(s["a_v"], s["b_v"], s["T_critical"]) - _integral(s["a_v"], s["b_v"], s["T_ref"])) + \
              (_integral(s["a_c"], s["b_c"], T) - _integral(s["a_c"], s["b_c"], s["T_critical"]))
    elif T > s["T_triple"]:
        tmp = _integral(s["a_v"], s["b_v"], T) - _integral(s["a_v"], s["b_v"], s["T_ref"])
    else:
        tmp = (_integral(s["a_v"], s["b_v"], s["T_triple"]) - _integral(s["a_v"], s["b_v"], s["T_ref"])) + \
              (_integral(s["a_s"], s["b_s"], T) - _integral(s["a_s"], s["b_s"], s["T_triple"]))
    return s["P_ref"] * math.exp((s["mu"] / RGAS) * tmp)


def latent_heat(s: Dict, T: float) -> float:
    """SaturationData%latent_heat (clima_saturationdata.f90:56-91) [erg/g]."""
    if T >= s["T_critical"]:
        return s["a_c"] + s["b_c"] * T
    if T > s["T_triple"]:
        return s["a_v"] + s["b_v"] * T
    return s["a_s"] + s["b_s"] * T


class AdiabatError(Exception):
    pass


# --------------------------------------------------------------------------------------------------
# make_profile (clima_adiabat_general.f90)
# --------------------------------------------------------------------------------------------------
class _Profile:
    """AdiabatProfileData (:18-71) with the procedures that work on it."""

    def __init__(self, sp, planet_mass, planet_radius, P_top, T_trop, RH, T_surf, rtol, atol, nz):
        self.sp, self.ng = sp, len(sp["mass"])
        self.planet_mass, self.planet_radius = planet_mass, planet_radius
        self.P_top, self.T_trop, self.RH, self.T_surf = P_top, T_trop, np.asarray(RH, float), T_surf
        self.rtol, self.atol = rtol, atol
        ne = 2 * nz + 1
        self.P = np.zeros(ne)
        self.z = np.zeros(ne)
        self.T = np.zeros(ne)
        self.f_i = np.zeros((ne, self.ng))
        self.P_trop = -1.0
        self.sp_type = np.zeros(self.ng, dtype=int)
        self.f_i_dry = np.zeros(self.ng)
        self.f_i_cur = np.zeros(self.ng)
        self.gout_old = np.zeros(self.ng + 1)
        self.j = 1
        self.stopping_reason = "ptop"
        self.nfcn = 0

    def psat(self, i, T):
        s = self.sp["sat"][i]
        return math.inf if s is None else self.RH[i] * sat_pressure(s, T)

    def update_f_i_dry(self, P, f_i_layer):                                             # :529-546
        P_i = np.asarray(f_i_layer) * P
        P_dry = float(sum(P_i[i] for i in range(self.ng) if self.sp_type[i] == DRY))
        with np.errstate(divide="ignore", invalid="ignore"):
            self.f_i_dry = P_i / P_dry

    def mixing_ratios(self, P, T):                                                      # :548-574
        f = np.zeros(self.ng)
        f_moist = 0.0
        for i in range(self.ng):
            if self.sp_type[i] == CONDENSING:
                f[i] = min(self.RH[i] * sat_pressure(self.sp["sat"][i], T) / P, 1.0)
                f_moist += f[i]
        f_dry = max(1.0 - f_moist, 1.0e-40)
        for i in range(self.ng):
            if self.sp_type[i] == DRY:
                f[i] = f_dry * self.f_i_dry[i]
        return f, f_dry

    def root_fcn(self, P, T):                                                           # :483-513
        f, _ = self.mixing_ratios(P, T)
        g = np.empty(self.ng + 1)
        g[0] = T - self.T_trop
        for i in range(self.ng):
            g[i + 1] = 1.0 if self.sp_type[i] == CONDENSING else self.psat(i, T) - f[i] * P
        return g

    def rhs(self, P, u):                                                                # :576-656
        T, z = float(u[0]), float(u[1])
        sp, ng = self.sp, self.ng
        Rgas_si = RGAS / 1.0e7
        f, f_dry = self.mixing_ratios(P, T)
        mubar = 0.0
        for i in range(ng):
            mubar += f[i] * sp["mass"][i]
        cp_dry = 2.2250738585072014e-308
        cp_i = np.zeros(ng)
        L_i = np.zeros(ng)
        for i in range(ng):
            if self.sp_type[i] == CONDENSING:
                L_i[i] = latent_heat(sp["sat"][i], T) * sp["mass"][i] * 1.0e-7
            cp = heat_capacity(sp["thermo"][i], T)
            if cp is None:
                raise AdiabatError("Failed to compute heat capacity")
            cp_i[i] = cp
            if self.sp_type[i] == DRY:
                cp_dry += self.f_i_dry[i] * cp
        first = second = 0.0
        for i in range(ng):
            if self.sp_type[i] == CONDENSING:
                beta = L_i[i] / (Rgas_si * T)
                first += f[i] * (cp_i[i] - Rgas_si * beta + Rgas_si * beta ** 2.0)
                second += beta * f[i]
        dlnT_dlnP = 1.0 / (f_dry * ((cp_dry * f_dry + first) / (Rgas_si * (f_dry + second))) + second)
        dT_dP = dlnT_dlnP * (T / P)
        grav = gravity(self.planet_radius, self.planet_mass, z)
        dz_dP = -(RGAS * T) / (grav * P * mubar)
        self.nfcn += 1
        return np.array([dT_dP, dz_dP])

    def solout(self, dop: Dop853, nr, P_old, P_cur, y):                                 # :355-449
        irtrn = 1
        gout = self.root_fcn(P_cur, float(y[0]))
        if nr == 1:
            self.gout_old = gout.copy()
        P_roots = np.zeros(self.ng + 1)
        found = np.zeros(self.ng + 1, dtype=bool)
        for i in range(self.ng + 1):
            go, gn = self.gout_old[i], gout[i]
            if (go < 0 and gn > 0) or (go > 0 and gn < 0):
                found[i] = True
                root, _, info = zeroin(lambda x, i=i: self.root_fcn(x, dop.contd8(0, x))[i], P_old, P_cur, 1.0e-8)
                if info != 0:
                    raise AdiabatError('brent failed in "find_root"')
                P_roots[i] = root
        if found.any():
            irtrn = -1
            ind = int(np.argmax(P_roots))
            self.P_root = float(P_roots[ind])
            self.u_root = np.array([dop.contd8(0, self.P_root), dop.contd8(1, self.P_root)])
            if ind == 0:
                self.stopping_reason = "tropopause"
            else:
                self.stopping_reason = "saturation"
                self.ind_sat = ind - 1
        self.gout_old = gout.copy()
        ne = self.P.shape[0]
        if self.j < ne:
            PP = self.P_root if irtrn == -1 else P_cur
            if self.P[self.j] <= P_old and self.P[self.j] >= PP:
                while self.P[self.j] >= PP:
                    j = self.j
                    self.T[j] = dop.contd8(0, self.P[j])
                    self.z[j] = dop.contd8(1, self.P[j])
                    self.f_i[j], _ = self.mixing_ratios(self.P[j], self.T[j])
                    self.j += 1
                    if self.j >= ne:
                        break
        return irtrn


def altitude_vs_P(P, T, mubar, P0, z0, planet_mass, planet_radius):                     # :659-669
    G_grav_cgs = 6.67e-8
    return ((N_AVO * K_BOLTZ * T) / (G_grav_cgs * planet_mass * mubar) * math.log(P / P0)
            + 1 / (planet_radius + z0)) ** (-1.0) - planet_radius


def make_profile(T_surf: float, P_i_surf: Sequence[float], sp: Dict, nz: int, planet_mass: float,
                 planet_radius: float, P_top: float, T_trop: float, RH: Sequence[float],
                 rtol: float = 1.0e-9, atol: float = 1.0e-12, ocean_fcns: Optional[Sequence] = None) -> Dict:
    """-> dict(P, z, T (2 nz + 1), f_i (2 nz + 1, ng), P_trop, N_surface (ng), N_ocean (ng, ng), nfcn)."""
    P_i_surf = np.asarray(P_i_surf, dtype=np.float64)
    ng = len(sp["mass"])
    if T_surf < T_trop:
        raise AdiabatError('make_profile: Input "T_surf" is less than input "T_trop"')
    if np.any(P_i_surf < 0.0):
        raise AdiabatError('make_profile: Surface pressures (input "P_i_surf") must be positive')
    if P_i_surf.shape[0] != ng:
        raise AdiabatError('make_profile: Input "P_i_surf" has the wrong shape')
    if T_trop < 0.0:
        raise AdiabatError('make_profile: Input "T_trop" is less than 0')
    d = _Profile(sp, planet_mass, planet_radius, P_top, T_trop, RH, T_surf, rtol, atol, nz)
    grav = gravity(planet_radius, planet_mass, 0.0)
    P_i_cur = np.zeros(ng)
    N_surface = np.zeros(ng)
    for i in range(ng):
        P_sat = d.psat(i, T_surf)
        if P_i_surf[i] > P_sat:
            P_i_cur[i] = P_sat
            N_surface[i] = (P_i_surf[i] - P_sat) / (sp["mass"][i] * grav)
            d.sp_type[i] = CONDENSING
        else:
            P_i_cur[i] = P_i_surf[i]
            d.sp_type[i] = DRY
    # gases dissolved in oceans (:226-246): ocean_fcns[j](T_surf, ng, P_i [bar]) -> mol/kg of every gas in ocean j
    N_ocean = np.zeros((ng, ng))
    for j in range(ng):
        if ocean_fcns is not None and ocean_fcns[j] is not None:
            m_i = np.asarray(ocean_fcns[j](T_surf, ng, P_i_cur / 1.0e6), dtype=np.float64)
            for i in range(ng):
                if i != j:
                    N_ocean[i, j] = m_i[i] * N_surface[j] * (sp["mass"][j] / 1.0e3)
    P_surf = 0.0
    for i in range(ng):
        P_surf += P_i_cur[i]
    if P_top > P_surf:
        raise AdiabatError('make_profile: "P_top" is bigger than the surface pressure')
    d.f_i_cur = P_i_cur / P_surf
    d.update_f_i_dry(P_surf, d.f_i_cur)
    ne = 2 * nz + 1
    l0, l1 = math.log10(P_surf), math.log10(P_top)
    d.P = np.array([10.0 ** (l0 + (l1 - l0) * k / (ne - 1)) for k in range(ne)])
    d.P[0], d.P[-1] = P_surf, P_top

    # integrate (:274-353)
    dop = Dop853(d.rhs, 2, None)
    dop.solout = lambda nr, xo, x, y: d.solout(dop, nr, xo, x, y)
    d.f_i[0] = d.f_i_cur
    d.T[0], d.z[0] = T_surf, 0.0
    u, Pn = np.array([T_surf, 0.0]), P_surf
    while True:
        d.stopping_reason = "ptop"
        Pn, u, idid = dop.integrate(Pn, u, P_top, rtol, atol)
        if idid < 0:
            raise AdiabatError("dop853 integration failed: %6d" % idid)
        if d.stopping_reason in ("ptop", "tropopause"):
            break
        Pn, u = d.P_root, d.u_root.copy()
        f_cur, _ = d.mixing_ratios(Pn, float(u[0]))
        d.sp_type[d.ind_sat] = CONDENSING
        d.update_f_i_dry(Pn, f_cur)
    if d.stopping_reason == "tropopause":
        f_cur, _ = d.mixing_ratios(d.P_root, float(d.u_root[0]))
        d.P_trop = d.P_root
        mubar = 0.0
        for i in range(ng):
            mubar += f_cur[i] * sp["mass"][i]
        for i in range(d.j, ne):
            d.T[i] = T_trop
            d.f_i[i] = f_cur
            d.z[i] = altitude_vs_P(d.P[i], T_trop, mubar, d.P_root, float(d.u_root[1]), planet_mass, planet_radius)
    if np.any(d.z[:ne - 1] < 0.0):
        raise AdiabatError('"make_profile" yielded negative altitudes. This may be caused by the lack of a '
                           'hydrostatic solution to the entered atmosphere.')
    return dict(P=d.P, z=d.z, T=d.T, f_i=d.f_i, P_trop=d.P_trop, N_surface=N_surface, N_ocean=N_ocean, nfcn=d.nfcn,
                sp_type=d.sp_type.copy())


# --------------------------------------------------------------------------------------------------
# compute_altitude (clima_adiabat_altitude.f90) and AdiabatClimate_make_profile (clima_adiabat.f90:401-472)
# --------------------------------------------------------------------------------------------------
def _interp_linear(xg: np.ndarray, fg: np.ndarray, x: float) -> float:
    """linear_interp_1d%evaluate (linear_interpolation_module.F90:236-262): the bracketing interval of an
    increasing grid, the end intervals outside it (linear extrapolation)."""
    n = xg.shape[0]
    i = int(np.searchsorted(xg, x, side="right")) - 1
    i = min(max(i, 0), n - 2)
    q = (x - xg[i]) / (xg[i + 1] - xg[i])
    return (1.0 - q) * fg[i] + q * fg[i + 1]


def solve_ivp_dop853(rhs: Callable, x_start: float, y_start: Sequence[float], x_eval: np.ndarray, rtol: float,
                     atol: float) -> np.ndarray:
    """clima_useful.f90:97-236: y at every x_eval (monotone, starting on the x_start side)."""
    n, m = len(y_start), len(x_eval)
    big = max(1.0, abs(x_start), float(np.max(np.abs(x_eval))))
    xtol = max(1.0e-12 * big, 10.0 * float(np.spacing(big)))
    y_eval = np.zeros((n, m))
    st = dict(j=0)
    while st["j"] < m and abs(x_eval[st["j"]] - x_start) <= xtol:
        y_eval[:, st["j"]] = y_start
        st["j"] += 1
    if st["j"] >= m:
        return y_eval
    dop = Dop853(rhs, n, None)

    def solout(nr, xold, x, y):
        x_lo, x_hi = min(xold, x) - xtol, max(xold, x) + xtol
        while st["j"] < m:
            xe = x_eval[st["j"]]
            if xe < x_lo or xe > x_hi:
                break
            y_eval[:, st["j"]] = [dop.contd8(i, xe) for i in range(n)]
            st["j"] += 1
        return 1
    dop.solout = solout
    _, _, idid = dop.integrate(x_start, np.asarray(y_start, float), float(x_eval[-1]), rtol, atol)
    if idid < 0:
        raise AdiabatError("compute_altitude: altitude segment integration failed.")
    if st["j"] < m:
        raise AdiabatError("solve_ivp_dop853: integration finished before sampling all output points.")
    return y_eval


def compute_altitude(P_surf, T_surf, f_i_surf, P, T, f_i, mass, P_top, planet_mass, planet_radius, rtol, atol,
                     reference_pressure=-1.0):
    """-> (z (nz), dz (nz)).  reference_pressure <= 0 (the default, clima_adiabat.f90:64): the planet radius is at the
    surface and z is integrated upward from it.  reference_pressure > 0 (:97-169): the radius is at that pressure,
    z is integrated from there upward and downward, then shifted so that it is again zero at the surface."""
    nz = P.shape[0]
    if nz < 2:
        raise AdiabatError("compute_altitude requires nz >= 2")
    P_e = np.zeros(2 * nz + 1)
    P_e[0] = P_surf
    for i in range(nz):
        P_e[2 * i + 1] = P[i]
        if i < nz - 1:
            P_e[2 * i + 2] = math.sqrt(P[i] * P[i + 1])
    P_e[2 * nz] = P_top
    if np.min(P_e[:-1] - P_e[1:]) <= 10.0 * np.spacing(P_e[0]):
        raise AdiabatError("compute_altitude: pressure grid is not strictly decreasing.")
    P_grid = np.concatenate([[P_surf], P])
    T_grid = np.concatenate([[T_surf], T])
    mub = np.zeros(nz + 1)
    for j in range(len(mass)):
        mub[0] += f_i_surf[j] * mass[j]
        mub[1:] += f_i[:, j] * mass[j]
    xg = np.log10(P_grid[::-1])
    Tg, mg = T_grid[::-1].copy(), mub[::-1].copy()

    def rhs(Pq, u):
        lp = math.log10(Pq)
        Tq, mq = _interp_linear(xg, Tg, lp), _interp_linear(xg, mg, lp)
        grav = gravity(planet_radius, planet_mass, float(u[0]))
        return np.array([-(RGAS * Tq) / (grav * Pq * mq)])
    z_e = np.zeros(2 * nz + 1)
    if reference_pressure <= 0.0:
        z_e[1:2 * nz] = solve_ivp_dop853(rhs, P_e[0], [0.0], P_e[1:2 * nz], rtol, atol)[0]
        z_e[2 * nz] = z_e[2 * nz - 1] + (z_e[2 * nz - 1] - z_e[2 * nz - 2])
    else:
        Pr = reference_pressure
        if Pr < P_top or Pr > P_surf:
            raise AdiabatError("compute_altitude: reference_pressure=%11.4E outside model domain [%11.4E, %11.4E]"
                               % (Pr, P_top, P_surf))
        ne = 2 * nz + 1
        k = next(kk for kk in range(ne - 1) if P_e[kk] >= Pr >= P_e[kk + 1])   # 0-based bracket [k, k + 1]
        if Pr == P_e[k]:
            top, bot = k + 1, k - 1
        elif Pr == P_e[k + 1]:
            top, bot = k + 2, k
        else:
            top, bot = k + 1, k
        if top <= ne - 2:   # upward, to the last edge but one
            z_e[top:ne - 1] = solve_ivp_dop853(rhs, Pr, [0.0], P_e[top:ne - 1], rtol, atol)[0]
        if bot >= 0:        # downward, to the surface
            z_e[bot::-1] = solve_ivp_dop853(rhs, Pr, [0.0], P_e[bot::-1], rtol, atol)[0]
        z_e[:ne - 1] -= z_e[0]
        z_e[ne - 1] = z_e[ne - 2] + (z_e[ne - 2] - z_e[ne - 3])
    z = z_e[1:2 * nz:2].copy()
    dz = z_e[2:2 * nz + 1:2] - z_e[0:2 * nz - 1:2]
    return z, dz


def adiabat_column(T_surf, P_i_surf, sp, nz, planet_mass, planet_radius, P_top=1.0, T_trop=180.0, RH=None,
                   rtol=1.0e-9, atol=1.0e-12, reference_pressure=-1.0) -> Dict:
    """What `AdiabatClimate%make_profile` leaves in the object: P, T (nz), f_i, densities (nz, ng), z, dz,
    P_surf, P_trop, N_surface, N_atmos -- the model-grid column that copy_atm_to_radiative_grid doubles."""
    ng = len(sp["mass"])
    RH = np.ones(ng) if RH is None else np.asarray(RH, float)
    e = make_profile(T_surf, P_i_surf, sp, nz, planet_mass, planet_radius, P_top, T_trop, RH, rtol, atol)
    P = e["P"][1:2 * nz:2].copy()
    T = e["T"][1:2 * nz:2].copy()
    f_i = e["f_i"][1:2 * nz:2].copy()
    z, dz = compute_altitude(e["P"][0], T_surf, e["f_i"][0], P, T, f_i, sp["mass"], P_top, planet_mass,
                             planet_radius, rtol, atol, reference_pressure)
    density = P / (K_BOLTZ * T)
    densities = f_i * density[:, None]
    N_atmos = np.array([np.sum(density * f_i[:, i] * dz) / N_AVO for i in range(ng)])
    return dict(P=P, T=T, f_i=f_i, densities=densities, z=z, dz=dz, P_surf=float(e["P"][0]),
                P_trop=e["P_trop"], N_surface=e["N_surface"], N_atmos=N_atmos, f_i_surf=e["f_i"][0].copy(),
                edges=e)


# --------------------------------------------------------------------------------------------------
# particles: set_particle_density_and_radii + interpolate_particles (clima_adiabat.f90:1022-1044, 1047-1123)
# --------------------------------------------------------------------------------------------------
def interpolate_particles(P_layers, P_set, pdensities, pradii):
    """-> (pdensities, pradii) at the layer pressures, each (nz, np): 10 ** (linear interpolation in log10 P of the log10
    of the values given at P_set (strictly decreasing), the table extended by `huge` / `tiny` with the end values held)."""
    tiny, huge = 2.2250738585072014e-308, 1.7976931348623157e308
    P_set = np.asarray(P_set, float)
    x = np.log10(np.concatenate([[huge], P_set, [tiny]])[::-1])
    if np.any(np.diff(x) <= 0.0):
        raise AdiabatError("Interpolation initialization in `set_particle_density_and_radii` failed")
    out = []
    for vals in (np.asarray(pdensities, float), np.asarray(pradii, float)):
        res = np.empty((len(P_layers), vals.shape[1]))
        for k in range(vals.shape[1]):
            tmp = np.concatenate([[vals[0, k]], vals[:, k], [vals[-1, k]]])
            tmp = np.log10(np.maximum(tmp[::-1], tiny))
            for j, Pj in enumerate(P_layers):
                res[j, k] = 10.0 ** _interp_linear(x, tmp, math.log10(Pj))
        out.append(res)
    return out[0], out[1]
