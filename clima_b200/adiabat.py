"""N3: the batched user-facing entry of `AdiabatClimate`.

`AdiabatClimate.TOA_fluxes(T_surf, P_i_surf)` (src/adiabat/clima_adiabat.f90:776-799) is three steps:
`make_profile` (the pseudo-adiabat on the model grid, :401-472), `copy_atm_to_radiative_grid` (every layer
twice + two ghost copies of the top layer, :729-773) and `rad%TOA_fluxes`.  A sweep
(examples/EarlyMars.ipynb, cell 7) calls it once per (T_surf, P_i_surf) pair.  Here the first step stays a
per-column host call -- the reference's own `make_profile` through its Python package where that exists, any
callable with the same outputs otherwise (batching it is row N4, not built) -- and the other two run ONCE for
the whole batch: the doubled grid is built for all columns and `Radtran.radiate_batch` does the RT on the GPU.
"""
from __future__ import annotations

from typing import Callable, Dict, Sequence, Tuple

import math

import numpy as np

from .radtran import ClimaException, Radtran


def copy_atm_to_radiative_grid(T, P, densities, dz, pdensities=None, pradii=None) -> Dict:
    """`AdiabatClimate_copy_atm_to_radiative_grid` (clima_adiabat.f90:729-773) for one column: model layer i ->
    radiative layers 2i, 2i+1 (dz halved), then two ghost layers that copy radiative layer 2 nz (the upper
    half of the top layer).  Pressures are passed through unchanged (the caller converts units)."""
    nz = len(T)
    nzr = 2 * nz + 2

    def rep(a, half=False):
        a = np.asarray(a, dtype=np.float64)
        out = np.empty((nzr,) + a.shape[1:])
        out[0:2 * nz:2] = a * (0.5 if half else 1.0)
        out[1:2 * nz:2] = a * (0.5 if half else 1.0)
        out[2 * nz] = out[2 * nz - 1]
        out[2 * nz + 1] = out[2 * nz - 1]
        return out
    r = dict(T=rep(T), P=rep(P), densities=rep(densities), dz=rep(dz, half=True))
    if pdensities is not None:
        r.update(pdensities=rep(pdensities), radii=rep(pradii))
    return r


class AdiabatClimateBatch:
    """`make_profile(T_surf, P_i_surf) -> dict(T(nz) [K], P(nz) [dynes/cm^2], densities(nz, ng)
    [molecules/cm^3], dz(nz) [cm])` on the model grid, ground first (the members `AdiabatClimate` holds after
    its own `make_profile`); `rad` the GPU `Radtran` built for nz_r = 2 nz + 2 layers."""

    def __init__(self, rad: Radtran, make_profile: Callable[[float, np.ndarray], Dict]):
        self.rad = rad
        self.make_profile = make_profile

    def columns(self, T_surf: Sequence[float], P_i_surf: np.ndarray) -> Dict:
        """make_profile + copy_atm_to_radiative_grid for every column -> the `cols` dict of radiate_batch."""
        T_surf = np.ascontiguousarray(T_surf, dtype=np.float64)
        P_i_surf = np.asarray(P_i_surf, dtype=np.float64)
        ncol = T_surf.shape[0]
        if P_i_surf.ndim != 2 or P_i_surf.shape[1] != ncol:
            raise ClimaException('"T_surf" and "P_i_surf" have inconsistent dimensions')
        if P_i_surf.shape[0] != self.rad.tables.ng:
            raise ClimaException('"P_i_surf" has the wrong input dimension.')
        parts = []
        for c in range(ncol):
            prof = self.make_profile(float(T_surf[c]), P_i_surf[:, c])
            parts.append(copy_atm_to_radiative_grid(prof["T"], np.asarray(prof["P"]) / 1.0e6, prof["densities"], prof["dz"]))
        nzr = parts[0]["T"].shape[0]
        if any(p["T"].shape[0] != nzr for p in parts):
            raise ClimaException("make_profile returned columns of different sizes")
        cols = dict(nz=nzr, ng=self.rad.tables.ng, np=0, ncol=ncol, T_surface=T_surf)
        for k in ("T", "P", "dz"):
            cols[k] = np.asfortranarray(np.stack([p[k] for p in parts], axis=-1))
        cols["densities"] = np.asfortranarray(np.stack([p["densities"] for p in parts], axis=-1))
        return cols

    def TOA_fluxes_batch(self, T_surf: Sequence[float], P_i_surf: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        """(ISR, OLR) [mW/m^2] for every (T_surf[c], P_i_surf[:, c]): AdiabatClimate.TOA_fluxes per column."""
        r = self.rad.radiate_batch(self.columns(T_surf, P_i_surf))
        return r["ISR"], r["OLR"]

    def TOA_fluxes(self, T_surf: float, P_i_surf: np.ndarray) -> Tuple[float, float]:
        isr, olr = self.TOA_fluxes_batch([T_surf], np.asarray(P_i_surf, dtype=np.float64)[:, None])
        return float(isr[0]), float(olr[0])


# ------------------------------------------------------------------------------------------------------
# N4: make_profile itself on the GPU (include/clima_rt.h `clima_rt_adiabat_set`, `clima_rt_make_profile_batch`,
# `clima_rt_adiabat_toa_fluxes_batch`; kernel clima_b200/csrc/adiabat_profile.cu)
# ------------------------------------------------------------------------------------------------------
import ctypes as C  # noqa: E402
from dataclasses import dataclass, field  # noqa: E402
from typing import List, Optional  # noqa: E402

from ._abi import ERR_LEN, alloc_out, c_double_p, c_int32_p  # noqa: E402
from .radtran import _check, _i  # noqa: E402

MAX_GASES, MAX_TEMPS = 16, 4
SAT_KEYS = ("mu", "T_ref", "P_ref", "T_triple", "T_critical", "a_v", "b_v", "a_s", "b_s", "a_c", "b_c")
PROFILE_STATUS = {1: "T_surf below T_trop", 2: "negative surface pressure", 3: "P_top above the surface pressure",
                  4: "dop853 failed", 5: "brent failed", 6: "heat capacity out of range", 7: "negative altitudes",
                  8: "altitude integration failed", 9: "pressure grid not decreasing", 10: "negative T_trop",
                  11: "reference_pressure outside the model domain"}


def species_from_yaml(path_or_dict, names: Optional[Sequence[str]] = None) -> Dict:
    """The `species:` list of a species.yaml (templates/AdiabatClimate/species.yaml) as the dict this module
    and the oracle use: names, mass [g/mol] (from the `atoms:` masses and each composition, as
    src/clima_types_create.f90 does), thermo (model, temps, data) and sat (None or the LinearLatentHeat
    parameters of src/clima_saturationdata.f90:186-283).  `names` selects and orders the gases."""
    if isinstance(path_or_dict, dict):
        y = path_or_dict
    else:
        import yaml
        with open(path_or_dict) as f:
            y = yaml.safe_load(f)
    atoms = {a["name"]: float(a["mass"]) for a in y["atoms"]}
    by_name = {s["name"]: s for s in y["species"]}
    order = list(names) if names is not None else [s["name"] for s in y["species"]]
    sp = dict(names=[], mass=[], thermo=[], sat=[])
    for n in order:
        if n not in by_name:
            raise ClimaException('species "%s" is not in the species file' % n)
        s = by_name[n]
        sp["names"].append(n)
        sp["mass"].append(sum(atoms[k] * float(v) for k, v in s["composition"].items()))
        t = s["thermo"]
        if t["model"] not in ("Shomate", "Nasa9"):
            raise ClimaException('thermo "model" must be Shomate or Nasa9 for species "%s"' % n)
        sp["thermo"].append(dict(model=t["model"], temps=[float(v) for v in t["temperature-ranges"]],
                                 data=[[float(v) for v in r] for r in t["data"]]))
        q = s.get("saturation")
        if q is None:
            sp["sat"].append(None)
            continue
        if q["model"] != "LinearLatentHeat":
            raise ClimaException('Saturation "model" must be "LinearLatentHeat" for species "%s"' % n)
        p = q["parameters"]
        vals = (p["mu"], p["T-ref"], p["P-ref"], p["T-triple"], p["T-critical"], q["vaporization"]["a"],
                q["vaporization"]["b"], q["sublimation"]["a"], q["sublimation"]["b"], q["super-critical"]["a"],
                q["super-critical"]["b"])
        sp["sat"].append({k: float(v) for k, v in zip(SAT_KEYS, vals)})
    return sp


class CAdiabat(C.Structure):
    _fields_ = [("ng", C.c_int32), ("nz", C.c_int32), ("double_radiative_grid", C.c_int32),
                ("mass", c_double_p), ("thermo_model", c_int32_p), ("thermo_ntemps", c_int32_p),
                ("thermo_temps", c_double_p), ("thermo_data", c_double_p), ("has_sat", c_int32_p),
                ("sat", c_double_p), ("RH", c_double_p),
                ("planet_mass", C.c_double), ("planet_radius", C.c_double), ("P_top", C.c_double),
                ("T_trop", C.c_double), ("rtol", C.c_double), ("atol", C.c_double), ("reference_pressure", C.c_double)]


PROFILE_FIELDS = ("P", "T", "z", "dz", "f_i", "f_i_surf", "N_surface", "P_surf", "P_trop")


class CProfileOut(C.Structure):
    _fields_ = [(k, c_double_p) for k in PROFILE_FIELDS] + [("status", c_int32_p), ("nfcn", c_int32_p)]


class AdiabatClimate:
    """The part of `AdiabatClimate` (src/adiabat/clima_adiabat.f90) on the path make_profile -> TOA_fluxes,
    with the column loop of a sweep inside the call.  Public members keep the reference's names: P_top,
    T_trop, RH, rtol, atol, double_radiative_grid; after `make_profile` P, T, f_i, z, dz, densities, P_surf,
    P_trop, N_surface, N_atmos, convecting_with_below, lapse_rate describe the (last) column."""

    def __init__(self, rad: Radtran, species: Dict, nz: int, planet_mass: float, planet_radius: float,
                 double_radiative_grid: bool = True):
        self.rad, self.sp, self.nz = rad, species, int(nz)
        self.planet_mass, self.planet_radius = float(planet_mass), float(planet_radius)
        self.double_radiative_grid = bool(double_radiative_grid)
        self.ng = len(species["mass"])
        self.P_top, self.T_trop = 1.0, 180.0
        self.RH = np.ones(self.ng)
        self.rtol, self.atol = 1.0e-9, 1.0e-12
        self.reference_pressure = -1.0  # > 0: the planet radius is at this pressure (clima_adiabat.f90:58-64)
        self._sent = None
        self.ocean_fcns = [None] * self.ng  # per ocean-forming species: fcn(T_surf, ng, P_i [bar]) -> mol/kg (ng)

    # -- the adiabat description goes to the device whenever a public member has changed
    def _sync(self):
        key = (self.P_top, self.T_trop, tuple(np.asarray(self.RH, float)), self.rtol, self.atol,
               self.double_radiative_grid, self.nz, self.reference_pressure)
        if key == self._sent:
            return
        sp, ng = self.sp, self.ng
        if ng > MAX_GASES:
            raise ClimaException("more than %d gases" % MAX_GASES)
        if len(self.RH) != ng:
            raise ClimaException('make_profile: Input "RH" has the wrong dimension.')
        mass = np.ascontiguousarray(sp["mass"], dtype=np.float64)
        model = np.zeros(ng, np.int32)
        ntemps = np.zeros(ng, np.int32)
        temps = np.zeros((ng, MAX_TEMPS + 1))
        data = np.zeros((ng, MAX_TEMPS, 9))
        has_sat = np.zeros(ng, np.int32)
        sat = np.zeros((ng, 11))
        for i in range(ng):
            t = sp["thermo"][i]
            model[i] = 1 if t["model"] == "Shomate" else 2
            nt = len(t["temps"]) - 1
            if nt < 1 or nt > MAX_TEMPS:
                raise ClimaException("1 to %d temperature ranges per gas" % MAX_TEMPS)
            ntemps[i] = nt
            temps[i, :nt + 1] = t["temps"]
            for k in range(nt):
                data[i, k, :len(t["data"][k])] = t["data"][k]
            if sp["sat"][i] is not None:
                has_sat[i] = 1
                sat[i] = [sp["sat"][i][k] for k in SAT_KEYS]
        RH = np.ascontiguousarray(self.RH, dtype=np.float64)
        a = CAdiabat(ng, self.nz, int(self.double_radiative_grid), mass.ctypes.data_as(c_double_p),
                     model.ctypes.data_as(c_int32_p), ntemps.ctypes.data_as(c_int32_p),
                     temps.ctypes.data_as(c_double_p), data.ctypes.data_as(c_double_p),
                     has_sat.ctypes.data_as(c_int32_p), sat.ctypes.data_as(c_double_p), RH.ctypes.data_as(c_double_p),
                     self.planet_mass, self.planet_radius, float(self.P_top), float(self.T_trop), float(self.rtol),
                     float(self.atol), float(self.reference_pressure))
        err = C.create_string_buffer(ERR_LEN + 1)
        self.rad._lib.clima_rt_adiabat_set(self.rad._h, C.byref(a), err)
        _check(err)
        self._sent = key

    def _inputs(self, T_surf, P_i_surf):
        T_surf = np.ascontiguousarray(T_surf, dtype=np.float64).reshape(-1)
        P_i = np.asfortranarray(P_i_surf, dtype=np.float64)
        if P_i.ndim == 1:
            P_i = np.asfortranarray(P_i[:, None])
        if P_i.shape[0] != self.ng:
            raise ClimaException("P_i_surf has the wrong dimension")
        if P_i.shape[1] != T_surf.shape[0]:
            raise ClimaException('"T_surf" and "P_i_surf" have inconsistent dimensions')
        return T_surf, P_i

    def _profile_out(self, ncol):
        nz, ng = self.nz, self.ng
        shapes = dict(P=(nz, ncol), T=(nz, ncol), z=(nz, ncol), dz=(nz, ncol), f_i=(nz, ng, ncol),
                      f_i_surf=(ng, ncol), N_surface=(ng, ncol), P_surf=(ncol,), P_trop=(ncol,))
        arrs = {k: np.zeros(s, order="F") for k, s in shapes.items()}
        arrs["status"] = np.zeros(ncol, np.int32)
        arrs["nfcn"] = np.zeros(ncol, np.int32)
        o = CProfileOut()
        for k in PROFILE_FIELDS:
            setattr(o, k, arrs[k].ctypes.data_as(c_double_p))
        o.status = arrs["status"].ctypes.data_as(c_int32_p)
        o.nfcn = arrs["nfcn"].ctypes.data_as(c_int32_p)
        return o, arrs

    def _derived(self, p: Dict) -> Dict:
        """densities, N_atmos, convecting_with_below, lapse_rate (clima_adiabat.f90:446-470) on the host."""
        k_boltz, N_avo = 1.380649e-16, 6.02214076e23
        density = p["P"] / (k_boltz * p["T"])
        p["densities"] = p["f_i"] * density[:, None, :]
        p["N_atmos"] = np.einsum("zc,zgc,zc->gc", density, p["f_i"], p["dz"]) / N_avo
        p["convecting_with_below"] = p["P"] > p["P_trop"][None, :]
        lnT = np.log(np.vstack([self._T_surf[None, :], p["T"]]))
        lnP = np.log(np.vstack([p["P_surf"][None, :], p["P"]]))
        p["lapse_rate"] = np.diff(lnT, axis=0) / np.diff(lnP, axis=0)
        return p

    def _column_T_trop(self, T_trop):
        """clima_rt_adiabat_set_column_T_trop: one tropopause temperature per column for the next batch calls
        (None: back to self.T_trop)."""
        err = C.create_string_buffer(ERR_LEN + 1)
        if T_trop is None:
            self.rad._lib.clima_rt_adiabat_set_column_T_trop(self.rad._h, _i(0), None, err)
        else:
            t = np.ascontiguousarray(T_trop, dtype=np.float64).reshape(-1)
            self.rad._lib.clima_rt_adiabat_set_column_T_trop(self.rad._h, _i(t.shape[0]), t.ctypes.data_as(c_double_p), err)
        _check(err)

    def make_profile_batch(self, T_surf, P_i_surf, check: bool = True, T_trop=None) -> Dict:
        """`make_profile` for every (T_surf[c], P_i_surf[:, c]) -> dict of (nz, ncol) / (nz, ng, ncol) arrays
        (+ status, nfcn per column).  check=False returns the per-column status instead of raising."""
        self._sync()
        T_surf, P_i = self._inputs(T_surf, P_i_surf)
        ncol = T_surf.shape[0]
        o, arrs = self._profile_out(ncol)
        err = C.create_string_buffer(ERR_LEN + 1)
        if T_trop is not None:   # a tropopause temperature per column (the solvers with solve_for_T_trop)
            self._column_T_trop(T_trop)
        try:
            self.rad._lib.clima_rt_make_profile_batch(self.rad._h, _i(ncol), T_surf.ctypes.data_as(c_double_p),
                                                      P_i.ctypes.data_as(c_double_p), C.byref(o), err)
        finally:
            if T_trop is not None:
                self._column_T_trop(None)
        if check or not arrs["status"].any():
            _check(err)
        self._T_surf = T_surf
        with np.errstate(all="ignore"):
            return self._oceans(self._derived(arrs))

    def make_profile(self, T_surf: float, P_i_surf):
        """One column; leaves the reference's member names on the object."""
        p = self.make_profile_batch([T_surf], np.asarray(P_i_surf, dtype=np.float64)[:, None])
        self.T_surf = float(T_surf)
        for k in ("P", "T", "z", "dz", "lapse_rate", "convecting_with_below"):
            setattr(self, k, p[k][:, 0].copy())
        for k in ("f_i", "densities"):
            setattr(self, k, p[k][:, :, 0].copy())
        for k in ("f_i_surf", "N_surface", "N_atmos"):
            setattr(self, k, p[k][:, 0].copy())
        self.N_ocean = p["N_ocean"][:, :, 0].copy()
        self.P_surf, self.P_trop = float(p["P_surf"][0]), float(p["P_trop"][0])

    def TOA_fluxes_batch(self, T_surf, P_i_surf, full: bool = False, T_trop=None):
        """(ISR, OLR) of every column: make_profile, the radiative grid and the RT in ONE device call
        (`AdiabatClimate%TOA_fluxes` per column in the reference).  full=True returns the whole result dict
        of `Radtran.radiate_batch` plus the profile arrays."""
        self._sync()
        T_surf, P_i = self._inputs(T_surf, P_i_surf)
        ncol = T_surf.shape[0]
        nzr = 2 * self.nz + 2 if self.double_radiative_grid else self.nz
        o, arrs = alloc_out(nzr, ncol, self.rad.tables.nw_ir, self.rad.tables.nw_sol, False)
        po, parrs = self._profile_out(ncol) if full else (None, None)
        err = C.create_string_buffer(ERR_LEN + 1)
        if T_trop is not None:
            self._column_T_trop(T_trop)
        try:
            self.rad._lib.clima_rt_adiabat_toa_fluxes_batch(self.rad._h, _i(ncol), T_surf.ctypes.data_as(c_double_p),
                                                            P_i.ctypes.data_as(c_double_p),
                                                            C.byref(po) if full else None, C.byref(o), err)
        finally:
            if T_trop is not None:
                self._column_T_trop(None)
        _check(err)
        if full:
            self._T_surf = T_surf
            arrs["profile"] = self._oceans(self._derived(parrs))
            return arrs
        return arrs["ISR"], arrs["OLR"]

    def TOA_fluxes(self, T_surf: float, P_i_surf):
        isr, olr = self.TOA_fluxes_batch([T_surf], np.asarray(P_i_surf, dtype=np.float64)[:, None])
        return float(isr[0]), float(olr[0])

    def set_ocean_solubility_fcn(self, species: str, fcn):
        """`AdiabatClimate%set_ocean_solubility_fcn` (clima_adiabat.f90:1125-1150): `fcn(T_surf, ng, P_i)` with P_i the
        surface partial pressures in bars returns the mol/kg of every gas dissolved in an ocean of `species`; None
        removes it.  The callbacks are host code and touch only N_ocean (clima_adiabat_general.f90:226-246), never the
        profile, so they run on the host after the device call, per column, exactly as the reference calls them."""
        if species not in self.sp["names"]:
            raise ClimaException('Gas "%s" is not in the list of species' % species)
        self.ocean_fcns[self.sp["names"].index(species)] = fcn

    def _oceans(self, p: Dict) -> Dict:
        ng, ncol = self.ng, p["P_surf"].shape[0]
        N_ocean = np.zeros((ng, ng, ncol), order="F")
        if any(f is not None for f in self.ocean_fcns):
            P_i_cur = p["f_i_surf"] * p["P_surf"][None, :]
            for c in np.where(p["status"] == 0)[0]:
                for j, fcn in enumerate(self.ocean_fcns):
                    if fcn is None:
                        continue
                    m_i = np.asarray(fcn(float(self._T_surf[c]), ng, P_i_cur[:, c] / 1.0e6), dtype=np.float64)
                    N_ocean[:, j, c] = m_i * p["N_surface"][j, c] * (self.sp["mass"][j] / 1.0e3)
                    N_ocean[j, j, c] = 0.0  # an ocean does not dissolve into itself
        p["N_ocean"] = N_ocean
        return p

    def set_particle_density_and_radii(self, P, pdensities, pradii):
        """`AdiabatClimate%set_particle_density_and_radii` (clima_adiabat.f90:1047-1123): P (n) [dynes/cm^2, strictly
        decreasing], pdensities (n, np) [particles/cm^3], pradii (n, np) [cm]; make_profile interpolates them onto its
        pressure grid (log-log, end values held) and TOA_fluxes passes them to the RT."""
        P = np.ascontiguousarray(P, dtype=np.float64).reshape(-1)
        pd = np.asfortranarray(pdensities, dtype=np.float64)
        pr = np.asfortranarray(pradii, dtype=np.float64)
        npart = len(self.rad.tables.particle_names)
        if P.shape[0] != pd.shape[0]:
            raise ClimaException("`P` and `pdensities` have incompatible shapes")
        if P.shape[0] != pr.shape[0]:
            raise ClimaException("`P` and `pradii` have incompatible shapes")
        if pd.ndim != 2 or pd.shape[1] != npart:
            raise ClimaException("The second dimension of `pdensities` does not match the number of particles")
        if pr.ndim != 2 or pr.shape[1] != npart:
            raise ClimaException("The second dimension of `pradii` does not match the number of particles")
        err = C.create_string_buffer(ERR_LEN + 1)
        self.rad._lib.clima_rt_adiabat_set_particles(self.rad._h, _i(P.shape[0]), P.ctypes.data_as(c_double_p),
                                                     pd.ctypes.data_as(c_double_p), pr.ctypes.data_as(c_double_p), err)
        _check(err)

    # ---- the two other column builders of the reference as lock-step Newton iterations over the whole batch: every
    # iteration evaluates the residuals of all columns and of all their finite-difference perturbations in ONE
    # make_profile_batch call (the reference runs MINPACK hybrd1 per column, 20-40 make_profile calls each)
    @staticmethod
    def _newton_batch(X0, resid, tol, max_iter=60, h=1.0e-5):
        """Solves F(X[:, c]) = 0 for every column c of X0 (n, ncol), X = log10 of the unknown pressures.
        `resid(idx, X) -> (F (n, m), ok (m))` evaluates m trial columns, idx[j] = the column X[:, j] belongs to.
        Damped Newton with forward differences; converged like hybrd1 (relative change of X below tol) or |F| < tol."""
        n, ncol = X0.shape
        X = X0.copy()
        allc = np.arange(ncol)
        F, ok = resid(allc, X)
        done, failed = np.zeros(ncol, bool), ~ok
        for _ in range(max_iter):
            act = np.where(~done & ~failed)[0]
            if act.size == 0:
                break
            m = act.size
            Xp = np.repeat(X[:, act], n, axis=1)                      # block q holds the n perturbations of column act[q]
            for k in range(n):
                Xp[k, k::n] += h
            Fp, okp = resid(np.repeat(act, n), Xp)
            step = np.zeros((n, m))
            live = np.ones(m, bool)
            for q, c in enumerate(act):
                J = (Fp[:, q * n:(q + 1) * n] - F[:, [c]]) / h        # J[i, k] = dF_i / dx_k
                try:
                    if not okp[q * n:(q + 1) * n].all() or not np.all(np.isfinite(J)):
                        raise np.linalg.LinAlgError
                    step[:, q] = np.linalg.solve(J, -F[:, c])
                except np.linalg.LinAlgError:
                    live[q] = False
            step = np.clip(np.nan_to_num(step), -1.0, 1.0)            # at most a decade per iteration
            # hybrd's test: a Newton step below tol relative to x means x is the root to that accuracy (the residual
            # itself bottoms out at the integration noise of make_profile, ~1e-7 at the default rtol)
            small = live & (np.max(np.abs(step) / np.maximum(np.abs(X[:, act]), 1.0), axis=0) < tol)
            X[:, act[small]] += step[:, small]                        # (the step is taken, as hybrd does)
            done[act[small]] = True
            live &= ~small
            if not live.any():
                continue
            lam = np.ones(m)
            f0 = np.max(np.abs(F[:, act]), axis=0)
            for _ls in range(10):                                      # halve the step until the residual drops
                Xn = X[:, act] + lam[None, :] * step
                Fn, okn = resid(act, Xn)
                fn = np.max(np.abs(np.nan_to_num(Fn, nan=np.inf)), axis=0)
                worse = live & (~okn | ~(fn < f0 + 10.0 * tol))
                if not worse.any():
                    break
                lam[worse] *= 0.5
            good = live & okn & (fn < f0 + 10.0 * tol)
            idx = act[good]
            dx = np.max(np.abs(lam[None, :] * step)[:, good] / np.maximum(np.abs(Xn[:, good]), 1.0), axis=0)
            X[:, idx] = Xn[:, good]
            F[:, idx] = Fn[:, good]
            done[idx] = (dx < tol) | (fn[good] < tol)
            failed[act[live & ~good]] = True
        return X, done

    def make_column_batch(self, T_surf, N_i_surf, tol: float = 1.0e-8) -> Dict:
        """`AdiabatClimate%make_column` (clima_adiabat.f90:476-574) for every (T_surf[c], N_i_surf[:, c]) [mol/cm^2]: the
        surface pressures whose profile holds N_atmos + N_surface + sum(N_ocean) = N_i_surf.  Initial guesses as in the
        reference (N m g times 1, 0.5, 2, 0.1, 5, 0.01, tried in turn for the columns that have not converged).
        Returns the make_profile_batch dict of the solutions plus "P_i_surf" (ng, ncol)."""
        T_surf = np.ascontiguousarray(T_surf, dtype=np.float64).reshape(-1)
        N_t = np.asarray(N_i_surf, dtype=np.float64)
        if N_t.ndim == 1:
            N_t = N_t[:, None]
        if N_t.shape != (self.ng, T_surf.shape[0]):
            raise ClimaException("N_i_surf has the wrong dimension")
        ncol = T_surf.shape[0]
        grav = 6.67430e-11 * (self.planet_mass / 1.0e3) / (self.planet_radius / 1.0e2) ** 2.0 * 1.0e2
        mass = np.asarray(self.sp["mass"], dtype=np.float64)
        X = np.zeros((self.ng, ncol))
        solved = np.zeros(ncol, bool)
        tiny = math.sqrt(2.2250738585072014e-308)
        for sf in (1.0, 0.5, 2.0, 0.1, 5.0, 0.01):
            todo = np.where(~solved)[0]
            if todo.size == 0:
                break

            def resid(idx, Xq, todo=todo):
                cols = todo[idx]
                with np.errstate(all="ignore"):
                    p = self.make_profile_batch(T_surf[cols], np.asfortranarray(10.0 ** Xq), check=False)
                    N = p["N_atmos"] + p["N_surface"] + p["N_ocean"].sum(axis=1)
                    den = np.where(N_t[:, cols] > 0.0, N_t[:, cols], 1.0)   # a gas with no inventory: absolute residual
                    return (N - N_t[:, cols]) / den, p["status"] == 0
            X0 = np.log10(np.maximum(N_t[:, todo] * mass[:, None] * grav * sf, tiny))
            Xs, done = self._newton_batch(X0, resid, tol)
            X[:, todo[done]] = Xs[:, done]
            solved[todo[done]] = True
        if not solved.all():
            raise ClimaException("column %d: hybrd1 root solve failed in make_column." % int(np.where(~solved)[0][0]))
        P_i = np.asfortranarray(10.0 ** X)
        p = self.make_profile_batch(T_surf, P_i)
        p["P_i_surf"] = P_i
        return p

    def make_profile_bg_gas_batch(self, T_surf, P_i_surf, P_surf, bg_gas: str, tol: float = 1.0e-8) -> Dict:
        """`AdiabatClimate%make_profile_bg_gas` (clima_adiabat.f90:579-650): the pressure of `bg_gas` is varied until the
        surface pressure of the profile is P_surf (scalar or (ncol)); guesses P_surf times 1 and 0.1 as in the reference."""
        T_surf, P_i = self._inputs(T_surf, P_i_surf)
        ncol = T_surf.shape[0]
        P_t = np.broadcast_to(np.asarray(P_surf, dtype=np.float64), (ncol,)).copy()
        if np.any(P_t <= 0.0):
            raise ClimaException("P_surf must be greater than zero.")
        if bg_gas not in self.sp["names"]:
            raise ClimaException('Gas "%s" is not in the list of species' % bg_gas)
        ind = self.sp["names"].index(bg_gas)
        X = np.zeros((1, ncol))
        solved = np.zeros(ncol, bool)
        for sf in (1.0, 0.1):
            todo = np.where(~solved)[0]
            if todo.size == 0:
                break

            def resid(idx, Xq, todo=todo):
                cols = todo[idx]
                Pq = np.asfortranarray(P_i[:, cols].copy())
                Pq[ind] = 10.0 ** Xq[0]
                with np.errstate(all="ignore"):
                    p = self.make_profile_batch(T_surf[cols], Pq, check=False)
                    return (p["P_surf"] / P_t[cols] - 1.0)[None, :], p["status"] == 0
            Xs, done = self._newton_batch(np.log10(P_t[todo] * sf)[None, :], resid, tol)
            X[:, todo[done]] = Xs[:, done]
            solved[todo[done]] = True
        if not solved.all():
            raise ClimaException("column %d: hybrd1 root solve failed in make_profile_bg_gas." % int(np.where(~solved)[0][0]))
        P_sol = np.asfortranarray(P_i.copy())
        P_sol[ind] = 10.0 ** X[0]
        p = self.make_profile_batch(T_surf, P_sol)
        p["P_i_surf"] = P_sol
        return p

    def make_column(self, T_surf: float, N_i_surf):
        """One column; leaves the reference's member names on the object (like make_profile)."""
        p = self.make_column_batch([T_surf], np.asarray(N_i_surf, dtype=np.float64)[:, None])
        self.make_profile(T_surf, p["P_i_surf"][:, 0])

    def make_profile_bg_gas(self, T_surf: float, P_i_surf, P_surf: float, bg_gas: str):
        p = self.make_profile_bg_gas_batch([T_surf], np.asarray(P_i_surf, dtype=np.float64)[:, None], P_surf, bg_gas)
        self.make_profile(T_surf, p["P_i_surf"][:, 0])

    def surface_temperature_batch(self, P_i_surf, T_guess=280.0, surface_heat_flow: float = 0.0, tol: float = 1.0e-8,
                                  solve_for_T_trop: bool = False):
        """`AdiabatClimate%surface_temperature` (clima_adiabat.f90:882-975) for every composition P_i_surf[:, c]: the
        surface temperature at which ISR - OLR + surface_heat_flow = 0, unknown x = log10(T_surf) as in the reference,
        every iteration ONE batched TOA_fluxes call over all columns and their perturbations.  With solve_for_T_trop the
        tropopause temperature of each column is a second unknown that must equal the skin temperature of its own bond
        albedo (:953-959; initial guess self.T_trop).  Returns (T_surf (ncol), ISR, OLR) and, with solve_for_T_trop,
        T_trop (ncol) as a fourth array."""
        P_i = np.asfortranarray(P_i_surf, dtype=np.float64)
        if P_i.ndim == 1:
            P_i = np.asfortranarray(P_i[:, None])
        if P_i.shape[0] != self.ng:
            raise ClimaException("P_i_surf has the wrong dimension")
        ncol = P_i.shape[1]
        nun = 2 if solve_for_T_trop else 1
        X0 = np.empty((nun, ncol))
        X0[0] = np.log10(np.broadcast_to(np.asarray(T_guess, dtype=np.float64), (ncol,)))
        if solve_for_T_trop:
            X0[1] = math.log10(self.T_trop)
        S = self.rad.bolometric_flux()   # W/m^2
        sigma_si = 5.670374419e-8

        def fluxes(T, cols, Tt):
            r = self.TOA_fluxes_batch(T, np.asfortranarray(P_i[:, cols]), full=solve_for_T_trop, T_trop=Tt)
            if not solve_for_T_trop:
                return r[0], r[1], None
            return r["ISR"], r["OLR"], r["fup_n_sol"][-1] / r["fdn_n_sol"][-1]   # bond albedo at the top (:955)

        def resid(idx, Xq):
            T = 10.0 ** Xq[0]
            Tt = 10.0 ** Xq[1] if solve_for_T_trop else None
            m = T.shape[0]
            try:
                isr, olr, alb = fluxes(T, idx, Tt)
            except ClimaException:
                # one failing column fails the batched call (like the reference's loop): find out which, column by column
                isr, olr, alb = np.full(m, np.nan), np.full(m, np.nan), np.full(m, np.nan)
                for j in range(m):
                    try:
                        a, b, c = fluxes(T[j:j + 1], idx[j:j + 1], None if Tt is None else Tt[j:j + 1])
                        isr[j], olr[j] = a[0], b[0]
                        if c is not None:
                            alb[j] = c[0]
                    except ClimaException:
                        pass
            with np.errstate(all="ignore"):
                F = np.empty((nun, m))
                F[0] = (isr - olr + surface_heat_flow) / olr
                if solve_for_T_trop:
                    skin = ((S * (1.0 - alb)) / (4.0 * sigma_si)) ** 0.25 * 0.5 ** 0.25   # clima_eqns.f90:248-261
                    F[1] = skin / Tt - 1.0
            return F, np.all(np.isfinite(F), axis=0)
        X, done = self._newton_batch(X0, resid, tol, h=1.0e-4)
        if not done.all():
            raise ClimaException("column %d: hybrd1 root solve failed." % int(np.where(~done)[0][0]))
        T = 10.0 ** X[0]
        if solve_for_T_trop:
            Tt = 10.0 ** X[1]
            isr, olr = self.TOA_fluxes_batch(T, P_i, T_trop=Tt)
            return T, isr, olr, Tt
        isr, olr = self.TOA_fluxes_batch(T, P_i)
        return T, isr, olr

    def TOA_fluxes_column_batch(self, T_surf, N_i_surf):
        """`AdiabatClimate%TOA_fluxes_column` (clima_adiabat.f90:802-825): make_column, then the RT -> (ISR, OLR)."""
        sol = self.make_column_batch(T_surf, N_i_surf)
        return self.TOA_fluxes_batch(T_surf, sol["P_i_surf"])

    def TOA_fluxes_bg_gas_batch(self, T_surf, P_i_surf, P_surf, bg_gas: str):
        """`AdiabatClimate%TOA_fluxes_bg_gas` (clima_adiabat.f90:828-853): make_profile_bg_gas, then the RT."""
        sol = self.make_profile_bg_gas_batch(T_surf, P_i_surf, P_surf, bg_gas)
        return self.TOA_fluxes_batch(T_surf, sol["P_i_surf"])

    def _joint_solve(self, X0, cols_of, extra_residuals, surface_heat_flow, tol):
        """Energy balance plus column-builder constraints as ONE Newton system per column (the reference nests a
        hybrd1 of the column builder inside a hybrd1 for T_surf; the root is the same).  X = [log10 T_surf; log10 of
        the unknown pressures]; cols_of(X, idx) -> P_i_surf (ng, m); extra_residuals(profile, idx) -> (n - 1, m)."""
        def resid(idx, Xq):
            m = Xq.shape[1]
            T = 10.0 ** Xq[0]
            P = np.asfortranarray(cols_of(Xq, idx))
            F = np.full((Xq.shape[0], m), np.nan)
            try:
                r = self.TOA_fluxes_batch(T, P, full=True)
                with np.errstate(all="ignore"):
                    F[0] = (r["ISR"] - r["OLR"] + surface_heat_flow) / r["OLR"]
                    F[1:] = extra_residuals(r["profile"], idx)
            except ClimaException:
                for j in range(m):
                    try:
                        r = self.TOA_fluxes_batch(T[j:j + 1], P[:, j:j + 1], full=True)
                        with np.errstate(all="ignore"):
                            F[0, j] = (r["ISR"][0] - r["OLR"][0] + surface_heat_flow) / r["OLR"][0]
                            F[1:, j] = extra_residuals(r["profile"], idx[j:j + 1])[:, 0]
                    except ClimaException:
                        pass
            return F, np.all(np.isfinite(F), axis=0)
        X, done = self._newton_batch(X0, resid, tol, h=1.0e-4)
        if not done.all():
            raise ClimaException("column %d: hybrd1 root solve failed." % int(np.where(~done)[0][0]))
        return X

    def surface_temperature_column_batch(self, N_i_surf, T_guess=280.0, surface_heat_flow: float = 0.0,
                                         tol: float = 1.0e-8):
        """`AdiabatClimate%surface_temperature_column` (clima_adiabat.f90:977-992): ISR = OLR with the atmosphere built
        from column inventories N_i_surf (ng, ncol) [mol/cm^2].  Returns (T_surf, ISR, OLR, P_i_surf)."""
        N_t = np.asarray(N_i_surf, dtype=np.float64)
        if N_t.ndim == 1:
            N_t = N_t[:, None]
        ncol = N_t.shape[1]
        Tg = np.broadcast_to(np.asarray(T_guess, dtype=np.float64), (ncol,)).copy()
        start = self.make_column_batch(Tg, N_t)                       # pressures that hold the inventories at T_guess
        X0 = np.vstack([np.log10(Tg)[None, :], np.log10(start["P_i_surf"])])
        den = np.where(N_t > 0.0, N_t, 1.0)

        def extra(p, idx):
            N = p["N_atmos"] + p["N_surface"] + p["N_ocean"].sum(axis=1)
            return (N - N_t[:, idx]) / den[:, idx]
        X = self._joint_solve(X0, lambda Xq, idx: 10.0 ** Xq[1:], extra, surface_heat_flow, tol)
        T, P_i = 10.0 ** X[0], np.asfortranarray(10.0 ** X[1:])
        isr, olr = self.TOA_fluxes_batch(T, P_i)
        return T, isr, olr, P_i

    def surface_temperature_bg_gas_batch(self, P_i_surf, P_surf, bg_gas: str, T_guess=280.0,
                                         surface_heat_flow: float = 0.0, tol: float = 1.0e-8):
        """`AdiabatClimate%surface_temperature_bg_gas` (clima_adiabat.f90:996-1020): ISR = OLR with the pressure of
        `bg_gas` adjusted so that the surface pressure is P_surf.  Returns (T_surf, ISR, OLR, P_i_surf)."""
        P_i = np.asfortranarray(P_i_surf, dtype=np.float64)
        if P_i.ndim == 1:
            P_i = np.asfortranarray(P_i[:, None])
        ncol = P_i.shape[1]
        P_t = np.broadcast_to(np.asarray(P_surf, dtype=np.float64), (ncol,)).copy()
        Tg = np.broadcast_to(np.asarray(T_guess, dtype=np.float64), (ncol,)).copy()
        start = self.make_profile_bg_gas_batch(Tg, P_i, P_t, bg_gas)
        ind = self.sp["names"].index(bg_gas)
        X0 = np.vstack([np.log10(Tg)[None, :], np.log10(start["P_i_surf"][ind])[None, :]])

        def cols_of(Xq, idx):
            Pq = P_i[:, idx].copy()
            Pq[ind] = 10.0 ** Xq[1]
            return Pq
        X = self._joint_solve(X0, cols_of, lambda p, idx: (p["P_surf"] / P_t[idx] - 1.0)[None, :], surface_heat_flow, tol)
        T = 10.0 ** X[0]
        P_sol = np.asfortranarray(cols_of(X, np.arange(ncol)))
        isr, olr = self.TOA_fluxes_batch(T, P_sol)
        return T, isr, olr, P_sol

    def surface_temperature(self, P_i_surf, T_guess: float = 280.0) -> float:
        T, _, _ = self.surface_temperature_batch(np.asarray(P_i_surf, dtype=np.float64)[:, None], T_guess)
        self.make_profile(float(T[0]), P_i_surf)
        return float(T[0])

    def kernel_ms(self) -> float:
        """Device time of the last make_profile kernel (CUDA events on the handle's stream)."""
        st = (C.c_double * 2)()
        self.rad._lib.clima_rt_get_adiabat_stats(self.rad._h, st)
        return float(st[0])

    NOALLOC_MAX_WORK_DOUBLES = 2 * 1024 * 1024  # CLIMA_RT_NOALLOC_MAX_WORK_DOUBLES of include/clima_rt.h

    def make_profile_batch_noalloc(self, ncol: int, T_surf_ptr: int, P_i_surf_ptr: int, out_ptrs: Dict[str, int]):
        """`clima_rt_make_profile_batch_noalloc`: the same kernel with its work arrays in the library's static
        device array; every argument is a raw DEVICE address of the caller (`tensor.data_ptr()`): T_surf (ncol),
        P_i_surf (ng, ncol) and, in `out_ptrs`, P, T, z, dz (nz, ncol), f_i (nz, ng, ncol), f_i_surf, N_surface
        (ng, ncol), P_surf, P_trop (ncol) as float64 and status, nfcn (ncol) as int32.  The call allocates nothing;
        per-column failures stay in `status`.  `_sync()` (clima_rt_adiabat_set, which allocates) must have run:
        any make_profile / TOA_fluxes call, or call it first."""
        self._sync()
        o = CProfileOut()
        for k in PROFILE_FIELDS:
            setattr(o, k, C.cast(C.c_void_p(out_ptrs.get(k, 0) or None), c_double_p))
        o.status = C.cast(C.c_void_p(out_ptrs.get("status", 0) or None), c_int32_p)
        o.nfcn = C.cast(C.c_void_p(out_ptrs.get("nfcn", 0) or None), c_int32_p)
        err = C.create_string_buffer(ERR_LEN + 1)
        self.rad._lib.clima_rt_make_profile_batch_noalloc(
            self.rad._h, _i(ncol), C.cast(C.c_void_p(T_surf_ptr or None), c_double_p),
            C.cast(C.c_void_p(P_i_surf_ptr or None), c_double_p), C.byref(o), err)
        _check(err)
