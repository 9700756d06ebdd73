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
