"""Host-side mirror of the reference's `Radtran` interface over the CUDA C-ABI library.

Same method names, argument meaning and error behaviour as `type Radtran`
(src/radtran/clima_radtran.f90:31-85: `radiate`, `TOA_fluxes`, `set_custom_optical_properties`,
`unset_custom_optical_properties`, public knobs `diurnal_fac`, `surface_albedo`, ...), plus the
batched-column entry point `radiate_batch` / `TOA_fluxes_batch` that the north star adds.
Errors follow the Cython layer (clima/cython/_clima.pyx:11,48-49): a non-empty `err` raises
`ClimaException`.

All compute goes through `libclima_rt.so` (include/clima_rt.h).  There is no CPU fallback: if the
library is missing or no CUDA device is usable, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import numpy as np

from ._abi import (COut, CParams, CTables, ERR_LEN, OpacityTables, RadtranParams, alloc_out,
                   c_double_p, dptr)

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libclima_rt.so")
_lib = None

EXPORTS = ["clima_rt_create", "clima_rt_destroy", "clima_rt_set_params",
           "clima_rt_set_custom_optical_properties", "clima_rt_unset_custom_optical_properties",
           "clima_rt_radiate_batch", "clima_rt_radiate_batch_device", "clima_rt_jacobian_batch",
           "clima_rt_get_opacity",
           "clima_rt_get_launch_count", "clima_rt_get_stage_ms", "clima_rt_get_stream", "clima_rt_get_rorr_stats", "clima_rt_get_rorr_cut_stats", "clima_rt_measure_fp64_peak",
           "clima_rt_version_get",
           "clima_rt_comm_unique_id", "clima_rt_comm_init", "clima_rt_comm_destroy", "clima_rt_allgather_device",
           "clima_rt_radiate_batch_sharded",
           "clima_rt_adiabat_set", "clima_rt_make_profile_batch", "clima_rt_adiabat_toa_fluxes_batch", "clima_rt_get_adiabat_stats",
           "clima_rt_make_profile_batch_noalloc", "clima_rt_adiabat_set_particles",
           "clima_rt_adiabat_set_column_T_trop"]
COMM_ID_BYTES = 128


class ClimaException(Exception):
    """clima/cython/_clima.pyx:11"""


def load_library() -> C.CDLL:
    """Loads the CUDA C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        path = os.environ.get("CLIMA_RT_LIBRARY", _LIB_PATH)  # developer knob: A/B builds of the same ABI
        if not os.path.exists(path):
            raise ClimaException(
                f"{path} is missing: build it with `python -m clima_b200.build` "
                "(there is no CPU fallback)")
        _lib = C.CDLL(path)
        for name in EXPORTS:
            getattr(_lib, name).restype = None
    return _lib


def _i(v):
    return C.byref(C.c_int32(int(v)))


def _check(err):
    if len(err.value.strip()) > 0:
        raise ClimaException(err.value.decode())


def measure_fp64_peak(device: int = 0) -> float:
    lib = load_library()
    out = C.c_double(0.0)
    err = C.create_string_buffer(ERR_LEN + 1)
    lib.clima_rt_measure_fp64_peak(_i(device), C.byref(out), err)
    _check(err)
    return out.value


def version() -> str:
    buf = C.create_string_buffer(101)
    load_library().clima_rt_version_get(buf)
    return buf.value.decode()


class ClimaRadtranWrk:
    """`type ClimaRadtranWrk` (clima_radtran.f90:11-25) of one column."""

    def __init__(self, fup_a, fdn_a, fup_n, fdn_n, amean, tau_band):
        self.fup_a, self.fdn_a, self.fup_n, self.fdn_n = fup_a, fdn_a, fup_n, fdn_n
        self.amean, self.tau_band = amean, tau_band


class Radtran:
    def __init__(self, tables: OpacityTables, params: RadtranParams, device: int = 0):
        self._lib = load_library()
        self.tables = tables
        self._params = params
        self._ct, self._keep_t = tables.to_c()
        cp, self._keep_p = params.to_c()
        self._h = C.c_void_p()
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_create(C.byref(self._ct), C.byref(cp), _i(device), C.byref(self._h), err)
        _check(err)
        self.device = device
        self.wrk_ir: Optional[ClimaRadtranWrk] = None
        self.wrk_sol: Optional[ClimaRadtranWrk] = None
        self.f_total: Optional[np.ndarray] = None

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.clima_rt_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # ---- run-time knobs (public members of type Radtran, clima_radtran.f90:51-68)
    @property
    def params(self) -> RadtranParams:
        return self._params

    @params.setter
    def params(self, p: RadtranParams):
        cp, keep = p.to_c()
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_set_params(self._h, C.byref(cp), err)
        _check(err)
        self._params, self._keep_p = p, keep

    def set_custom_optical_properties(self, log10P, dtau_dz, w0, g0):
        """State after OpticalProperties_set_custom_optical_properties
        (clima_radtran_types.f90:504-531): tables already on the opacity grid."""
        arrs = [np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="F"))
                for a in (log10P, dtau_dz, w0, g0)]
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_set_custom_optical_properties(self._h, _i(len(arrs[0])),
                                                         *(dptr(a) for a in arrs), err)
        _check(err)

    def unset_custom_optical_properties(self):
        self._lib.clima_rt_unset_custom_optical_properties(self._h)

    # ---- the batched-column entry point
    def radiate_batch(self, cols: Dict, compute_solar=True, compute_opacity=True, spectral=False,
                      out=None):
        """`radiate` + `TOA_fluxes` for cols['ncol'] columns.  `cols` holds Fortran-ordered
        T_surface(ncol), T(nz,ncol), P(nz,ncol) [bar], densities(nz,ng,ncol), dz(nz,ncol) and
        optionally pdensities/radii(nz,np,ncol).  Returns a dict of result arrays."""
        nz, ng, npp, ncol = cols["nz"], cols["ng"], cols["np"], cols["ncol"]
        if out is None:
            out = alloc_out(nz, ncol, self.tables.nw_ir, self.tables.nw_sol, spectral)
        o, arrs = out
        err = C.create_string_buffer(ERR_LEN + 1)
        pd = dptr(cols["pdensities"]) if "pdensities" in cols else c_double_p()
        rd = dptr(cols["radii"]) if "radii" in cols else c_double_p()
        self._lib.clima_rt_radiate_batch(self._h, _i(ncol), _i(nz), _i(ng), _i(npp),
                                         dptr(cols["T_surface"]), dptr(cols["T"]), dptr(cols["P"]),
                                         dptr(cols["densities"]), dptr(cols["dz"]), pd, rd,
                                         _i(compute_solar), _i(compute_opacity), C.byref(o), err)
        _check(err)
        return arrs

    def radiate_batch_device(self, ncol, nz, ng, npp, inp: Dict[str, int], outp: Dict[str, int],
                             compute_solar=True, compute_opacity=True):
        """Same call with raw DEVICE addresses (integers, e.g. `tensor.data_ptr()`): `inp` has
        T_surface, T, P, densities, dz (and pdensities, radii); `outp` any member of
        `clima_rt_out` (the seven integrated results are mandatory)."""
        from ._abi import OUT_FIELDS, c_int32_p
        o = COut()
        for n in OUT_FIELDS:
            setattr(o, n, C.cast(C.c_void_p(outp.get(n, 0) or None), c_double_p))
        o.status = C.cast(C.c_void_p(outp.get("status", 0) or None), c_int32_p)

        def dp(k):
            return C.cast(C.c_void_p(inp.get(k, 0) or None), c_double_p)
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_radiate_batch_device(self._h, _i(ncol), _i(nz), _i(ng), _i(npp),
                                                dp("T_surface"), dp("T"), dp("P"), dp("densities"),
                                                dp("dz"), dp("pdensities"), dp("radii"),
                                                _i(compute_solar), _i(compute_opacity), C.byref(o), err)
        _check(err)

    def TOA_fluxes_batch(self, cols: Dict):
        r = self.radiate_batch(cols)
        return r["ISR"], r["OLR"]

    # ---- multi-GPU: the NCCL entry points of the C ABI (include/clima_rt.h)
    @staticmethod
    def comm_unique_id() -> bytes:
        """ncclGetUniqueId through the C ABI: rank 0 calls it and hands the bytes to the other ranks."""
        buf = C.create_string_buffer(COMM_ID_BYTES)
        err = C.create_string_buffer(ERR_LEN + 1)
        load_library().clima_rt_comm_unique_id(buf, err)
        _check(err)
        return buf.raw

    def comm_init(self, nranks: int, rank: int, uid: bytes):
        if len(uid) != COMM_ID_BYTES:
            raise ClimaException("comm_init: the unique id must be %d bytes" % COMM_ID_BYTES)
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_comm_init(self._h, _i(nranks), _i(rank), C.create_string_buffer(uid, COMM_ID_BYTES), err)
        _check(err)
        self._nranks = nranks

    def comm_destroy(self):
        self._lib.clima_rt_comm_destroy(self._h)

    def allgather_device(self, send_ptr: int, recv_ptr: int, count: int):
        """ncclAllGather of `count` doubles per rank between device buffers, on the handle's stream."""
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_allgather_device(self._h, C.cast(C.c_void_p(send_ptr), c_double_p),
                                            C.cast(C.c_void_p(recv_ptr), c_double_p),
                                            C.byref(C.c_int64(int(count))), err)
        _check(err)

    def radiate_batch_sharded(self, cols: Dict, ncol_max: int, compute_solar=True, compute_opacity=True):
        """This rank's columns (`cols`, possibly empty) through `radiate_batch`, then the all-gather:
        returns the dict of integrated results of ALL ranks, arrays (nz+1, ncol_max, nranks) /
        (ncol_max, nranks), padding columns zero."""
        nz, ng, npp, ncol = cols["nz"], cols["ng"], cols["np"], cols["ncol"]
        nr = self._nranks
        L = nz + 1
        o = COut()
        arrs = {}
        for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"):
            arrs[k] = np.zeros((L, ncol_max, nr), order="F")
        for k in ("ISR", "OLR"):
            arrs[k] = np.zeros((ncol_max, nr), order="F")
        for k, a in arrs.items():
            setattr(o, k, dptr(a))
        err = C.create_string_buffer(ERR_LEN + 1)
        nul = c_double_p()
        g = lambda k: dptr(cols[k]) if ncol > 0 else nul  # noqa: E731
        pd = dptr(cols["pdensities"]) if "pdensities" in cols and ncol > 0 else nul
        rd = dptr(cols["radii"]) if "radii" in cols and ncol > 0 else nul
        self._lib.clima_rt_radiate_batch_sharded(self._h, _i(ncol), _i(ncol_max), _i(nz), _i(ng), _i(npp),
                                                 g("T_surface"), g("T"), g("P"), g("densities"), g("dz"), pd, rd,
                                                 _i(compute_solar), _i(compute_opacity), C.byref(o), err)
        _check(err)
        return arrs

    def jacobian_batch(self, T_surface, T):
        """Perturbed-temperature batch of AdiabatClimate_jacobian_from_base
        (src/adiabat/clima_adiabat_solve.f90:798-820) in one call: T_surface(ncol), T(nz,ncol) on the
        MODEL grid; the handle must hold the base column (radiate on the doubled grid, ncol=1).
        Returns f_total(nz+1,ncol) = rad%f_total(2i-1) of every perturbed column."""
        Ts = np.ascontiguousarray(T_surface, dtype=np.float64)
        T = np.asfortranarray(T, dtype=np.float64)
        nz, ncol = T.shape
        if Ts.shape != (ncol,):
            raise ClimaException("T_surface and T disagree on the number of columns")
        out = np.zeros((nz + 1, ncol), order="F")
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_jacobian_batch(self._h, _i(ncol), _i(nz), dptr(Ts), dptr(T), dptr(out), err)
        _check(err)
        return out

    # ---- the reference's single-column API (clima_radtran.f90:221-342)
    def radiate(self, T_surface, T, P, densities, dz, pdensities=None, radii=None,
                compute_solar=True, compute_opacity=True):
        T = np.ascontiguousarray(T, dtype=np.float64)
        nz = T.shape[0]
        dens = np.asfortranarray(densities, dtype=np.float64)
        cols = dict(ncol=1, nz=nz, ng=dens.shape[1] if dens.ndim > 1 else 1,
                    np=0 if pdensities is None else np.asarray(pdensities).reshape(nz, -1).shape[1],
                    T_surface=np.array([float(T_surface)]), T=T,
                    P=np.ascontiguousarray(P, dtype=np.float64),
                    densities=np.ascontiguousarray(dens.ravel(order="F")),
                    dz=np.ascontiguousarray(dz, dtype=np.float64))
        if pdensities is not None:
            cols["pdensities"] = np.ascontiguousarray(np.asfortranarray(pdensities, dtype=np.float64).ravel(order="F"))
        if radii is not None:
            cols["radii"] = np.ascontiguousarray(np.asfortranarray(radii, dtype=np.float64).ravel(order="F"))
        if (pdensities is None) != (radii is None):
            raise ClimaException("Both pdensities and radii must be arguments.")
        r = self.radiate_batch(cols, compute_solar, compute_opacity, spectral=True)
        self.wrk_ir = ClimaRadtranWrk(r["fup_a_ir"][..., 0], r["fdn_a_ir"][..., 0], r["fup_n_ir"][:, 0],
                                      r["fdn_n_ir"][:, 0], np.zeros_like(r["fup_a_ir"][..., 0]),
                                      r["tau_band_ir"][..., 0])
        self.wrk_sol = ClimaRadtranWrk(r["fup_a_sol"][..., 0], r["fdn_a_sol"][..., 0], r["fup_n_sol"][:, 0],
                                       r["fdn_n_sol"][:, 0], r["amean_sol"][..., 0], r["tau_band_sol"][..., 0])
        self.f_total = r["f_total"][:, 0]
        self._last = r
        return r

    def TOA_fluxes(self, T_surface, T, P, densities, dz, pdensities=None, radii=None):
        r = self.radiate(T_surface, T, P, densities, dz, pdensities, radii)
        return float(r["ISR"][0]), float(r["OLR"][0])

    # ---- host-side helpers of type Radtran that do not touch the device path
    def bolometric_flux(self) -> float:
        """Bolometric stellar flux at the planet in W/m^2 (Radtran_bolometric_flux,
        clima_radtran.f90:353-364): sum of photons_sol * bin width in frequency, scaled."""
        i0, i1 = self.tables.sol_ind
        freq = self.tables.freq[i0 - 1:i1 + 1]
        flux = 0.0
        for i in range(self.tables.nw_sol):
            flux = flux + float(self.tables.photons_sol[i]) * float(freq[i] - freq[i + 1])
        return self._params.photon_scale_factor * flux / 1.0e3

    def set_bolometric_flux(self, flux: float):
        """Radtran_set_bolometric_flux (clima_radtran.f90:345-350): rescales photon_scale_factor."""
        import dataclasses
        p1 = dataclasses.replace(self._params, photon_scale_factor=1.0)
        self._params = p1
        self.params = dataclasses.replace(p1, photon_scale_factor=float(flux) / self.bolometric_flux())

    def equilibrium_temperature(self, bond_albedo: float) -> float:
        """clima_eqns.f90:248-254 through Radtran_equilibrium_temperature (clima_radtran.f90:376-382)"""
        sigma_si = 5.670374419e-8  # clima_const.f90:16
        return ((self.bolometric_flux() * (1.0 - bond_albedo)) / (4.0 * sigma_si)) ** 0.25

    def skin_temperature(self, bond_albedo: float) -> float:
        """clima_eqns.f90:256-261 through Radtran_skin_temperature (clima_radtran.f90:367-373)"""
        return self.equilibrium_temperature(bond_albedo) * 0.5 ** 0.25

    @staticmethod
    def apply_radiation_enhancement(r: Dict, rad_enhancement: float) -> Dict:
        """Radtran_apply_radiation_enhancement (clima_radtran.f90:402-411) on a result dict of
        radiate / radiate_batch: scales the solar fluxes in place and rebuilds f_total."""
        for k in ("fdn_n_sol", "fup_n_sol", "fdn_a_sol", "fup_a_sol"):
            if k in r and r[k] is not None:
                r[k] *= rad_enhancement
        r["f_total"][...] = (r["fdn_n_sol"] - r["fup_n_sol"]) + (r["fdn_n_ir"] - r["fup_n_ir"])
        return r

    # ---- introspection / test hooks
    def get_opacity(self, nz: int, icol: int = 0):
        G, nw = self.tables.ngauss, self.tables.nw
        tau = np.zeros((nz, G, nw), order="F")
        w0 = np.zeros((nz, G, nw), order="F")
        g = np.zeros((nz, nw), order="F")
        tb = np.zeros((nz, nw), order="F")
        err = C.create_string_buffer(ERR_LEN + 1)
        self._lib.clima_rt_get_opacity(self._h, _i(icol), dptr(tau), dptr(w0), dptr(g), dptr(tb), err)
        _check(err)
        return tau, w0, g, tb

    def launch_count(self) -> int:
        n = C.c_int64(0)
        self._lib.clima_rt_get_launch_count(self._h, C.byref(n))
        return n.value

    def stream(self) -> int:
        """cudaStream_t (as an integer) the handle launches on."""
        p = C.c_void_p()
        self._lib.clima_rt_get_stream(self._h, C.byref(p))
        return p.value or 0

    def rorr_stats(self):
        """(units processed, units through the exact kernel, worklist overflows)."""
        a = (C.c_int64 * 3)()
        self._lib.clima_rt_get_rorr_stats(self._h, a)
        return tuple(a)

    def rorr_cut_stats(self):
        """dict of the staircase-kernel counters (clima_rt_get_rorr_cut_stats)."""
        a = (C.c_int64 * 8)()
        self._lib.clima_rt_get_rorr_cut_stats(self._h, a)
        k = ("warp_merges", "warp_passes", "restarts", "units_to_exact", "cold_warp_merges", "last_kernel",
             "closed_form_warp_merges", "jacobian_delta_calls")
        return dict(zip(k, list(a)[:8]))

    def stage_ms(self):
        a = (C.c_double * 6)()
        self._lib.clima_rt_get_stage_ms(self._h, a)
        return list(a)
