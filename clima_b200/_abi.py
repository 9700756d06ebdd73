"""ctypes mirror of include/clima_rt.h (the C-ABI structs) and the Python-side table container.

`OpacityTables` mirrors what `type OpticalProperties` + the two `RTChannel`s hold after
`create_OpticalProperties` / `create_RTChannel` have run
(reference: src/radtran/clima_radtran_types.f90:23-141,263-269;
src/radtran/clima_radtran_types_create.f90:226-270,272-645).  All arrays are float64,
column-major (Fortran order) where multi-dimensional, species indices 1-based like the Fortran.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

ERR_LEN = 1024  # clima/fortran/clima_c_api.f90:5

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_double_pp = C.POINTER(c_double_p)


class CTables(C.Structure):
    _fields_ = [
        ("nw", C.c_int32), ("wavl", c_double_p),
        ("ng", C.c_int32), ("np", C.c_int32),
        ("nk", C.c_int32), ("ngauss", C.c_int32), ("weights", c_double_p),
        ("k_sp_ind", c_int32_p), ("k_npress", c_int32_p), ("k_ntemp", c_int32_p),
        ("k_log10P", c_double_pp), ("k_temp", c_double_pp), ("k_log10k", c_double_pp),
        ("ncia", C.c_int32), ("cia_sp_ind", c_int32_p), ("cia_dim", c_int32_p),
        ("cia_ntemp", c_int32_p), ("cia_temp", c_double_pp), ("cia_data", c_double_pp),
        ("nray", C.c_int32), ("ray_sp_ind", c_int32_p), ("ray_xs", c_double_p),
        ("npxs", C.c_int32), ("pxs_sp_ind", c_int32_p), ("pxs_dim", c_int32_p),
        ("pxs_ntemp", c_int32_p), ("pxs_temp", c_double_pp), ("pxs_data", c_double_pp),
        ("has_cont", C.c_int32), ("cont_LH2O", C.c_int32), ("cont_ntemp", C.c_int32),
        ("cont_temp", c_double_p), ("cont_log10_H2O", c_double_p), ("cont_log10_foreign", c_double_p),
        ("npart", C.c_int32), ("part_p_ind", c_int32_p), ("part_nrad", c_int32_p),
        ("part_radii", c_double_pp), ("part_w0", c_double_pp), ("part_qext", c_double_pp),
        ("part_gt", c_double_pp),
        ("ir_ind_start", C.c_int32), ("ir_ind_end", C.c_int32),
        ("sol_ind_start", C.c_int32), ("sol_ind_end", C.c_int32),
        ("photons_sol", c_double_p),
    ]


class CParams(C.Structure):
    _fields_ = [
        ("diurnal_fac", C.c_double), ("photon_scale_factor", C.c_double),
        ("ir_tau_min", C.c_double), ("has_hard_surface", C.c_int32), ("nzenith", C.c_int32),
        ("zenith_u", c_double_p), ("zenith_weights", c_double_p),
        ("surface_albedo", c_double_p), ("surface_emissivity", c_double_p),
    ]


OUT_FIELDS = ["fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR",
              "fup_a_ir", "fdn_a_ir", "tau_band_ir", "fup_a_sol", "fdn_a_sol", "amean_sol",
              "tau_band_sol"]


class COut(C.Structure):
    _fields_ = [(n, c_double_p) for n in OUT_FIELDS] + [("status", c_int32_p)]


def _f64(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="K"))


def dptr(a: Optional[np.ndarray]):
    if a is None:
        return c_double_p()
    assert a.dtype == np.float64
    return a.ctypes.data_as(c_double_p)


def iptr(a: Optional[np.ndarray]):
    if a is None:
        return c_int32_p()
    assert a.dtype == np.int32
    return a.ctypes.data_as(c_int32_p)


def _pp(arrs: List[np.ndarray]):
    n = max(len(arrs), 1)
    arr = (c_double_p * n)()
    for i, a in enumerate(arrs):
        arr[i] = dptr(a)
    return arr


@dataclass
class KTable:  # type Ktable, clima_radtran_types.f90:23-38
    sp_ind: int
    log10P: np.ndarray  # (npress) log10 bar
    temp: np.ndarray    # (ntemp)
    log10k: np.ndarray  # (ngauss, npress, ntemp, nw) Fortran-ordered


@dataclass
class XS1D:  # type Xsection (CIA / photolysis), clima_radtran_types.f90:44-55
    sp_ind: List[int]
    dim: int
    temp: Optional[np.ndarray]  # (ntemp) for dim 1
    data: np.ndarray            # dim 1: log10 xs (ntemp, nw) F-order; dim 0: xs_0d (nw)


@dataclass
class Particle:  # type ParticleXsection, clima_radtran_types.f90:57-66
    p_ind: int
    radii: np.ndarray  # (nrad) cm
    w0: np.ndarray     # (nrad, nw) F-order
    qext: np.ndarray
    gt: np.ndarray


@dataclass
class OpacityTables:
    wavl: np.ndarray            # (nw+1) nm
    species_names: List[str]
    particle_names: List[str]
    weights: np.ndarray         # (ngauss)
    k: List[KTable]
    cia: List[XS1D] = field(default_factory=list)
    ray_sp_ind: List[int] = field(default_factory=list)
    ray_xs: Optional[np.ndarray] = None  # (nw, nray) F-order
    pxs: List[XS1D] = field(default_factory=list)
    cont_LH2O: int = 0                   # 0 = no continuum
    cont_temp: Optional[np.ndarray] = None
    cont_log10_H2O: Optional[np.ndarray] = None      # (ntemp, nw) F-order
    cont_log10_foreign: Optional[np.ndarray] = None
    part: List[Particle] = field(default_factory=list)
    ir_ind: tuple = (1, 1)   # (ind_start, ind_end) 1-based inclusive
    sol_ind: tuple = (1, 1)
    photons_sol: Optional[np.ndarray] = None  # (nw_sol)

    @property
    def nw(self): return len(self.wavl) - 1
    @property
    def ng(self): return len(self.species_names)
    @property
    def np_(self): return len(self.particle_names)
    @property
    def ngauss(self): return len(self.weights)
    @property
    def nw_ir(self): return self.ir_ind[1] - self.ir_ind[0] + 1
    @property
    def nw_sol(self): return self.sol_ind[1] - self.sol_ind[0] + 1
    @property
    def freq(self):  # types_create.f90:360-361
        return 299792458.0 / (self.wavl * 1.0e-9)

    def to_c(self):
        """Returns (CTables, keepalive)."""
        keep = []

        def F(a):  # flatten in Fortran order
            b = np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="F"))
            keep.append(b)
            return b

        def I(a):
            b = np.ascontiguousarray(np.asarray(a, dtype=np.int32).ravel(order="F"))
            keep.append(b)
            return b

        t = CTables()
        t.nw = self.nw
        t.wavl = dptr(F(self.wavl))
        t.ng = self.ng
        t.np = self.np_
        t.nk = len(self.k)
        t.ngauss = self.ngauss
        t.weights = dptr(F(self.weights))
        t.k_sp_ind = iptr(I([k.sp_ind for k in self.k]))
        t.k_npress = iptr(I([len(k.log10P) for k in self.k]))
        t.k_ntemp = iptr(I([len(k.temp) for k in self.k]))
        for name, arrs in (("k_log10P", [F(k.log10P) for k in self.k]),
                           ("k_temp", [F(k.temp) for k in self.k]),
                           ("k_log10k", [F(k.log10k) for k in self.k])):
            pp = _pp(arrs)
            keep.append(pp)
            setattr(t, name, C.cast(pp, c_double_pp))

        def xs_group(prefix, lst, two_species):
            setattr(t, "n" + prefix, len(lst))
            sp = []
            for x in lst:
                sp.extend(x.sp_ind[:2] if two_species else x.sp_ind[:1])
            setattr(t, prefix + "_sp_ind", iptr(I(sp if sp else [0])))
            setattr(t, prefix + "_dim", iptr(I([x.dim for x in lst] or [0])))
            setattr(t, prefix + "_ntemp", iptr(I([len(x.temp) if x.dim == 1 else 0 for x in lst] or [0])))
            temps = _pp([F(x.temp) if x.dim == 1 else F([0.0]) for x in lst])
            datas = _pp([F(x.data) for x in lst])
            keep.extend([temps, datas])
            setattr(t, prefix + "_temp", C.cast(temps, c_double_pp))
            setattr(t, prefix + "_data", C.cast(datas, c_double_pp))

        xs_group("cia", self.cia, True)
        xs_group("pxs", self.pxs, False)
        t.nray = len(self.ray_sp_ind)
        t.ray_sp_ind = iptr(I(self.ray_sp_ind or [0]))
        t.ray_xs = dptr(F(self.ray_xs if self.ray_xs is not None else [0.0]))
        t.has_cont = 1 if self.cont_LH2O > 0 else 0
        t.cont_LH2O = self.cont_LH2O
        if t.has_cont:
            t.cont_ntemp = len(self.cont_temp)
            t.cont_temp = dptr(F(self.cont_temp))
            t.cont_log10_H2O = dptr(F(self.cont_log10_H2O))
            t.cont_log10_foreign = dptr(F(self.cont_log10_foreign))
        t.npart = len(self.part)
        t.part_p_ind = iptr(I([p.p_ind for p in self.part] or [0]))
        t.part_nrad = iptr(I([len(p.radii) for p in self.part] or [0]))
        for name, arrs in (("part_radii", [F(p.radii) for p in self.part]),
                           ("part_w0", [F(p.w0) for p in self.part]),
                           ("part_qext", [F(p.qext) for p in self.part]),
                           ("part_gt", [F(p.gt) for p in self.part])):
            pp = _pp(arrs)
            keep.append(pp)
            setattr(t, name, C.cast(pp, c_double_pp))
        t.ir_ind_start, t.ir_ind_end = self.ir_ind
        t.sol_ind_start, t.sol_ind_end = self.sol_ind
        t.photons_sol = dptr(F(self.photons_sol))
        return t, keep


@dataclass
class RadtranParams:
    """Public run-time members of `type Radtran` (src/radtran/clima_radtran.f90:51-68)."""
    zenith_u: np.ndarray
    zenith_weights: np.ndarray
    surface_albedo: np.ndarray      # (nw_sol)
    surface_emissivity: np.ndarray  # (nw_ir)
    diurnal_fac: float = 0.5
    photon_scale_factor: float = 1.0
    ir_tau_min: float = 1.0e-6
    has_hard_surface: bool = True

    def to_c(self):
        keep = [np.ascontiguousarray(np.asarray(a, dtype=np.float64)) for a in
                (self.zenith_u, self.zenith_weights, self.surface_albedo, self.surface_emissivity)]
        p = CParams()
        p.diurnal_fac = self.diurnal_fac
        p.photon_scale_factor = self.photon_scale_factor
        p.ir_tau_min = self.ir_tau_min
        p.has_hard_surface = 1 if self.has_hard_surface else 0
        p.nzenith = len(keep[0])
        p.zenith_u, p.zenith_weights, p.surface_albedo, p.surface_emissivity = (dptr(a) for a in keep)
        return p, keep


def zenith_angles_and_weights(n: int):
    """cos(zenith) nodes and weights: Gauss-Legendre on mu in [0,1], weights w/2
    (src/clima_eqns.f90:26-41, src/radtran/clima_radtran.f90:162-165; futils `gauss_legendre`
    == numpy.polynomial.legendre.leggauss)."""
    x, w = np.polynomial.legendre.leggauss(n)
    mu = x / 2.0 + 1.0 / 2.0
    ang = np.arccos(mu) * 180.0 / np.pi
    return np.cos(ang * np.pi / 180.0), w / 2.0


def alloc_out(nz: int, ncol: int, nw_ir: int, nw_sol: int, spectral: bool):
    """Allocate result arrays (Fortran order, column index last) and the COut struct."""
    L = nz + 1
    arrs = {
        "fup_n_ir": np.zeros((L, ncol), order="F"), "fdn_n_ir": np.zeros((L, ncol), order="F"),
        "fup_n_sol": np.zeros((L, ncol), order="F"), "fdn_n_sol": np.zeros((L, ncol), order="F"),
        "f_total": np.zeros((L, ncol), order="F"),
        "ISR": np.zeros(ncol), "OLR": np.zeros(ncol),
    }
    if spectral:
        arrs.update({
            "fup_a_ir": np.zeros((L, nw_ir, ncol), order="F"),
            "fdn_a_ir": np.zeros((L, nw_ir, ncol), order="F"),
            "tau_band_ir": np.zeros((nz, nw_ir, ncol), order="F"),
            "fup_a_sol": np.zeros((L, nw_sol, ncol), order="F"),
            "fdn_a_sol": np.zeros((L, nw_sol, ncol), order="F"),
            "amean_sol": np.zeros((L, nw_sol, ncol), order="F"),
            "tau_band_sol": np.zeros((nz, nw_sol, ncol), order="F"),
        })
    status = np.zeros(ncol, dtype=np.int32)
    o = COut()
    for n in OUT_FIELDS:
        setattr(o, n, dptr(arrs[n]) if n in arrs else c_double_p())
    o.status = iptr(status)
    arrs["status"] = status
    return o, arrs
