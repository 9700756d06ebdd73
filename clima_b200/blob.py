"""Flat binary case file for callers of the C ABI that are not Python (tests/c_driver.c; a Fortran
`Radtran_pack_tables` would fill `clima_rt_tables` directly from `self%op`, fortran/clima_radtran_pack.f90).

Layout: the magic "CRTB1\\0\\0\\0", then records in the fixed order of `write_case`; a record is an
int32 kind (0 = int32 data, 1 = float64 data), an int32 count n and n values.  Arrays are Fortran
ordered exactly as `clima_rt_tables` / `clima_rt_params` / the batched inputs take them."""
from __future__ import annotations

import struct
from typing import Dict, Optional

import numpy as np

from ._abi import OpacityTables, RadtranParams

MAGIC = b"CRTB1\0\0\0"


def _w(f, a, kind):
    a = np.asarray(a, dtype=np.int32 if kind == 0 else np.float64).ravel(order="F")
    f.write(struct.pack("<ii", kind, a.size))
    f.write(a.tobytes())


def write_case(path: str, tb: OpacityTables, pr: RadtranParams, cols: Dict, golden: Dict,
               jac: Optional[Dict] = None):
    """tables, params, one batch of columns, the expected integrated results (`golden`: the oracle's
    fup_n_ir ... OLR) and optionally a Jacobian case (`jac`: T_surface(ncol), T(nzm,ncol), f(nzm+1,ncol))."""
    I = lambda f, a: _w(f, a, 0)  # noqa: E731
    D = lambda f, a: _w(f, a, 1)  # noqa: E731
    with open(path, "wb") as f:
        f.write(MAGIC)
        has_cont = 1 if tb.cont_LH2O > 0 else 0
        I(f, [tb.nw, tb.ng, tb.np_, len(tb.k), tb.ngauss, len(tb.cia), len(tb.ray_sp_ind), len(tb.pxs), has_cont,
              tb.cont_LH2O, len(tb.cont_temp) if has_cont else 0, len(tb.part), tb.ir_ind[0], tb.ir_ind[1],
              tb.sol_ind[0], tb.sol_ind[1]])
        D(f, tb.wavl); D(f, tb.weights)
        I(f, [k.sp_ind for k in tb.k]); I(f, [len(k.log10P) for k in tb.k]); I(f, [len(k.temp) for k in tb.k])
        for k in tb.k:
            D(f, k.log10P); D(f, k.temp); D(f, k.log10k)
        for lst, two in ((tb.cia, True), (tb.pxs, False)):
            sp = []
            for x in lst:
                sp.extend(x.sp_ind[:2] if two else x.sp_ind[:1])
            I(f, sp); I(f, [x.dim for x in lst]); I(f, [len(x.temp) if x.dim == 1 else 0 for x in lst])
            for x in lst:
                D(f, x.temp if x.dim == 1 else [0.0]); D(f, x.data)
            if two:  # Rayleigh sits between CIA and photolysis in clima_rt_tables
                I(f, tb.ray_sp_ind); D(f, tb.ray_xs if tb.ray_xs is not None else [0.0])
        D(f, tb.cont_temp if has_cont else [0.0]); D(f, tb.cont_log10_H2O if has_cont else [0.0])
        D(f, tb.cont_log10_foreign if has_cont else [0.0])
        I(f, [p.p_ind for p in tb.part]); I(f, [len(p.radii) for p in tb.part])
        for p in tb.part:
            D(f, p.radii); D(f, p.w0); D(f, p.qext); D(f, p.gt)
        D(f, tb.photons_sol)
        D(f, [pr.diurnal_fac, pr.photon_scale_factor, pr.ir_tau_min])
        I(f, [1 if pr.has_hard_surface else 0, len(pr.zenith_u)])
        D(f, pr.zenith_u); D(f, pr.zenith_weights); D(f, pr.surface_albedo); D(f, pr.surface_emissivity)
        I(f, [cols["ncol"], cols["nz"], cols["ng"], cols["np"]])
        for k in ("T_surface", "T", "P", "densities", "dz"):
            D(f, cols[k])
        if cols["np"] > 0:
            D(f, cols["pdensities"]); D(f, cols["radii"])
        for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR"):
            D(f, golden[k])
        if jac is None:
            I(f, [0, 0, 0])
        else:
            I(f, [1, jac["T"].shape[1], jac["T"].shape[0]])
            D(f, jac["T_surface"]); D(f, jac["T"]); D(f, jac["f"])
