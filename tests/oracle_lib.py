"""ctypes loader for the CPU oracle (oracle/oracle.cpp).  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from clima_b200._abi import (CParams, COut, CTables, ERR_LEN, OpacityTables, RadtranParams, alloc_out,
                             dptr, iptr)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_libs = {}


def build_oracle():
    subprocess.run(["make", "-C", ORACLE_DIR, "-s"], check=True)


def load(kind="parity"):
    if kind in _libs:
        return _libs[kind]
    path = os.path.join(ORACLE_DIR, f"liboracle_{kind}.so")
    if not os.path.exists(path):
        build_oracle()
    lib = C.CDLL(path)
    lib.oracle_max_threads.restype = C.c_int32
    _libs[kind] = lib
    return lib


def _i(v):
    return C.byref(C.c_int32(int(v)))


def _d(v):
    return C.byref(C.c_double(float(v)))


class OracleRadtran:
    """Single-column `Radtran` (src/radtran/clima_radtran.f90:31-85) on the CPU oracle, plus the
    batched loop used as CPU baseline."""

    def __init__(self, tables: OpacityTables, params: RadtranParams, kind="parity", nthreads_bins=1):
        self.lib = load(kind)
        self.tables = tables
        self.params = params
        self._ct, self._keep_t = tables.to_c()
        self._cp, self._keep_p = params.to_c()
        self._h = C.c_void_p()
        err = C.create_string_buffer(ERR_LEN + 1)
        self.lib.oracle_create(C.byref(self._ct), C.byref(self._cp), _i(nthreads_bins), C.byref(self._h), err)
        assert not err.value, err.value

    def __del__(self):
        try:
            if self._h:
                self.lib.oracle_destroy(self._h)
        except Exception:
            pass

    def set_params(self, params: RadtranParams):
        self.params = params
        self._cp, self._keep_p = params.to_c()
        err = C.create_string_buffer(ERR_LEN + 1)
        self.lib.oracle_set_params(self._h, C.byref(self._cp), err)
        assert not err.value, err.value

    def set_custom(self, log10P, dtau_dz, w0, g0):
        arrs = [np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="F"))
                for a in (log10P, dtau_dz, w0, g0)]
        err = C.create_string_buffer(ERR_LEN + 1)
        self.lib.oracle_set_custom_optical_properties(self._h, _i(len(arrs[0])), *(dptr(a) for a in arrs), err)
        assert not err.value, err.value

    def unset_custom(self):
        self.lib.oracle_unset_custom_optical_properties(self._h)

    def radiate_batch(self, cols, compute_solar=True, compute_opacity=True, spectral=False, nthreads=1):
        nz, ng, npp, ncol = cols["nz"], cols["ng"], cols["np"], cols["ncol"]
        o, arrs = alloc_out(nz, ncol, self.tables.nw_ir, self.tables.nw_sol, spectral)
        err = C.create_string_buffer(ERR_LEN + 1)
        pd = dptr(cols["pdensities"]) if "pdensities" in cols else dptr(None)
        rd = dptr(cols["radii"]) if "radii" in cols else dptr(None)
        self.lib.oracle_radiate_batch(self._h, _i(ncol), _i(nz), _i(ng), _i(npp), dptr(cols["T_surface"]),
                                      dptr(cols["T"]), dptr(cols["P"]), dptr(cols["densities"]),
                                      dptr(cols["dz"]), pd, rd, _i(compute_solar), _i(compute_opacity),
                                      _i(nthreads), C.byref(o), err)
        if err.value.strip():
            raise RuntimeError(err.value.decode())
        return arrs

    def radiate(self, cols, icol=0, compute_solar=True, compute_opacity=True, spectral=True):
        """One column through the stateful single-column path (keeps opacity / stale solar)."""
        nz, ng, npp = cols["nz"], cols["ng"], cols["np"]
        o, arrs = alloc_out(nz, 1, self.tables.nw_ir, self.tables.nw_sol, spectral)
        err = C.create_string_buffer(ERR_LEN + 1)

        def col(a):
            return np.ascontiguousarray(np.asarray(a)[..., icol].ravel(order="F"))
        T, P, D, dz = col(cols["T"]), col(cols["P"]), col(cols["densities"]), col(cols["dz"])
        pdv = col(cols["pdensities"]) if "pdensities" in cols else None
        rdv = col(cols["radii"]) if "radii" in cols else None
        self.lib.oracle_radiate(self._h, _i(nz), _i(ng), _i(npp), _d(cols["T_surface"][icol]), dptr(T), dptr(P),
                                dptr(D), dptr(dz), dptr(pdv), dptr(rdv), _i(compute_solar),
                                _i(compute_opacity), C.byref(o), err)
        if err.value.strip():
            raise RuntimeError(err.value.decode())
        return arrs

    def get_opacity(self, nz):
        G, nw = self.tables.ngauss, self.tables.nw
        tau = np.zeros((nz, G, nw), order="F")
        w0 = np.zeros((nz, G, nw), order="F")
        g = np.zeros((nz, nw), order="F")
        tb = np.zeros((nz, nw), order="F")
        self.lib.oracle_get_opacity(self._h, dptr(tau), dptr(w0), dptr(g), dptr(tb))
        return tau, w0, g, tb
