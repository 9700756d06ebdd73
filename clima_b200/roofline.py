"""Algorithmic work of the RT hot path per column (restated in DESIGN.md).

Counting convention: every +,-,*,/,sqrt,exp,compare counts 1; an FMA-able a*b+c counts 2; a
comparison sort of n keys counts n*log2(n); work that is provably invariant is counted once
(duplicated layers of the doubled grid, zenith-independent two-stream terms)."""
from __future__ import annotations

import math
from dataclasses import dataclass


# DRAM traffic of the dominant kernel (rorr_fast::optics_rorr_kernel<16,1,1>) per column, from
# profiles/r1p_ncu_full_summary.txt (ncu --set full, 256 columns of the cfg2 workload):
# (73,949,000 + 622,805,000) bytes / 256 columns.  Almost all of it is the kernel's OUTPUT
# (tau and w0 of every unit, 2 x 8 x G bytes x nw x unique layers = 2.59 MB per column).
RORR_FAST_DRAM_BYTES_PER_COLUMN = (73949000 + 622805000) / 256.0
# The same for rorr_cut::rorr_cut_kernel<16,true> (the default kernel on sweeps), ncu --set full at
# 1024 columns of the cfg2 workload, column chains (profiles/r2b_ncu_rorr_cut_summary.txt):
# dram__bytes_read.sum 443.0 MB + dram__bytes_write.sum 2742.2 MB; the writes are the kernel's
# output (tau, w0 of every unit: 2.59 MB per column), which the two-stream kernels read back.
RORR_CUT_DRAM_BYTES_PER_COLUMN = (443.00032e6 + 2742.222e6) / 1024.0
RORR_DRAM_SOURCE = {
    "fast": "ncu r1p capture at 256 columns (profiles/r1p_ncu_full_summary.txt), scaled per column",
    "cut": "ncu r2b capture at 1024 columns (profiles/r2b_ncu_rorr_cut_summary.txt), scaled per column",
}


@dataclass
class Dims:
    nw: int      # opacity-grid bins
    L_ir: int    # IR channel bins
    L_sol: int   # solar channel bins
    G: int       # ngauss
    K: int       # k-distribution gases
    C: int       # CIA pairs
    R: int       # Rayleigh species
    X: int       # photolysis cross sections
    Z: int       # zenith angles
    nz: int      # RT layers (nz_r)
    U: int       # unique layers
    ng: int      # gases in the model


def stage_ops(d: Dims) -> dict:
    G2 = d.G * d.G
    ops = {
        "k_interp": d.nw * d.U * d.K * d.G * 11,
        "continua": d.nw * d.U * (10 * d.C + 2 * d.R + 2 * d.X + 18) + d.nw * d.nz * 10 * d.G,
        "rorr": d.nw * d.U * (d.K - 1) * (G2 * (5 + 2 * math.log2(d.G)) + 2 * d.G),
        "ir": d.L_ir * d.G * d.nz * 88 + d.L_ir * (d.nz + 1) * 5,
        "solar": d.L_sol * d.G * d.nz * (55 + 70 * d.Z),
        "integrate": (4 * d.L_ir + 6 * d.L_sol * d.Z) * d.G * (d.nz + 1)
                     + 4 * (d.L_ir + d.L_sol) * (d.nz + 1),
    }
    ops["total"] = sum(ops.values())
    return ops


def column_bytes(d: Dims) -> dict:
    """Algorithmic HBM bytes per column: inputs once in, integrated flux profiles once out."""
    inp = d.nz * (3 + d.ng) * 8 + 8
    out = 5 * (d.nz + 1) * 8 + 16
    return {"in": inp, "out": out, "total": inp + out}


def dims_of(tables, params, cols) -> Dims:
    import numpy as np
    nz = cols["nz"]
    # unique layers of column 0 (consecutive duplicates share a slot, csrc/prep.cu)
    T, P, dz = cols["T"][:, 0], cols["P"][:, 0], cols["dz"][:, 0]
    dens = cols["densities"][:, :, 0]
    same = (T[1:] == T[:-1]) & (P[1:] == P[:-1]) & (dz[1:] == dz[:-1]) & np.all(dens[1:] == dens[:-1], axis=1)
    U = int(nz - np.count_nonzero(same))
    return Dims(nw=tables.nw, L_ir=tables.nw_ir, L_sol=tables.nw_sol, G=tables.ngauss, K=len(tables.k),
                C=len(tables.cia), R=len(tables.ray_sp_ind), X=len(tables.pxs), Z=len(params.zenith_u),
                nz=nz, U=U, ng=tables.ng)
