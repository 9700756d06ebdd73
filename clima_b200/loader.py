"""Load-time table packer (SURVEY.md §8 "next" row N2): turns the reference's data files into the
`OpacityTables` the device handle is created from.

Host code, runs once per handle; nothing here is on the timed path.  It mirrors, routine by routine,
  * `read_stellar_flux`      src/radtran/clima_radtran_types_create.f90:9-78
  * `create_RTChannel`       src/radtran/clima_radtran_types_create.f90:226-270
  * `create_RayleighXsection`src/radtran/clima_radtran_types_create.f90:1048-1088 + `rayleigh_vardavas`
                             src/clima_eqns.f90:240-246
  * `create_Ktable`          src/radtran/clima_radtran_types_create.f90:1265-1378
and the two futils routines the reader depends on (`addpnt`, `inter2`; futils v0.1.14 is a CPM
dependency that is not under /root/reference -- restated from the published TUV algorithm they
implement: Madronich's `addpnt`/`inter2`, grid insertion and bin averaging of a piecewise-linear
function).

HDF5: neither h5py nor libhdf5 exists in this image, so the HDF5 readers take any mapping-like object
with the h5py dataset interface (`f[name][...]`); `open_hdf5()` imports h5py lazily and fails loudly
when it is missing.  The tests drive the readers with in-memory stand-ins of the documented schema.
"""
from __future__ import annotations

from typing import Dict, Mapping, Sequence, Tuple

import numpy as np

from ._abi import KTable

C_LIGHT = 299792458.0  # src/clima_const.f90:13


class LoaderError(Exception):
    """The reference returns `err` strings from these routines; the packer raises them."""


# ----------------------------------------------------------------------------- atmosphere files
K_BOLTZ = 1.380649e-16  # erg/K, src/clima_const.f90:10


def read_atmosphere_file(path: str) -> Tuple[Sequence[str], np.ndarray]:
    """`AtmosphereFile` (src/clima_types_create.f90: create_AtmosphereFile): a header line of column
    labels, then one row per level.  Returns (labels, columns[nlabels, nz])."""
    with open(path) as f:
        lines = [ln for ln in f.read().splitlines() if ln.strip()]
    labels = lines[0].split()
    rows = [[float(x) for x in ln.split()] for ln in lines[1:]]
    if any(len(r) != len(labels) for r in rows):
        raise LoaderError(f'Number of columns is not consistent in "{path}"')
    return labels, np.asarray(rows, dtype=np.float64).T.copy()


def vertical_grid(bottom: float, top: float, nz: int) -> Tuple[np.ndarray, np.ndarray]:
    """src/clima_eqns.f90:172-184: uniform layers, z at the layer centres (cm)."""
    dz = np.full(nz, (top - bottom) / nz)
    z = np.empty(nz)
    z[0] = dz[0] / 2.0
    for i in range(1, nz):
        z[i] = z[i - 1] + dz[i]
    return z, dz


def unpack_atmospherefile(labels: Sequence[str], columns: np.ndarray, species_names: Sequence[str],
                          z: np.ndarray, filename: str = "atmosphere.txt"):
    """src/clima_types_create.f90:424-514: log10(mixing ratio), temperature and log10(pressure) of the
    file interpolated linearly in altitude (file `alt` is in km) onto the layer centres z (cm).
    futils `interp` is restated as linear interpolation that holds the end values outside the file's
    range (immaterial for the ModernEarth template: its 200 levels ARE vertical_grid(0, 1e7, 200))."""
    labels = list(labels)
    if "alt" not in labels:
        raise LoaderError(f'"alt" was not found in input file "{filename}"')
    alt = columns[labels.index("alt")] * 1.0e5
    mix = np.empty((len(z), len(species_names)), order="F")
    for i, sp in enumerate(species_names):
        if sp not in labels:
            raise LoaderError(f'Species "{sp}" was not found in "{filename}"')
        mix[:, i] = 10.0 ** np.interp(z, alt, np.log10(columns[labels.index(sp)]))
    if not np.all(np.isclose(mix.sum(1), 1.0, rtol=1.0e-2, atol=0.0)):
        raise LoaderError(f'mixing ratios do not sum to close to 1 in "{filename}"')
    for need in ("temp", "press"):
        if need not in labels:
            raise LoaderError(f'"{need}" was not found in input file "{filename}"')
    T = np.interp(z, alt, columns[labels.index("temp")])
    P = 10.0 ** np.interp(z, alt, np.log10(columns[labels.index("press")]))
    return mix, T, P


def column_from_atmosphere(labels, columns, species_names, nz: int, top: float = 1.0e7) -> Dict:
    """The single column of tests/test_radtran.f90:44-58: vertical_grid(0, top, nz), the file unpacked
    onto it, number densities P*1e6/(k T) times the mixing ratios.  Returns the `cols` dict of
    Radtran.radiate_batch with ncol = 1 (ground-first)."""
    z, dz = vertical_grid(0.0, top, nz)
    mix, T, P = unpack_atmospherefile(labels, columns, species_names, z)
    density = (P * 1.0e6) / (K_BOLTZ * T)
    dens = np.asfortranarray((mix * density[:, None])[:, :, None])
    return dict(nz=nz, ng=len(species_names), np=0, ncol=1, T_surface=np.array([T[0]]), T=np.asfortranarray(T[:, None]),
                P=np.asfortranarray(P[:, None]), densities=dens, dz=np.asfortranarray(dz[:, None]))


# ----------------------------------------------------------------------------- futils: addpnt, inter2
def addpnt(x: np.ndarray, y: np.ndarray, n: int, xnew: float, ynew: float) -> int:
    """Insert (xnew, ynew) into the ascending grid x[:n] (arrays have spare capacity); returns n+1.
    A point equal to an existing abscissa is an error, like the original (ierr = 3... family)."""
    if n + 1 > len(x):
        raise LoaderError("addpnt: not enough space in the arrays")
    if np.any(x[:n] == xnew):
        raise LoaderError("addpnt: the point is already on the grid")
    if n > 1 and np.any(np.diff(x[:n]) <= 0.0):
        raise LoaderError("addpnt: grid is not strictly ascending")
    ins = int(np.searchsorted(x[:n], xnew))
    x[ins + 1:n + 1] = x[ins:n].copy()
    y[ins + 1:n + 1] = y[ins:n].copy()
    x[ins], y[ins] = xnew, ynew
    return n + 1


def inter2(xg: np.ndarray, x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """Bin averages over the bins xg[i]..xg[i+1] of the piecewise-linear function through (x, y):
    yg[i] = (1/(xg[i+1]-xg[i])) * integral of y over the bin (trapezoids between data points, partial
    trapezoids at the bin edges).  Bins outside [x[0], x[-1]] are 0."""
    xg = np.asarray(xg, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    n, ng = len(x), len(xg)
    if np.any(np.diff(x) <= 0.0) or np.any(np.diff(xg) <= 0.0):
        raise LoaderError("inter2: grids must be strictly ascending")
    yg = np.zeros(ng - 1)
    jstart = 0
    for i in range(ng - 1):
        lo, hi = xg[i], xg[i + 1]
        if hi <= x[0] or lo >= x[n - 1]:
            continue
        # first data interval that reaches into the bin
        k = jstart
        while k < n - 1 and x[k + 1] <= lo:
            k += 1
        jstart = k
        area = 0.0
        while k < n - 1 and x[k] < hi:
            a1 = max(x[k], lo)
            a2 = min(x[k + 1], hi)
            if x[k + 1] == x[k]:
                darea = 0.0
            else:
                slope = (y[k + 1] - y[k]) / (x[k + 1] - x[k])
                b1 = y[k] + slope * (a1 - x[k])
                b2 = y[k] + slope * (a2 - x[k])
                darea = (a2 - a1) * (b2 + b1) / 2.0
            area += darea
            k += 1
        yg[i] = area / (hi - lo)
    return yg


# --------------------------------------------------------------------------------- stellar flux
def read_stellar_flux(star_file: str, wavl: np.ndarray) -> np.ndarray:
    """Stellar spectrum file (one header line, then `wavelength[nm] flux[mW/m2/nm]` rows) -> photon
    flux per bin of `wavl` in mW/m2/Hz (types_create.f90:9-78): zero padding points just outside the
    file's range and at 0 / huge, bin averages by inter2, nm -> Hz with the bin-centre wavelength."""
    try:
        with open(star_file) as f:
            lines = f.read().splitlines()
    except OSError:
        raise LoaderError(f"The input file {star_file} does not exist.")
    fw, ff = [], []
    for ln in lines[1:]:
        p = ln.replace(",", " ").split()
        if not p:
            break  # list-directed read ends at the first empty record
        try:
            fw.append(float(p[0].replace("d", "e").replace("D", "e")))
            ff.append(float(p[1].replace("d", "e").replace("D", "e")))
        except (ValueError, IndexError):
            raise LoaderError(f"Problem reading {star_file}")
    n = len(fw)
    file_wav = np.zeros(n + 4)
    file_flux = np.zeros(n + 4)
    file_wav[:n], file_flux[:n] = fw, ff
    rdelta = 1.0e-4
    try:
        i = n
        i = addpnt(file_wav, file_flux, i, file_wav[0] * (1.0 - rdelta), 0.0)
        i = addpnt(file_wav, file_flux, i, 0.0, 0.0)
        i = addpnt(file_wav, file_flux, i, file_wav[i - 1] * (1.0 + rdelta), 0.0)
        i = addpnt(file_wav, file_flux, i, np.finfo(np.float64).max, 0.0)
        flux = inter2(wavl, file_wav, file_flux)
    except LoaderError:
        raise LoaderError(f"Problem interpolating {star_file}")
    wavl = np.asarray(wavl, dtype=np.float64)
    wav_av = 0.5 * (wavl[:-1] + wavl[1:])
    return flux * (((wav_av * 1.0e-9) * wav_av) / C_LIGHT)


# ------------------------------------------------------------------------------------ Rayleigh
def rayleigh_vardavas(A: float, B: float, Delta: float, lam_nm):
    """Vardavas & Carver cross section [cm2] at wavelength lam_nm (clima_eqns.f90:240-246)."""
    lam_um = np.asarray(lam_nm, dtype=np.float64) * 1.0e-3
    l2 = lam_um * lam_um  # x**2.0_dp, x**4.0_dp: compilers expand these to products
    t = A * (1.0 + B / l2)
    return 4.577e-21 * ((6.0 + 3.0 * Delta) / (6.0 - 7.0 * Delta)) * (t * t) * (1.0 / (l2 * l2))


def rayleigh_table(rayleigh: Mapping, species_names: Sequence[str], wavl: np.ndarray,
                   filename: str = "rayleigh.yaml") -> Tuple[list, np.ndarray]:
    """`rayleigh` is the parsed rayleigh.yaml ({species: {data: {Delta, A, B}}}); returns
    (ray_sp_ind 1-based, ray_xs (nw, nray) Fortran-ordered), each cross section evaluated at the LOWER
    edge of its bin like create_RayleighXsection (types_create.f90:1083-1086)."""
    wavl = np.asarray(wavl, dtype=np.float64)
    inds, cols = [], []
    for i, sp in enumerate(species_names):
        if sp not in rayleigh:
            continue
        try:
            d = rayleigh[sp]["data"]
            Delta, A, B = float(d["Delta"]), float(d["A"]), float(d["B"])
        except (KeyError, TypeError) as e:
            raise LoaderError(f"{filename}: species {sp}: missing {e}")
        inds.append(i + 1)
        cols.append(rayleigh_vardavas(A, B, Delta, wavl[:-1]))
    if not cols:
        return [], np.zeros((len(wavl) - 1, 0), order="F")
    return inds, np.asfortranarray(np.stack(cols, axis=1))


# ------------------------------------------------------------------------------------ channels
def channel_indices(channel_wavl: np.ndarray, op_wavl: np.ndarray, filename: str = "bins.h5") -> Tuple[int, int]:
    """(ind_start, ind_end), 1-based inclusive bin range of a channel inside the opacity grid
    (create_RTChannel, types_create.f90:250-268): nearest edges, sizes must agree, edges within 1e-7
    relative."""
    cw = np.asarray(channel_wavl, dtype=np.float64)
    ow = np.asarray(op_wavl, dtype=np.float64)
    ind1 = int(np.argmin(np.abs(cw[0] - ow))) + 1
    ind2 = int(np.argmin(np.abs(cw[-1] - ow))) + 1
    msg = f'The wavelength bins "{filename}" are not compatible with the k-distribution wavelength bins.'
    if len(cw) != ind2 - ind1 + 1:
        raise LoaderError(msg)
    if not np.allclose(cw, ow[ind1 - 1:ind2], rtol=1.0e-7, atol=0.0):
        raise LoaderError(msg)
    return ind1, ind2 - 1


# ------------------------------------------------------------------------------------ k-tables
def open_hdf5(path: str):
    try:
        import h5py  # noqa: PLC0415  (absent from the build image; present wherever the data are)
    except ImportError as e:
        raise LoaderError(f'Failed to read "{path}": h5py is not installed') from e
    return h5py.File(path, "r")


def _dataset(h: Mapping, name: str, ndim: int, filename: str) -> np.ndarray:
    if name not in h:
        raise LoaderError(f'"{filename}" does not contain the dataset "{name}"')
    a = np.asarray(h[name][...], dtype=np.float64)
    if a.ndim != ndim:
        raise LoaderError(f'Dataset "{name}" in "{filename}" has the wrong number of dimensions')
    return a


def ktable_from_hdf5(h: Mapping, sp_ind: int, filename: str = "<memory>") -> Tuple[KTable, np.ndarray, np.ndarray]:
    """create_Ktable (types_create.f90:1265-1378): datasets `weights` (G), `log10P`, `T`, `wavelengths`
    (micron, nw+1) and `log10k`.  h5fortran hands Fortran the dims (G, nP, nT, nw); h5py shows the same
    bytes C-ordered, i.e. with shape (nw, nT, nP, G), so the Fortran array is its transpose.
    Returns (KTable, weights, wavl[nm])."""
    weights = _dataset(h, "weights", 1, filename)
    log10P = _dataset(h, "log10P", 1, filename)
    temp = _dataset(h, "T", 1, filename)
    wavl = _dataset(h, "wavelengths", 1, filename) * 1.0e3
    lk = _dataset(h, "log10k", 4, filename)
    G, nP, nT, nw = len(weights), len(log10P), len(temp), len(wavl) - 1
    if lk.shape != (nw, nT, nP, G):
        raise LoaderError(f'Dataset "log10k" in "{filename}" has shape {lk.shape}, expected {(nw, nT, nP, G)}')
    if np.any(np.diff(log10P) <= 0.0) or np.any(np.diff(temp) <= 0.0):
        # the reference's interpolator initialisation fails on non-monotone axes (:1358-1366)
        raise LoaderError(f'Failed to initialize interpolator for "{filename}"')
    return KTable(sp_ind=sp_ind, log10P=log10P, temp=temp, log10k=np.asfortranarray(lk.transpose(3, 2, 1, 0))), weights, wavl


def check_same_grid(tables: Dict[str, Tuple[KTable, np.ndarray, np.ndarray]]) -> Tuple[np.ndarray, np.ndarray]:
    """All k-tables of a handle share the g-weights and the wavelength grid
    (create_OpticalProperties, types_create.f90:140-189)."""
    names = list(tables)
    _, w0, wl0 = tables[names[0]]
    for nm in names[1:]:
        _, w, wl = tables[nm]
        if len(w) != len(w0) or not np.allclose(w, w0, rtol=1e-10, atol=0.0):
            raise LoaderError(f'k-distribution "{nm}" has different weights than "{names[0]}"')
        if len(wl) != len(wl0) or not np.allclose(wl, wl0, rtol=1e-10, atol=0.0):
            raise LoaderError(f'k-distribution "{nm}" has different wavelength bins than "{names[0]}"')
    return w0, wl0
