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


# =====================================================================================================
# N2 continued: the remaining readers of clima_radtran_types_create.f90 and the selection logic of
# create_OpticalProperties.  Every HDF5 reader takes a mapping with the h5py dataset interface.
# =====================================================================================================
LOG10TINY = float(np.log10(np.sqrt(np.finfo(np.float64).tiny)))  # src/clima_const.f90:21


def read_wavl(h: Mapping, channel: str, filename: str = "bins.h5") -> np.ndarray:
    """`read_wavl` (types_create.f90:647-687): channel bin edges "sol_wavl" / "ir_wavl" in micron -> nm."""
    name = {"solar": "sol_wavl", "sol": "sol_wavl", "ir": "ir_wavl"}.get(channel)
    if name is None:
        raise LoaderError("read_wavl received an invalid channel_type")
    return _dataset(h, name, 1, f"{filename}/{name}") * 1.0e3


def _rebin_rows(wav_um: np.ndarray, rows: np.ndarray, wavl: np.ndarray, filename: str) -> np.ndarray:
    """The block repeated in read_h5_Xsection (:1181-1236) and create_WaterContinuum (:945-1040): for
    every row (one temperature) of log10 values on the file's wavelength grid, pad with `log10tiny`
    just outside the range and at 0 / huge (four `addpnt`), then bin-average onto the opacity grid
    with `inter2`.  rows: (n, nwf) -> (n, nw)."""
    wav = np.asarray(wav_um, dtype=np.float64) * 1.0e3  # micron -> nm
    n, nwf = rows.shape
    out = np.empty((n, len(wavl) - 1))
    rdelta = 1.0e-4
    for i in range(n):
        x = np.zeros(nwf + 4)
        y = np.zeros(nwf + 4)
        x[:nwf], y[:nwf] = wav, rows[i]
        try:
            k = nwf
            k = addpnt(x, y, k, x[0] * (1.0 - rdelta), LOG10TINY)
            k = addpnt(x, y, k, 0.0, LOG10TINY)
            k = addpnt(x, y, k, x[k - 1] * (1.0 + rdelta), LOG10TINY)
            k = addpnt(x, y, k, np.finfo(np.float64).max, LOG10TINY)
            out[i] = inter2(wavl, x, y)
        except LoaderError:
            raise LoaderError(f'Problem interpolating data in "{filename}"')
    return out


def xsection_from_hdf5(h: Mapping, sp_ind: Sequence[int], wavl: np.ndarray, filename: str = "<memory>"):
    """`read_h5_Xsection` (types_create.f90:1106-1263; used by create_CIAXsection :1090-1104 and the
    absorption cross sections): datasets "wavelengths" [micron], "log10xs" (nwf) or (ntemp, nwf),
    "T" (ntemp).  dim 0 -> xs_0d(nw) = 10**rebinned; dim 1 -> log10 xs (ntemp, nw), interpolated
    linearly in T at run time."""
    from ._abi import XS1D
    if "log10xs" not in h:
        raise LoaderError(f'{filename}: dataset "log10xs" does not exist')
    data = np.asarray(h["log10xs"][...], dtype=np.float64)
    dim = data.ndim - 1
    if dim not in (0, 1):
        raise LoaderError(f"Issue reading {filename}")
    wav = _dataset(h, "wavelengths", 1, filename)
    if dim == 0:
        if data.shape[0] != wav.shape[0]:
            raise LoaderError(f'"log10xs" has a bad dimension in "{filename}"')
        xs0 = 10.0 ** _rebin_rows(wav, data[None, :], wavl, filename)[0]
        return XS1D(list(sp_ind), 0, None, xs0)
    temp = _dataset(h, "T", 1, filename)
    if data.shape[0] != temp.shape[0]:
        raise LoaderError(f'"log10xs" has a bad dimension in "{filename}"')
    return XS1D(list(sp_ind), 1, temp, np.asfortranarray(_rebin_rows(wav, data, wavl, filename)))


def water_continuum_from_hdf5(h: Mapping, species_names: Sequence[str], wavl: np.ndarray, model: str = "MT_CKD",
                              filename: str = "<memory>") -> Dict:
    """`create_WaterContinuum` (types_create.f90:868-1046): "wavelengths" [micron], "T", "log10xs_H2O" and
    "log10xs_foreign" (ntemp, nwf).  Returns the keyword arguments of OpacityTables."""
    species_names = list(species_names)
    if "H2O" not in species_names:
        raise LoaderError('"H2O" must be a species to include the "continuum" opacity')
    if not len(species_names) > 1:
        raise LoaderError('There must be more than 1 species in order to use the "continuum" opacity')
    wav = _dataset(h, "wavelengths", 1, filename)
    temp = _dataset(h, "T", 1, filename)
    out = dict(cont_LH2O=species_names.index("H2O") + 1, cont_temp=temp)
    for ds, key in (("log10xs_H2O", "cont_log10_H2O"), ("log10xs_foreign", "cont_log10_foreign")):
        d = _dataset(h, ds, 2, filename)
        if d.shape[0] != temp.shape[0]:
            raise LoaderError(f'"{ds}" has a bad dimension in "{filename}"')
        out[key] = np.asfortranarray(_rebin_rows(wav, d, wavl, filename))
    return out


def interp_discrete_to_bins(wavl: np.ndarray, wv: np.ndarray, f: np.ndarray, mode: str, fill: float = 0.0) -> np.ndarray:
    """futils `interp_discrete_to_bins` (v0.1.14, not under /root/reference; restated from its name, its
    arguments and its two call sites, types_create.f90:830-838 and :1461): discrete samples (wv, f) ->
    one value per bin of `wavl` = the mean over the bin of the piecewise-linear function through the
    samples; outside the sampled range the function is held at its end value ('Constant') or set to
    `fill` ('FillValue').  PARITY UNPINNED like every futils routine."""
    wavl = np.asarray(wavl, dtype=np.float64)
    wv = np.asarray(wv, dtype=np.float64)
    f = np.asarray(f, dtype=np.float64)
    if np.any(np.diff(wv) <= 0.0):
        raise LoaderError("interp_discrete_to_bins: wavelengths must be strictly ascending")
    lo, hi = (f[0], f[-1]) if mode == "Constant" else (fill, fill)
    if mode not in ("Constant", "FillValue"):
        raise LoaderError(f'interp_discrete_to_bins: unknown extrapolation "{mode}"')
    big = np.finfo(np.float64).max
    x = np.concatenate(([-big, np.nextafter(wv[0], -np.inf)], wv, [np.nextafter(wv[-1], np.inf), big]))
    y = np.concatenate(([lo, lo], f, [hi, hi]))
    out = np.empty(len(wavl) - 1)
    for i in range(len(wavl) - 1):
        a, b = wavl[i], wavl[i + 1]
        inner = x[(x > a) & (x < b)]
        xs = np.concatenate(([a], inner, [b]))
        ys = np.interp(xs, x, y)
        out[i] = np.sum(0.5 * (ys[1:] + ys[:-1]) * np.diff(xs)) / (b - a)
    return out


def particle_from_hdf5(h: Mapping, p_ind: int, wavl: np.ndarray, filename: str = "<memory>"):
    """`create_ParticleXsection` (types_create.f90:734-866): "wavelengths" (on the opacity grid's unit,
    nm), "radii" [micron -> cm], "w0", "qext", "g0" (nrad, nwf), each row put on the bins with
    interp_discrete_to_bins(..., 'Constant')."""
    from ._abi import Particle
    wv = _dataset(h, "wavelengths", 1, filename)
    radii = _dataset(h, "radii", 1, filename) / 1.0e4
    arrs = {}
    for ds in ("w0", "qext", "g0"):
        d = _dataset(h, ds, 2, filename)
        if d.shape != (radii.shape[0], wv.shape[0]):
            raise LoaderError(f'"{ds}" has the wrong shape in "{filename}"')
        arrs[ds] = np.asfortranarray(np.stack([interp_discrete_to_bins(wavl, wv, d[i], "Constant")
                                                for i in range(radii.shape[0])]))
    return Particle(p_ind, radii, arrs["w0"], arrs["qext"], arrs["g0"])


def photolysis_from_hdf5(h: Mapping, sp_ind: int, wavl: np.ndarray, filename: str = "<memory>"):
    """`create_PhotolysisXsection` (types_create.f90:1407-1468): "wavelengths", "photoabsorption";
    log10(max(xs, tiny)) on the bins with interp_discrete_to_bins(..., 'FillValue', log10tiny), then 10**."""
    from ._abi import XS1D
    wv = _dataset(h, "wavelengths", 1, filename)
    xs = _dataset(h, "photoabsorption", 1, filename)
    lx = np.log10(np.maximum(xs, np.finfo(np.float64).tiny))
    return XS1D([sp_ind], 0, None, 10.0 ** interp_discrete_to_bins(wavl, wv, lx, "FillValue", LOG10TINY))


def parse_optical_properties(settings: Mapping) -> Dict:
    """The `optical-properties` block of settings.yaml as `unpack_settingsopacity` reads it
    (src/clima_types_create.f90:760-960): species lists, k-method, and per opacity either `true`/`false`
    (everything the data directory offers) or an explicit list."""
    op = settings.get("optical-properties")
    if op is None:
        raise LoaderError('"optical-properties" is required in the settings file')
    sp = op.get("species", {})
    out = dict(gases=list(sp.get("gases", [])), particles=list(sp.get("particles", []) or []),
               k_method=op.get("k-method", "RandomOverlapResortRebin"), number_of_bins=op.get("number-of-bins"),
               wavelength_bins_file=op.get("wavelength-bins-file"))
    if out["k_method"] not in ("RandomOverlapResortRebin", "AdaptiveEquivalentExtinction"):
        raise LoaderError(f'k-method "{out["k_method"]}" is not an option.')
    oc = op.get("opacities", {})
    for key, name in (("k-distributions", "k"), ("CIA", "cia"), ("rayleigh", "rayleigh"), ("photolysis-xs", "pxs")):
        v = oc.get(key)
        if isinstance(v, bool):
            out[name] = v
        elif v is None:
            out[name] = False
        else:
            v = list(v)
            if len(set(v)) != len(v):
                raise LoaderError(f'duplicate in optical-properties/opacities/{key}')
            out[name] = v
    wc = oc.get("water-continuum")
    out["water_continuum"] = wc if isinstance(wc, str) else None
    out["particle_xs"] = [dict(name=p["name"], dat=p.get("data", p["name"])) for p in (oc.get("particle-xs") or [])]
    return out


def build_tables(settings: Mapping, data: Mapping, star_file: str):
    """`create_OpticalProperties` (types_create.f90:272-645) + the two `create_RTChannel` + `read_stellar_flux`
    of `create_Radtran_2` (clima_radtran.f90:128-219): settings dict (parsed YAML) + data directory ->
    OpacityTables.  `data` maps a path relative to the data directory ("kdistributions/H2O.h5",
    "CIA/CO2-CO2.h5", "rayleigh/rayleigh.yaml", "kdistributions/bins.h5", ...) to an opened HDF5-like
    mapping (or, for the YAML file, its parsed dict); a missing key is a missing file."""
    from ._abi import OpacityTables
    so = parse_optical_properties(settings)
    species, particles = so["gases"], so["particles"]

    def listing(kind, folder, names):
        if kind is True:
            return [n for n in names if f"{folder}/{n}.h5" in data]
        return list(kind) if kind else []

    # k-distributions (:300-380)
    knames = listing(so["k"], "kdistributions", species)
    if so["k"] is True and not knames:
        raise LoaderError("No k-distribution data was found, but at least one k-distribution is needed.")
    if not knames:
        raise LoaderError("You must specify at least one k-distribution in the settings file.")
    tabs = {}
    for n in knames:
        if n not in species:
            raise LoaderError(f'Species "{n}" in optical property "k-distributions" is not in the list of species.')
        tabs[n] = ktable_from_hdf5(data[f"kdistributions/{n}.h5"], species.index(n) + 1, f"kdistributions/{n}.h5")
    weights, wavl = check_same_grid(tabs)
    # CIA (:384-460): with `true` every pair the directory offers, minus the H2O pairs when the water
    # continuum is on (:431-432)
    if so["cia"] is True:
        pairs = [f"{a}-{b}" for a in species for b in species if f"CIA/{a}-{b}.h5" in data
                 and not (so["water_continuum"] and "H2O" in (a, b))]
    else:
        pairs = list(so["cia"]) if so["cia"] else []
    cia = []
    for pr in pairs:
        a, _, b = pr.partition("-")
        if a not in species or b not in species:
            raise LoaderError(f'CIA pair "{pr}" contains a species that is not in the list of species.')
        if so["water_continuum"] and "H2O" in (a, b):
            raise LoaderError(f'Optical property "water-continuum" is set, but CIA "{pr}" is also set.')
        cia.append(xsection_from_hdf5(data[f"CIA/{pr}.h5"], [species.index(a) + 1, species.index(b) + 1], wavl, f"CIA/{pr}.h5"))
    # Rayleigh (:470-530)
    ray_ind, ray_xs = [], None
    if so["rayleigh"]:
        ry = data.get("rayleigh/rayleigh.yaml")
        if ry is None:
            raise LoaderError('There is an issue with formatting in "rayleigh/rayleigh.yaml"')
        if so["rayleigh"] is not True:
            for n in so["rayleigh"]:
                if n not in species:
                    raise LoaderError(f'Species "{n}" in optical property "rayleigh" is not in the list of species.')
                if n not in ry:
                    raise LoaderError(f'There is an issue with formatting in "rayleigh/rayleigh.yaml"')
            ry = {n: ry[n] for n in so["rayleigh"]}
        ray_ind, ray_xs = rayleigh_table(ry, species, wavl, "rayleigh/rayleigh.yaml")
    # photolysis cross sections (:540-580)
    pnames = listing(so["pxs"], "xsections", species)
    pxs = []
    for n in pnames:
        if n not in species:
            raise LoaderError(f'Species "{n}" in optical property "photolysis-xs" is not in the list of species.')
        if f"xsections/{n}.h5" not in data:
            raise LoaderError(f'Species "{n}" does not have photolysis xsection data')
        pxs.append(photolysis_from_hdf5(data[f"xsections/{n}.h5"], species.index(n) + 1, wavl, f"xsections/{n}.h5"))
    # particles (:586-600)
    part = []
    for p in so["particle_xs"]:
        if p["name"] not in particles:
            raise LoaderError(f'Species "{p["name"]}" in optical property "particle-xs" is not in the list of particles.')
        fn = f'aerosol_xsections/{p["dat"]}/mie_{p["dat"]}.h5'
        if fn not in data:
            raise LoaderError(f"Was unable to open mie data file {fn}")
        part.append(particle_from_hdf5(data[fn], particles.index(p["name"]) + 1, wavl, fn))
    kw = {}
    if so["water_continuum"]:
        fn = f'water_continuum/{so["water_continuum"]}.h5'
        if fn not in data:
            raise LoaderError(f'Continuum "{so["water_continuum"]}" is not avaliable.')
        kw.update(water_continuum_from_hdf5(data[fn], species, wavl, so["water_continuum"], fn))
    # channels (:226-270) and the stellar flux on the solar channel's bins (clima_radtran.f90:172-180)
    bins = so["wavelength_bins_file"] or "kdistributions/bins.h5"
    if bins not in data:
        raise LoaderError(f'Failed to read "{bins}".')
    ir_ind = channel_indices(read_wavl(data[bins], "ir", bins), wavl, bins)
    sol_wavl = read_wavl(data[bins], "solar", bins)
    sol_ind = channel_indices(sol_wavl, wavl, bins)
    photons = read_stellar_flux(star_file, sol_wavl)
    return OpacityTables(wavl=wavl, species_names=list(species), particle_names=list(particles), weights=weights,
                         k=[tabs[n][0] for n in knames], cia=cia, ray_sp_ind=ray_ind, ray_xs=ray_xs, pxs=pxs, part=part,
                         ir_ind=ir_ind, sol_ind=sol_ind, photons_sol=photons, **kw)
