"""Load-time packer (N2): addpnt/inter2, stellar flux reader, Rayleigh, channel indices, k-table schema.
CPU only.  The checks are independent restatements (exact integration of the piecewise-linear
function on the merged grid), conservation properties and a literature value."""
import numpy as np
import pytest

from clima_b200 import loader, synth
from clima_b200.loader import LoaderError


def _exact_bin_average(xg, x, y):
    """integral of the piecewise-linear interpolant (0 outside the data) on the merged grid"""
    out = np.zeros(len(xg) - 1)
    for i in range(len(xg) - 1):
        lo, hi = max(xg[i], x[0]), min(xg[i + 1], x[-1])
        if hi <= lo:
            continue
        pts = np.unique(np.concatenate(([lo, hi], x[(x > lo) & (x < hi)])))
        out[i] = np.trapezoid(np.interp(pts, x, y), pts) / (xg[i + 1] - xg[i])
    return out


def test_inter2_is_the_bin_average_of_the_piecewise_linear_function():
    rng = np.random.default_rng(3)
    x = np.sort(rng.uniform(100.0, 900.0, 57))
    y = rng.uniform(0.0, 5.0, 57)
    for xg in (np.linspace(50.0, 1000.0, 23), np.sort(rng.uniform(0.0, 1200.0, 40)), x[5:30].copy()):
        got = loader.inter2(xg, x, y)
        np.testing.assert_allclose(got, _exact_bin_average(xg, x, y), rtol=1e-13, atol=1e-15)
    # bins entirely outside the data are zero; integral over a covering grid is conserved
    xg = np.array([0.0, 50.0, 99.0, 500.0, 950.0, 2000.0])
    got = loader.inter2(xg, x, y)
    assert got[0] == 0.0 and got[1] == 0.0
    assert np.sum(got * np.diff(xg)) == pytest.approx(np.trapezoid(y, x), rel=1e-13)
    with pytest.raises(LoaderError):
        loader.inter2(np.array([0.0, 2.0, 1.0]), x, y)


def test_addpnt_keeps_the_grid_sorted_and_rejects_duplicates():
    x = np.array([1.0, 2.0, 4.0, 0.0, 0.0])
    y = np.array([10.0, 20.0, 40.0, 0.0, 0.0])
    n = loader.addpnt(x, y, 3, 3.0, 30.0)
    n = loader.addpnt(x, y, n, 0.5, 5.0)
    assert n == 5 and list(x) == [0.5, 1.0, 2.0, 3.0, 4.0] and list(y) == [5.0, 10.0, 20.0, 30.0, 40.0]
    with pytest.raises(LoaderError):
        loader.addpnt(x, y, 5, 9.0, 0.0)  # full
    with pytest.raises(LoaderError):
        loader.addpnt(np.array([1.0, 2.0, 0.0]), np.zeros(3), 2, 2.0, 0.0)  # duplicate abscissa


def test_read_stellar_flux_conserves_energy_and_converts_units(tmp_path):
    lam = np.linspace(120.0, 3000.0, 400)
    flux = 3.0 * np.exp(-((lam - 600.0) / 400.0) ** 2) + 0.01
    lam = np.array([float(f"{v:.10e}") for v in lam]); flux = np.array([float(f"{v:.10e}") for v in flux])  # as written
    f = tmp_path / "star.txt"
    f.write_text("Wavelength (nm)   Solar flux (mW/m^2/nm)\n" + "\n".join(f"{a:.10e} {b:.10e}" for a, b in zip(lam, flux)) + "\n")
    wavl = np.concatenate(([50.0], np.geomspace(100.0, 4000.0, 60), [1.0e5]))
    pf = loader.read_stellar_flux(str(f), wavl)
    assert pf.shape == (len(wavl) - 1,)
    wav_av = 0.5 * (wavl[:-1] + wavl[1:])
    per_nm = pf / (((wav_av * 1.0e-9) * wav_av) / loader.C_LIGHT)  # undo :75
    # energy: file integral + the two thin ramps to zero just outside its range (rdelta = 1e-4, :55-58)
    ramps = 0.5 * flux[0] * lam[0] * 1e-4 + 0.5 * flux[-1] * lam[-1] * 1e-4
    assert np.sum(per_nm * np.diff(wavl)) == pytest.approx(np.trapezoid(flux, lam) + ramps, rel=1e-12)
    assert pf[0] == 0.0 and pf[-1] == 0.0  # bins outside the file
    with pytest.raises(LoaderError, match="does not exist"):
        loader.read_stellar_flux(str(tmp_path / "nope.txt"), wavl)
    bad = tmp_path / "bad.txt"
    bad.write_text("header\n1.0 2.0\nabc def\n")
    with pytest.raises(LoaderError, match="Problem reading"):
        loader.read_stellar_flux(str(bad), wavl)


def test_rayleigh_vardavas_scaling_and_literature_value():
    # N2 (Vardavas & Carver 1984): A = 29.06e-5, B = 7.70e-3, Delta = 0.0305; Sneep & Ubachs 2005
    # measure 5.1e-27 cm2 at 532 nm
    s = loader.rayleigh_vardavas(29.06e-5, 7.70e-3, 0.0305, 532.0)
    assert s == pytest.approx(5.1e-27, rel=0.08)
    # B = 0: pure lambda^-4
    a, b = loader.rayleigh_vardavas(1e-4, 0.0, 0.0, np.array([400.0, 800.0]))
    assert a / b == pytest.approx(16.0, rel=1e-14)
    wavl = np.array([200.0, 300.0, 500.0, 1000.0])
    ray = {"N2": {"data": {"Delta": 0.0305, "A": 29.06e-5, "B": 7.70e-3}}, "CO2": {"data": {"Delta": 0.0805, "A": 43.9e-5, "B": 6.4e-3}}}
    inds, xs = loader.rayleigh_table(ray, ["H2O", "CO2", "N2"], wavl)
    assert inds == [2, 3] and xs.shape == (3, 2) and xs.flags.f_contiguous
    assert xs[1, 1] == pytest.approx(loader.rayleigh_vardavas(29.06e-5, 7.70e-3, 0.0305, 300.0), rel=1e-15)  # lower bin edge (:1085)
    with pytest.raises(LoaderError):
        loader.rayleigh_table({"N2": {"data": {"A": 1.0}}}, ["N2"], wavl)


def test_channel_indices_follow_create_rtchannel():
    op = np.geomspace(100.0, 1.0e5, 41)
    assert loader.channel_indices(op, op) == (1, 40)
    assert loader.channel_indices(op[10:31] * (1 + 1e-9), op) == (11, 30)
    with pytest.raises(LoaderError, match="not compatible"):
        loader.channel_indices(op[10:31:2], op)
    w = op[10:31].copy(); w[5] *= 1.001
    with pytest.raises(LoaderError, match="not compatible"):
        loader.channel_indices(w, op)


class _H5(dict):
    """stand-in with the h5py interface the reader uses: `name in f`, `f[name][...]`"""


def _as_h5(k, weights, wavl_nm):
    return _H5(weights=np.asarray(weights), log10P=k.log10P, T=k.temp, wavelengths=np.asarray(wavl_nm) * 1e-3,
               log10k=np.ascontiguousarray(np.transpose(k.log10k, (3, 2, 1, 0))))  # what h5py shows of Fortran (G,nP,nT,nw)


def test_ktable_schema_roundtrip_and_errors():
    tb = synth.make_tables(nw=7, G=8, nk=2, npress=5, ntemp=4, n_sol=5, ir_start=2)
    loaded = {}
    for i, k in enumerate(tb.k):
        kt, w, wl = loader.ktable_from_hdf5(_as_h5(k, tb.weights, tb.wavl), k.sp_ind, f"k{i}.h5")
        assert kt.log10k.flags.f_contiguous and kt.log10k.shape == k.log10k.shape
        assert np.array_equal(kt.log10k, k.log10k) and np.array_equal(kt.log10P, k.log10P) and np.array_equal(kt.temp, k.temp)
        np.testing.assert_allclose(wl, tb.wavl, rtol=1e-15)
        loaded[f"k{i}.h5"] = (kt, w, wl)
    w, wl = loader.check_same_grid(loaded)
    assert np.array_equal(w, tb.weights)
    h = _as_h5(tb.k[0], tb.weights, tb.wavl)
    del h["log10P"]
    with pytest.raises(LoaderError, match='dataset "log10P"'):
        loader.ktable_from_hdf5(h, 1, "x.h5")
    h = _as_h5(tb.k[0], tb.weights, tb.wavl)
    h["log10k"] = h["log10k"][:, :, :, :-1]
    with pytest.raises(LoaderError, match="shape"):
        loader.ktable_from_hdf5(h, 1, "x.h5")
    h = _as_h5(tb.k[0], tb.weights, tb.wavl)
    h["T"] = h["T"][::-1].copy()
    with pytest.raises(LoaderError, match="interpolator"):
        loader.ktable_from_hdf5(h, 1, "x.h5")
    bad = dict(loaded)
    kt, w, wl = loaded["k1.h5"]
    bad["k1.h5"] = (kt, w * 1.01, wl)
    with pytest.raises(LoaderError, match="different weights"):
        loader.check_same_grid(bad)
    with pytest.raises(LoaderError, match="h5py"):
        loader.open_hdf5("/nonexistent.h5")
