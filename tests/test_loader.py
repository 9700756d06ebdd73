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


# ------------------------------------------------------------------ N2: the remaining readers
class _DS:
    """stand-in for an h5py dataset: `f[name][...]`"""
    def __init__(self, a): self.a = np.asarray(a)
    def __getitem__(self, k): return self.a


def _h5(**kw):
    return {k: _DS(v) for k, v in kw.items()}


def _fake_datadir(tb, extra_cia=("H2O-H2O",)):
    """An in-memory data directory in the reference's schema, generated FROM synthetic tables so that
    build_tables can be checked against them: k-tables as h5py would show them (C-ordered transpose),
    CIA / continuum / photolysis / Mie on finer file grids, rayleigh.yaml, bins.h5."""
    G, nw = tb.ngauss, tb.nw
    data = {}
    wl_um = tb.wavl / 1.0e3
    for k in tb.k:
        name = tb.species_names[k.sp_ind - 1]
        data[f"kdistributions/{name}.h5"] = _h5(weights=tb.weights, log10P=k.log10P, T=k.temp, wavelengths=wl_um,
                                               log10k=np.ascontiguousarray(np.asarray(k.log10k).transpose(3, 2, 1, 0)))
    fine = np.geomspace(wl_um[0] * 1.001, wl_um[-1] * 0.999, 900)
    rng = np.random.default_rng(4)
    for a, b in (("CO2", "CO2"), ("N2", "N2"), ("H2O", "H2O")):
        T = np.linspace(150.0, 800.0, 7)
        data[f"CIA/{a}-{b}.h5"] = _h5(wavelengths=fine, T=T, log10xs=-46.0 + np.sin(3.0 * np.log10(fine))[None, :] + 1e-3 * T[:, None])
    data["CIA/N2-H2.h5"] = _h5(wavelengths=fine, log10xs=-47.0 + 0.5 * np.cos(2.0 * np.log10(fine)))
    Tc = np.linspace(200.0, 700.0, 9)
    data["water_continuum/MT_CKD.h5"] = _h5(wavelengths=fine, T=Tc, log10xs_H2O=-42.0 - 1e-3 * Tc[:, None] + 0.0 * fine[None, :],
                                            log10xs_foreign=-44.0 + 0.3 * np.sin(np.log10(fine))[None, :] + 0.0 * Tc[:, None])
    wv_nm = np.geomspace(150.0, 400.0, 60)
    data["xsections/O2.h5"] = _h5(wavelengths=wv_nm, photoabsorption=1e-19 * np.exp(-(wv_nm - 150.0) / 40.0))
    r_um = np.geomspace(0.01, 10.0, 8)
    wv_mie = np.geomspace(tb.wavl[0] * 1.01, tb.wavl[-1] * 0.99, 120)
    data["aerosol_xsections/khare1984/mie_khare1984.h5"] = _h5(
        wavelengths=wv_mie, radii=r_um, w0=rng.uniform(0.1, 0.9, (8, 120)), qext=rng.uniform(0.5, 3.0, (8, 120)),
        g0=rng.uniform(0.0, 0.8, (8, 120)))
    data["rayleigh/rayleigh.yaml"] = {"N2": {"data": {"Delta": 0.0305, "A": 29.06e-5, "B": 7.7e-3}},
                                      "CO2": {"data": {"Delta": 0.0805, "A": 43.9e-5, "B": 6.4e-3}},
                                      "Ar": {"data": {"Delta": 0.0006, "A": 27.92e-5, "B": 5.6e-3}}}
    data["kdistributions/bins.h5"] = _h5(sol_wavl=wl_um[tb.sol_ind[0] - 1:tb.sol_ind[1] + 1],
                                         ir_wavl=wl_um[tb.ir_ind[0] - 1:tb.ir_ind[1] + 1])
    return data


def test_xsection_reader_rebins_log10_values_and_pads_with_log10tiny():
    wavl = np.geomspace(200.0, 20000.0, 25)
    wf = np.geomspace(0.5, 10.0, 300)  # micron: covers 500 nm .. 10 um only
    xs = loader.xsection_from_hdf5(_h5(wavelengths=wf, log10xs=np.full(300, -45.0)), [1, 2], wavl)
    assert xs.dim == 0 and xs.data.shape == (24,)
    inside = (wavl[:-1] >= 500.0) & (wavl[1:] <= 10000.0)
    np.testing.assert_allclose(xs.data[inside], 1.0e-45, rtol=1e-12)
    outside = (wavl[1:] < 499.0) | (wavl[:-1] > 10002.0)
    np.testing.assert_allclose(np.log10(xs.data[outside]), loader.LOG10TINY, rtol=1e-12)  # the padding value
    T = np.array([100.0, 300.0, 900.0])
    x1 = loader.xsection_from_hdf5(_h5(wavelengths=wf, T=T, log10xs=-44.0 - 1e-3 * T[:, None] + 0.0 * wf[None, :]), [3, 3], wavl)
    assert x1.dim == 1 and x1.data.shape == (3, 24) and np.isfortran(x1.data)
    np.testing.assert_allclose(x1.data[:, inside], (-44.0 - 1e-3 * T)[:, None] * np.ones((1, inside.sum())), rtol=1e-12)
    with pytest.raises(LoaderError, match="log10xs"):
        loader.xsection_from_hdf5(_h5(wavelengths=wf), [1, 1], wavl)
    with pytest.raises(LoaderError, match="bad dimension"):
        loader.xsection_from_hdf5(_h5(wavelengths=wf, T=T, log10xs=np.zeros((2, 300))), [1, 1], wavl)


def test_interp_discrete_to_bins_modes():
    wavl = np.array([0.0, 1.0, 2.0, 4.0, 8.0])
    wv = np.array([1.5, 2.5, 3.5])
    f = np.array([10.0, 20.0, 30.0])
    c = loader.interp_discrete_to_bins(wavl, wv, f, "Constant")
    assert c[0] == pytest.approx(10.0) and c[3] == pytest.approx(30.0)       # held at the end values
    assert c[2] == pytest.approx((0.5 * (15.0 + 20.0) * 0.5 + 0.5 * (20.0 + 30.0) * 1.0 + 30.0 * 0.5) / 2.0)
    z = loader.interp_discrete_to_bins(wavl, wv, f, "FillValue", -5.0)
    assert z[0] == pytest.approx(-5.0, abs=1e-9) and z[3] == pytest.approx(-5.0, abs=1e-9)
    with pytest.raises(LoaderError):
        loader.interp_discrete_to_bins(wavl, wv[::-1], f, "Constant")


def test_parse_optical_properties_of_the_reference_templates():
    """The `optical-properties` blocks of the reference's own templates (inline copies, and the files
    themselves where /root/reference exists)."""
    import os
    import yaml
    modern_earth = yaml.safe_load("""
optical-properties:
  species:
    gases: [H2O, CO2, O2, N2, O3, CH4]
    particles: [HCaer1]
  k-method: RandomOverlapResortRebin
  opacities: {k-distributions: true, CIA: true, rayleigh: true, photolysis-xs: true,
    water-continuum: MT_CKD}
""")
    so = loader.parse_optical_properties(modern_earth)
    assert so["gases"] == ["H2O", "CO2", "O2", "N2", "O3", "CH4"] and so["particles"] == ["HCaer1"]
    assert so["k"] is True and so["cia"] is True and so["rayleigh"] is True and so["pxs"] is True
    assert so["water_continuum"] == "MT_CKD" and so["particle_xs"] == []
    for rel in ("templates/ModernEarth/settings.yaml", "templates/AdiabatClimate/Earth/settings.yaml",
                "tests/settings_RCE_test.yaml"):
        path = os.path.join("/root/reference", rel)
        if os.path.exists(path):
            s2 = loader.parse_optical_properties(yaml.safe_load(open(path)))
            assert s2["k_method"] == "RandomOverlapResortRebin" and s2["k"] is True and s2["water_continuum"] == "MT_CKD"
            # AdiabatClimate settings take their species from species.yaml, not from this block
            assert rel.startswith("templates/ModernEarth") == (len(s2["gases"]) > 0)
    with pytest.raises(LoaderError, match="k-method"):
        loader.parse_optical_properties({"optical-properties": {"k-method": "Nope", "species": {"gases": ["H2O"]}}})
    with pytest.raises(LoaderError, match="required"):
        loader.parse_optical_properties({})


def test_build_tables_from_a_data_directory(tmp_path):
    """create_OpticalProperties + create_RTChannel + read_stellar_flux on an in-memory data directory:
    `true` selects what the directory offers, the H2O CIA pair is dropped because the water continuum is
    on (types_create.f90:431-432), explicit lists are honoured, the k-tables come back bit for bit, and
    the result drives the CPU oracle."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from oracle_lib import OracleRadtran
    tb0 = synth.make_tables(nw=14, G=8, nk=3, npress=6, ntemp=5, n_sol=9, ir_start=5)
    data = _fake_datadir(tb0)
    star = tmp_path / "star.txt"
    lam = np.geomspace(100.0, 6.0e5, 500)
    star.write_text("Wavelength (nm)  flux\n" + "\n".join(f"{a:.8e} {1.0e3 * np.exp(-a / 2000.0) + 1e-3:.8e}" for a in lam) + "\n")
    settings = {"optical-properties": {"species": {"gases": list(tb0.species_names), "particles": ["haze"]},
                                       "k-method": "RandomOverlapResortRebin",
                                       "opacities": {"k-distributions": True, "CIA": True, "rayleigh": True, "photolysis-xs": True,
                                                     "water-continuum": "MT_CKD",
                                                     "particle-xs": [{"name": "haze", "data": "khare1984"}]}}}
    tb = loader.build_tables(settings, data, str(star))
    assert [tb.species_names[k.sp_ind - 1] for k in tb.k] == [tb0.species_names[k.sp_ind - 1] for k in tb0.k]
    for k, k0 in zip(tb.k, tb0.k):
        assert np.array_equal(k.log10k, k0.log10k) and np.array_equal(k.log10P, k0.log10P)
    np.testing.assert_allclose(tb.wavl, tb0.wavl, rtol=1e-14)
    pairs = [(tb.species_names[c.sp_ind[0] - 1], tb.species_names[c.sp_ind[1] - 1]) for c in tb.cia]
    assert ("CO2", "CO2") in pairs and ("N2", "N2") in pairs and ("N2", "H2") in pairs
    assert not any("H2O" in p for p in pairs)
    assert [tb.species_names[i - 1] for i in tb.ray_sp_ind] == ["CO2", "N2"]  # species order; Ar is not a species
    assert len(tb.pxs) == 1 and tb.species_names[tb.pxs[0].sp_ind[0] - 1] == "O2"
    assert tb.cont_LH2O == tb.species_names.index("H2O") + 1 and tb.cont_log10_H2O.shape == (9, 14)
    assert len(tb.part) == 1 and tb.part[0].w0.shape == (8, 14) and tb.part[0].radii[0] == pytest.approx(1.0e-6)
    assert tb.ir_ind == tb0.ir_ind and tb.sol_ind == tb0.sol_ind and tb.photons_sol.shape == (tb0.nw_sol,)
    # explicit lists
    s2 = {"optical-properties": {"species": {"gases": list(tb0.species_names)},
                                 "opacities": {"k-distributions": ["CO2"], "CIA": ["CO2-CO2"], "rayleigh": ["N2"]}}}
    t2 = loader.build_tables(s2, data, str(star))
    assert len(t2.k) == 1 and len(t2.cia) == 1 and t2.ray_sp_ind == [tb0.species_names.index("N2") + 1] and t2.cont_LH2O == 0
    with pytest.raises(LoaderError, match="water-continuum"):
        loader.build_tables({"optical-properties": {"species": {"gases": list(tb0.species_names)},
                                                    "opacities": {"k-distributions": True, "CIA": ["H2O-H2O"],
                                                                  "water-continuum": "MT_CKD"}}}, data, str(star))
    with pytest.raises(LoaderError, match="at least one k-distribution"):
        loader.build_tables({"optical-properties": {"species": {"gases": ["Xe"]}, "opacities": {"k-distributions": True}}},
                            data, str(star))
    # the packed tables drive the checker end to end (1 particle species in the model)
    pr = synth.default_params(tb, nzenith=2)
    cols = synth.make_columns(2, nz=6, seed=3, doubled=False, npart=1)
    r = OracleRadtran(tb, pr).radiate_batch(cols)
    assert np.all(np.isfinite(r["fup_n_ir"])) and np.all(r["OLR"] > 0)
