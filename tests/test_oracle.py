"""CPU tests of the oracle (oracle/oracle.cpp): unit-level semantics of the restated routines,
closed-form limits, agreement with the independent NumPy/SciPy restatement (oracle/oracle_np.py)
and with the frozen golden vectors (tests/golden/).

The reference's own tests hold no numeric assertions for this path (tests/test_radtran.f90:67-81
only prints), so the oracle is "parity unpinned" against the Fortran binary; these tests are what
pins it instead."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from cases import compare, small_case
from oracle_lib import OracleRadtran, load

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import oracle_np  # noqa: E402

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int32)


def _d(a):
    return a.ctypes.data_as(dp)


def _i(v):
    return C.byref(C.c_int32(int(v)))


def _f(v):
    return C.byref(C.c_double(float(v)))


@pytest.fixture(scope="module")
def lib():
    return load("parity")


# ------------------------------------------------------------------------------ unit semantics
def test_dintrv_bracket_semantics(lib):
    """linear_interpolation_module.F90:348-350"""
    xt = np.array([1.0, 2.0, 4.0, 8.0])
    out = C.c_int32(0)
    for x, want in [(0.5, 1), (1.0, 1), (1.5, 1), (2.0, 2), (3.99, 2), (4.0, 3), (8.0, 3), (9.0, 3)]:
        lib.oracle_dintrv(_d(xt), _i(4), _f(x), C.byref(out))
        assert out.value == want, (x, out.value, want)


def test_interp_1d_2d_are_bilinear_and_extrapolate(lib):
    x = np.array([0.0, 1.0, 3.0]); y = np.array([10.0, 20.0])
    f1 = 2.0 * x + 1.0
    o = C.c_double(0)
    for xx in (-1.0, 0.25, 2.0, 5.0):  # interp_1d extrapolates linearly (:233-271)
        lib.oracle_interp_1d(_d(x), _d(f1), _i(3), _f(xx), C.byref(o))
        assert abs(o.value - (2.0 * xx + 1.0)) < 1e-13
    F = np.asfortranarray(3.0 * x[:, None] - 0.5 * y[None, :] + 0.1 * x[:, None] * y[None, :])
    for xx, yy in [(0.5, 12.0), (2.5, 19.0), (3.0, 20.0)]:
        lib.oracle_interp_2d(_d(x), _i(3), _d(y), _i(2), _d(F), _f(xx), _f(yy), C.byref(o))
        assert abs(o.value - (3.0 * xx - 0.5 * yy + 0.1 * xx * yy)) < 1e-12


def test_planck_and_ten2power(lib):
    o = C.c_double(0)
    lib.oracle_planck(_f(3.0e13), _f(288.0), C.byref(o))
    assert abs(o.value - oracle_np.planck(3.0e13, 288.0)) <= 1e-15 * o.value
    lib.oracle_ten2power(_f(-20.5), C.byref(o))
    assert abs(o.value / 10.0 ** -20.5 - 1.0) < 1e-14


def test_mrgrnk_is_a_stable_ascending_rank(lib):
    rng = np.random.default_rng(0)
    for n in (1, 2, 7, 64, 256, 400):
        x = rng.integers(0, 12, n).astype(np.float64)  # many ties
        inds = np.zeros(n, dtype=np.int32)
        lib.oracle_mrgrnk(_d(x), _i(n), inds.ctypes.data_as(ip))
        assert np.array_equal(inds - 1, np.argsort(x, kind="stable"))


def test_rebin_conserves_the_integral_and_is_identity_on_equal_bins(lib):
    rng = np.random.default_rng(1)
    n_old, n_new = 256, 16
    w = rng.uniform(0.1, 1.0, n_old); w /= w.sum()
    old_e = np.concatenate(([0.0], np.cumsum(w))); old_e[-1] = 1.0
    old_v = np.sort(10.0 ** rng.uniform(-8, 2, n_old))
    new_w = rng.uniform(0.5, 1.0, n_new); new_w /= new_w.sum()
    new_e = np.concatenate(([0.0], np.cumsum(new_w))); new_e[-1] = 1.0
    new_v = np.zeros(n_new)
    lib.oracle_rebin(_d(old_e), _d(old_v), _i(n_old), _d(new_e), _d(new_v), _i(n_new))
    assert abs(np.sum(new_v * np.diff(new_e)) / np.sum(old_v * np.diff(old_e)) - 1.0) < 1e-13
    assert np.allclose(new_v, oracle_np.rebin_cumulative(old_e, old_v, new_e), rtol=1e-11, atol=0)
    same = np.zeros(n_old)
    lib.oracle_rebin(_d(old_e), _d(old_v), _i(n_old), _d(old_e), _d(same), _i(n_old))
    assert np.allclose(same, old_v, rtol=1e-14, atol=0)


def test_rebin_against_brute_force_integration(lib):
    """The oracle's `rebin` (a reconstruction: futils is not on the box) against numerical integration of
    the piecewise-constant function it stands for -- a midpoint sum on a grid 200x finer than the finest
    bin -- on random, unaligned, tie-laden bins: zero-width old bins (repeated edges), new edges that
    coincide with old ones, and new bins narrower than an old bin."""
    rng = np.random.default_rng(7)
    for trial in range(12):
        n_old = int(rng.integers(3, 40))
        n_new = int(rng.integers(1, 12))
        w = rng.uniform(0.0, 1.0, n_old)
        w[rng.uniform(size=n_old) < 0.2] = 0.0          # zero-width old bins
        if w.sum() == 0.0:
            w[0] = 1.0
        old_e = np.concatenate(([0.0], np.cumsum(w / w.sum()))); old_e[-1] = 1.0
        old_v = 10.0 ** rng.uniform(-3, 3, n_old)
        inner = np.sort(rng.uniform(0.0, 1.0, n_new - 1))
        take = rng.uniform(size=n_new - 1) < 0.3          # some new edges sit exactly on old edges
        inner[take] = rng.choice(old_e[1:-1], size=int(take.sum())) if n_old > 2 else inner[take]
        new_e = np.concatenate(([0.0], np.sort(inner), [1.0]))
        if np.any(np.diff(new_e) <= 0.0):
            continue
        new_v = np.zeros(n_new)
        lib.oracle_rebin(_d(old_e), _d(old_v), _i(n_old), _d(new_e), _d(new_v), _i(n_new))
        # brute force: f(x) = old_v[j] on [old_e[j], old_e[j+1]); exact integral per new bin by splitting
        # at every old edge, cross-checked by a fine midpoint sum
        for i in range(n_new):
            a, b = new_e[i], new_e[i + 1]
            cuts = np.unique(np.concatenate(([a, b], old_e[(old_e > a) & (old_e < b)])))
            mids = 0.5 * (cuts[1:] + cuts[:-1])
            j = np.clip(np.searchsorted(old_e, mids, side="right") - 1, 0, n_old - 1)
            exact = np.sum(old_v[j] * np.diff(cuts)) / (b - a)
            assert new_v[i] == pytest.approx(exact, rel=1e-12), (trial, i)
            xs = a + (np.arange(4000) + 0.5) * (b - a) / 4000.0
            jj = np.clip(np.searchsorted(old_e, xs, side="right") - 1, 0, n_old - 1)
            assert new_v[i] == pytest.approx(np.mean(old_v[jj]), rel=2e-2 if n_old > 1 else 1e-12)


def test_weights_to_bins(lib):
    w = np.array([0.1, 0.2, 0.3, 0.4]); b = np.zeros(5)
    lib.oracle_weights_to_bins(_d(w), _i(4), _d(b))
    assert np.allclose(b, [0.0, 0.1, 0.3, 0.6, 1.0], rtol=0, atol=1e-15) and b[0] == 0.0


def test_tridiag_residual(lib):
    rng = np.random.default_rng(2)
    n = 60
    a = rng.uniform(-1, 1, n); c = rng.uniform(-1, 1, n); b = 3.0 + rng.uniform(0, 1, n); d = rng.uniform(-1, 1, n)
    a[0] = 0.0; c[-1] = 0.0
    M = np.diag(b) + np.diag(a[1:], -1) + np.diag(c[:-1], 1)
    aa, bb, cc, dd = a.copy(), b.copy(), c.copy(), d.copy()
    lib.oracle_tridiag(_i(n), _d(aa), _d(bb), _d(cc), _d(dd))
    assert np.max(np.abs(M @ dd - d)) < 1e-13


# ------------------------------------------------------------------------------ closed-form limits
def _ir(lib, tau, w0, gt, emis, hard, B, tau_min=1e-6):
    nz = len(tau)
    fup = np.zeros(nz + 1); fdn = np.zeros(nz + 1)
    lib.oracle_two_stream_ir(_i(nz), _d(tau), _d(w0), _d(gt), _f(emis), _i(1 if hard else 0), _f(tau_min),
                             _d(B), _d(fup), _d(fdn))
    return fup, fdn


def test_ir_isothermal_pure_absorption_closed_form(lib):
    """w0 = 0, isothermal, T_surface = T, emissivity 1: fup = pi B at every level and
    fdn(i) = pi B (1 - exp(-2 tauc_i)) (hemispheric mean, u1 = 0.5), independent of the layering."""
    rng = np.random.default_rng(3)
    for nz in (1, 5, 40):
        tau = 10.0 ** rng.uniform(-3, 0.5, nz)
        B = np.full(nz + 1, 0.37)
        fup, fdn = _ir(lib, tau, np.zeros(nz), np.zeros(nz), 1.0, True, B)
        tauc = np.concatenate(([0.0], np.cumsum(tau)))
        assert np.allclose(fup, np.pi * 0.37, rtol=1e-12)
        assert np.allclose(fdn, np.pi * 0.37 * (1.0 - np.exp(-2.0 * tauc)), rtol=1e-11, atol=1e-15)


def test_ir_thin_layer_guard_is_continuous(lib):
    """tau <= ir_tau_min switches to the zero-slope source (twostream.f90:216-227)."""
    B = np.array([0.2, 0.3, 0.5])
    w0 = np.array([0.3, 0.3]); gt = np.array([0.1, 0.1])
    f1 = _ir(lib, np.array([0.999e-6, 0.5]), w0, gt, 0.9, True, B)
    f2 = _ir(lib, np.array([1.001e-6, 0.5]), w0, gt, 0.9, True, B)
    assert np.allclose(f1[0], f2[0], rtol=1e-5) and np.allclose(f1[1], f2[1], rtol=1e-5, atol=1e-8)


def test_solar_conservative_scattering_energy_balance(lib):
    """w0 -> 1 (no absorption): net flux is constant with height (Toon 89 conserves energy)."""
    nz = 12
    tau = np.full(nz, 0.3); w0 = np.full(nz, 1.0 - 1e-12); gt = np.full(nz, 0.4)
    am = np.zeros(nz + 1); fup = np.zeros(nz + 1); fdn = np.zeros(nz + 1); sr = C.c_double(0)
    lib.oracle_two_stream_solar(_i(nz), _d(tau), _d(w0), _d(gt), _f(0.6), _f(0.3), _d(am), C.byref(sr),
                                _d(fup), _d(fdn))
    net = fdn - fup
    assert np.max(np.abs(net - net[0])) < 1e-8 * abs(net[0])
    assert abs(fup[nz] - 0.3 * fdn[nz]) < 1e-12  # surface albedo boundary condition
    assert fdn[0] == 0.6  # direct(1) = u0


def test_two_stream_matches_numpy_banded_solve(lib):
    rng = np.random.default_rng(4)
    nz = 30
    tau = 10.0 ** rng.uniform(-7, 1.5, nz); w0 = rng.uniform(0, 0.99999, nz); gt = rng.uniform(0, 0.9, nz)
    B = np.sort(rng.uniform(0.1, 1.0, nz + 1))
    for hard in (True, False):
        fu, fd = _ir(lib, tau, w0, gt, 0.85, hard, B)
        nu, nd = oracle_np.two_stream_ir(tau, w0, gt, 0.85, hard, 1e-6, B)
        assert np.allclose(fu, nu, rtol=1e-10) and np.allclose(fd, nd, rtol=1e-10, atol=1e-14)
    am = np.zeros(nz + 1); fup = np.zeros(nz + 1); fdn = np.zeros(nz + 1); sr = C.c_double(0)
    lib.oracle_two_stream_solar(_i(nz), _d(tau), _d(w0), _d(gt), _f(0.35), _f(0.2), _d(am), C.byref(sr),
                                _d(fup), _d(fdn))
    a2, u2, d2 = oracle_np.two_stream_solar(tau, w0, gt, 0.35, 0.2)
    assert np.allclose(fup, u2, rtol=1e-10, atol=1e-16) and np.allclose(fdn, d2, rtol=1e-10, atol=1e-16)
    assert np.allclose(am, a2, rtol=1e-10, atol=1e-16)


# ------------------------------------------------------------------------------ whole path
def _np_batch(tb, pr, cols, custom=None):
    outs = []
    for c in range(cols["ncol"]):
        outs.append(oracle_np.radiate(
            tb, pr, cols["T_surface"][c], cols["T"][:, c], cols["P"][:, c], cols["densities"][:, :, c],
            cols["dz"][:, c], cols["pdensities"][:, :, c] if "pdensities" in cols else None,
            cols["radii"][:, :, c] if "radii" in cols else None, custom))
    keys = [k for k in outs[0] if k != "opacity"]
    return {k: np.stack([np.asarray(o[k]) for o in outs], axis=-1) for k in keys}, outs


@pytest.mark.parametrize("kw", [
    dict(G=16, nw=10, nz=8, ncol=2),
    dict(G=8, nw=9, nz=9, ncol=2, doubled=False, has_hard_surface=False, nzen=3),
    dict(G=20, nw=8, nz=6, ncol=1, nzen=1),
    dict(G=16, nw=10, nz=8, ncol=2, npart=2),
])
def test_cpp_oracle_agrees_with_numpy_restatement(kw):
    tb, pr, cols = small_case(**kw)
    ro = OracleRadtran(tb, pr).radiate_batch(cols, spectral=True)
    rn, _ = _np_batch(tb, pr, cols)
    # element-wise tolerance is looser: the two restatements solve the tridiagonal system with
    # different algorithms (pivoted LAPACK vs Thomas) and tiny downward IR fluxes near the top of
    # the atmosphere are differences of large terms
    compare(rn, ro, tol=1e-10, elem_tol=1e-6)


def test_opacity_agrees_with_numpy_restatement():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=1)
    o = OracleRadtran(tb, pr)
    o.radiate(cols, 0)
    _, outs = _np_batch(tb, pr, cols)
    for a, b, name in zip(o.get_opacity(cols["nz"]), outs[0]["opacity"], ("tau", "w0", "g", "tau_band")):
        m = np.abs(b) > 0
        assert np.array_equal(a[~m], b[~m]), name
        if m.any():
            assert np.max(np.abs(a[m] - b[m]) / np.abs(b[m])) < 1e-11, name


def test_single_kgas_makes_rorr_the_identity():
    """nk = 1: k_rorr skips the merge loop (clima_radtran_types.f90:823): tau = continua + k1*col."""
    tb, pr, cols = small_case(G=16, nw=6, nz=6, ncol=1, nk=1)
    o = OracleRadtran(tb, pr)
    o.radiate(cols, 0)
    tau = o.get_opacity(cols["nz"])[0]
    _, outs = _np_batch(tb, pr, cols)
    assert np.allclose(tau, outs[0]["opacity"][0], rtol=1e-12)
    assert np.all(np.diff(tau, axis=1) >= 0)  # ascending in g


def test_zero_custom_opacity_does_not_change_fluxes():
    """tests/test_radtran.f90:86-104"""
    tb, pr, cols = small_case(G=16, nw=8, nz=6, ncol=2)
    o = OracleRadtran(tb, pr)
    r0 = o.radiate_batch(cols, spectral=True)
    z = np.zeros((5, tb.nw), order="F")
    o.set_custom(np.linspace(0.0, 7.0, 5), z, z, z)
    r1 = o.radiate_batch(cols, spectral=True)
    compare(r1, r0, tol=1e-13)


def test_doubled_grid_dedupe_equals_undoubled_physics():
    """Splitting every layer in two identical halves (clima_adiabat.f90:729-773) must leave the
    fluxes at the original level boundaries unchanged to rounding for the IR channel's band tau."""
    from clima_b200 import synth
    tb, pr, cols = small_case(G=16, nw=8, nz=6, ncol=1, doubled=False)
    T, P, D, dz = synth.double_grid(cols["T"], cols["P"], cols["densities"], cols["dz"], ghosts=False)
    c2 = dict(cols, T=T, P=P, densities=D, dz=dz, nz=T.shape[0])
    r1 = OracleRadtran(tb, pr).radiate_batch(cols, spectral=True)
    r2 = OracleRadtran(tb, pr).radiate_batch(c2, spectral=True)
    tb1 = r1["tau_band_ir"][:, :, 0]
    tb2 = r2["tau_band_ir"][:, :, 0]
    assert np.allclose(tb2[0::2] + tb2[1::2], tb1, rtol=1e-12)


def test_frozen_opacity_and_stale_solar_semantics():
    """compute_opacity=.false. / compute_solar=.false. (clima_radtran.f90:248-289)."""
    tb, pr, cols = small_case(G=16, nw=8, nz=6, ncol=1, nzen=1)
    o = OracleRadtran(tb, pr)
    r0 = o.radiate(cols, 0, spectral=False)
    hot = dict(cols, T=cols["T"] * 1.02, T_surface=cols["T_surface"] * 1.02)
    r1 = o.radiate(hot, 0, compute_solar=False, compute_opacity=False, spectral=False)
    assert np.array_equal(r1["fup_n_sol"], r0["fup_n_sol"])  # stale solar
    assert np.all(r1["fup_n_ir"][-1] > r0["fup_n_ir"][-1])    # warmer -> more OLR with same opacity
    assert np.allclose(r1["f_total"], (r0["fdn_n_sol"] - r0["fup_n_sol"]) + (r1["fdn_n_ir"] - r1["fup_n_ir"]),
                       rtol=0, atol=1e-9 * np.max(np.abs(r1["fup_n_ir"])))


# ------------------------------------------------------------------------------ golden vectors
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["small_g16", "small_g8_nohard", "small_particles_custom", "scanon_2col"])
def test_oracle_reproduces_golden_vectors(name):
    """Frozen outputs of the parity oracle (tests/golden/make_golden.py); guards against drift of
    the oracle itself."""
    import make_golden
    case = make_golden.CASES[name]
    tb, pr, cols, custom = case()
    ref = np.load(os.path.join(GOLDEN, name + ".npz"))
    o = OracleRadtran(tb, pr)
    if custom is not None:
        o.set_custom(*custom)
    r = o.radiate_batch(cols, spectral=False, nthreads=4)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR"):
        s = np.max(np.abs(ref[k]))
        assert np.max(np.abs(r[k] - ref[k])) <= 1e-12 * s, k
