"""Row N4: batched `AdiabatClimate%make_profile` / `TOA_fluxes` (clima_b200/csrc/adiabat_profile.cu behind
clima_rt_make_profile_batch / clima_rt_adiabat_toa_fluxes_batch) against the numpy restatement of the reference
(oracle/adiabat_np.py: pinned for make_profile against outputs the reference printed, see its header).  Inputs are the cases of the reference's own
tests/test_adiabat.f90 (which holds no numbers) on the reference's species data (tests/golden/adiabat_species.json).
Tolerances: see PROFILE_TOL / TIGHT_TOL below (the integration error of the reference's default settings bounds
what any two builds can agree on); 1e-9 on ISR / OLR like the rest of the path, on converged profiles."""
import json
import math
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import adiabat_np as A  # noqa: E402

EARTH_MASS, EARTH_RADIUS = 5.972e27, 6.371e8  # templates/AdiabatClimate/Earth/settings.yaml
BAR = 1.0e6
# tests/test_adiabat.f90:31, :99: [H2O, CO2, N2, H2, CH4, CO, O2] in bars
P_I_OCEAN = np.array([270.0, 400e-6, 1.0, 1e-10, 1e-10, 1e-10, 1e-10]) * BAR
P_I_CO2 = np.array([270.0, 10.0, 1.0, 1e-10, 1e-10, 1e-10, 1e-10]) * BAR


def species():
    with open(os.path.join(HERE, "golden", "adiabat_species.json")) as f:
        return json.load(f)


def one_gas(cp=29.0, mass=28.0):
    """A single non-condensing gas with a constant heat capacity (Shomate, only the A coefficient)."""
    return dict(names=["X"], mass=[mass], sat=[None],
                thermo=[dict(model="Shomate", temps=[0.0, 6000.0], data=[[cp, 0, 0, 0, 0, 0, 0]])])


# ------------------------------------------------------------------------------------------------------------
# the oracle against closed forms and an independent integrator (CPU)
# ------------------------------------------------------------------------------------------------------------
def test_zeroin_finds_roots_to_the_tolerance():
    r, fr, info = A.zeroin(lambda x: x ** 3 - 2.0 * x - 5.0, 2.0, 3.0, 1e-12)
    assert info == 0 and abs(r - 2.0945514815423265) < 1e-11
    r, _, info = A.zeroin(lambda x: math.cos(x) - x, 1.0, 0.0, 1e-10)  # descending bracket, as find_root uses it
    assert info == 0 and abs(r - 0.7390851332151607) < 1e-9
    assert A.zeroin(lambda x: x * x + 1.0, -1.0, 1.0, 1e-8)[2] == -1


def test_dop853_against_scipy_and_a_closed_form():
    from scipy.integrate import solve_ivp
    f = lambda x, y: np.array([y[1], -y[0]])  # noqa: E731
    d = A.Dop853(f, 2)
    x, y, idid = d.integrate(0.0, np.array([0.0, 1.0]), 10.0, 1e-11, 1e-13)
    assert idid == 1 and x == 10.0
    assert abs(y[0] - math.sin(10.0)) < 1e-9 and abs(y[1] - math.cos(10.0)) < 1e-9
    assert abs(d.contd8(0, 9.9) - math.sin(9.9)) < 1e-8          # dense output inside the last step
    s = solve_ivp(f, (0.0, 10.0), [0.0, 1.0], method="DOP853", rtol=1e-11, atol=1e-13)
    assert np.max(np.abs(s.y[:, -1] - y)) < 1e-9
    # backward integration (pressure decreases along the adiabat) and a solout that stops it
    stop = dict(n=0)

    def solout(nr, xo, xn, yn):
        stop["n"] = nr
        return -1 if xn < 5.0 else 1
    d2 = A.Dop853(lambda x, y: np.array([y[0] / x]), 1, solout)   # y = x
    x, y, idid = d2.integrate(10.0, np.array([10.0]), 1.0, 1e-10, 1e-12)
    assert idid == 2 and x < 5.0 and abs(y[0] - x) < 1e-8 and stop["n"] > 1


def test_dry_adiabat_of_a_constant_cp_gas():
    """T = T_s (P / P_s)^(R / cp) below the tropopause, T_trop above it; P_trop from the same law."""
    cp, nz, Ts, Ps, Tt = 29.0, 40, 300.0, 1.0e6, 200.0
    e = A.make_profile(Ts, [Ps], one_gas(cp), nz, EARTH_MASS, EARTH_RADIUS, 1.0, Tt, [1.0], 1e-11, 1e-13)
    kappa = (A.RGAS / 1e7) / cp
    want = np.maximum(Ts * (e["P"] / Ps) ** kappa, Tt)
    assert np.max(np.abs(e["T"][:-1] / want[:-1] - 1.0)) < 1e-9
    assert abs(e["P_trop"] / (Ps * (Tt / Ts) ** (1.0 / kappa)) - 1.0) < 1e-8
    assert np.all(np.diff(e["z"][:-1]) > 0.0) and np.all(e["f_i"] == 1.0)
    assert e["P"][0] == Ps and e["P"][-1] == 1.0 and np.allclose(np.diff(np.log10(e["P"])), -6.0 / (2 * nz))


def test_condensing_gases_follow_their_saturation_curve():
    sp = species()
    c = A.adiabat_column(280.0, P_I_CO2, sp, 30, EARTH_MASS, EARTH_RADIUS, T_trop=150.0)
    e = c["edges"]
    assert list(e["sp_type"][:3]) == [1, 1, 0]          # H2O from the surface, CO2 aloft (tests/test_adiabat.f90:95)
    assert c["N_surface"][0] > 0.0 and c["N_surface"][1] == 0.0
    below = c["P"] > c["P_trop"]
    for i in np.where(below)[0]:
        assert abs(c["f_i"][i, 0] * c["P"][i] / A.sat_pressure(sp["sat"][0], c["T"][i]) - 1.0) < 1e-12
    assert np.allclose(c["f_i"].sum(axis=1), 1.0, rtol=0, atol=1e-12)
    # CO2 is dry at the surface (its mixing ratio relative to N2 is the surface one) and saturated higher up
    assert abs(c["f_i"][0, 1] / c["f_i"][0, 2] - 10.0) < 1e-9
    top = np.where(below)[0][-1]
    assert abs(c["f_i"][top, 1] * c["P"][top] / A.sat_pressure(sp["sat"][1], c["T"][top]) - 1.0) < 1e-12
    assert 150.0 <= c["T"].min() and c["T"][-1] == 150.0 and c["P_trop"] > 0.0
    assert np.all(c["dz"] > 0.0) and np.all(np.diff(c["z"]) > 0.0)


def test_compute_altitude_isothermal_closed_form():
    """Constant T and mubar: 1/(R+z) = 1/R + (Rgas T)/(G M mubar) ln(P/P0) with the G of `gravity`."""
    nz, T0, mu = 20, 250.0, 28.0
    P = np.logspace(5.9, 0.2, nz)
    z, dz = A.compute_altitude(1.0e6, T0, [1.0], P, np.full(nz, T0), np.ones((nz, 1)), [mu], 1.0, EARTH_MASS,
                               EARTH_RADIUS, 1e-11, 1e-13)
    GM = A.G_GRAV * 1.0e3 * EARTH_MASS   # cm^3/s^2
    want = 1.0 / (1.0 / EARTH_RADIUS + (A.RGAS * T0) / (GM * mu) * np.log(P / 1.0e6)) - EARTH_RADIUS
    assert np.max(np.abs(z / want - 1.0)) < 1e-9
    assert np.all(dz > 0.0) and abs(dz.sum() / (2.0 * z[-1] - z[-2] + 0.5 * (z[-1] - z[-2])) - 1.0) < 0.2


def tutorial_case():
    with open(os.path.join(HERE, "golden", "tutorial_adiabat_outputs.json")) as f:
        ref = json.load(f)
    i = ref["inputs"]
    return ref, i["T_surf"], np.array(i["P_i_bar"]) * BAR, i


def test_oracle_against_the_outputs_the_reference_printed():
    """THE PIN of this path: examples/Tutorial_AdiabatClimate.ipynb stores what the reference printed after
    `c.make_profile(280, [260, 10, 1, ...] bar)` with P_top = 1e-2, T_trop = 150 (fixture + generator in tests/golden).
      * P_surf = 11.010013835789936 bar: all 17 digits (saturation pressure of H2O at 280 K + the dry gases);
      * N_surface(H2O) = 14696.15911658 mol/cm^2: all 13 printed digits (masses, gravity);
      * N_atmos (9 printed digits): sum of density * f_i * dz / N_avo over the layers, i.e. T(P), z(P), the mixing
        ratios, the CO2 saturation event and the tropopause of the pseudo-adiabat.  The notebook's version of the
        reference took dz from make_profile's own z; today's takes it from `compute_altitude` (interpolating T and mubar
        linearly in log10 P between the layer centres), which moves N_atmos by 3e-4 -- so the printed numbers pin the
        adiabat integration (checked with dz from its own z) and not compute_altitude."""
    ref, T_surf, P_i, inp = tutorial_case()
    sp = species()
    want = np.array(ref["N_atmos"])
    for rtol, tol in ((1e-9, 1e-7), (1e-12, 2e-9)):
        c = A.adiabat_column(T_surf, P_i, sp, inp["nz"], inp["planet_mass"], inp["planet_radius"], P_top=inp["P_top"],
                             T_trop=inp["T_trop"], rtol=rtol)
        assert repr(c["P_surf"] / 1e6) == ref["P_surf_text"]
        assert "%.8f" % c["N_surface"][0] == ref["N_surface_text"].split()[0].lstrip("[") and not c["N_surface"][1:].any()
        ze = c["edges"]["z"]
        dz_own = ze[2::2] - ze[0:-1:2]
        dens = c["P"] / (A.K_BOLTZ * c["T"])
        N_own = np.array([np.sum(dens * c["f_i"][:, i] * dz_own) / A.N_AVO for i in range(7)])
        assert np.max(np.abs(N_own / want - 1.0)) < tol, (rtol, N_own / want - 1.0)
        # today's dz (compute_altitude): a different discretisation of the same column, 3e-4 away
        assert 1e-5 < np.max(np.abs(c["N_atmos"] / want - 1.0)) < 1e-3
    assert c["edges"]["sp_type"][1] == 1 and c["P_trop"] > 0        # "CO2 condensation starts at about 3 bars"
    k = int(np.argmax(c["f_i"][:, 1] < c["f_i"][0, 1] * 0.999))
    assert 2.0e6 < c["P"][k] < 4.0e6
    assert abs(c["P_trop"] / 1e6 / 1e-2 - 1.0) < 0.5                # "the tropopause is at 1e-2 bars"


def test_converged_oracle_against_an_independent_integration():
    """The pseudo-adiabat of the reference's CO2-condensing case (tests/test_adiabat.f90:95-104) integrated a second
    way: scipy's DOP853 with its own step control and its own event location (terminal events for "CO2 reaches
    saturation" and "T reaches T_trop"), the right-hand side written from Graham et al. (2021) eq. 1 with numpy
    vectors instead of the oracle's loops.  Both at tight tolerance: T, z and the tropopause agree to 1e-9."""
    from scipy.integrate import solve_ivp
    sp = species()
    nz, T_surf, T_trop = 30, 280.0, 150.0
    e = A.make_profile(T_surf, P_I_CO2, sp, nz, EARTH_MASS, EARTH_RADIUS, 1.0, T_trop, np.ones(7), 1e-13, 1e-12)
    mass = np.array(sp["mass"])
    R = A.RGAS / 1.0e7
    psat = [(lambda T, q=q: A.sat_pressure(q, T)) if q is not None else None for q in sp["sat"]]
    cond = np.zeros(7, bool)
    P_i = P_I_CO2.copy()
    for i in range(7):
        if psat[i] is not None and P_i[i] > psat[i](T_surf):
            P_i[i], cond[i] = psat[i](T_surf), True
    P0 = P_i.sum()
    assert abs(P0 / e["P"][0] - 1.0) < 1e-14
    f_dry_share = np.where(cond, 0.0, P_i) / P_i[~cond].sum()

    def mixing(P, T):
        f = np.zeros(7)
        for i in np.where(cond)[0]:
            f[i] = min(psat[i](T) / P, 1.0)
        fd = max(1.0 - f.sum(), 1e-40)
        f[~cond] = fd * f_dry_share[~cond]
        return f, fd

    def rhs(P, u):
        T, z = u
        f, fd = mixing(P, T)
        cp = np.array([A.heat_capacity(t, T) for t in sp["thermo"]])
        L = np.array([A.latent_heat(sp["sat"][i], T) * mass[i] * 1e-7 if cond[i] else 0.0 for i in range(7)])
        beta = L / (R * T)
        s1 = np.sum(f[cond] * (cp[cond] - R * beta[cond] + R * beta[cond] ** 2))
        s2 = np.sum(beta[cond] * f[cond])
        cpd = np.sum(f_dry_share[~cond] * cp[~cond])
        dlnT = 1.0 / (fd * ((cpd * fd + s1) / (R * (fd + s2))) + s2)
        g = A.gravity(EARTH_RADIUS, EARTH_MASS, z)
        return [dlnT * T / P, -(A.RGAS * T) / (g * P * np.dot(f, mass))]

    def ev_trop(P, u):
        return u[0] - T_trop

    def ev_co2(P, u):
        return psat[1](u[0]) - mixing(P, u[0])[0][1] * P if not cond[1] else 1.0
    ev_trop.terminal = ev_co2.terminal = True
    P_start, u = P0, [T_surf, 0.0]
    pieces = []
    for _ in range(3):
        s = solve_ivp(rhs, (P_start, 1.0), u, method="DOP853", rtol=1e-12, atol=1e-10, dense_output=True,
                      events=[ev_trop, ev_co2])
        pieces.append((P_start, s.t[-1], s.sol))
        if s.status != 1:
            break
        if s.t_events[0].size:             # tropopause
            P_trop = float(s.t_events[0][0])
            break
        f, _ = mixing(s.t[-1], s.y[0, -1])  # CO2 starts to condense: restart with it on its saturation curve
        cond[1] = True
        Pd = f * s.t[-1]
        f_dry_share = np.where(cond, 0.0, Pd) / Pd[~cond].sum()
        P_start, u = float(s.t[-1]), [float(s.y[0, -1]), float(s.y[1, -1])]
    assert cond[1] and abs(P_trop / e["P_trop"] - 1.0) < 1e-9
    for k, Pk in enumerate(e["P"][:-1]):
        if Pk < P_trop:
            assert e["T"][k] == T_trop
            continue
        hi, lo, sol = next(p for p in pieces if p[0] >= Pk >= p[1])
        T, z = sol(Pk)
        assert abs(T / e["T"][k] - 1.0) < 1e-9 and abs(z - e["z"][k]) < 1e-9 * max(e["z"][k], 1e5)


def test_make_profile_input_errors():
    sp = species()
    with pytest.raises(A.AdiabatError, match="less than input"):
        A.make_profile(150.0, P_I_OCEAN, sp, 10, EARTH_MASS, EARTH_RADIUS, 1.0, 180.0, np.ones(7))
    with pytest.raises(A.AdiabatError, match="must be positive"):
        A.make_profile(300.0, -P_I_OCEAN, sp, 10, EARTH_MASS, EARTH_RADIUS, 1.0, 180.0, np.ones(7))
    with pytest.raises(A.AdiabatError, match="bigger than the surface"):
        A.make_profile(300.0, P_I_OCEAN * 1e-12, sp, 10, EARTH_MASS, EARTH_RADIUS, 1.0, 180.0, np.ones(7))
    with pytest.raises(A.AdiabatError, match="wrong shape"):
        A.make_profile(300.0, P_I_OCEAN[:3], sp, 10, EARTH_MASS, EARTH_RADIUS, 1.0, 180.0, np.ones(7))


def test_species_file_reader_on_the_reference_template():
    from clima_b200.adiabat import species_from_yaml
    path = "/root/reference/templates/AdiabatClimate/species.yaml"
    if not os.path.exists(path):
        pytest.skip("the reference tree is not on this machine")
    sp = species_from_yaml(path)
    assert sp == species()
    assert sp["names"][:3] == ["H2O", "CO2", "N2"] and abs(sp["mass"][0] - 18.01534) < 1e-12
    assert sp["sat"][2] is None and sp["sat"][0]["T_critical"] == 647.0 and sp["sat"][0]["P_ref"] == 1.0142e6
    sub = species_from_yaml(path, ["N2", "H2O"])
    assert sub["names"] == ["N2", "H2O"] and sub["sat"][0] is None


# ------------------------------------------------------------------------------------------------------------
# the device path against the oracle
# ------------------------------------------------------------------------------------------------------------
def sweep(n, seed=0):
    """Columns around the reference's test cases: warm oceans, the CO2-condensing case, dry thin atmospheres."""
    rng = np.random.default_rng(seed)
    T = rng.uniform(240.0, 360.0, n)
    P = np.empty((7, n), order="F")
    for c in range(n):
        base = (P_I_OCEAN, P_I_CO2, np.array([1e-4, 0.05, 0.8, 1e-3, 1e-6, 1e-7, 0.2]) * BAR)[c % 3]
        P[:, c] = base * rng.uniform(0.5, 2.0, 7)
    return T, P


def gpu_model(nz=24, **kw):
    from cases import small_case
    from clima_b200.adiabat import AdiabatClimate
    from clima_b200.radtran import Radtran
    tb, pr, _ = small_case(G=16, nw=10, nz=4, ncol=1, nzen=2)
    ac = AdiabatClimate(Radtran(tb, pr), species(), nz, EARTH_MASS, EARTH_RADIUS, **kw)
    return tb, pr, ac


def oracle_column(ac, T, Pi):
    return A.adiabat_column(T, Pi, ac.sp, ac.nz, ac.planet_mass, ac.planet_radius, ac.P_top, ac.T_trop, ac.RH,
                            ac.rtol, ac.atol, ac.reference_pressure)


# What two builds of this algorithm can agree on.  At the reference's default tolerances (rtol 1e-9, atol 1e-12 on
# T [K] and z [cm]) the oracle is itself 4e-8 (T), 4e-7 (z), 4e-6 (dz) away from its converged solution (rtol
# 1e-13) on these columns: z starts at 0 with an absolute tolerance of 1e-12 cm, where the error estimate is
# rounding noise, so accept/reject decisions -- and from there the whole step sequence -- differ between any two
# builds (two compilations of the device code differ from each other by as much).  compute_altitude adds a floor
# that no rtol removes: it integrates over T and mubar interpolated LINEARLY in log10 P, the step that straddles a
# kink of the interpolant picks up an error its estimator does not see (4e-3 cm of 4e5 at rtol 1e-12; 1.3e-7
# relative was seen), and which step straddles it is again build-dependent (experiments/adiabat_noise.py: a 1e-13
# perturbation of T_surf moves the oracle's own altitude error between 1e-9 and 2e-6).  So: at the default settings
# device and oracle must agree within that band (test_make_profile_batch_matches_the_oracle also measures both
# against the converged solution), and within 1e-9 on T and the mixing ratios / 5e-7 on the altitudes at rtol 1e-12.
# Full-size finding (4096 columns, test_full_size_sweep_properties_of_make_profile): 1.7 % of the columns differ between
# two builds by more than 1e-4 in dz, up to 8e-4, always at the layers around the TROPOPAUSE, where T(log10 P) of the
# altitude integration has its sharpest kink; the oracle does exactly the same under a 1e-13 perturbation of T_surf
# (column 3672 of that sweep: dz error against the converged solution between 1.8e-6 and 8.0e-4 over 60 perturbations,
# median 8.5e-6, experiments/adiabat_noise.py).  It is a property of the reference's algorithm at its default
# tolerances, not of a build, hence the dz bound below.
PROFILE_TOL = dict(T=1e-6, z=2e-5, dz=2e-3, f=1e-6)
TIGHT_TOL = dict(T=1e-9, z=5e-7, dz=5e-7, f=1e-9)


def same_profile(p, c, o, tol=PROFILE_TOL):
    assert np.max(np.abs(p["P"][:, c] / o["P"] - 1.0)) < 1e-14
    for k in ("T", "z", "dz"):
        assert np.max(np.abs(p[k][:, c] / o[k] - 1.0)) < tol[k], k
    assert np.max(np.abs(p["f_i"][:, :, c] - o["f_i"])) < tol["f"]
    assert np.max(np.abs(p["densities"][:, :, c] - o["densities"]) / o["densities"].max()) < tol["f"]
    assert abs(p["P_surf"][c] / o["P_surf"] - 1.0) < 1e-13
    assert (p["P_trop"][c] < 0) == (o["P_trop"] < 0)
    if o["P_trop"] > 0:
        assert abs(p["P_trop"][c] / o["P_trop"] - 1.0) < 10 * tol["T"]
    assert np.allclose(p["N_surface"][:, c], o["N_surface"], rtol=1e-12, atol=0)
    assert np.allclose(p["N_atmos"][:, c], o["N_atmos"], rtol=tol["dz"], atol=0)


@pytest.mark.gpu
def test_make_profile_batch_matches_the_oracle():
    _, _, ac = gpu_model(nz=24)
    ac.T_trop = 150.0
    T, P = sweep(12)
    T[1], P[:, 1] = 280.0, P_I_CO2        # tests/test_adiabat.f90:95-104 exactly
    p = ac.make_profile_batch(T, P)
    assert not p["status"].any() and np.all(p["nfcn"] > 50)
    e_dev = dict(T=0.0, z=0.0, dz=0.0)
    e_ora = dict(T=0.0, z=0.0, dz=0.0)
    for c in range(T.shape[0]):
        o = oracle_column(ac, T[c], P[:, c])
        same_profile(p, c, o)
        conv = A.adiabat_column(T[c], P[:, c], ac.sp, ac.nz, ac.planet_mass, ac.planet_radius, ac.P_top, ac.T_trop,
                                ac.RH, 1e-13, ac.atol)
        for k in e_dev:
            e_dev[k] = max(e_dev[k], float(np.max(np.abs(p[k][:, c] / conv[k] - 1.0))))
            e_ora[k] = max(e_ora[k], float(np.max(np.abs(o[k] / conv[k] - 1.0))))
    # the device is as close to the converged solution as the reference's algorithm is: on T directly, on the
    # altitudes within the band that a 1e-13 perturbation of the input moves the oracle itself through
    # (experiments/adiabat_noise.py: 1e-9 ... 2e-6 on z)
    assert e_dev["T"] < 5.0 * e_ora["T"] + 1e-9, (e_dev, e_ora)
    assert e_dev["z"] < PROFILE_TOL["z"] and e_dev["dz"] < PROFILE_TOL["dz"], (e_dev, e_ora)
    assert (p["P_trop"] > 0).any()


@pytest.mark.gpu
def test_device_on_the_tutorial_case_of_the_reference():
    """The column the reference's notebook printed numbers for (the oracle is pinned against them above): device against
    oracle at nz = 100, P_top = 1e-2, both kernels, and the printed P_surf / N_surface directly."""
    ref, T_surf, P_i, inp = tutorial_case()
    _, _, ac = gpu_model(nz=inp["nz"])
    ac.T_trop, ac.P_top, ac.rtol = inp["T_trop"], inp["P_top"], 1e-12
    o = oracle_column(ac, T_surf, P_i)
    for mode in ("group", "thread"):
        os.environ["CLIMA_RT_ADIABAT"] = mode
        try:
            p = ac.make_profile_batch([T_surf], P_i[:, None])
        finally:
            del os.environ["CLIMA_RT_ADIABAT"]
        same_profile(p, 0, o, tol=TIGHT_TOL)
        assert abs(p["P_surf"][0] / 1e6 / ref["P_surf_bar"] - 1.0) < 1e-15
        assert abs(p["N_surface"][0, 0] / ref["N_surface"][0] - 1.0) < 1e-12
        # N_atmos with today's dz is 3e-4 from the printed numbers (see the oracle test); the mixing ratios and T that
        # enter it are the pinned ones
        assert np.max(np.abs(p["N_atmos"][:, 0] / np.array(ref["N_atmos"]) - 1.0)) < 1e-3


@pytest.mark.gpu
def test_make_profile_tight_tolerance_and_knobs():
    """rtol 1e-12: device and oracle agree to 1e-9 however their steps fall; RH, T_trop, P_top reach the device."""
    _, _, ac = gpu_model(nz=16)
    ac.rtol, ac.T_trop, ac.P_top = 1e-12, 200.0, 10.0
    ac.RH = np.array([0.6, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0])
    T, P = sweep(6, seed=3)
    p = ac.make_profile_batch(T, P)
    for c in range(6):
        o = oracle_column(ac, T[c], P[:, c])
        same_profile(p, c, o, tol=TIGHT_TOL)
        assert o["P"][-1] > 10.0 and p["T"][:, c].min() >= 200.0
    q = ac.make_profile_batch(T, P)
    ac.RH = np.ones(7)
    r = ac.make_profile_batch(T, P)
    assert np.array_equal(p["T"], q["T"]) and not np.array_equal(p["T"], r["T"])


@pytest.mark.gpu
def test_make_profile_column_does_not_depend_on_the_batch():
    _, _, ac = gpu_model(nz=20)
    T, P = sweep(70, seed=5)
    p = ac.make_profile_batch(T, P)
    for c in (0, 33, 69):
        ac.make_profile(T[c], P[:, c])
        assert np.array_equal(ac.T, p["T"][:, c]) and np.array_equal(ac.dz, p["dz"][:, c])
        assert np.array_equal(ac.f_i, p["f_i"][:, :, c]) and ac.P_trop == p["P_trop"][c]
        assert ac.lapse_rate.shape == (20,) and ac.convecting_with_below.dtype == bool


@pytest.mark.gpu
def test_both_make_profile_kernels_agree(monkeypatch):
    """One lane per gas (default up to 64 columns per SM) and one thread per column (beyond) restate the same
    algorithm: same status, same tropopause, profiles within the reproducibility band of the default tolerances and
    1e-9 / 5e-7 at rtol 1e-12; each against the oracle as well."""
    _, _, ac = gpu_model(nz=20)
    ac.T_trop = 150.0
    T, P = sweep(40, seed=9)
    T[3] = 100.0
    res = {}
    for rtol, tol in ((1e-9, PROFILE_TOL), (1e-12, TIGHT_TOL)):
        ac.rtol = rtol
        for mode in ("group", "thread"):
            monkeypatch.setenv("CLIMA_RT_ADIABAT", mode)
            res[mode] = ac.make_profile_batch(T, P, check=False)
        a, b = res["group"], res["thread"]
        assert np.array_equal(a["status"], b["status"]) and a["status"][3] == 1 and a["status"].sum() == 1
        ok = a["status"] == 0
        # the surface pressure is a sum over the gases: a butterfly over the lanes against a running sum
        assert np.max(np.abs(a["P"][:, ok] / b["P"][:, ok] - 1.0)) < 1e-14
        assert np.array_equal(a["N_surface"][:, ok], b["N_surface"][:, ok])
        for k in ("T", "z", "dz"):
            assert np.max(np.abs(a[k][:, ok] / b[k][:, ok] - 1.0)) < tol[k], (k, rtol)
        assert np.max(np.abs(a["f_i"][:, :, ok] - b["f_i"][:, :, ok])) < tol["f"]
        assert np.array_equal(a["P_trop"] > 0, b["P_trop"] > 0)
        for c in (0, 17, 39):
            o = oracle_column(ac, T[c], P[:, c])
            same_profile(a, c, o, tol)
            same_profile(b, c, o, tol)


@pytest.mark.gpu
def test_eleven_gases_nasa9_and_three_condensers(monkeypatch):
    """More than 8 gases (groups of 16 lanes per column), NASA-9 heat capacities (clima_eqns.f90:94-103) next to
    Shomate ones, and a third condensable: both kernels against the oracle."""
    import copy
    from clima_b200 import synth
    from clima_b200.adiabat import AdiabatClimate
    from clima_b200.radtran import Radtran
    sp = species()
    extra = ["NH3", "He", "Ar", "SO2"]
    for k, (n, m) in enumerate(zip(extra, (17.03, 4.0026, 39.948, 64.06))):
        sp["names"].append(n)
        sp["mass"].append(m)
        # cp / R = a3 + a4 T + a5 T^2 (+ a1 / T^2 for the last range): two temperature ranges
        sp["thermo"].append(dict(model="Nasa9", temps=[0.0, 400.0, 6000.0],
                                 data=[[0.0, 0.0, 3.5 + 0.2 * k, 1.0e-3, -2.0e-7, 0.0, 0.0, 0.0, 0.0],
                                       [1.0e3, 0.0, 3.6 + 0.2 * k, 8.0e-4, -1.0e-7, 0.0, 0.0, 0.0, 0.0]]))
        sp["sat"].append(None)
    nh3 = copy.deepcopy(sp["sat"][1])       # an NH3-like condensable from CO2's fit: lower vapour pressure
    nh3.update(mu=17.03, P_ref=nh3["P_ref"] * 0.05)
    sp["sat"][7] = nh3
    names = sp["names"]
    tb = synth.make_tables(nw=10, G=16, nk=3, ncia=6, nray=3, npxs=2, cont=True, npress=8, ntemp=6, n_sol=6, ir_start=3,
                           seed=1, species=names)
    assert tb.ng == 11
    ac = AdiabatClimate(Radtran(tb, synth.default_params(tb, nzenith=2)), sp, 18, EARTH_MASS, EARTH_RADIUS)
    ac.T_trop, ac.rtol = 140.0, 1e-12
    rng = np.random.default_rng(4)
    n = 10
    T = rng.uniform(250.0, 330.0, n)
    P = np.empty((11, n), order="F")
    for c in range(n):
        P[:7, c] = (P_I_OCEAN, P_I_CO2)[c % 2] * rng.uniform(0.5, 2.0, 7)
        P[7:, c] = np.array([0.3, 1e-3, 0.01, 1e-5]) * BAR * rng.uniform(0.5, 2.0, 4)
    got = {}
    for mode in ("group", "thread"):
        monkeypatch.setenv("CLIMA_RT_ADIABAT", mode)
        got[mode] = ac.make_profile_batch(T, P)
    cond_nh3 = 0
    for c in range(n):
        o = oracle_column(ac, T[c], P[:, c])
        cond_nh3 += int(o["edges"]["sp_type"][7] == 1)
        for mode in got:
            same_profile(got[mode], c, o, tol=TIGHT_TOL)
    assert cond_nh3 > 0                       # the third condensable did condense somewhere
    monkeypatch.delenv("CLIMA_RT_ADIABAT")
    isr, olr = ac.TOA_fluxes_batch(T, P)      # and the columns go through the RT with 11 gases
    assert np.all(np.isfinite(olr)) and np.all(olr > 0.0) and np.all(isr > 0.0)


def test_particle_interpolation_oracle():
    """interpolate_particles: log-log between the given levels, end values held outside them, zero density -> tiny."""
    Pset = np.array([1.0e6, 1.0e4, 1.0e2])
    pd = np.array([[100.0], [1.0], [0.0]])
    pr = np.array([[1.0e-4], [4.0e-4], [1.0e-5]])
    d, r = A.interpolate_particles([1.0e7, 1.0e6, 1.0e5, 1.0e4, 1.0], Pset, pd, pr)
    assert np.allclose(d[:4, 0], [100.0, 100.0, 10.0, 1.0], rtol=1e-13) and d[4, 0] < 1e-300
    assert np.allclose(r[:, 0], [1.0e-4, 1.0e-4, 2.0e-4, 4.0e-4, 1.0e-5], rtol=1e-13)
    with pytest.raises(A.AdiabatError, match="Interpolation initialization"):
        A.interpolate_particles([1.0], Pset[::-1], pd, pr)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["group", "thread"])
def test_adiabat_toa_fluxes_with_particles(mode, monkeypatch):
    """Tables with two particle species: `set_particle_density_and_radii` + `interpolate_particles` on the device and
    the particle arrays of copy_atm_to_radiative_grid, checked through the fluxes against the oracle chain; the
    constructor's default (no particles, 1 micron) before any profile is set."""
    from cases import small_case
    from clima_b200.adiabat import AdiabatClimate, copy_atm_to_radiative_grid
    from clima_b200.radtran import ClimaException, Radtran
    from oracle_lib import OracleRadtran
    monkeypatch.setenv("CLIMA_RT_ADIABAT", mode)
    tb, pr, _ = small_case(G=16, nw=10, nz=4, ncol=1, nzen=2, npart=2)
    nz = 12
    ac = AdiabatClimate(Radtran(tb, pr), species(), nz, EARTH_MASS, EARTH_RADIUS)
    ac.T_trop, ac.rtol = 160.0, 1e-12
    T, P = sweep(5, seed=8)
    o = OracleRadtran(tb, pr)
    Pset = np.array([3.0e6, 1.0e5, 1.0e3, 10.0])
    pdens = np.array([[50.0, 0.0], [20.0, 5.0], [1.0, 8.0], [0.0, 0.1]])
    prad = np.array([[2.0e-4, 1.0e-5], [1.0e-4, 2.0e-5], [5.0e-5, 3.0e-5], [2.0e-5, 3.0e-5]])
    for given in (False, True):
        if given:
            ac.set_particle_density_and_radii(Pset, pdens, prad)
        r = ac.TOA_fluxes_batch(T, P, full=True)
        for c in range(5):
            col = oracle_column(ac, T[c], P[:, c])
            if given:
                pd_l, pr_l = A.interpolate_particles(col["P"], Pset, pdens, prad)
            else:
                pd_l, pr_l = np.full((nz, 2), 2.2250738585072014e-308), np.full((nz, 2), 1.0e-4)
            g = copy_atm_to_radiative_grid(col["T"], col["P"] / 1.0e6, col["densities"], col["dz"], pd_l, pr_l)
            cols = dict(nz=2 * nz + 2, ng=7, np=2, ncol=1, T_surface=np.array([T[c]]))
            for k in ("T", "P", "dz"):
                cols[k] = np.asfortranarray(g[k][:, None])
            for k in ("densities", "pdensities", "radii"):
                cols[k] = np.asfortranarray(g[k][:, :, None])
            ro = o.radiate(cols, 0, spectral=False)
            assert abs(r["ISR"][c] / ro["ISR"][0] - 1.0) < 1e-7 and abs(r["OLR"][c] / ro["OLR"][0] - 1.0) < 1e-7
            assert np.max(np.abs(r["f_total"][:, c] - ro["f_total"][:, 0])) < 1e-7 * np.max(np.abs(ro["f_total"]))
        if not given:
            olr0 = r["OLR"].copy()
    assert np.max(np.abs(r["OLR"] / olr0 - 1.0)) > 1e-3       # the particles do change the fluxes
    # the particle profile may also be set before the first profile call of a fresh handle
    ac2 = AdiabatClimate(Radtran(tb, pr), species(), nz, EARTH_MASS, EARTH_RADIUS)
    ac2.T_trop, ac2.rtol = 160.0, 1e-12
    ac2.set_particle_density_and_radii(Pset, pdens, prad)
    assert np.array_equal(ac2.TOA_fluxes_batch(T, P)[1], r["OLR"])
    with pytest.raises(ClimaException, match="Interpolation initialization"):
        ac.set_particle_density_and_radii(Pset[::-1], pdens, prad)
    with pytest.raises(ClimaException, match="`pradii` must be larger than zero"):
        ac.set_particle_density_and_radii(Pset, pdens, -prad)
    with pytest.raises(ClimaException, match="does not match the number of particles"):
        ac.set_particle_density_and_radii(Pset, pdens[:, :1], prad[:, :1])


def test_reference_pressure_oracle():
    """reference_pressure at the surface pressure is the default case bit for bit; higher up the radius of the planet is
    reached at lower pressure, gravity below it is stronger and the column is thinner."""
    sp = species()
    a = A.adiabat_column(300.0, P_I_OCEAN, sp, 20, EARTH_MASS, EARTH_RADIUS, T_trop=200.0)
    b = A.adiabat_column(300.0, P_I_OCEAN, sp, 20, EARTH_MASS, EARTH_RADIUS, T_trop=200.0, reference_pressure=a["P_surf"])
    assert np.array_equal(a["z"], b["z"]) and np.array_equal(a["dz"], b["dz"])
    c = A.adiabat_column(300.0, P_I_OCEAN, sp, 20, EARTH_MASS, EARTH_RADIUS, T_trop=200.0, reference_pressure=50.0)
    assert np.all(c["z"] < a["z"]) and 0.97 < c["z"][-1] / a["z"][-1] < 0.99 and np.array_equal(c["T"], a["T"])
    with pytest.raises(A.AdiabatError, match="outside model domain"):
        A.adiabat_column(300.0, P_I_OCEAN, sp, 20, EARTH_MASS, EARTH_RADIUS, T_trop=200.0, reference_pressure=0.5)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["group", "thread"])
def test_reference_pressure_on_the_device(mode, monkeypatch):
    """compute_altitude with reference_pressure > 0 (clima_adiabat_altitude.f90:97-169): between two edges, exactly on a
    layer centre of one column (the is_on_k / is_on_kp1 branches), at the top; outside the domain = status 11."""
    from clima_b200.radtran import ClimaException
    monkeypatch.setenv("CLIMA_RT_ADIABAT", mode)
    _, _, ac = gpu_model(nz=16)
    ac.T_trop, ac.rtol = 170.0, 1e-12
    T, P = sweep(5, seed=17)
    base = ac.make_profile_batch(T, P)
    for Pr in (1.0e3, float(base["P"][5, 2]), 1.0):
        ac.reference_pressure = Pr
        p = ac.make_profile_batch(T, P)
        for c in range(5):
            same_profile(p, c, oracle_column(ac, T[c], P[:, c]), tol=TIGHT_TOL)
        assert np.all(p["z"] < base["z"]) and np.array_equal(p["T"], base["T"])
    ac.reference_pressure = 1.0e9
    q = ac.make_profile_batch(T, P, check=False)
    assert np.all(q["status"] == 11)
    with pytest.raises(ClimaException, match="reference_pressure outside model domain"):
        ac.TOA_fluxes_batch(T, P)


@pytest.mark.gpu
def test_ocean_solubility_functions_fill_N_ocean():
    """N_ocean (clima_adiabat_general.f90:226-246) from host callbacks after the device call, against the oracle; a
    Henry's-law style function as in the reference's tutorial (cells 50-52)."""
    _, _, ac = gpu_model(nz=10)

    def henry(T_surf, ng, P_i):
        k = np.array([0.0, 3.4e-2, 6.1e-4, 7.8e-4, 1.4e-3, 9.5e-4, 1.3e-3]) * np.exp(-(T_surf - 298.0) / 50.0)
        return k * P_i
    ac.set_ocean_solubility_fcn("H2O", henry)
    T, P = sweep(6, seed=13)
    p = ac.make_profile_batch(T, P)
    assert p["N_ocean"].shape == (7, 7, 6) and (p["N_ocean"][:, 0, :] > 0).any()
    for c in range(6):
        e = A.make_profile(T[c], P[:, c], ac.sp, ac.nz, ac.planet_mass, ac.planet_radius, ac.P_top, ac.T_trop, ac.RH,
                           ocean_fcns=[henry] + [None] * 6)
        assert np.allclose(p["N_ocean"][:, :, c], e["N_ocean"], rtol=1e-12, atol=0.0)
        assert p["N_ocean"][0, 0, c] == 0.0 and not p["N_ocean"][:, 1:, c].any()
    ac.make_profile(T[0], P[:, 0])
    assert np.array_equal(ac.N_ocean, p["N_ocean"][:, :, 0])
    ac.set_ocean_solubility_fcn("H2O", None)
    assert not ac.make_profile_batch(T, P)["N_ocean"].any()
    from clima_b200.radtran import ClimaException
    with pytest.raises(ClimaException, match="not in the list of species"):
        ac.set_ocean_solubility_fcn("Xe", henry)


@pytest.mark.gpu
def test_full_size_sweep_properties_of_make_profile():
    """BASELINE's sweep size (4096 columns, nz = 50), where the oracle is too slow to be the checker: size-independent
    properties of every column, both kernels on the same batch, a subset bit-identical to the batch, 8 columns against
    the oracle."""
    _, _, ac = gpu_model(nz=50)
    n = 4096
    T, P = sweep(n, seed=31)
    p = ac.make_profile_batch(T, P)                      # one lane per gas at this size
    assert not p["status"].any() and np.all(np.isfinite(p["T"])) and np.all(np.isfinite(p["dz"]))
    assert np.max(np.abs(p["f_i"].sum(axis=1) - 1.0)) < 1e-12            # mixing ratios sum to one
    assert np.all(np.diff(p["z"], axis=0) > 0.0) and np.all(p["dz"] > 0.0)
    assert np.all(np.diff(p["P"], axis=0) < 0.0) and np.all(p["P"][0] < p["P_surf"][None, :])
    assert np.all(p["T"] >= ac.T_trop) and np.all(np.diff(p["T"], axis=0) <= 1e-9)  # cooling upward, then isothermal
    strat = p["P"] < p["P_trop"][None, :]
    assert np.all(p["T"][strat] == ac.T_trop) and (p["P_trop"] > 0).sum() > n // 2
    # hydrostatic consistency: the column mass above the surface from densities * dz against P_surf / g
    mass = np.array(ac.sp["mass"])
    col_mass = np.einsum("zgc,g,zc->c", p["densities"], mass, p["dz"]) / 6.02214076e23
    g0 = A.gravity(EARTH_RADIUS, EARTH_MASS, 0.0)
    assert np.all(np.abs(col_mass * g0 / (p["P_surf"] - ac.P_top) - 1.0) < 0.03)   # layer-mean discretisation, g(z)
    # a water-saturated column stays on the saturation curve
    wet = np.where(p["N_surface"][0] > 0)[0][:50]
    for c in wet:
        k = int(np.argmax(~(p["P"][:, c] > p["P_trop"][c]))) or 50
        f_sat = np.array([A.sat_pressure(ac.sp["sat"][0], t) for t in p["T"][:k, c]]) / p["P"][:k, c]
        assert np.max(np.abs(p["f_i"][:k, 0, c] / np.minimum(f_sat, 1.0) - 1.0)) < 1e-12
    sub = np.arange(0, n, 97)
    q = ac.make_profile_batch(T[sub], np.asfortranarray(P[:, sub]))
    for k in ("T", "z", "dz", "f_i", "P_trop"):
        assert np.array_equal(q[k], p[k][..., sub]), k
    os.environ["CLIMA_RT_ADIABAT"] = "thread"
    try:
        r = ac.make_profile_batch(T, P)
    finally:
        del os.environ["CLIMA_RT_ADIABAT"]
    assert np.max(np.abs(r["T"] / p["T"] - 1.0)) < PROFILE_TOL["T"]
    ddz = np.max(np.abs(r["dz"] / p["dz"] - 1.0), axis=0)       # per column: see the note at PROFILE_TOL
    assert ddz.max() < PROFILE_TOL["dz"] and np.median(ddz) < 1e-5 and (ddz > 1e-4).mean() < 0.05
    worst = int(np.argmax(ddz))
    k = int(np.argmax(np.abs(r["dz"][:, worst] / p["dz"][:, worst] - 1.0)))
    assert abs(np.log(p["P"][k, worst] / p["P_trop"][worst])) < 1.0      # ... and it sits at the tropopause
    for c in range(0, n, 512):
        same_profile(p, c, oracle_column(ac, T[c], P[:, c]))


@pytest.mark.gpu
def test_make_column_and_bg_gas_batches():
    """`make_column` / `make_profile_bg_gas` (clima_adiabat.f90:476-650) as lock-step Newton iterations over the batch:
    the inventories / the surface pressure are met, one column agrees with MINPACK hybrd (scipy fsolve, the reference's
    solver) run on the oracle, and the numbers the reference printed in its tutorial (cells 16-20) come out."""
    from scipy.optimize import fsolve
    _, _, ac = gpu_model(nz=12)
    ac.T_trop = 200.0
    rng = np.random.default_rng(5)
    n = 7
    T = rng.uniform(260.0, 330.0, n)
    N_i = np.empty((7, n))
    for c in range(n):
        N_i[:, c] = np.array([15.0e3, 23.0, 36.0, 1e-10, 1e-10, 1e-10, 1e-10]) * rng.uniform(0.3, 3.0, 7)
    N_i[0, 3] = 1.0e-3                                    # a column whose water does not condense at the surface
    p = ac.make_column_batch(T, N_i)
    tot = p["N_atmos"] + p["N_surface"] + p["N_ocean"].sum(axis=1)
    # x is converged to 1e-8 (hybrd's test); the inventories themselves carry the integration noise of make_profile
    assert np.max(np.abs(tot / N_i - 1.0)) < 1e-6 and p["N_surface"][0, 3] == 0.0 and (p["N_surface"][0] > 0).sum() == n - 1
    # MINPACK hybrd (mode 2, unit scaling = hybrd1) on the oracle, same residual as the reference's fcn
    c = 1
    grav = A.gravity(EARTH_RADIUS, EARTH_MASS, 0.0)

    def fcn(x):
        col = A.adiabat_column(T[c], 10.0 ** x, ac.sp, ac.nz, ac.planet_mass, ac.planet_radius, ac.P_top, ac.T_trop)
        return col["N_atmos"] + col["N_surface"] - N_i[:, c]
    x0 = np.log10(N_i[:, c] * np.array(ac.sp["mass"]) * grav)
    x, _, ier, _ = fsolve(fcn, x0, xtol=1e-8, diag=np.ones(7), full_output=True)
    # hybrd's residual is absolute and its stopping test a norm over all unknowns, so it leaves the 1e-10 mol/cm^2 trace
    # gases wherever they are once the three major inventories are met (H2 ends 95 % off here); the batch solver works
    # on relative residuals and meets every inventory.  The major gases agree:
    assert ier == 1 and np.max(np.abs(10.0 ** x[:3] / p["P_i_surf"][:3, c] - 1.0)) < 1e-6
    assert np.max(np.abs(fcn(x)[3:] / N_i[3:, c])) > 1e-3 > 1e-6 > np.max(np.abs(tot[3:, c] / N_i[3:, c] - 1.0))
    # background gas: the surface pressure is met, the other gases keep their pressures
    T2, P2 = sweep(5, seed=19)
    P2[0], P2[1] = 1.0e4, 1.0e5                            # little water and CO2, so that 2 bar are reachable with N2
    q = ac.make_profile_bg_gas_batch(T2, P2, 2.0e6, "N2")
    assert np.max(np.abs(q["P_surf"] / 2.0e6 - 1.0)) < 1e-7
    assert np.array_equal(np.delete(q["P_i_surf"], 2, axis=0), np.delete(P2, 2, axis=0))
    from clima_b200.radtran import ClimaException
    with pytest.raises(ClimaException, match="not in the list of species"):
        ac.make_profile_bg_gas_batch(T2, P2, 2.0e6, "Xe")
    with pytest.raises(ClimaException, match="root solve failed in make_profile_bg_gas"):
        ac.make_profile_bg_gas_batch(T2, P2 * 100.0, 2.0e6, "N2")   # the other gases alone exceed 2 bar
    # the reference's tutorial: T_trop = 200, P_top = 1e-2, nz = 100; N_i = [15e3, 23, 36, 1e-10 ...]
    _, _, tc = gpu_model(nz=100)
    tc.T_trop, tc.P_top = 200.0, 1.0e-2
    tc.make_column(280.0, np.array([15.0e3, 23.0, 36.0, 1e-10, 1e-10, 1e-10, 1e-10]))
    assert abs(tc.N_surface[0] / 14999.93596532 - 1.0) < 1e-8          # cell 18: printed N_surface
    assert abs(tc.N_atmos[0] / 6.40346849e-02 - 1.0) < 2e-3             # printed N_atmos (its dz predates compute_altitude)
    assert np.max(np.abs((tc.N_atmos + tc.N_surface) / np.array([1.5e4, 23.0, 36.0, 1e-10, 1e-10, 1e-10, 1e-10]) - 1.0)) < 1e-6
    tc.make_profile_bg_gas(280.0, np.array([260.0, 1.0, 1.0, 1e-10, 1e-10, 1e-10, 1e-10]) * BAR, 2.0e6, "N2")
    assert abs(tc.P_surf / 2000000.0 - 1.0) < 1e-7                      # cell 20: printed P_surf


@pytest.mark.gpu
def test_surface_temperature_batch():
    """`AdiabatClimate%surface_temperature` (clima_adiabat.f90:882-975) for a batch of compositions: ISR = OLR at every
    solution; one column against MINPACK hybrd (the reference's solver) driving the same residual; the residual of the
    oracle chain (adiabat oracle + RT oracle) at the device's answer."""
    import dataclasses
    from scipy.optimize import fsolve
    from clima_b200.adiabat import copy_atm_to_radiative_grid
    from oracle_lib import OracleRadtran
    tb, pr, ac = gpu_model(nz=14)
    ac.T_trop = 160.0
    ac.rad.params = dataclasses.replace(ac.rad.params, photon_scale_factor=0.3)   # puts the balance near 330 K
    n = 6
    P = np.asfortranarray(np.repeat(P_I_OCEAN[:, None], n, 1))
    P[1] = np.logspace(-4, 0, n) * BAR                                   # 1e-4 ... 1 bar of CO2
    T, isr, olr = ac.surface_temperature_batch(P, T_guess=300.0)
    assert np.all((T > 250.0) & (T < 400.0)) and np.max(np.abs(isr / olr - 1.0)) < 1e-6
    assert np.all(np.diff(T) > 0.0)                                     # more CO2, warmer surface
    c = 2

    def f(x):
        a, b = ac.TOA_fluxes(float(10.0 ** x[0]), P[:, c])
        return [a - b]
    x, _, ier, _ = fsolve(f, [np.log10(300.0)], xtol=1e-8, diag=[1.0], full_output=True)
    assert ier == 1 and abs(10.0 ** x[0] / T[c] - 1.0) < 1e-6
    ac.rtol = 1e-12
    Tc, _, _ = ac.surface_temperature_batch(P[:, c:c + 1], T_guess=float(T[c]))
    col = oracle_column(ac, float(Tc[0]), P[:, c])
    g = copy_atm_to_radiative_grid(col["T"], col["P"] / 1.0e6, col["densities"], col["dz"])
    cols = dict(nz=2 * 14 + 2, ng=7, np=0, ncol=1, T_surface=np.array([float(Tc[0])]))
    for k in ("T", "P", "dz"):
        cols[k] = np.asfortranarray(g[k][:, None])
    cols["densities"] = np.asfortranarray(g["densities"][:, :, None])
    ro = OracleRadtran(tb, dataclasses.replace(pr, photon_scale_factor=0.3)).radiate(cols, 0, spectral=False)
    assert abs(ro["ISR"][0] / ro["OLR"][0] - 1.0) < 1e-6
    assert abs(ac.surface_temperature(P[:, c], T_guess=310.0) / Tc[0] - 1.0) < 1e-6 and abs(ac.T_surf / Tc[0] - 1.0) < 1e-6
    from clima_b200.radtran import ClimaException
    with pytest.raises(ClimaException, match="root solve failed"):
        ac.rad.params = dataclasses.replace(ac.rad.params, photon_scale_factor=1.0e-4)   # no balance above T_trop
        ac.surface_temperature_batch(P[:, :2], T_guess=300.0)


@pytest.mark.gpu
def test_per_column_T_trop_and_solve_for_T_trop():
    """A tropopause temperature per column (clima_rt_adiabat_set_column_T_trop): each column equals the single-value
    call with its own T_trop; then the reference's solve_for_T_trop (clima_adiabat.f90:926-959, on in its own
    tests/test_adiabat.f90:27): T_trop of every column ends at the skin temperature of that column's bond albedo."""
    import dataclasses
    from clima_b200.radtran import ClimaException
    _, _, ac = gpu_model(nz=14)
    T, P = sweep(6, seed=23)
    Tt = np.array([150.0, 165.0, 180.0, 195.0, 210.0, 160.0])
    p = ac.make_profile_batch(T, P, T_trop=Tt)
    r = ac.TOA_fluxes_batch(T, P, T_trop=Tt)
    for c in range(6):
        ac.T_trop = float(Tt[c])
        q = ac.make_profile_batch(T[c:c + 1], P[:, c:c + 1])
        for k in ("T", "z", "dz", "f_i", "P_trop"):
            assert np.array_equal(q[k][..., 0], p[k][..., c]), k
        assert ac.TOA_fluxes(T[c], P[:, c])[1] == r[1][c]
        assert p["T"][:, c].min() == Tt[c] or p["P_trop"][c] < 0
    ac.T_trop = 180.0
    assert np.array_equal(ac.make_profile_batch(T, P)["T"], ac.make_profile_batch(T, P, T_trop=np.full(6, 180.0))["T"])
    with pytest.raises(ClimaException, match="tropopause temperatures, the batch has"):
        ac.make_profile_batch(T, P, T_trop=Tt[:4])
    # solve_for_T_trop
    ac.rad.params = dataclasses.replace(ac.rad.params, photon_scale_factor=0.3)
    n = 4
    Pc = np.asfortranarray(np.repeat(P_I_OCEAN[:, None], n, 1))
    Pc[1] = np.logspace(-3, 0, n) * BAR
    Ts, isr, olr, Tts = ac.surface_temperature_batch(Pc, T_guess=300.0, solve_for_T_trop=True)
    assert np.max(np.abs(isr / olr - 1.0)) < 1e-6 and np.all((Tts > 100.0) & (Tts < Ts))
    full = ac.TOA_fluxes_batch(Ts, Pc, full=True, T_trop=Tts)
    alb = full["fup_n_sol"][-1] / full["fdn_n_sol"][-1]
    skin = np.array([ac.rad.skin_temperature(a) for a in alb])
    assert np.max(np.abs(skin / Tts - 1.0)) < 1e-6
    has = full["profile"]["P_trop"] > 0            # a hot column may reach P_top before it cools to its T_trop
    assert has.any() and np.all(np.abs(full["profile"]["T"].min(axis=0)[has] / Tts[has] - 1.0) < 1e-12)


@pytest.mark.gpu
def test_surface_temperature_column_and_bg_gas_batches():
    """The two other climate solves of the reference (clima_adiabat.f90:977-1020): energy balance together with the
    inventories / with a background gas at a fixed surface pressure.  At the solution all constraints hold, and the
    answer equals the nested route (column builder inside, energy balance outside) the reference takes."""
    import dataclasses
    _, _, ac = gpu_model(nz=12)
    ac.T_trop = 160.0
    ac.rad.params = dataclasses.replace(ac.rad.params, photon_scale_factor=0.3)
    n = 3
    N_i = np.repeat(np.array([15.0e3, 1.0, 36.0, 1e-10, 1e-10, 1e-10, 1e-10])[:, None], n, 1)
    N_i[1] = [0.1, 1.0, 10.0]                                           # CO2 inventories
    T, isr, olr, P_i = ac.surface_temperature_column_batch(N_i, T_guess=310.0)
    assert np.max(np.abs(isr / olr - 1.0)) < 1e-6 and np.all(np.diff(T) > 0.0)
    p = ac.make_profile_batch(T, P_i)
    tot = p["N_atmos"] + p["N_surface"] + p["N_ocean"].sum(axis=1)
    assert np.max(np.abs(tot / N_i - 1.0)) < 1e-6
    a, b = ac.TOA_fluxes_column_batch(T, N_i)                            # nested: make_column at the answer, then the RT
    assert np.max(np.abs(a / isr - 1.0)) < 1e-6 and np.max(np.abs(b / olr - 1.0)) < 1e-6
    Pg = np.asfortranarray(np.repeat(P_I_OCEAN[:, None], n, 1))
    Pg[0], Pg[1] = 1.0e4, [1.0e3, 1.0e4, 1.0e5]                          # a dry-ish atmosphere topped up with N2 to 1.5 bar
    T2, isr2, olr2, P2 = ac.surface_temperature_bg_gas_batch(Pg, 1.5e6, "N2", T_guess=300.0)
    assert np.max(np.abs(isr2 / olr2 - 1.0)) < 1e-6
    q = ac.make_profile_batch(T2, P2)
    assert np.max(np.abs(q["P_surf"] / 1.5e6 - 1.0)) < 1e-6 and np.array_equal(np.delete(P2, 2, 0), np.delete(Pg, 2, 0))
    a2, b2 = ac.TOA_fluxes_bg_gas_batch(T2, Pg, 1.5e6, "N2")
    assert np.max(np.abs(a2 / isr2 - 1.0)) < 1e-6 and np.max(np.abs(b2 / olr2 - 1.0)) < 1e-6


@pytest.mark.gpu
def test_make_profile_failures_are_reported_per_column():
    from clima_b200.radtran import ClimaException
    _, _, ac = gpu_model(nz=12)
    T, P = sweep(5, seed=2)
    T[1] = 100.0                       # below T_trop
    P[2, 2] = -1.0                     # negative pressure
    P[:, 3] = 1e-9                     # surface pressure below P_top
    p = ac.make_profile_batch(T, P, check=False)
    assert list(p["status"]) == [0, 1, 2, 3, 0]
    same_profile(p, 4, oracle_column(ac, T[4], P[:, 4]))
    with pytest.raises(ClimaException, match='column 1: make_profile: Input "T_surf" is less than input "T_trop"'):
        ac.make_profile_batch(T, P)
    with pytest.raises(ClimaException, match="column 1"):
        ac.TOA_fluxes_batch(T, P)
    with pytest.raises(ClimaException, match="inconsistent dimensions"):
        ac.TOA_fluxes_batch(T[:2], P)
    # a heat capacity outside its temperature ranges (clima_adiabat_general.f90:618-621)
    sp = species()
    sp["thermo"][2]["temps"] = [0.0, 250.0]
    from clima_b200.adiabat import AdiabatClimate
    bad = AdiabatClimate(ac.rad, sp, 12, EARTH_MASS, EARTH_RADIUS)
    with pytest.raises(ClimaException, match="Failed to compute heat capacity"):
        bad.make_profile_batch(T[:1], P[:, :1])


@pytest.mark.gpu
@pytest.mark.parametrize("doubled", [True, False])
def test_adiabat_toa_fluxes_batch_matches_the_per_column_reference_loop(doubled):
    """AdiabatClimate%TOA_fluxes per column (make_profile, copy_atm_to_radiative_grid, rad%TOA_fluxes) restated
    with the oracles, against the one device call."""
    from clima_b200.adiabat import copy_atm_to_radiative_grid
    from oracle_lib import OracleRadtran
    tb, pr, ac = gpu_model(nz=14, double_radiative_grid=doubled)
    ac.T_trop = 160.0
    ac.rtol = 1e-12                       # converged profiles on both sides, so that 1e-9 on the fluxes is meaningful
    T, P = sweep(9, seed=7)
    r = ac.TOA_fluxes_batch(T, P, full=True)
    o = OracleRadtran(tb, pr)
    nzr = 2 * 14 + 2 if doubled else 14
    assert r["f_total"].shape == (nzr + 1, 9)
    pg = r["profile"]
    for c in range(9):
        col = oracle_column(ac, T[c], P[:, c])
        same_profile(pg, c, col, tol=TIGHT_TOL)
        # (a) the RT on the column the device built: the 1e-9 bar of the path
        # (b) the whole chain restated on the CPU: bounded by the altitude floor above
        dev = dict(T=pg["T"][:, c], P=pg["P"][:, c], densities=pg["densities"][:, :, c], dz=pg["dz"][:, c])
        for src, tol in ((dev, 1e-9), (col, 1e-7)):
            if doubled:
                g = copy_atm_to_radiative_grid(src["T"], src["P"] / 1.0e6, src["densities"], src["dz"])
            else:
                g = dict(T=src["T"], P=src["P"] / 1.0e6, densities=src["densities"], dz=src["dz"])
            cols = dict(nz=nzr, ng=7, np=0, ncol=1, T_surface=np.array([T[c]]))
            for k in ("T", "P", "dz"):
                cols[k] = np.asfortranarray(g[k][:, None])
            cols["densities"] = np.asfortranarray(g["densities"][:, :, None])
            ro = o.radiate(cols, 0, spectral=False)
            assert abs(r["ISR"][c] / ro["ISR"][0] - 1.0) < tol and abs(r["OLR"][c] / ro["OLR"][0] - 1.0) < tol
            assert np.max(np.abs(r["f_total"][:, c] - ro["f_total"][:, 0])) < tol * np.max(np.abs(ro["f_total"]))
    isr, olr = ac.TOA_fluxes(T[3], P[:, 3])
    assert isr == r["ISR"][3] or abs(isr / r["ISR"][3] - 1.0) < 1e-12
    assert abs(olr / r["OLR"][3] - 1.0) < 1e-12


@pytest.mark.gpu
def test_adiabat_needs_matching_tables():
    from cases import small_case
    from clima_b200.adiabat import AdiabatClimate
    from clima_b200.radtran import ClimaException, Radtran
    tb, pr, _ = small_case(G=16, nw=10, nz=4, ncol=1, nzen=2)
    sp = species()
    few = {k: v[:3] for k, v in sp.items()}
    with pytest.raises(ClimaException, match="differs from the radiative tables"):
        AdiabatClimate(Radtran(tb, pr), few, 10, EARTH_MASS, EARTH_RADIUS).make_profile_batch([300.0], P_I_OCEAN[:3, None])
    g = Radtran(tb, pr)
    with pytest.raises(ClimaException, match="clima_rt_adiabat_set has not been called"):
        import ctypes as C
        from clima_b200._abi import ERR_LEN
        from clima_b200.radtran import _check, _i
        err = C.create_string_buffer(ERR_LEN + 1)
        z = np.zeros(8)
        g._lib.clima_rt_make_profile_batch(g._h, _i(1), z.ctypes.data_as(C.POINTER(C.c_double)),
                                           z.ctypes.data_as(C.POINTER(C.c_double)), None, err)
        _check(err)
