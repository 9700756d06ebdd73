"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs.  Tolerance 1e-9 relative (north_star), FP64."""
import numpy as np
import pytest

from cases import TOL, compare, rel_err, small_case
from clima_b200 import synth
from clima_b200.radtran import ClimaException, Radtran
from oracle_lib import OracleRadtran

pytestmark = pytest.mark.gpu

# Tests of the SORTING RORR kernels (rorr_impl.cuh: quantised keys, certificate, worklist, closed-form
# merges, unit order).  The default RORR path is the sort-free staircase kernel (optics_rorr_cut.cu);
# these select the sorting kernels with CLIMA_RT_RORR=sort.
SORT_KERNEL_TESTS = {
    "test_rorr_fast_path_is_bit_identical_to_exact_path", "test_rorr_near_ties_go_through_the_exact_kernel",
    "test_rorr_worklist_overflow_falls_back_to_exact", "test_rorr_closed_form_merges_match_the_sorted_path",
    "test_bit_reproducible_with_worklist_and_closed_form_merges", "test_rorr_general_walk_matches_the_single_edge_walk",
    "test_full_size_fast_keys_match_exact_kernel", "test_rorr_unit_order_does_not_change_results",
}


@pytest.fixture(autouse=True)
def _rorr_flavour(request, monkeypatch):
    if request.node.name.split("[")[0] in SORT_KERNEL_TESTS:
        monkeypatch.setenv("CLIMA_RT_RORR", "sort")
    yield


def _same(a, b, bitwise, tol=1e-12):
    """Results of one column computed in different batches: bit-identical with the sorting kernels; the
    staircase kernel starts a unit from the staircases of its chain predecessor, and where sums tie
    exactly ANY staircase through the tie group is a valid sorted prefix (same integral, different
    summation order), so a column's result depends on its neighbours at the rounding level."""
    if bitwise:
        return np.array_equal(a, b)
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    s = np.max(np.abs(b)) if b.size else 0.0
    return bool(np.all(np.abs(a - b) <= tol * max(s, 1e-300)))


FLAVOURS = pytest.mark.parametrize("flavour", ["cut", "sort"])


def _set_flavour(monkeypatch, flavour):
    if flavour == "sort":
        monkeypatch.setenv("CLIMA_RT_RORR", "sort")
    else:  # the staircase kernel also for batches smaller than its automatic threshold
        monkeypatch.delenv("CLIMA_RT_RORR", raising=False)
        monkeypatch.setenv("CLIMA_RT_RORR_CHAIN", "col")


def run_both(tb, pr, cols, spectral=True, **kw):
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr, kind="parity")
    rg = g.radiate_batch(cols, spectral=spectral, **kw)
    ro = o.radiate_batch(cols, spectral=spectral, nthreads=8, **kw)
    return g, o, rg, ro


@pytest.mark.parametrize("G", [16, 8, 20, 32, 5])
def test_small_parity_ngauss(G):
    tb, pr, cols = small_case(G=G, nw=10, nz=8, ncol=3)
    g, o, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)
    assert np.all(rg["status"] == 0)


def test_opacity_parity():
    """tau, w0, g, tau_band of OpticalPropertiesResult (clima_radtran_types.f90:242-247)."""
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2)
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    g.radiate_batch(cols)
    o.radiate(cols, icol=1)
    tg = g.get_opacity(cols["nz"], icol=1)
    to = o.get_opacity(cols["nz"])
    for a, b, name in zip(tg, to, ("tau", "w0", "g", "tau_band")):
        m = np.abs(b) > 0
        e = np.max(np.abs(a[m] - b[m]) / np.abs(b[m])) if m.any() else 0.0
        assert e < 1e-11, (name, e)


@pytest.mark.parametrize("nzen", [1, 3, 8])
def test_zenith_passes(nzen):
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2, nzen=nzen)
    _, _, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)


def test_undoubled_odd_nz_and_no_hard_surface():
    tb, pr, cols = small_case(G=16, nw=10, nz=9, ncol=2, doubled=False, has_hard_surface=False)
    _, _, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)


def test_emissivity_albedo_vectors():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2)
    rng = np.random.default_rng(3)
    pr.surface_albedo = rng.uniform(0.05, 0.6, tb.nw_sol)
    pr.surface_emissivity = rng.uniform(0.7, 1.0, tb.nw_ir)
    pr.photon_scale_factor = 0.4286
    _, _, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)


def test_particles_and_custom_opacity():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2, npart=2)
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    compare(g.radiate_batch(cols, spectral=True), o.radiate_batch(cols, spectral=True))
    # custom optical properties (clima_radtran_types.f90:504-572)
    rng = np.random.default_rng(5)
    nP = 7
    log10P = np.linspace(0.5, 6.5, nP)
    dtau = np.asfortranarray(10.0 ** rng.uniform(-9, -6, (nP, tb.nw)))
    w0 = np.asfortranarray(rng.uniform(0.1, 0.9, (nP, tb.nw)))
    g0 = np.asfortranarray(rng.uniform(0.0, 0.8, (nP, tb.nw)))
    g.set_custom_optical_properties(log10P, dtau, w0, g0)
    o.set_custom(log10P, dtau, w0, g0)
    r1 = g.radiate_batch(cols, spectral=True)
    compare(r1, o.radiate_batch(cols, spectral=True))
    # all-zero custom opacity must not change fluxes (tests/test_radtran.f90:86-104)
    z = np.zeros((nP, tb.nw), order="F")
    g.set_custom_optical_properties(log10P, z, z, z)
    o.set_custom(log10P, z, z, z)
    r0 = g.radiate_batch(cols, spectral=True)
    compare(r0, o.radiate_batch(cols, spectral=True))
    g.unset_custom_optical_properties()
    o.unset_custom()
    r2 = g.radiate_batch(cols, spectral=True)
    compare(r2, o.radiate_batch(cols, spectral=True))
    assert rel_err(r0["fup_n_ir"], r2["fup_n_ir"]) < 1e-12


def test_particle_radius_out_of_range_fails_like_reference():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2, npart=1)
    cols["radii"][:] = 1.0  # 1 cm: outside the Mie table
    g = Radtran(tb, pr)
    with pytest.raises(ClimaException, match="Opacity computation failed"):
        g.radiate_batch(cols)


def test_particle_radius_out_of_range_in_an_early_chunk(monkeypatch):
    """The radius check looks at every chunk, not only at the last one (ncol=7 in chunks of 3, the
    bad radius in column 0)."""
    monkeypatch.setenv("CLIMA_RT_CHUNK_COLS", "3")
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=7, npart=1)
    cols["radii"][:, :, 0] = 1.0
    g = Radtran(tb, pr)
    with pytest.raises(ClimaException, match="Opacity computation failed"):
        g.radiate_batch(cols)
    # the same batch without the bad column is fine
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=7, npart=1)
    Radtran(tb, pr).radiate_batch(cols)


def test_stale_solar_state_is_refused_when_it_is_not_the_callers(monkeypatch):
    """compute_solar=.false. (clima_radtran.f90:286-289) reuses the previous call's solar fluxes.
    A fresh handle has the reference's all-zero wrk_sol; after a multi-chunk solar call only the
    last chunk's fluxes remain and the call must fail instead of mixing columns."""
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=7)
    o = OracleRadtran(tb, pr)
    g = Radtran(tb, pr)
    r0 = g.radiate_batch(cols, compute_solar=False)  # fresh handle: zeros, like the reference
    assert np.all(r0["fup_n_sol"] == 0.0) and np.all(r0["fdn_n_sol"] == 0.0)
    ro = o.radiate(cols, 2, compute_solar=False, spectral=False)
    assert rel_err(r0["f_total"][:, 2], ro["f_total"][:, 0]) < TOL
    r1 = g.radiate_batch(cols)
    r2 = g.radiate_batch(cols, compute_solar=False)  # same single-chunk batch: allowed
    assert np.array_equal(r1["f_total"], r2["f_total"])
    monkeypatch.setenv("CLIMA_RT_CHUNK_COLS", "3")
    g3 = Radtran(tb, pr)
    g3.radiate_batch(cols)  # three chunks
    with pytest.raises(ClimaException, match="solar"):
        g3.radiate_batch(cols, compute_solar=False)


def test_null_handle_is_an_error_string_not_a_crash():
    import ctypes as C
    from clima_b200.radtran import load_library
    from clima_b200._abi import ERR_LEN
    lib = load_library()
    err = C.create_string_buffer(ERR_LEN + 1)
    lib.clima_rt_set_params(None, None, err)
    assert b"NULL" in err.value
    n = C.c_int64(-1)
    lib.clima_rt_get_launch_count(None, C.byref(n))
    assert n.value == 0
    lib.clima_rt_unset_custom_optical_properties(None)
    ms = (C.c_double * 6)()
    lib.clima_rt_get_stage_ms(None, ms)


def test_frozen_opacity_jacobian_mode():
    """compute_opacity=.false., compute_solar=.false. (clima_adiabat_solve.f90:798-820): perturbed
    temperatures re-use ONE column's opacity and the stale solar fluxes."""
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=1, nzen=1)
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    rg0 = g.radiate_batch(cols, spectral=False)
    ro0 = o.radiate(cols, 0, spectral=False)
    compare(rg0, ro0, spectral=False)
    nz = cols["nz"]
    npert = 5
    pert = {k: v for k, v in cols.items()}
    pert["ncol"] = npert
    pert["T_surface"] = np.repeat(cols["T_surface"], npert) * (1 + 1e-2 * np.arange(npert))
    T = np.repeat(cols["T"], npert, axis=1).copy(order="F")
    for c in range(npert):
        T[c % nz:, c] *= 1.01
    pert["T"] = T
    for k in ("P", "dz"):
        pert[k] = np.asfortranarray(np.repeat(cols[k], npert, axis=1))
    pert["densities"] = np.asfortranarray(np.repeat(cols["densities"], npert, axis=2))
    rg = g.radiate_batch(pert, compute_solar=False, compute_opacity=False)
    for c in range(npert):
        ro = o.radiate(pert, c, compute_solar=False, compute_opacity=False, spectral=False)
        sub = {k: (v[..., c:c + 1] if v.ndim > 1 else v[c:c + 1]) for k, v in rg.items()}
        compare(sub, ro, spectral=False)


def test_same_batch_frozen_opacity():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=3)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols)
    r2 = g.radiate_batch(cols, compute_opacity=False)
    for k in ("fup_n_ir", "fup_n_sol", "f_total"):
        assert np.array_equal(r1[k], r2[k])


@FLAVOURS
def test_single_column_api_matches_batch(flavour, monkeypatch):
    _set_flavour(monkeypatch, flavour)
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2)
    g = Radtran(tb, pr)
    rb = g.radiate_batch(cols)
    isr, olr = g.TOA_fluxes(cols["T_surface"][1], cols["T"][:, 1], cols["P"][:, 1],
                            cols["densities"][:, :, 1], cols["dz"][:, 1])
    bw = flavour == "sort"
    assert _same(isr, rb["ISR"][1], bw) and _same(olr, rb["OLR"][1], bw)
    assert _same(g.f_total, rb["f_total"][:, 1], bw)
    assert g.wrk_ir.fup_a.shape == (cols["nz"] + 1, tb.nw_ir)


def test_error_convention():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2)
    g = Radtran(tb, pr)
    bad = dict(cols)
    bad["ng"] = cols["ng"] - 1
    with pytest.raises(ClimaException, match="densities"):
        g.radiate_batch(bad)
    with pytest.raises(ClimaException, match="retained"):
        Radtran(tb, pr).radiate_batch(cols, compute_opacity=False)


@FLAVOURS
def test_chunking_is_transparent(flavour, monkeypatch):
    _set_flavour(monkeypatch, flavour)
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=7)
    r1 = Radtran(tb, pr).radiate_batch(cols, spectral=True)
    monkeypatch.setenv("CLIMA_RT_CHUNK_COLS", "3")
    r2 = Radtran(tb, pr).radiate_batch(cols, spectral=True)
    for k in r1:
        assert _same(r1[k], r2[k], flavour == "sort"), k


def test_bit_reproducible():
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=4)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols, spectral=True)
    r2 = g.radiate_batch(cols, spectral=True)
    for k in r1:
        assert np.array_equal(r1[k], r2[k]), k


def test_scanon_full_size_parity():
    """S-canon dimensions (nw=100, G=16, K=5, nz_r=202, Z=4), a few columns of the cfg2 sweep."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(6, nz=100)
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr, kind="parity")
    rg = g.radiate_batch(cols, spectral=True)
    ro = o.radiate_batch(cols, spectral=True, nthreads=8)
    w = compare(rg, ro, verbose=True)
    assert w["OLR"] < TOL and w["ISR"] < TOL


def test_modern_earth_like_single_column():
    """cfg1 shape: nz=200 undoubled, 8 zenith angles, albedo 0.15 (tests/test_radtran.f90:33-38)."""
    tb = synth.make_tables()
    pr = synth.default_params(tb, nzenith=8, albedo=0.15)
    cols = synth.make_columns(1, nz=200, doubled=False, T_surf_range=(288.0, 288.0))
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr, kind="parity")
    compare(g.radiate_batch(cols, spectral=True), o.radiate_batch(cols, spectral=True), verbose=True)


def test_rorr_fast_path_is_bit_identical_to_exact_path(monkeypatch):
    """The 32-bit-key fast RORR kernel must return the permutation of the stable 64-bit sort (or
    hand the unit to the exact kernel): results are bit-identical with CLIMA_RT_RORR=exact."""
    tb, pr, cols = small_case(G=16, nw=12, nz=10, ncol=3)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols, spectral=True)
    units, exact, ovf = g.rorr_stats()
    assert units > 0 and exact < units and ovf == 0
    monkeypatch.setenv("CLIMA_RT_RORR", "exact")
    g2 = Radtran(tb, pr)
    r2 = g2.radiate_batch(cols, spectral=True)
    assert g2.rorr_stats()[1] == g2.rorr_stats()[0]
    for k in r1:
        assert np.array_equal(r1[k], r2[k]), k


@pytest.mark.parametrize("G", [8, 16, 32])
def test_rorr_near_ties_go_through_the_exact_kernel(G):
    """k-distributions made of two tight clusters (steps of 1e-13 in log10 k, six decades apart):
    the sums a_i + b_j collide in the 32-bit quantisation, the fast kernel must notice and defer to
    the exact kernel, and the result must still match the oracle."""
    tb, pr, cols = small_case(G=G, nw=6, nz=6, ncol=2)
    for kt in tb.k[:3]:
        gidx = np.arange(G)
        step = np.where(gidx < G // 2, 0.0, 6.0) + 1e-13 * gidx
        kt.log10k[...] = kt.log10k[0:1, :, :, :] + step[:, None, None, None]
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    compare(g.radiate_batch(cols, spectral=True), o.radiate_batch(cols, spectral=True))
    units, exact, ovf = g.rorr_stats()
    assert exact > 0


def test_rorr_worklist_overflow_falls_back_to_exact(monkeypatch):
    """Every unit uncertified (all sums tied in the quantisation) -> worklist overflow -> the chunk
    is redone by the exact kernel; results still match the oracle."""
    tb, pr, cols = small_case(G=16, nw=40, nz=60, ncol=40, doubled=False)
    for kt in tb.k:
        gidx = np.arange(16)
        step = np.where(gidx < 8, 0.0, 6.0) + 1e-13 * gidx
        kt.log10k[...] = kt.log10k[0:1, :, :, :] + step[:, None, None, None]
    g = Radtran(tb, pr)
    rg = g.radiate_batch(cols)
    units, exact, ovf = g.rorr_stats()
    assert ovf >= 1
    sub = {k: (v[..., :2] if v.ndim > 1 else v[:2]) for k, v in cols.items() if isinstance(v, np.ndarray)}
    c2 = dict(cols, **sub); c2["ncol"] = 2
    ro = OracleRadtran(tb, pr).radiate_batch(c2)
    rg2 = {k: (v[..., :2] if v.ndim > 1 else v[:2]) for k, v in rg.items()}
    compare(rg2, ro, spectral=False)


def test_rorr_closed_form_merges_match_the_sorted_path(monkeypatch):
    """Merges whose rows (or columns) do not interleave skip the sort and use the closed form
    a_i + <b> (rorr_impl.cuh).  With CLIMA_RT_RORR_NOSKIP=1 every merge is sorted and rebinned; the
    two must agree to rounding (the sorted path carries the rounding of its 256-term cumulative
    weights), far inside the 1e-9 contract, on a batch where both kinds of merge occur."""
    tb, pr, cols = small_case(G=16, nw=12, nz=10, ncol=3, nk=3)
    # make one gas negligible so that its merge is row-major for sure
    tb.k[2].log10k[...] -= 12.0
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols, spectral=True)
    t1 = g.get_opacity(cols["nz"], icol=1)[0]
    monkeypatch.setenv("CLIMA_RT_RORR_NOSKIP", "1")
    g2 = Radtran(tb, pr)
    r2 = g2.radiate_batch(cols, spectral=True)
    t2 = g2.get_opacity(cols["nz"], icol=1)[0]
    assert np.max(np.abs(t1 - t2) / np.abs(t2)) < 1e-12
    assert not np.array_equal(t1, t2)  # the closed form was actually taken somewhere
    compare(r1, r2, tol=1e-12)
    compare(r1, OracleRadtran(tb, pr).radiate_batch(cols, spectral=True))


@FLAVOURS
def test_sharded_matches_single(flavour, monkeypatch):
    """clima_b200/shard.py with the production compute callable: the blocks of a 2-way partition,
    computed one after the other on this GPU and re-assembled, equal the single-batch result bit for
    bit (columns are independent units; the multi-process gather itself is covered on CPU by
    tests/test_shard_gloo.py)."""
    from clima_b200 import shard
    _set_flavour(monkeypatch, flavour)
    bw = flavour == "sort"
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=5)
    g = Radtran(tb, pr)
    whole = g.radiate_batch(cols)
    parts = [g.radiate_batch(shard.slice_columns(cols, b, e)) for b, e in shard.partition(5, 2)]
    for k in shard.RESULT_PROFILES:
        assert _same(np.concatenate([p[k] for p in parts], axis=1), whole[k], bw), k
    for k in shard.RESULT_SCALARS:
        assert _same(np.concatenate([p[k] for p in parts]), whole[k], bw), k
    one = shard.radiate_sharded(g.radiate_batch, cols, dist=None)
    for k in shard.RESULT_PROFILES + shard.RESULT_SCALARS:
        assert np.array_equal(one[k], whole[k]), k


def test_bit_reproducible_with_worklist_and_closed_form_merges():
    """Run-to-run bit reproducibility on a batch large enough that the exact kernel regroups
    uncertified units (worklist order comes from atomics) and closed-form merges occur: the result
    of a unit must not depend on which other units share its warp."""
    tb = synth.make_tables(nw=30, n_sol=20, ir_start=10)
    pr = synth.default_params(tb, nzenith=2)
    cols = synth.make_columns(48, nz=40, seed=7, doubled=True)
    g = Radtran(tb, pr)
    r1 = {k: v.copy() for k, v in g.radiate_batch(cols, spectral=True).items()}
    units, exact, ovf = g.rorr_stats()
    assert 0 < exact < units
    for _ in range(3):
        r2 = g.radiate_batch(cols, spectral=True)
        for k in r1:
            assert np.array_equal(r1[k], r2[k]), k


def test_rorr_general_walk_matches_the_single_edge_walk(monkeypatch):
    """Gauss-Legendre weight sets take the single-edge rebin walk (rt_rebin1); the general looping
    walk (forced here) must give bit-identical opacities."""
    tb, pr, cols = small_case(G=16, nw=12, nz=10, ncol=3)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols, spectral=True)
    monkeypatch.setenv("CLIMA_RT_RORR_GENERAL_WALK", "1")
    g2 = Radtran(tb, pr)
    r2 = g2.radiate_batch(cols, spectral=True)
    for k in r1:
        assert np.array_equal(r1[k], r2[k]), k


@pytest.mark.parametrize("flavour", ["default", "staircase"])
@pytest.mark.parametrize("G", [5, 16])
def test_rorr_weights_with_several_edges_per_element(G, flavour, monkeypatch):
    """A weight set in which one sorted element (weight w_i*w_j) spans several new bins -- one
    dominant g-point next to very narrow ones -- selects the general walk at table load
    (capi.cu, DevTables::pad0 bit 1) and must still match the oracle."""
    if flavour == "staircase":  # the staircase kernel has no single-edge assumption to fall back from
        monkeypatch.setenv("CLIMA_RT_RORR_CHAIN", "col")
    tb, pr, cols = small_case(G=G, nw=8, nz=8, ncol=2)
    w = np.full(G, 0.02 / (G - 1))
    w[G // 2] = 0.98
    tb.weights[...] = w / w.sum()
    assert (tb.weights.max() ** 2) > tb.weights.min()
    g, o, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)


def test_jacobian_batch_entry_matches_the_reference_loop():
    """clima_rt_jacobian_batch == the loop of AdiabatClimate_jacobian_from_base
    (clima_adiabat_solve.f90:798-820): for each perturbed model-grid temperature profile, copy to the
    doubled radiative grid (clima_adiabat.f90:729-773), radiate with compute_opacity=.false.,
    compute_solar=.false., take f_total(2i-1).  Checked against the oracle run column by column."""
    nzm = 7
    tb, pr, base = small_case(G=16, nw=10, nz=nzm, ncol=1, nzen=1, doubled=True)
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    g.radiate_batch(base, spectral=False)   # objective at the base state (retains opacity + solar)
    o.radiate(base, 0, spectral=False)
    nzr = base["nz"]
    assert nzr == 2 * nzm + 2
    Tm = base["T"][0:2 * nzm:2, 0].copy()
    Ts0 = float(base["T_surface"][0])
    ncol = nzm + 1
    T = np.repeat(Tm[:, None], ncol, axis=1).copy(order="F")
    Ts = np.full(ncol, Ts0)
    eps = 1.0e-2
    Ts[0] += eps * Ts0
    for i in range(nzm):
        T[i, i + 1] += eps * Tm[i]
    f = g.jacobian_batch(Ts, T)
    assert f.shape == (nzm + 1, ncol)
    for c in range(ncol):
        Tr = np.empty(nzr)
        Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
        pert = dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None]))
        ro = o.radiate(pert, 0, compute_solar=False, compute_opacity=False, spectral=False)
        ref = ro["f_total"][0:2 * nzm + 1:2, 0]
        assert rel_err(f[:, c], ref) < TOL
    # the base column itself is reproduced bit for bit by an unperturbed entry
    f0 = g.jacobian_batch(np.array([Ts0]), Tm[:, None])
    r0 = g.radiate_batch(base, compute_solar=False, compute_opacity=False)
    assert np.array_equal(f0[:, 0], r0["f_total"][0:2 * nzm + 1:2, 0])


def test_jacobian_batch_without_base_fails_loudly():
    tb, pr, cols = small_case(G=16, nw=10, nz=6, ncol=2, nzen=1)
    g = Radtran(tb, pr)
    with pytest.raises(ClimaException, match="base column"):
        g.jacobian_batch(np.array([300.0]), np.full((6, 1), 250.0))


def test_early_mars_like_columns_cfg3():
    """cfg3 shape: CO2-H2O-N2 columns with CIA on a pCO2 x T_surf grid, Mars gravity, photon scale
    0.4286, albedo 0.2, doubled grid (examples/EarlyMars)."""
    tb = synth.make_tables()
    pr = synth.default_params(tb, nzenith=4, albedo=0.2, photon_scale_factor=0.4286)
    cols = synth.make_columns(6, nz=40, T_surf_range=(200.0, 320.0), pco2_range=(1.0e-2, 10.0 ** 0.5),
                              gravity=371.0, seed=7)
    g, o, rg, ro = run_both(tb, pr, cols, spectral=False)
    compare(rg, ro, spectral=False)


def test_randomised_composition_cfg5():
    """cfg5 shape: log-uniform mixing ratios per column (no similarity between neighbouring columns:
    the column-grouped RORR unit order must not matter)."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(7, nz=30, randomize=True, seed=11)
    g, o, rg, ro = run_both(tb, pr, cols, spectral=False)
    compare(rg, ro, spectral=False)


def test_full_size_fast_keys_match_exact_kernel(monkeypatch):
    """Size-independent property at the bench dimensions: with saturated (plateau) k-distributions
    most units go through the two-window fast keys; the fluxes must agree with a run that sends
    every unit through the 64-bit-key kernel.  Not bitwise at this size: the 64-bit keys carry the
    element index in their low 8 mantissa bits, so sums closer than 2^-44 relative are ranked by
    index there, while the certified fast path ranks them by exact value (what mrgrnk does); the
    difference is bounded by 2^-44 * w/dE < 1e-13 relative per rebinned value."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(48, nz=100)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols)
    units, exact, ovf = g.rorr_stats()
    assert 0 < exact < 0.05 * units, (units, exact)
    monkeypatch.setenv("CLIMA_RT_RORR", "exact")
    r2 = Radtran(tb, pr).radiate_batch(cols)
    compare(r1, r2, spectral=False, tol=1e-12)


def test_absent_gases_give_exact_ties():
    """A k-gas with zero density contributes b_j = 0 for every g-point: all sums of a row tie exactly
    (the stable order of mrgrnk is the index order) and the merge must return the running mixture
    unchanged; with the FIRST k-gas absent the mixture starts from zeros."""
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=3, nk=3)
    for sp in (tb.k[1].sp_ind - 1, tb.k[0].sp_ind - 1):
        c2 = dict(cols)
        d = cols["densities"].copy(order="F")
        d[:, sp, :] = 0.0
        c2["densities"] = d
        g, o, rg, ro = run_both(tb, pr, c2)
        compare(rg, ro)



@pytest.mark.parametrize("scale", [1.0e6, 1.0e-10])
def test_extreme_optical_depths(scale):
    """Columns a million times denser (layer optical depths far beyond the point where exp(-lambda*tau)
    underflows: the solver's lean exp holds its argument at -700) and ten orders thinner (every layer
    under the thin-layer guard of two_stream_ir :216-227 and the tau clamps of radiate.f90).

    In the thin limit the Toon matrix entries cancel (B = e2*e2n - e4*e4n ~ 2*lambda*tau), so ANY two
    evaluation orders of the same recurrence differ by ~1e-16/(lambda*tau) of the flux scale; the
    downward IR flux itself is ~1e-5 of the upward one there.  The contract is therefore checked
    against the common flux scale of the channel (what f_total sees), not per array."""
    tb, pr, cols = small_case(G=16, nw=12, nz=12, ncol=3, nk=3)
    c2 = dict(cols)
    c2["densities"] = (cols["densities"] * scale).copy(order="F")
    g, o, rg, ro = run_both(tb, pr, c2)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol"):
        assert np.all(np.isfinite(rg[k])), k
    if scale > 1.0:
        compare(rg, ro)
        return
    for ch in ("ir", "sol"):
        s_ch = max(np.max(np.abs(ro[f"fup_n_{ch}"])), np.max(np.abs(ro[f"fdn_n_{ch}"])))
        for k in (f"fup_n_{ch}", f"fdn_n_{ch}"):
            assert np.max(np.abs(rg[k] - ro[k])) <= TOL * s_ch, k
    for k in ("OLR", "ISR"):
        assert np.max(np.abs(rg[k] - ro[k]) / np.abs(ro[k])) <= TOL, k


def test_tables_from_the_load_time_packer(tmp_path):
    """N2: Rayleigh cross sections and the solar photon flux produced by clima_b200.loader (the
    restatement of create_RayleighXsection / read_stellar_flux) drive the device path."""
    from clima_b200 import loader
    tb, pr, cols = small_case(G=16, nw=12, nz=10, ncol=3, nk=3)
    ray = {"N2": {"data": {"Delta": 0.0305, "A": 29.06e-5, "B": 7.70e-3}},
           "CO2": {"data": {"Delta": 0.0805, "A": 43.9e-5, "B": 6.4e-3}}}
    tb.ray_sp_ind, tb.ray_xs = loader.rayleigh_table(ray, tb.species_names, tb.wavl)
    assert len(tb.ray_sp_ind) == 2
    lam = np.geomspace(0.5 * tb.wavl[0], 2.0 * tb.wavl[-1], 300)
    star = tmp_path / "star.txt"
    star.write_text("nm mW/m2/nm\n" + "\n".join(f"{a:.8e} {1.0e3 * np.exp(-((np.log(a / 550.0)) ** 2)):.8e}" for a in lam))
    i0, i1 = tb.sol_ind
    tb.photons_sol = loader.read_stellar_flux(str(star), tb.wavl[i0 - 1:i1 + 1])
    assert tb.photons_sol.shape == (i1 - i0 + 1,) and np.all(tb.photons_sol > 0.0)
    g, o, rg, ro = run_both(tb, pr, cols)
    compare(rg, ro)


def test_bolometric_flux_helpers_and_radiation_enhancement():
    """Host helpers of type Radtran (clima_radtran.f90:345-411): set_bolometric_flux rescales
    photon_scale_factor so that bolometric_flux() returns the request, the solar fluxes follow
    linearly (and match the oracle run with the same scale factor), skin / equilibrium temperatures
    follow clima_eqns.f90:248-261, apply_radiation_enhancement rebuilds f_total."""
    import dataclasses
    tb, pr, cols = small_case(G=16, nw=12, nz=10, ncol=2, nk=3)
    g = Radtran(tb, pr)
    f0 = g.bolometric_flux()
    freq = tb.freq[tb.sol_ind[0] - 1:tb.sol_ind[1] + 1]
    assert f0 == pytest.approx(pr.photon_scale_factor * np.sum(tb.photons_sol * (freq[:-1] - freq[1:])) / 1.0e3, rel=1e-14)
    r0 = {k: v.copy() for k, v in g.radiate_batch(cols).items()}
    g.set_bolometric_flux(1361.0)
    assert g.bolometric_flux() == pytest.approx(1361.0, rel=1e-14)
    scale = g.params.photon_scale_factor / pr.photon_scale_factor
    r1 = g.radiate_batch(cols)
    np.testing.assert_allclose(r1["ISR"], r0["ISR"] * scale, rtol=1e-13)
    np.testing.assert_allclose(r1["fdn_n_sol"], r0["fdn_n_sol"] * scale, rtol=1e-13)
    np.testing.assert_array_equal(r1["fup_n_ir"], r0["fup_n_ir"])
    ro = OracleRadtran(tb, dataclasses.replace(pr, photon_scale_factor=g.params.photon_scale_factor)).radiate_batch(cols)
    compare(r1, ro, spectral=False)
    teq = g.equilibrium_temperature(0.3)
    assert teq == pytest.approx((1361.0 * 0.7 / (4.0 * 5.670374419e-8)) ** 0.25, rel=1e-14)
    assert g.skin_temperature(0.3) == pytest.approx(teq * 0.5 ** 0.25, rel=1e-15)
    r2 = {k: v.copy() for k, v in r1.items()}
    Radtran.apply_radiation_enhancement(r2, 1.25)
    np.testing.assert_allclose(r2["fdn_n_sol"], 1.25 * r1["fdn_n_sol"], rtol=1e-15)
    np.testing.assert_allclose(r2["f_total"], (r2["fdn_n_sol"] - r2["fup_n_sol"]) + (r1["fdn_n_ir"] - r1["fup_n_ir"]), rtol=1e-15, atol=0)


def test_rorr_unit_order_does_not_change_results(monkeypatch):
    """The RORR kernels group 4 units per warp either as the same (bin, slot) of 4 adjacent columns
    (chosen when neighbouring columns are alike) or as 4 adjacent slots of one column; the order is a
    scheduling choice only.  Alike columns (a sweep) and unrelated ones (randomised), both orders
    forced, against the automatic choice and the oracle."""
    tb, pr, _ = small_case(G=16, nw=10, nz=12, ncol=9, nk=3)
    for kw in (dict(), dict(randomize=True, seed=5)):
        cols = synth.make_columns(9, nz=12, **kw)
        monkeypatch.delenv("CLIMA_RT_RORR_GROUP", raising=False)
        g, o, r_auto, ro = run_both(tb, pr, cols, spectral=False)
        compare(r_auto, ro, spectral=False)
        res = {}
        for mode in ("layers", "columns"):
            monkeypatch.setenv("CLIMA_RT_RORR_GROUP", mode)
            res[mode] = Radtran(tb, pr).radiate_batch(cols)
        compare(res["layers"], res["columns"], spectral=False, tol=1e-12)
        compare(r_auto, res["columns"], spectral=False, tol=1e-12)


# ------------------------------------------------------------------ sort-free (staircase) RORR kernel
def _opacity_err(g1, g2, nz, icols):
    e = 0.0
    for i in icols:
        for a, b in zip(g1.get_opacity(nz, icol=i), g2.get_opacity(nz, icol=i)):
            m = np.abs(b) > 0
            if m.any():
                e = max(e, float(np.max(np.abs(a[m] - b[m]) / np.abs(b[m]))))
    return e


@pytest.mark.parametrize("G", [16, 8, 11, 32])
def test_staircase_kernel_matches_the_sorting_kernels(G, monkeypatch):
    """optics_rorr_cut.cu (no sort: the rebin from staircase cuts, chains over columns or over layer
    slots) against the sorting kernels (mrgrnk order + rebin walk) on the same columns: tau, w0, tau_band
    to 1e-12 elementwise, fluxes to 1e-11; and against the oracle."""
    tb, pr, cols = small_case(G=G, nw=10, nz=14, ncol=7, nk=3)
    monkeypatch.setenv("CLIMA_RT_RORR", "sort")
    gs = Radtran(tb, pr)
    rs = gs.radiate_batch(cols, spectral=True)
    monkeypatch.delenv("CLIMA_RT_RORR")
    ro = OracleRadtran(tb, pr).radiate_batch(cols, spectral=True)
    for chain in ("slot", "col"):
        monkeypatch.setenv("CLIMA_RT_RORR_CHAIN", chain)
        gc = Radtran(tb, pr)
        rc = gc.radiate_batch(cols, spectral=True)
        units, exact, ovf = gc.rorr_stats()
        assert units > 0 and exact == 0 and ovf == 0, (units, exact, ovf)
        assert _opacity_err(gc, gs, cols["nz"], range(7)) < 1e-12
        compare(rc, rs, tol=1e-11)
        compare(rc, ro)


@pytest.mark.parametrize("chain", ["col", "slot"])
@pytest.mark.parametrize("kw", [dict(), dict(randomize=True, seed=11),
                                dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0)])
def test_staircase_kernel_full_size_columns(kw, chain, monkeypatch):
    """cfg2 / cfg5 / cfg3 columns at S-canon size (nz_r=202, nw=100, G=16, K=5), 16 columns each: the
    staircase kernel (both chain directions) against the oracle, and tau against the sorting kernels."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(16, nz=100, **kw)
    monkeypatch.setenv("CLIMA_RT_RORR_CHAIN", chain)
    g = Radtran(tb, pr)
    rg = g.radiate_batch(cols, spectral=True)
    units, exact, ovf = g.rorr_stats()
    assert exact <= 1e-4 * units and ovf == 0, (units, exact, ovf)
    ro = OracleRadtran(tb, pr, kind="parity").radiate_batch(cols, spectral=True, nthreads=8)
    compare(rg, ro)
    assert g.rorr_cut_stats()["last_kernel"] == (1 if chain == "col" else 0)
    monkeypatch.delenv("CLIMA_RT_RORR_CHAIN")
    monkeypatch.setenv("CLIMA_RT_RORR", "sort")
    gs = Radtran(tb, pr)
    gs.radiate_batch(cols)
    assert gs.rorr_cut_stats()["last_kernel"] == 2
    assert _opacity_err(g, gs, cols["nz"], (0, 7, 15)) < 1e-11


@pytest.mark.parametrize("chain", ["col", "slot"])
def test_staircase_kernel_ties_plateaus_and_absent_gases(chain, monkeypatch):
    """Exact ties (two clusters of identical k values, an absent gas: b = 0) and near ties (steps of
    1e-13 in log10 k): any staircase through a tie group is a valid sorted prefix, the rebin does not
    depend on the order inside it."""
    tb, pr, cols = small_case(G=16, nw=8, nz=8, ncol=5, nk=3)
    gidx = np.arange(16)
    tb.k[0].log10k[...] = tb.k[0].log10k[0:1] + (np.where(gidx < 8, 0.0, 6.0) + 1e-13 * gidx)[:, None, None, None]
    tb.k[1].log10k[...] = tb.k[1].log10k[0:1] + np.where(gidx < 5, 0.0, 3.0)[:, None, None, None]
    d = cols["densities"].copy(order="F")
    d[:, tb.k[2].sp_ind - 1, 1::2] = 0.0
    c2 = dict(cols, densities=d)
    monkeypatch.setenv("CLIMA_RT_RORR_CHAIN", chain)
    g, o, rg, ro = run_both(tb, pr, c2)
    compare(rg, ro)
    assert g.rorr_stats()[1] == 0


def test_cfg1_modern_earth_template_column():
    """BASELINE.json configs[0] from the reference's own template files (tests/test_radtran.f90:33-65):
    atmosphere.txt (200 levels) + Sun_now.txt, 8 zenith angles, albedo 0.15; synthetic k-tables."""
    from cases import cfg1_template_case
    tb, pr, cols = cfg1_template_case()
    assert cols["nz"] == 200 and abs(cols["T_surface"][0] - 288.15) < 0.01
    g, o, rg, ro = run_both(tb, pr, cols)
    w = compare(rg, ro, verbose=True)
    assert w["OLR"] < TOL and w["ISR"] < TOL
    assert 2.0e5 < rg["OLR"][0] < 4.5e5  # mW/m2: an Earth-like OLR


@pytest.mark.parametrize("eps", [1e-3, 1e-6, 1e-9, 1e-12])
def test_solar_resonance_lambda_u0(eps):
    """two_stream_solar divides by denom = lambda**2 - 1/u0**2 without a guard
    (clima_radtran_twostream.f90:80; SURVEY Appendix C.10 "preserve").  The zenith cosine is put a
    relative distance eps from 1/lambda of one (layer, bin, g-point) of a Rayleigh-dominated column
    (g = 0: lambda**2 = 3 (1 - w0)).  Near the pole C+ and C- grow like 1/eps while the boundary
    conditions subtract them again, so ANY two evaluation orders differ by ~1e-16/eps in that term;
    the GPU (reciprocal-multiply) must stay within that of the oracle (IEEE division) and both must
    stay finite."""
    from cases import rayleigh_dominated_case
    tb, pr, cols = rayleigh_dominated_case()
    g0 = Radtran(tb, pr)
    g0.radiate_batch(cols)
    tau, w0, gt, _ = g0.get_opacity(cols["nz"], icol=0)
    lam = np.sqrt(3.0 * (1.0 - w0)) if np.all(gt == 0) else None
    assert lam is not None
    sol = slice(tb.sol_ind[0] - 1, tb.sol_ind[1])
    cand = np.argwhere((1.0 / lam[:, :, sol] > 0.6) & (1.0 / lam[:, :, sol] < 0.95) & (tau[:, :, sol] > 1e-3))
    assert len(cand) > 0
    n, gg, ll = cand[len(cand) // 2]
    u0 = (1.0 / lam[n, gg, sol][ll]) * (1.0 + eps)
    pr.zenith_u = np.array([u0])
    pr.zenith_weights = np.array([1.0])
    g, o, rg, ro = run_both(tb, pr, cols)
    for k in ("fup_n_sol", "fdn_n_sol", "fup_a_sol", "fdn_a_sol", "amean_sol"):
        assert np.all(np.isfinite(rg[k])) == np.all(np.isfinite(ro[k])), k
    tol = max(TOL, 1.0e-14 / eps)
    for k in ("fup_n_sol", "fdn_n_sol", "fup_a_sol", "fdn_a_sol"):
        assert rel_err(rg[k], ro[k]) <= tol, (k, rel_err(rg[k], ro[k]), tol)


def test_rorr_kernel_choice_follows_the_columns():
    """A sweep of alike columns takes the staircase kernel with column chains; columns with randomised
    composition keep the sorting kernels; a tiny batch keeps them too (too few chains)."""
    tb = synth.make_tables(nw=20, n_sol=12, ir_start=8)
    pr = synth.default_params(tb, nzenith=1)
    g = Radtran(tb, pr)
    g.radiate_batch(synth.make_columns(64, nz=20, T_surf_range=(280.0, 290.0)))
    st = g.rorr_cut_stats()
    assert st["last_kernel"] == 1 and st["warp_merges"] > 0 and st["warp_passes"] < 1.5 * st["warp_merges"], st
    g.radiate_batch(synth.make_columns(64, nz=20, randomize=True, seed=3))
    assert g.rorr_cut_stats()["last_kernel"] == 2
    g.radiate_batch(synth.make_columns(4, nz=20, T_surf_range=(280.0, 281.0)))
    assert g.rorr_cut_stats()["last_kernel"] == 2


def test_nccl_entry_points_single_rank():
    """The multi-GPU entry points of the C ABI with a communicator of ONE rank (the only size a single
    GPU can host): clima_rt_comm_unique_id / _init, clima_rt_radiate_batch_sharded == radiate_batch
    padded to ncol_max, clima_rt_allgather_device, and the error strings without a communicator."""
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=5)
    g = Radtran(tb, pr)
    ref = g.radiate_batch(cols)
    with pytest.raises(ClimaException, match="no communicator"):
        g._nranks = 1
        g.radiate_batch_sharded(cols, 7)
    uid = Radtran.comm_unique_id()
    assert len(uid) == 128
    g.comm_init(1, 0, uid)
    with pytest.raises(ClimaException, match="already has a communicator"):
        g.comm_init(1, 0, uid)
    out = g.radiate_batch_sharded(cols, 7)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"):
        assert out[k].shape == (cols["nz"] + 1, 7, 1)
        assert np.array_equal(out[k][:, :5, 0], ref[k]), k
        assert np.all(out[k][:, 5:, 0] == 0.0)
    for k in ("ISR", "OLR"):
        assert np.array_equal(out[k][:5, 0], ref[k]), k
    with pytest.raises(ClimaException, match="ncol must be in"):
        g.radiate_batch_sharded(cols, 3)
    g.comm_destroy()
    g.comm_init(1, 0, Radtran.comm_unique_id())  # a handle can be given a new communicator
    out2 = g.radiate_batch_sharded(cols, 5)
    assert np.array_equal(out2["OLR"][:, 0], ref["OLR"])


def test_c_driver_runs_the_abi_without_python(tmp_path):
    """tests/c_driver.c: plain C against include/clima_rt.h, linked with libclima_rt.so -- the closest
    thing to the Fortran bind(c) caller this box can compile.  It reads a case file (tables, params,
    columns, the ORACLE's fluxes, a frozen-opacity Jacobian case from the oracle loop), calls
    clima_rt_create / radiate_batch / jacobian_batch / destroy and checks 1e-9 itself."""
    import os
    import shutil
    import subprocess
    from clima_b200 import blob
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    nzm = 6
    tb, pr, cols = small_case(G=16, nw=10, nz=nzm, ncol=3, nzen=2, doubled=True)
    o = OracleRadtran(tb, pr)
    gold = o.radiate_batch(cols, spectral=False)
    # Jacobian case on column 0 (the loop of clima_adiabat_solve.f90:798-820 run by the oracle)
    base = {k: (np.asfortranarray(v[..., :1]) if isinstance(v, np.ndarray) and v.shape[-1] == 3 else v) for k, v in cols.items()}
    base["ncol"] = 1
    o.radiate(base, 0, spectral=False)
    nzr = cols["nz"]
    Tm = base["T"][0:2 * nzm:2, 0].copy()
    Ts0 = float(base["T_surface"][0])
    ncj = nzm + 1
    T = np.repeat(Tm[:, None], ncj, axis=1).copy(order="F")
    Ts = np.full(ncj, Ts0)
    Ts[0] += 1.0e-2 * Ts0
    for i in range(nzm):
        T[i, i + 1] += 1.0e-2 * Tm[i]
    f = np.zeros((nzm + 1, ncj), order="F")
    for c in range(ncj):
        Tr = np.empty(nzr)
        Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
        pert = dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None]))
        ro = o.radiate(pert, 0, compute_solar=False, compute_opacity=False, spectral=False)
        f[:, c] = ro["f_total"][0:2 * nzm + 1:2, 0] if ro["f_total"].ndim == 2 else ro["f_total"][0:2 * nzm + 1:2]
    case = str(tmp_path / "case.crtb")
    blob.write_case(case, tb, pr, cols, gold, jac=dict(T_surface=Ts, T=T, f=f))
    exe = str(tmp_path / "c_driver")
    libdir = os.path.join(root, "clima_b200")
    r = subprocess.run([shutil.which("gcc"), "-std=c99", "-O1", "-I", os.path.join(root, "include"),
                        os.path.join(root, "tests", "c_driver.c"), "-o", exe, "-L", libdir, "-lclima_rt", "-lm",
                        "-Wl,-rpath," + libdir, "-Wl,--allow-shlib-undefined"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe, case], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "PASS" in r.stdout, r.stdout + r.stderr


def test_cfg4_jacobian_batch_at_nz50():
    """BASELINE.json configs[3] at its real size: nz = 50 model layers (nz_r = 102), Z = 1, albedo 0.3,
    51 perturbed columns (dT = 1e-2 T, `epsj`, clima_adiabat.f90:139), S-canon tables.  The batched entry
    against the reference loop run by the oracle, column by column."""
    nzm = 50
    tb = synth.make_tables()
    pr = synth.default_params(tb, nzenith=1, albedo=0.3)
    base = synth.make_columns(1, nz=nzm, doubled=True, T_surf_range=(288.0, 288.0))
    g = Radtran(tb, pr)
    o = OracleRadtran(tb, pr)
    g.radiate_batch(base)
    o.radiate(base, 0, spectral=False)
    nzr = base["nz"]
    assert nzr == 2 * nzm + 2
    Tm = base["T"][0:2 * nzm:2, 0].copy()
    Ts0 = float(base["T_surface"][0])
    ncol = nzm + 1
    T = np.repeat(Tm[:, None], ncol, axis=1).copy(order="F")
    Ts = np.full(ncol, Ts0)
    Ts[0] += 1.0e-2 * Ts0
    for i in range(nzm):
        T[i, i + 1] += 1.0e-2 * Tm[i]
    f = g.jacobian_batch(Ts, T)
    assert f.shape == (nzm + 1, ncol)
    scale = 0.0
    ref = np.zeros_like(f)
    for c in range(ncol):
        Tr = np.empty(nzr)
        Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
        pert = dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None]))
        ro = o.radiate(pert, 0, compute_solar=False, compute_opacity=False, spectral=False)
        ref[:, c] = ro["f_total"][0:2 * nzm + 1:2, 0]
        scale = max(scale, float(np.max(np.abs(ro["fup_n_ir"]))))
    assert np.max(np.abs(f - ref)) <= TOL * scale


@pytest.mark.parametrize("hard", [True, False])
def test_jacobian_delta_path_matches_the_full_solves(hard, monkeypatch):
    """clima_rt_jacobian_batch, delta path (jacobian_delta.cu: one factorisation of the two-stream matrix
    shared by all columns, per column a window solve + two chains) against the round-1 path that re-solves
    every column (CLIMA_RT_JAC=full), and against the oracle loop: surface, every layer, the top layer
    (4 radiative levels incl. the ghosts), an unchanged column, and a column that differs everywhere (too
    wide for the window: the call falls back to the full solves)."""
    nzm = 9
    tb, pr, base = small_case(G=16, nw=10, nz=nzm, ncol=1, nzen=1, doubled=True, has_hard_surface=hard)
    Tm = base["T"][0:2 * nzm:2, 0].copy()
    Ts0 = float(base["T_surface"][0])
    ncol = nzm + 2
    T = np.repeat(Tm[:, None], ncol, axis=1).copy(order="F")
    Ts = np.full(ncol, Ts0)
    Ts[0] += 1.0e-2 * Ts0
    for i in range(nzm):
        T[i, i + 1] += 1.0e-2 * Tm[i]
    # column ncol-1 stays equal to the base
    res = {}
    for mode in ("delta", "full"):
        if mode == "full":
            monkeypatch.setenv("CLIMA_RT_JAC", "full")
        g = Radtran(tb, pr)
        g.radiate_batch(base)
        res[mode] = g.jacobian_batch(Ts, T)
        st = g.rorr_cut_stats()
        assert (st["jacobian_delta_calls"] == 1) == (mode == "delta"), st
        if mode == "delta":
            fb = g.radiate_batch(base)["f_total"][0:2 * nzm + 1:2, 0]
            assert np.array_equal(res[mode][:, ncol - 1], fb)  # unchanged column: exactly the base
            # a column that differs in every level: falls back, same answer as the full path below
            Tall = np.asfortranarray((Tm * 1.01)[:, None]); Tsall = np.array([Ts0 * 1.01])
            f_all = g.jacobian_batch(Tsall, Tall)
            assert g.rorr_cut_stats()["jacobian_delta_calls"] == 1
            g.radiate_batch(base)
            res["wide"] = f_all
    scale = np.max(np.abs(res["full"]))
    assert np.max(np.abs(res["delta"] - res["full"])) <= 1e-12 * scale
    o = OracleRadtran(tb, pr)
    o.radiate(base, 0, spectral=False)
    nzr = base["nz"]
    for c in (0, 1, nzm, ncol - 1):
        Tr = np.empty(nzr)
        Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
        pert = dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None]))
        ro = o.radiate(pert, 0, compute_solar=False, compute_opacity=False, spectral=False)
        assert np.max(np.abs(res["delta"][:, c] - ro["f_total"][0:2 * nzm + 1:2, 0])) <= TOL * scale
    Tr = np.empty(nzr)
    Tr[0:2 * nzm:2] = Tm * 1.01; Tr[1:2 * nzm:2] = Tm * 1.01; Tr[2 * nzm:] = Tm[-1] * 1.01
    ro = o.radiate(dict(base, T_surface=np.array([Ts0 * 1.01]), T=np.asfortranarray(Tr[:, None])), 0,
                   compute_solar=False, compute_opacity=False, spectral=False)
    assert np.max(np.abs(res["wide"][:, 0] - ro["f_total"][0:2 * nzm + 1:2, 0])) <= TOL * scale


def test_jacobian_after_set_params_takes_the_new_parameters():
    """Changing the emissivity after the base call invalidates the Jacobian delta path's base state (its
    f_total was computed with the old emissivity): the next clima_rt_jacobian_batch re-solves every column
    with the new parameters, like the reference would (frozen opacity, stale solar, new `surface_emissivity`)."""
    import copy
    nzm = 6
    tb, pr, base = small_case(G=16, nw=10, nz=nzm, ncol=1, nzen=1, doubled=True)
    g = Radtran(tb, pr)
    g.radiate_batch(base)
    pr2 = copy.deepcopy(pr)
    pr2.surface_emissivity = np.full(tb.nw_ir, 0.8)
    g.params = pr2
    Tm = base["T"][0:2 * nzm:2, 0].copy()
    Ts = np.array([float(base["T_surface"][0]) * 1.01, float(base["T_surface"][0])])
    T = np.repeat(Tm[:, None], 2, axis=1).copy(order="F")
    T[2, 1] *= 1.01
    f = g.jacobian_batch(Ts, T)
    assert g.rorr_cut_stats()["jacobian_delta_calls"] == 0
    o = OracleRadtran(tb, pr2)
    o.radiate(base, 0, spectral=False)
    nzr = base["nz"]
    for c in range(2):
        Tr = np.empty(nzr)
        Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
        ro = o.radiate(dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None])), 0,
                       compute_solar=False, compute_opacity=False, spectral=False)
        ref = ro["f_total"][0:2 * nzm + 1:2, 0]
        assert np.max(np.abs(f[:, c] - ref)) <= TOL * np.max(np.abs(ro["fup_n_ir"]))
    # a new base call with the new parameters re-arms the delta path
    g.radiate_batch(base)
    f2 = g.jacobian_batch(Ts, T)
    assert g.rorr_cut_stats()["jacobian_delta_calls"] == 1
    assert np.max(np.abs(f2 - f)) <= 1e-12 * np.max(np.abs(f))


def test_staircase_kernel_backs_off_when_its_chains_do_not_carry():
    """Columns that are alike by their gas amounts (what similarity_kernel looks at) but have unrelated
    temperatures: the staircase kernel is chosen, reports > 1.5 passes per merge, and the following chunks
    take the sorting kernels; results match the oracle either way."""
    tb = synth.make_tables(nw=16, n_sol=10, ir_start=6)
    pr = synth.default_params(tb, nzenith=1)
    cols = synth.make_columns(48, nz=16, T_surf_range=(260.0, 260.0), doubled=False)
    rng = np.random.default_rng(9)
    T = np.asfortranarray(cols["T"] * rng.uniform(0.7, 2.5, (1, cols["T"].shape[1])))  # one factor per column
    cols = dict(cols, T=T)
    g = Radtran(tb, pr)
    r1 = g.radiate_batch(cols)
    st = g.rorr_cut_stats()
    assert st["last_kernel"] == 1 and st["warp_passes"] > 1.5 * st["warp_merges"], st
    r2 = g.radiate_batch(cols)
    assert g.rorr_cut_stats()["last_kernel"] == 2
    ro = OracleRadtran(tb, pr).radiate_batch(cols, nthreads=8)
    compare(r1, ro, spectral=False)
    compare(r2, ro, spectral=False)


def test_full_size_sweep_properties():
    """BASELINE config 2 at its full size (4096 columns, nz_r = 202, S-canon tables), where the oracle is too
    slow to be the checker: size-independent properties.  (a) every column finite, status 0, the staircase
    kernel with no unit handed to the exact kernel; (b) the same batch in REVERSED column order (every chain
    runs the other way, every staircase has another predecessor) gives the same per-column fluxes to 1e-11;
    (c) 48 columns picked across the sweep, radiated as their own batch, agree to 1e-11; (d) a second call is
    bit-identical; (e) 16 of the columns against the oracle."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    n = 4096
    cols = synth.make_columns(n, nz=100)
    g = Radtran(tb, pr)
    r = {k: v.copy() for k, v in g.radiate_batch(cols).items()}
    st = g.rorr_cut_stats()
    assert st["last_kernel"] == 1 and st["units_to_exact"] == 0 and g.rorr_stats()[2] == 0, st
    assert st["warp_passes"] < 1.2 * st["warp_merges"], st
    assert np.all(r["status"] == 0)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR"):
        assert np.all(np.isfinite(r[k])), k
    r2 = g.radiate_batch(cols)
    for k in r:
        assert np.array_equal(r[k], r2[k]), k  # (d)
    rev = {k: (np.asfortranarray(v[..., ::-1]) if isinstance(v, np.ndarray) and v.shape[-1] == n else v) for k, v in cols.items()}
    rr = g.radiate_batch(rev)
    scale = max(np.max(np.abs(r[k])) for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol"))
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"):
        assert np.max(np.abs(rr[k][:, ::-1] - r[k])) <= 1e-11 * scale, k  # (b)
    assert np.max(np.abs(rr["OLR"][::-1] / r["OLR"] - 1.0)) <= 1e-11
    pick = np.linspace(0, n - 1, 48).astype(int)
    sub = {k: (np.asfortranarray(v[..., pick]) if isinstance(v, np.ndarray) and v.shape[-1] == n else v) for k, v in cols.items()}
    sub["ncol"] = 48
    rs = g.radiate_batch(sub)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"):
        assert np.max(np.abs(rs[k] - r[k][:, pick])) <= 1e-11 * scale, k  # (c)
    ro = OracleRadtran(tb, pr, kind="parity").radiate_batch(
        {k: (np.asfortranarray(v[..., :16]) if isinstance(v, np.ndarray) and v.shape[-1] == 48 else v) for k, v in sub.items()} | {"ncol": 16},
        nthreads=8)
    compare({k: (v[..., :16] if v.ndim > 1 else v[:16]) for k, v in rs.items()}, ro, spectral=False)  # (e)


def test_full_size_randomised_columns_chunking(monkeypatch):
    """BASELINE config 5 at the per-GPU size of the 8-GPU run (8192 randomised columns = two chunks of 4096;
    sorting kernels): every column finite, and the same batch cut into chunks of 2048 is bit-identical
    (columns are independent units: a result must not depend on which columns share a chunk)."""
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(8192, nz=100, randomize=True, seed=11)
    g = Radtran(tb, pr)
    r1 = {k: v.copy() for k, v in g.radiate_batch(cols).items()}
    assert g.rorr_cut_stats()["last_kernel"] == 2
    assert np.all(r1["status"] == 0) and all(np.all(np.isfinite(r1[k])) for k in ("fup_n_ir", "fdn_n_ir", "f_total", "OLR"))
    monkeypatch.setenv("CLIMA_RT_CHUNK_COLS", "2048")
    r2 = Radtran(tb, pr).radiate_batch(cols)
    for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR"):
        assert np.array_equal(r1[k], r2[k]), k


def test_adiabat_climate_batch_entry_matches_the_per_column_loop():
    """N3 at the Python level: AdiabatClimateBatch.TOA_fluxes_batch (make_profile per column on the host, ONE
    doubled-grid batch on the GPU) against the loop the reference's users write (TOA_fluxes once per column,
    examples/EarlyMars.ipynb cell 7), here with the oracle doing each column; the column builder is an
    analytic stand-in for `make_profile` (row N4 is not built)."""
    from clima_b200.adiabat import AdiabatClimateBatch, copy_atm_to_radiative_grid
    tb, pr, _ = small_case(G=16, nw=10, nz=6, ncol=1, nzen=2)
    nzm, ng = 6, tb.ng
    kB = 1.380649e-16

    def make_profile(T_surf, P_i_surf):  # dry adiabat down to 200 K on log-spaced pressure layers
        Ps = float(np.sum(P_i_surf))
        edges = Ps * np.logspace(0.0, -4.0, nzm + 1)
        P = np.sqrt(edges[:-1] * edges[1:])
        T = np.maximum(200.0, T_surf * (P / Ps) ** 0.2)
        n = P / (kB * T)
        dens = n[:, None] * (P_i_surf / Ps)[None, :]
        dz = 8.0e5 * np.log(edges[:-1] / edges[1:]) * T / 250.0
        return dict(T=T, P=P, densities=dens, dz=dz)

    # the doubled grid: layers twice, dz halved, two ghost copies of the last radiative layer (:729-773)
    d = copy_atm_to_radiative_grid(np.arange(1.0, 4.0), np.arange(1.0, 4.0), np.ones((3, 2)), np.full(3, 2.0))
    assert list(d["T"]) == [1, 1, 2, 2, 3, 3, 3, 3] and list(d["dz"]) == [1, 1, 1, 1, 1, 1, 1, 1]
    rng = np.random.default_rng(2)
    ncol = 5
    T_surf = np.linspace(260.0, 320.0, ncol)
    P_i = np.asfortranarray(1.0e6 * np.abs(rng.uniform(1e-4, 0.6, (ng, ncol))))
    g = Radtran(tb, pr)
    acb = AdiabatClimateBatch(g, make_profile)
    isr, olr = acb.TOA_fluxes_batch(T_surf, P_i)
    o = OracleRadtran(tb, pr)
    cols = acb.columns(T_surf, P_i)
    assert cols["nz"] == 2 * nzm + 2
    for c in range(ncol):
        ro = o.radiate(cols, c, spectral=False)
        assert abs(isr[c] / ro["ISR"][0] - 1.0) < TOL and abs(olr[c] / ro["OLR"][0] - 1.0) < TOL
    i1, o1 = acb.TOA_fluxes(float(T_surf[2]), P_i[:, 2])
    assert abs(i1 / isr[2] - 1.0) < 1e-12 and abs(o1 / olr[2] - 1.0) < 1e-12
    with pytest.raises(ClimaException, match="inconsistent"):
        acb.TOA_fluxes_batch(T_surf[:3], P_i)
