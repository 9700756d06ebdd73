"""Shared parity cases: small synthetic tables/columns that the oracle finishes in seconds."""
import numpy as np

from clima_b200 import synth

TOL = 1.0e-9  # north_star: 1e-9 relative on layer fluxes and OLR (FP64)


def small_case(G=16, nw=12, nz=10, ncol=3, nk=3, npart=0, nzen=4, seed=1, doubled=True, **kw):
    n_sol = max(2, (2 * nw) // 3)
    ir_start = max(1, nw // 3)
    tb = synth.make_tables(nw=nw, G=G, nk=nk, ncia=6, nray=3, npxs=2, cont=True, npart=npart,
                           npress=8, ntemp=6, n_sol=n_sol, ir_start=ir_start, seed=seed)
    pr = synth.default_params(tb, nzenith=nzen, **kw)
    cols = synth.make_columns(ncol, nz=nz, seed=seed, doubled=doubled, npart=npart)
    return tb, pr, cols


def rel_err(a, b):
    """max |a-b| / max|b| (flux arrays span many decades; the scale is the array maximum)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    s = np.max(np.abs(b))
    if s == 0.0:
        return float(np.max(np.abs(a)))
    return float(np.max(np.abs(a - b)) / s)


def elem_rel_err(a, b, floor=1e-12):
    """max elementwise relative error over elements with |b| > floor*max|b| (the floor only keeps
    0/0 out: fluxes 12 decades below the column maximum -- deep fdn_n_sol under an opaque
    atmosphere -- are below one ulp of every sum they enter)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    s = np.max(np.abs(b))
    m = np.abs(b) > floor * s
    if not m.any():
        return 0.0
    return float(np.max(np.abs(a[m] - b[m]) / np.abs(b[m])))


FLUX_KEYS = ["fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol"]
SPEC_KEYS = ["fup_a_ir", "fdn_a_ir", "tau_band_ir", "fup_a_sol", "fdn_a_sol", "amean_sol", "tau_band_sol"]


def compare(r_gpu, r_ref, spectral=True, tol=TOL, verbose=False, elem_tol=None):
    worst = {}
    for k in FLUX_KEYS + (SPEC_KEYS if spectral else []):
        worst[k] = rel_err(r_gpu[k], r_ref[k])
    for k in FLUX_KEYS:
        worst[k + "_elem"] = elem_rel_err(r_gpu[k], r_ref[k])
    for k in ("ISR", "OLR"):
        worst[k] = float(np.max(np.abs(r_gpu[k] - r_ref[k]) / np.abs(r_ref[k])))
    # f_total is a difference of four large numbers: absolute tolerance vs max |flux|
    scale = max(np.max(np.abs(r_ref[k])) for k in FLUX_KEYS)
    worst["f_total"] = float(np.max(np.abs(r_gpu["f_total"] - r_ref["f_total"])) / scale)
    if verbose:
        for k, v in worst.items():
            print(f"  {k:16s} {v:.3e}")
    etol = tol if elem_tol is None else elem_tol
    bad = {k: v for k, v in worst.items() if not (v <= (etol if k.endswith('_elem') else tol))}
    assert not bad, f"parity failure (tol {tol}): {bad}"
    return worst


def cfg1_template_case():
    """BASELINE.json configs[0]: the ModernEarth template column of the reference's own test
    (/root/reference/tests/test_radtran.f90:33-65): atmosphere.txt on vertical_grid(0, 1e7, 200),
    optical-properties species of templates/ModernEarth/settings.yaml, 8 zenith angles, surface albedo
    0.15, the Sun_now.txt spectrum.  The k-tables / CIA / continuum are S-canon SYNTHETIC tables (the
    photochem_clima_data files are not on this box); the template data come from the fixture written by
    tests/golden/make_cfg1_template.py."""
    import os
    from clima_b200 import loader
    fx = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cfg1_modern_earth.npz"))
    species = ["H2O", "CO2", "O2", "N2", "O3", "CH4"]  # settings.yaml optical-properties: species: gases
    tb = synth.make_tables(species=species)
    assert np.array_equal(tb.wavl, fx["wavl"])
    tb.photons_sol = np.ascontiguousarray(fx["stellar_flux_on_scanon_grid"][tb.sol_ind[0] - 1:tb.sol_ind[1]])
    pr = synth.default_params(tb, nzenith=8, albedo=0.15)
    cols = loader.column_from_atmosphere([str(x) for x in fx["labels"]], fx["columns"], species, nz=200)
    return tb, pr, cols


def rayleigh_dominated_case(nz=12, ncol=1, seed=3):
    """Solar-resonance probe (clima_radtran_twostream.f90:80, `denom = lambda**2 - 1/u0**2`, no guard):
    small tables in which UV bins are Rayleigh dominated (w0 up to max_w0, g = 0), one zenith angle."""
    tb, pr, cols = small_case(G=16, nw=10, nz=nz, ncol=ncol, nk=3, nzen=1, seed=seed, doubled=False)
    return tb, pr, cols
