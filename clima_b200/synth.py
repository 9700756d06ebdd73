"""Synthetic opacity tables and atmospheric columns ("S-canon").

The real data (`photochem_clima_data` v0.3.1, fetched by the reference's CMake at configure time,
CMakeLists.txt:5) is not available offline, so every measurement uses synthetic tables that follow
the reference's HDF5 schema semantics (src/radtran/clima_radtran_types_create.f90:1265-1378 for
k-tables, :1090-1263 CIA, :1048-1088 Rayleigh, :868-1046 continuum, :734-866 particles).
All randomness: numpy default_rng(seed), default seed 20261104.
"""
from __future__ import annotations

import numpy as np

from ._abi import KTable, OpacityTables, Particle, RadtranParams, XS1D, zenith_angles_and_weights

SEED = 20261104
K_BOLTZ = 1.380649e-16   # erg/K, src/clima_const.f90:10
C_LIGHT = 299792458.0
PLANK = 6.62607004e-34
K_BOLTZ_SI = 1.380649e-23
N_AVO = 6.02214076e23

SPECIES = ["H2O", "CO2", "N2", "H2", "CH4", "CO", "O2"]
MU = {"H2O": 18.015, "CO2": 44.01, "N2": 28.013, "H2": 2.016, "CH4": 16.043, "CO": 28.01, "O2": 31.999}


def gauss_weights(G: int) -> np.ndarray:
    """k-table `weights`: Gauss-Legendre on [0,1]."""
    _, w = np.polynomial.legendre.leggauss(G)
    return w / 2.0


def rayleigh_vardavas(A, B, Delta, lam_nm):
    """src/clima_eqns.f90:240-246"""
    return 4.577e-21 * ((6.0 + 3.0 * Delta) / (6.0 - 7.0 * Delta)) * \
        (A * (1.0 + B / (lam_nm * 1.0e-3) ** 2.0)) ** 2.0 * (1.0 / (lam_nm * 1.0e-3) ** 4.0)


def planck_nu(nu, T):
    """src/clima_eqns.f90:64-73 (mW sr^-1 m^-2 Hz^-1)"""
    return 1.0e3 * ((2.0 * PLANK * nu ** 3.0) / (C_LIGHT ** 2.0)) * (1.0 / (np.exp((PLANK * nu) / (K_BOLTZ_SI * T)) - 1.0))


def make_tables(nw=100, G=16, nk=5, ncia=10, nray=5, npxs=3, cont=True, npart=0,
                npress=20, ntemp=16, n_sol=64, ir_start=37, seed=SEED,
                species=None) -> OpacityTables:
    """S-canon tables.  nw bins log-spaced 0.2-500 um; solar channel = bins 1..n_sol, IR channel =
    bins ir_start..nw (overlap exercises the shared-opacity logic, types_create.f90:250-268)."""
    rng = np.random.default_rng(seed)
    species = list(species or SPECIES)
    ng = len(species)
    wavl = np.logspace(np.log10(200.0), np.log10(5.0e5), nw + 1)  # nm
    lam_c = np.sqrt(wavl[:-1] * wavl[1:])
    x = np.log10(lam_c)
    weights = gauss_weights(G)
    gq = np.cumsum(weights) - 0.5 * weights  # g at bin centres

    k_species = [n for n in ["H2O", "CO2", "CH4", "CO", "O2", "O3"] if n in species][:nk]
    ktabs = []
    for i, name in enumerate(k_species):
        # per-gas grids differ a little so the per-table clamps are exercised
        log10P = np.linspace(-7.0 + 0.1 * i, 2.5 - 0.05 * i, npress)
        temp = np.linspace(50.0 + 5.0 * i, 2000.0 - 20.0 * i, ntemp)
        # smooth band structure in wavelength
        nb = 6
        centres = rng.uniform(x.min(), x.max(), nb)
        widths = rng.uniform(0.08, 0.35, nb)
        amp_max = {"H2O": 7.0, "CO2": 7.5, "CH4": 8.0, "CO": 8.0, "O2": 3.5, "O3": 7.0}[name]
        amps = rng.uniform(2.0, amp_max, nb)
        band = np.full(nw, -28.5 + rng.uniform(-0.5, 0.5))
        for c, wd, a in zip(centres, widths, amps):
            band = band + a * np.exp(-0.5 * ((x - c) / wd) ** 2)
        span = rng.uniform(4.0, 6.0, nw)              # decades between g=0 and g=1
        shape = gq[:, None] ** rng.uniform(1.5, 3.0, nw)[None, :]  # monotone increasing in g
        lk = np.empty((G, npress, ntemp, nw), order="F")
        pfac = 0.35 * (log10P[None, :, None, None] - 0.0)            # pressure broadening
        tfac = 0.6 * np.log10(temp[None, None, :, None] / 300.0) * (1.0 - gq[:, None, None, None])
        lk[...] = band[None, None, None, :] - 0.5 * span[None, None, None, :] \
            + span[None, None, None, :] * shape[:, None, None, :] \
            + pfac * (1.0 - gq[:, None, None, None]) + tfac
        lk += 0.02 * rng.standard_normal((1, npress, ntemp, nw))  # keeps monotonicity in g
        lk = np.clip(lk, -32.0, -16.5)
        lk = np.sort(lk, axis=0)  # k-distributions are ascending in g by construction
        ktabs.append(KTable(sp_ind=species.index(name) + 1, log10P=log10P, temp=temp,
                            log10k=np.asfortranarray(lk)))

    # CIA pairs (no H2O pairs when the water continuum is on, types_create.f90:431-432)
    pairs_all = [("CO2", "CO2"), ("N2", "N2"), ("H2", "H2"), ("CO2", "H2"), ("CO2", "CH4"),
                 ("N2", "H2"), ("CH4", "CH4"), ("N2", "CH4"), ("O2", "O2"), ("N2", "O2"),
                 ("H2", "CH4"), ("O2", "CO2")]
    cia = []
    for c in range(ncia):
        a, b = pairs_all[c % len(pairs_all)]
        if a not in species or b not in species:
            continue
        if c % 5 == 4:  # a few temperature-independent (dim 0) tables
            xs0 = 10.0 ** rng.uniform(-50.0, -44.0, nw)
            cia.append(XS1D([species.index(a) + 1, species.index(b) + 1], 0, None, xs0))
        else:
            nt = 10
            temp = np.linspace(100.0 + 10 * c, 900.0 + 40 * c, nt)
            base = -46.5 + 3.0 * np.sin(2.5 * x + c) + rng.uniform(-0.3, 0.3, nw)
            data = base[None, :] + 0.8 * np.log10(temp[:, None] / 300.0) * np.cos(x + 0.3 * c)[None, :]
            data = np.clip(data, -50.0, -43.0)
            cia.append(XS1D([species.index(a) + 1, species.index(b) + 1], 1, temp, np.asfortranarray(data)))

    # Rayleigh (left bin edge, types_create.f90:1084-1086)
    ray_names = ["N2", "CO2", "O2", "H2O", "H2"][:nray]
    ray_names = [n for n in ray_names if n in species]
    ray_par = {"N2": (29.06e-5, 7.7e-3, 0.0305), "CO2": (43.9e-5, 6.4e-3, 0.0805),
               "O2": (26.63e-5, 5.07e-3, 0.054), "H2O": (26.0e-5, 6.0e-3, 0.17),
               "H2": (13.58e-5, 7.52e-3, 0.0221)}
    ray_xs = np.zeros((nw, len(ray_names)), order="F")
    for i, n in enumerate(ray_names):
        A, B, D = ray_par[n]
        ray_xs[:, i] = rayleigh_vardavas(A, B, D, wavl[:-1])

    # photolysis cross sections (0-D, UV only)
    pxs = []
    for n in ["O2", "CO2", "H2O"][:npxs]:
        if n not in species:
            continue
        xs0 = np.where(lam_c < 320.0, 10.0 ** (-19.0 - 6.0 * (lam_c - 200.0) / 120.0), 0.0)
        xs0 = xs0 * rng.uniform(0.5, 2.0)
        pxs.append(XS1D([species.index(n) + 1], 0, None, xs0))

    kw = {}
    if cont and "H2O" in species and ng > 1:
        nt = 12
        ctemp = np.linspace(200.0, 750.0, nt)
        base = -42.5 - 1.2 * (x - 3.0) ** 2 / 4.0 + 0.8 * np.sin(3.0 * x)
        h2o = base[None, :] - 1.5 * np.log10(ctemp[:, None] / 296.0)
        foreign = base[None, :] - 2.2 - 0.4 * np.log10(ctemp[:, None] / 296.0)
        kw.update(cont_LH2O=species.index("H2O") + 1, cont_temp=ctemp,
                  cont_log10_H2O=np.asfortranarray(h2o), cont_log10_foreign=np.asfortranarray(foreign))

    particle_names = [f"aer{i + 1}" for i in range(npart)]
    part = []
    for i in range(npart):
        nr = 12
        radii = np.logspace(-6.0, -3.0, nr)  # cm
        size = 2.0 * np.pi * radii[:, None] / (lam_c[None, :] * 1.0e-7)
        qext = 2.0 * size ** 2 / (1.0 + size ** 2) + 0.05 * i
        w0 = np.clip(0.95 * size / (0.3 + size) - 0.1 * i, 0.0, 0.999)
        gt = np.clip(0.85 * size / (1.0 + size), 0.0, 0.95)
        part.append(Particle(i + 1, radii, np.asfortranarray(w0), np.asfortranarray(qext),
                             np.asfortranarray(gt)))

    sol_ind = (1, n_sol)
    ir_ind = (ir_start, nw)
    freq = C_LIGHT / (wavl * 1.0e-9)
    fs = 0.5 * (freq[sol_ind[0] - 1:sol_ind[1]] + freq[sol_ind[0]:sol_ind[1] + 1])
    photons_sol = np.pi * planck_nu(fs, 5772.0) * (6.957e8 / 1.496e11) ** 2  # mW/m^2/Hz at 1 AU

    return OpacityTables(wavl=wavl, species_names=species, particle_names=particle_names,
                         weights=weights, k=ktabs, cia=cia, ray_sp_ind=[species.index(n) + 1 for n in ray_names],
                         ray_xs=ray_xs, pxs=pxs, part=part, ir_ind=ir_ind, sol_ind=sol_ind,
                         photons_sol=photons_sol, **kw)


def default_params(tables: OpacityTables, nzenith=4, albedo=0.25, emissivity=1.0, **kw) -> RadtranParams:
    u, w = zenith_angles_and_weights(nzenith)
    return RadtranParams(zenith_u=u, zenith_weights=w,
                         surface_albedo=np.full(tables.nw_sol, albedo),
                         surface_emissivity=np.full(tables.nw_ir, emissivity), **kw)


# ------------------------------------------------------------------------------------------------
# columns
# ------------------------------------------------------------------------------------------------
def double_grid(T, P, dens, dz, ghosts=True):
    """Layer doubling + 2 ghost layers exactly like AdiabatClimate_copy_atm_to_radiative_grid
    (src/adiabat/clima_adiabat.f90:729-773).  Arrays are (nz, ncol) / (nz, ng, ncol), F-order."""
    nz = T.shape[0]
    nzr = 2 * nz + (2 if ghosts else 0)

    def rep(a, half=False):
        out = np.empty((nzr,) + a.shape[1:], order="F")
        out[0:2 * nz:2] = a * (0.5 if half else 1.0)
        out[1:2 * nz:2] = a * (0.5 if half else 1.0)
        if ghosts:
            out[2 * nz] = out[2 * nz - 1]
            out[2 * nz + 1] = out[2 * nz - 1]
        return out

    return rep(T), rep(P), rep(dens), rep(dz, half=True)


def make_columns(ncol, nz=100, species=None, seed=SEED, T_surf_range=(200.0, 400.0), T_trop=215.0,
                 P_surf=1.0, P_top=1.0e-5, doubled=True, randomize=False, pco2_range=None,
                 gravity=981.0, npart=0):
    """Analytic columns: dry-ish adiabat T = T_surf (P/P_surf)^0.19 down to an isothermal
    stratosphere at T_trop, hydrostatic layer thickness, ideal-gas number densities.
    Returns dict with T_surface(ncol), T(nz_r,ncol), P [bar], densities(nz_r,ng,ncol), dz [cm]
    (and pdensities/radii if npart>0), all Fortran-ordered, ground-first."""
    rng = np.random.default_rng(seed)
    species = list(species or SPECIES)
    ng = len(species)
    if ncol == 1:
        T_surf = np.array([0.5 * (T_surf_range[0] + T_surf_range[1])])
    else:
        T_surf = np.linspace(T_surf_range[0], T_surf_range[1], ncol)
    Ps = np.full(ncol, P_surf)
    # base dry mixing ratios
    f = {"CO2": 4.0e-4, "N2": 0.78, "H2": 1.0e-6, "CH4": 1.8e-6, "CO": 1.0e-7, "O2": 0.21}
    fmat = np.zeros((ng, ncol))
    for i, s in enumerate(species):
        if s in f:
            fmat[i] = f[s]
    if pco2_range is not None:
        pco2 = np.logspace(np.log10(pco2_range[0]), np.log10(pco2_range[1]), ncol)
        rng.shuffle(pco2)
        Ps = Ps + pco2
        fmat[species.index("CO2")] = pco2 / Ps
    if randomize:
        for i, s in enumerate(species):
            if s in ("N2", "H2O"):
                continue
            fmat[i] = 10.0 ** rng.uniform(-7.0, -1.0, ncol)
    edges = np.logspace(0.0, np.log10(P_top / P_surf), nz + 1)  # P/P_surf at layer edges
    pc = np.sqrt(edges[:-1] * edges[1:])
    P = np.asfortranarray(pc[:, None] * Ps[None, :])  # bar, layer centres
    T = np.maximum(T_trop, T_surf[None, :] * pc[:, None] ** 0.19)
    T = np.asfortranarray(T)
    # water: saturation-limited mixing ratio with relative humidity 0.5, cold-trapped aloft
    psat = 6.1e-3 * np.exp(5423.0 * (1.0 / 273.15 - 1.0 / T))  # bar
    fh2o = np.minimum(0.5 * psat / P, 0.5)
    fh2o = np.minimum.accumulate(fh2o, axis=0)
    dens = np.zeros((nz, ng, ncol), order="F")
    ntot = (P * 1.0e6) / (K_BOLTZ * T)
    iw = species.index("H2O") if "H2O" in species else -1
    dry = 1.0 - (fh2o if iw >= 0 else 0.0)
    fsum = fmat.sum(axis=0) - (fmat[iw] if iw >= 0 else 0.0)
    for i in range(ng):
        if i == iw:
            dens[:, i, :] = fh2o * ntot
        else:
            dens[:, i, :] = (fmat[i] / fsum)[None, :] * dry * ntot
    mubar = np.zeros((nz, ncol))
    for i, s in enumerate(species):
        mubar += MU[s] * dens[:, i, :] / ntot
    H = (N_AVO * K_BOLTZ * T) / (mubar * gravity)  # cm
    dlnp = np.log(edges[:-1] / edges[1:])
    dz = np.asfortranarray(H * dlnp[:, None])
    out = {}
    if doubled:
        T, P, dens, dz = double_grid(T, P, dens, dz)
    nzr = T.shape[0]
    out.update(T_surface=np.ascontiguousarray(T_surf), T=np.asfortranarray(T), P=np.asfortranarray(P),
               densities=np.asfortranarray(dens), dz=np.asfortranarray(dz), nz=nzr, ng=ng, np=npart,
               ncol=ncol)
    if npart > 0:
        pd = np.empty((nzr, npart, ncol), order="F")
        rd = np.empty((nzr, npart, ncol), order="F")
        for i in range(npart):
            pd[:, i, :] = 50.0 * (out["P"] / out["P"][0]) ** 0.5
            rd[:, i, :] = 1.0e-5 * (1.0 + i) * (1.0 + 0.5 * (out["P"] / out["P"][0]))
        out.update(pdensities=pd, radii=rd)
    return out
