"""N4 measurement: batched `make_profile` and the batched `AdiabatClimate%TOA_fluxes` on the device, with the
numpy oracle (the reference's algorithm, one column at a time on one host core) timed beside them on a sample.

    python tools/bench_adiabat.py [--ncol 4096] [--nz 50] [--reps 5] [--cpu-cols 8]

Prints one JSON line.  Workload: cfg2's radiative tables (100 bins, 16 g-points, 7 gases) with the model grid of
BASELINE.json (nz = 50 -> nz_r = 102), surface states spread around the cases of the reference's
tests/test_adiabat.f90.  Timed with CUDA events on the handle's stream (make_profile kernel) and with the host
clock around the public call (host inputs, host results)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ncol", type=int, default=4096)
    ap.add_argument("--nz", type=int, default=50)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cpu-cols", type=int, default=8)
    ap.add_argument("--solve", type=int, default=0, help="also solve ISR = OLR for this many compositions")
    ap.add_argument("--photon-scale", type=float, default=1.0)
    a = ap.parse_args()
    import torch
    import test_adiabat as TA
    from clima_b200 import synth
    from clima_b200.adiabat import AdiabatClimate
    from clima_b200.radtran import Radtran
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    rad = Radtran(tb, pr, device=0)
    ac = AdiabatClimate(rad, TA.species(), a.nz, TA.EARTH_MASS, TA.EARTH_RADIUS)
    T, P = TA.sweep(a.ncol, seed=11)
    ac.make_profile_batch(T[:64], P[:, :64])
    ac.TOA_fluxes_batch(T[:64], P[:, :64])
    t_prof, t_toa, t_kern = [], [], []
    for _ in range(a.reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        p = ac.make_profile_batch(T, P)
        t_prof.append(time.perf_counter() - t0)
        t_kern.append(ac.kernel_ms())
        t0 = time.perf_counter()
        isr, olr = ac.TOA_fluxes_batch(T, P)
        t_toa.append(time.perf_counter() - t0)
    # the oracle: one column at a time, one core
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    t0 = time.perf_counter()
    for c in range(a.cpu_cols):
        TA.oracle_column(ac, T[c], P[:, c])
    t_cpu = (time.perf_counter() - t0) / a.cpu_cols
    # the climate solve a user runs on top (AdiabatClimate%surface_temperature): `--solve N` compositions at once
    solve = None
    if a.solve > 0:
        import dataclasses
        rad.params = dataclasses.replace(rad.params, photon_scale_factor=a.photon_scale)
        Ps = np.asfortranarray(np.repeat(TA.P_I_OCEAN[:, None], a.solve, 1))
        Ps[1] = np.logspace(-4, 0, a.solve) * 1.0e6
        try:
            n0 = rad.launch_count()
            t0 = time.perf_counter()
            Ts, isr_s, olr_s = ac.surface_temperature_batch(Ps, T_guess=300.0)
            dt = time.perf_counter() - t0
            solve = dict(compositions=a.solve, seconds=dt, climates_per_s=a.solve / dt, T_surf_range=[float(Ts.min()), float(Ts.max())],
                         max_rel_imbalance=float(np.max(np.abs(isr_s / olr_s - 1.0))), kernel_launches=int(rad.launch_count() - n0),
                         photon_scale_factor=a.photon_scale)
        except Exception as e:
            solve = dict(error=str(e)[:200])
    line = dict(metric="profiles_per_second", ncol=a.ncol, nz=a.nz, ng=7,
                make_profile_kernel_ms=min(t_kern), make_profile_batch_ms=1e3 * min(t_prof), profiles_per_s=a.ncol / min(t_prof),
                toa_fluxes_batch_ms=1e3 * min(t_toa), toa_columns_per_s=a.ncol / min(t_toa),
                rhs_evaluations_per_column=float(np.mean(p["nfcn"])), rhs_evaluations_max=int(np.max(p["nfcn"])),
                with_tropopause=int((p["P_trop"] > 0).sum()),
                cpu_oracle_ms_per_column=1e3 * t_cpu, cpu_kind="port (numpy, 1 core)", cpu_sample_columns=a.cpu_cols,
                OLR_range=[float(olr.min()), float(olr.max())], surface_temperature_solve=solve)
    print(json.dumps(line))


if __name__ == "__main__":
    main()
