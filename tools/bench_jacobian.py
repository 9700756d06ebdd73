#!/usr/bin/env python
"""cfg4 of BASELINE.json: the frozen-opacity finite-difference Jacobian batch of the RCE solver
(src/adiabat/clima_adiabat_solve.f90:798-820) through clima_rt_jacobian_batch.
Base column nz=50 (nz_r=102), Z=1, S-canon tables; 51 perturbed temperature profiles per Jacobian;
timed as ONE Jacobian (51 columns) and as `--stack` stacked Jacobians in one call.
Prints one JSON line; the CPU figure is the oracle port running the same frozen-opacity columns
serially (the reference's own loop structure)."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from clima_b200 import synth
from clima_b200.radtran import Radtran


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stack", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    nzm = 50
    tb = synth.make_tables()
    pr = synth.default_params(tb, nzenith=1, albedo=0.3)
    base = synth.make_columns(1, nz=nzm, doubled=True)
    g = Radtran(tb, pr)
    r0 = g.radiate_batch(base)
    Tm = base["T"][0:2 * nzm:2, 0].copy(); Ts0 = float(base["T_surface"][0])
    def batch(nj):
        ncol = 51 * nj
        T = np.repeat(Tm[:, None], ncol, axis=1).copy(order="F"); Ts = np.full(ncol, Ts0)
        for j in range(nj):
            Ts[51 * j] += 1e-2 * Ts0 * (1 + 1e-3 * j)
            for i in range(nzm):
                T[i, 51 * j + i + 1] += 1e-2 * Tm[i] * (1 + 1e-3 * j)
        return Ts, T
    out = {}
    for name, nj in (("one_jacobian", 1), ("stacked", a.stack)):
        Ts, T = batch(nj)
        g.jacobian_batch(Ts, T)
        t0 = time.perf_counter()
        for _ in range(a.reps):
            f = g.jacobian_batch(Ts, T)
        dt = (time.perf_counter() - t0) / a.reps
        out[name] = {"jacobians": nj, "columns": 51 * nj, "ms_per_call": 1e3 * dt,
                     "jacobians_per_s": nj / dt, "columns_per_s": 51 * nj / dt,
                     "kernel_ms": g.stage_ms()}
    if not a.no_cpu:
        from oracle_lib import OracleRadtran
        o = OracleRadtran(tb, pr, kind="fast")
        o.radiate(base, 0, spectral=False)
        Ts, T = batch(1)
        nzr = base["nz"]
        t0 = time.perf_counter()
        for c in range(51):
            Tr = np.empty(nzr); Tr[0:2 * nzm:2] = T[:, c]; Tr[1:2 * nzm:2] = T[:, c]; Tr[2 * nzm:] = T[nzm - 1, c]
            pert = dict(base, T_surface=np.array([Ts[c]]), T=np.asfortranarray(Tr[:, None]))
            ro = o.radiate(pert, 0, compute_solar=False, compute_opacity=False, spectral=False)
        dt = time.perf_counter() - t0
        ref = ro["f_total"][0:2 * nzm + 1:2, 0]
        out["cpu_port"] = {"jacobians_per_s": 1.0 / dt, "columns_per_s": 51 / dt, "cores": 1,
                           "kind": "port", "max_rel_err_last_column": float(np.max(np.abs(f[:, 50] - ref)) / np.max(np.abs(ref))) if a.stack >= 1 else None}
    print(json.dumps({"workload": "cfg4: frozen-opacity IR-only Jacobian batch, nz=50 (nz_r=102), Z=1, 51 columns per Jacobian", **out}))


if __name__ == "__main__":
    main()
