#!/usr/bin/env python
"""GPU check of the sort-free (staircase) RORR kernel against the sorting kernels on the same columns:
max relative difference of tau / w0 / tau_band / fluxes, units sent to the exact kernel, and the stage
time of both, for cfg2 / cfg3 / cfg5 columns at full size (nz=100 doubled, S-canon tables).
usage (GPU box): tools/cut_check.py [ncol] [cfgs...]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
cfgs = sys.argv[2:] or ["cfg2", "cfg3", "cfg5"]
KW = dict(cfg2={}, cfg3=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0),
          cfg5=dict(randomize=True, seed=11))
tb = synth.make_tables()
pr = synth.default_params(tb)
FL = ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol")


def run(cols, env):
    for k, v in env.items():
        os.environ[k] = v
    g = Radtran(tb, pr)
    for k in env:
        del os.environ[k]
    g.radiate_batch(cols)
    s0 = g.stage_ms()
    t0 = time.perf_counter(); r = g.radiate_batch(cols); dt = time.perf_counter() - t0
    s1 = g.stage_ms()
    units, exact, ovf = g.rorr_stats()
    op = [g.get_opacity(cols["nz"], icol=i) for i in (0, ncol // 3, ncol - 1)]
    return dict(r=r, op=op, rorr_ms=s1[0], total_ms=s1[4], wall_ms=dt * 1e3,
                exact_share=exact / max(units, 1), ovf=int(ovf), cut=g.rorr_cut_stats())


out = {}
for name in cfgs:
    cols = synth.make_columns(ncol, nz=100, **KW[name])
    ref = run(cols, {"CLIMA_RT_RORR": "sort"})
    res = {"sort": dict(rorr_ms=ref["rorr_ms"], total_ms=ref["total_ms"], exact_share=ref["exact_share"])}
    for tag, env in (("cut_auto", {}), ("cut_slot", {"CLIMA_RT_RORR_CHAIN": "slot"}), ("cut_col", {"CLIMA_RT_RORR_CHAIN": "col"})):
        x = run(cols, env)
        d = {}
        for k in FL + ("OLR", "ISR"):
            d[k] = float(np.max(np.abs(x["r"][k] - ref["r"][k])) / np.max(np.abs(ref["r"][k])))
        for j, nm in enumerate(("tau", "w0", "g", "tau_band")):
            e = 0.0
            for a, b in zip(x["op"], ref["op"]):
                m = np.abs(b[j]) > 0
                if m.any():
                    e = max(e, float(np.max(np.abs(a[j][m] - b[j][m]) / np.abs(b[j][m]))))
            d[nm] = e
        res[tag] = dict(rorr_ms=x["rorr_ms"], total_ms=x["total_ms"], exact_share=x["exact_share"], ovf=x["ovf"], cut=x["cut"],
                        max_rel_vs_sort=max(d.values()), worst=max(d, key=d.get), tau=d["tau"],
                        finite=bool(all(np.all(np.isfinite(x["r"][k])) for k in FL)))
    out[name] = res
    print(name, json.dumps(res), flush=True)
print(json.dumps(dict(ncol=ncol, **out)))
