#!/usr/bin/env python
"""cfg3 / cfg5 at their full per-GPU sizes: throughput, share of units the 32-bit-key
certificate sends to the exact kernel, worklist overflows, and the fast-vs-exact property on a slice.
usage (GPU box): experiments/cfg35_fullsize.py [ncol]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
tb = synth.make_tables()
pr = synth.default_params(tb)
out = {}
for name, kw in (("cfg5_randomised", dict(randomize=True, seed=11)),
                 ("cfg3_early_mars", dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0))):
    cols = synth.make_columns(ncol, nz=100, **kw)
    g = Radtran(tb, pr)
    g.radiate_batch(cols)
    t0 = time.perf_counter(); r = g.radiate_batch(cols); dt = time.perf_counter() - t0
    units, exact, ovf = g.rorr_stats()
    ok = all(np.all(np.isfinite(r[k])) for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol"))
    out[name] = dict(columns=ncol, columns_per_s_host_to_host=ncol / dt, exact_kernel_share=exact / units,
                     worklist_overflows=int(ovf), finite=bool(ok), stage_ms=g.stage_ms())
    # property on a slice: all units through the 64-bit-key kernel
    sl = {k: (np.asfortranarray(v[..., :64]) if isinstance(v, np.ndarray) and v.shape[-1] == ncol else v) for k, v in cols.items()}
    r1 = Radtran(tb, pr).radiate_batch(sl)
    os.environ["CLIMA_RT_RORR"] = "exact"
    r2 = Radtran(tb, pr).radiate_batch(sl)
    del os.environ["CLIMA_RT_RORR"]
    out[name]["fast_vs_exact_max_rel"] = max(float(np.max(np.abs(r1[k] - r2[k])) / np.max(np.abs(r2[k])))
                                             for k in ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol"))
print(json.dumps(out))
