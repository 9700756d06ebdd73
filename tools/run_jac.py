#!/usr/bin/env python
"""One stacked clima_rt_jacobian_batch call (for ncu captures).  usage: tools/run_jac.py [njac]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran
nj = int(sys.argv[1]) if len(sys.argv) > 1 else 64
nzm = 50
tb = synth.make_tables()
g = Radtran(tb, synth.default_params(tb, nzenith=1, albedo=0.3))
base = synth.make_columns(1, nz=nzm, doubled=True)
g.radiate_batch(base)
Tm = base["T"][0:2 * nzm:2, 0].copy(); Ts0 = float(base["T_surface"][0])
ncol = 51 * nj
T = np.repeat(Tm[:, None], ncol, axis=1).copy(order="F"); Ts = np.full(ncol, Ts0)
for j in range(nj):
    Ts[51 * j] += 1e-2 * Ts0
    for i in range(nzm):
        T[i, 51 * j + i + 1] += 1e-2 * Tm[i]
f = g.jacobian_batch(Ts, T)
print(nj, g.stage_ms(), g.rorr_cut_stats()["jacobian_delta_calls"], float(f.mean()))
