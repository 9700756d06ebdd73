#!/usr/bin/env python
"""One call of the hot path on synthetic columns (for ncu captures).  usage: tools/run_once.py [ncol] [cfg2|cfg3|cfg5] [ncalls]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 256
cfg = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
ncalls = int(sys.argv[3]) if len(sys.argv) > 3 else 1
KW = dict(cfg2={}, cfg3=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0),
          cfg5=dict(randomize=True, seed=11))
tb = synth.make_tables()
g = Radtran(tb, synth.default_params(tb))
cols = synth.make_columns(ncol, nz=100, **KW[cfg])
for _ in range(ncalls):
    r = g.radiate_batch(cols)
print(ncol, cfg, g.stage_ms(), g.rorr_stats(), g.rorr_cut_stats(), float(r["OLR"].mean()))
