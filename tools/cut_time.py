#!/usr/bin/env python
"""Stage time of the default RORR path on synthetic columns, 3 timed calls.  usage: tools/cut_time.py ncol cfg"""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran
ncol = int(sys.argv[1]); cfg = sys.argv[2]
KW = dict(cfg2={}, cfg3=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0),
          cfg5=dict(randomize=True, seed=11))
tb = synth.make_tables()
g = Radtran(tb, synth.default_params(tb))
cols = synth.make_columns(ncol, nz=100, **KW[cfg])
g.radiate_batch(cols)
ms = []
for _ in range(3):
    g.radiate_batch(cols); ms.append(g.stage_ms()[0])
st = g.rorr_cut_stats()
print(json.dumps(dict(lib=os.environ.get("CLIMA_RT_LIBRARY", "default"), chain=os.environ.get("CLIMA_RT_RORR_CHAIN", "auto"),
                      len=os.environ.get("CLIMA_RT_RORR_CHAIN_LEN", "-"), ncol=ncol, cfg=cfg, rorr_ms=min(ms),
                      passes_per_merge=st["warp_passes"] / max(st["warp_merges"], 1), closed=st["closed_form_warp_merges"] / max(st["warp_merges"] + st["closed_form_warp_merges"], 1),
                      to_exact=st["units_to_exact"], restarts=st["restarts"], kernel=st["last_kernel"])))
