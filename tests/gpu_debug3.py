"""Run-to-run reproducibility probe at bench size."""
import os, sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from clima_b200 import synth
from clima_b200.radtran import Radtran
ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
tb = synth.make_tables(); pr = synth.default_params(tb, nzenith=4, albedo=0.25)
cols = synth.make_columns(ncol, nz=100, seed=synth.SEED, doubled=True)
g = Radtran(tb, pr)
r1 = {k: v.copy() for k, v in g.radiate_batch(cols).items()}
t1 = g.get_opacity(cols["nz"], 3)[0].copy()
for it in range(3):
    r2 = g.radiate_batch(cols)
    for k in ("fup_n_ir", "fup_n_sol", "OLR", "ISR"):
        d = np.abs(r2[k] - r1[k])
        if d.max() > 0:
            idx = np.unravel_index(np.argmax(d), d.shape)
            print(it, k, "max abs diff", d.max(), "rel", d.max() / np.abs(r1[k]).max(), "at", idx,
                  "ncols differing", np.count_nonzero(d.reshape(-1, ncol).max(axis=0) if d.ndim > 1 else d))
        else:
            print(it, k, "identical")
    t2 = g.get_opacity(cols["nz"], 3)[0]
    print(it, "tau col3 identical", np.array_equal(t1, t2))
print(g.rorr_stats())
