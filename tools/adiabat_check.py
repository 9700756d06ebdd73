"""Developer check: device make_profile against the oracle, column by column (max relative differences and
right-hand-side evaluation counts).  python tools/adiabat_check.py [ncol] [nz] [rtol]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_adiabat as TA  # noqa: E402

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 6
nz = int(sys.argv[2]) if len(sys.argv) > 2 else 24
_, _, ac = TA.gpu_model(nz=nz)
ac.T_trop = 150.0
if len(sys.argv) > 3:
    ac.rtol = float(sys.argv[3])
T, P = TA.sweep(ncol)
p = ac.make_profile_batch(T, P)
for c in range(ncol):
    o = TA.oracle_column(ac, T[c], P[:, c])
    d = {k: float(np.max(np.abs(p[k][:, c] / o[k] - 1.0))) for k in ("P", "T", "z", "dz")}
    d["f_i"] = float(np.max(np.abs(p["f_i"][:, :, c] - o["f_i"])))
    print(c, "nfcn gpu/oracle", p["nfcn"][c], o["edges"]["nfcn"], "P_trop", p["P_trop"][c], o["P_trop"],
          " ".join("%s %.2e" % kv for kv in d.items()))
