import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
from clima_b200 import synth
from clima_b200.radtran import Radtran
from oracle_lib import OracleRadtran
from cases import elem_rel_err, rel_err
tb = synth.make_tables(nw=16, n_sol=10, ir_start=6)
pr = synth.default_params(tb, nzenith=1)
cols = synth.make_columns(48, nz=16, T_surf_range=(260.0, 260.0), doubled=False)
rng = np.random.default_rng(9)
cols = dict(cols, T=np.asfortranarray(cols["T"] * rng.uniform(0.6, 3.0, cols["T"].shape)))
ro = OracleRadtran(tb, pr).radiate_batch(cols, nthreads=8)
ro2 = OracleRadtran(tb, pr, kind="fast").radiate_batch(cols, nthreads=8)
print("oracle parity vs fast build:", {k: elem_rel_err(ro2[k], ro[k]) for k in ("fup_n_ir","fdn_n_ir")})
for name, env in (("cut", {"CLIMA_RT_RORR_CHAIN": "col"}), ("sort", {"CLIMA_RT_RORR": "sort"}), ("exact", {"CLIMA_RT_RORR": "exact"})):
    os.environ.update(env)
    g = Radtran(tb, pr)
    for k in env: del os.environ[k]
    r = g.radiate_batch(cols)
    print(name, {k: (round(rel_err(r[k], ro[k]),16), elem_rel_err(r[k], ro[k])) for k in ("fup_n_ir","fdn_n_ir","fup_n_sol","fdn_n_sol")})
    k="fdn_n_ir"; a=r[k]; b=ro[k]; m=np.abs(b)>1e-12*np.abs(b).max(); e=np.where(m, np.abs(a-b)/np.abs(b), 0); i=np.unravel_index(e.argmax(), e.shape); print("  worst at", i, a[i], b[i], "max", np.abs(b).max())
