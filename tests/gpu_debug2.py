import sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from cases import small_case
from clima_b200.radtran import Radtran
from oracle_lib import OracleRadtran
np.set_printoptions(linewidth=200, precision=4)
nk = int(sys.argv[1]) if len(sys.argv) > 1 else 2
tb, pr, cols = small_case(G=16, nw=4, nz=4, ncol=1, nk=nk, doubled=False)
g = Radtran(tb, pr); o = OracleRadtran(tb, pr)
g.radiate_batch(cols); o.radiate(cols, 0)
tg = g.get_opacity(cols["nz"], 0)[0]; to = o.get_opacity(cols["nz"])[0]
for l in range(4):
    for n in range(4):
        r = tg[n, :, l] / to[n, :, l] - 1
        print(l, n, np.array2string(r, formatter={'float_kind': lambda x: "%8.1e" % x}))
print(g.rorr_stats())
