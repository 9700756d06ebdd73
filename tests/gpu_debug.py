"""Debug driver for the first GPU runs: prints the parity error table."""
import sys, time
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from cases import compare, small_case
from clima_b200 import synth
from clima_b200.radtran import Radtran, measure_fp64_peak
from oracle_lib import OracleRadtran

print("fp64 peak TFLOP/s:", measure_fp64_peak(0))
for G in (16, 8, 20):
    tb, pr, cols = small_case(G=G, nw=10, nz=8, ncol=3)
    g = Radtran(tb, pr); o = OracleRadtran(tb, pr)
    rg = g.radiate_batch(cols, spectral=True); ro = o.radiate_batch(cols, spectral=True, nthreads=8)
    print("G", G)
    try:
        compare(rg, ro, verbose=True)
    except AssertionError as e:
        print("FAIL", e)
    tg = g.get_opacity(cols["nz"], 0); o.radiate(cols, 0); to = o.get_opacity(cols["nz"])
    for a, b, n in zip(tg, to, ("tau", "w0", "g", "tband")):
        m = np.abs(b) > 0
        print("   opacity", n, np.max(np.abs(a[m]-b[m])/np.abs(b[m])) if m.any() else 0.0)
tb = synth.make_tables(); pr = synth.default_params(tb)
for ncol in (8, 256, 1024):
    cols = synth.make_columns(ncol, nz=100)
    g = Radtran(tb, pr)
    g.radiate_batch(cols)
    t = time.time(); r = g.radiate_batch(cols); dt = time.time() - t
    print(f"ncol {ncol}: {dt*1e3:.1f} ms  {ncol/dt:.0f} col/s  stages {g.stage_ms()}")
    if ncol == 8:
        o = OracleRadtran(tb, pr)
        ro = o.radiate_batch(cols, nthreads=8)
        try:
            compare(r, ro, spectral=False, verbose=True)
        except AssertionError as e:
            print("FAIL", e)
