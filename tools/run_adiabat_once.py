"""One batched AdiabatClimate TOA_fluxes call (for ncu launch lists): python tools/run_adiabat_once.py [ncol] [nz]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_adiabat as TA  # noqa: E402
from clima_b200 import synth  # noqa: E402
from clima_b200.adiabat import AdiabatClimate  # noqa: E402
from clima_b200.radtran import Radtran  # noqa: E402

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nz = int(sys.argv[2]) if len(sys.argv) > 2 else 50
tb = synth.make_tables()
ac = AdiabatClimate(Radtran(tb, synth.default_params(tb), device=0), TA.species(), nz, TA.EARTH_MASS, TA.EARTH_RADIUS)
T, P = TA.sweep(ncol, seed=11)
isr, olr = ac.TOA_fluxes_batch(T, P)
print("OLR", olr.min(), olr.max(), "stage ms", ac.rad.stage_ms() if hasattr(ac.rad, "stage_ms") else "")
