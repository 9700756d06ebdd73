"""Generates the golden vectors under tests/golden/ from the PARITY build of the CPU oracle
(oracle/oracle.cpp, -O2 -ffp-contract=off).  The reference's own tests hold no numeric vectors for
this path and the Fortran reference cannot be built in this image, so these are frozen outputs of
the oracle (after it agreed with the independent NumPy restatement oracle/oracle_np.py to <=1e-10),
not outputs of the reference: "parity unpinned".

    python tests/golden/make_golden.py        # rewrites tests/golden/*.npz

Inputs are regenerated from seeds by the CASES functions below (also imported by the tests), so only
the integrated outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


def _small_g16():
    from cases import small_case
    return small_case(G=16, nw=12, nz=10, ncol=3) + (None,)


def _small_g8_nohard():
    from cases import small_case
    return small_case(G=8, nw=9, nz=9, ncol=2, doubled=False, has_hard_surface=False, nzen=3, seed=7) + (None,)


def _small_particles_custom():
    from cases import small_case
    tb, pr, cols = small_case(G=16, nw=10, nz=8, ncol=2, npart=2, seed=11)
    rng = np.random.default_rng(5)
    nP = 7
    custom = (np.linspace(0.5, 6.5, nP), np.asfortranarray(10.0 ** rng.uniform(-9, -6, (nP, tb.nw))),
              np.asfortranarray(rng.uniform(0.1, 0.9, (nP, tb.nw))),
              np.asfortranarray(rng.uniform(0.0, 0.8, (nP, tb.nw))))
    return tb, pr, cols, custom


def _scanon_2col():
    from clima_b200 import synth
    tb = synth.make_tables()
    pr = synth.default_params(tb)
    cols = synth.make_columns(2, nz=100)
    return tb, pr, cols, None


CASES = {"small_g16": _small_g16, "small_g8_nohard": _small_g8_nohard,
         "small_particles_custom": _small_particles_custom, "scanon_2col": _scanon_2col}
KEYS = ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total", "ISR", "OLR")

if __name__ == "__main__":
    from oracle_lib import OracleRadtran
    for name, fn in CASES.items():
        tb, pr, cols, custom = fn()
        o = OracleRadtran(tb, pr, kind="parity")
        if custom is not None:
            o.set_custom(*custom)
        r = o.radiate_batch(cols, spectral=False, nthreads=1)
        np.savez(os.path.join(HERE, name + ".npz"), **{k: r[k] for k in KEYS})
        print(name, "OLR", r["OLR"], "ISR", r["ISR"])
