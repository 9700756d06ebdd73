"""Generates tests/golden/cfg1_modern_earth.npz from the reference's ModernEarth template
(BASELINE.json configs[0]; /root/reference/tests/test_radtran.f90:33-65):

  * the 200-level table of templates/ModernEarth/atmosphere.txt (labels + columns, verbatim), and
  * templates/ModernEarth/Sun_now.txt put on the S-canon wavelength grid by loader.read_stellar_flux
    (src/radtran/clima_radtran_types_create.f90:9-78) -- 100 numbers instead of the 800 KB file.

Runs only where /root/reference exists (this container); the GPU box reads the .npz.
    python tests/golden/make_cfg1_template.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from clima_b200 import loader, synth  # noqa: E402

REF = "/root/reference/templates/ModernEarth"


def main():
    labels, columns = loader.read_atmosphere_file(os.path.join(REF, "atmosphere.txt"))
    tb = synth.make_tables()
    photons = loader.read_stellar_flux(os.path.join(REF, "Sun_now.txt"), tb.wavl)
    np.savez_compressed(os.path.join(HERE, "cfg1_modern_earth.npz"), labels=np.array(labels), columns=columns,
                        wavl=tb.wavl, stellar_flux_on_scanon_grid=photons)
    print(labels, columns.shape, photons.shape, float(photons.sum()))


if __name__ == "__main__":
    main()
