"""Writes tests/golden/adiabat_species.json: the gases of the reference's own AdiabatClimate template
(templates/AdiabatClimate/species.yaml: masses from the atom table, heat-capacity polynomials, LinearLatentHeat
saturation data of H2O and CO2), parsed by clima_b200.adiabat.species_from_yaml, so that the adiabat tests
(tests/test_adiabat.py) use the reference's species data on machines without /root/reference.

    python tests/golden/make_adiabat_species.py
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from clima_b200.adiabat import species_from_yaml  # noqa: E402

sp = species_from_yaml("/root/reference/templates/AdiabatClimate/species.yaml")
with open(os.path.join(HERE, "adiabat_species.json"), "w") as f:
    json.dump(sp, f, indent=1)
print(sp["names"])
