"""Writes tests/golden/tutorial_adiabat_outputs.json: the numbers the REFERENCE itself printed for `make_profile` in its
tutorial notebook (examples/Tutorial_AdiabatClimate.ipynb, stored cell outputs: cells 6, 10, 11, 14), with the inputs of
the cell that produced them.  They are the only outputs of the real code for row N4 that exist on this machine: the oracle
(oracle/adiabat_np.py) is pinned against them in tests/test_adiabat.py.

    python tests/golden/make_tutorial_outputs.py
"""
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
nb = json.load(open("/root/reference/examples/Tutorial_AdiabatClimate.ipynb"))
cells = nb["cells"]


def out(i):
    return "".join("".join(o.get("text", "")) for o in cells[i].get("outputs", []))


def numbers(s):
    return [float(x) for x in re.findall(r"[-+]?\d+\.?\d*(?:[eE][-+]?\d+)?", s)]


src6 = "".join(cells[6]["source"])
assert "c.make_profile(T_surf, P_i)" in src6 and "c.P_top = 1.0e-2" in src6 and "c.T_trop = 150" in src6
assert "print(c.N_atmos)" in "".join(cells[10]["source"]) and "print(c.N_surface)" in "".join(cells[11]["source"])
assert "print(c.P_surf/1e6)" in "".join(cells[14]["source"])
fix = dict(
    source="examples/Tutorial_AdiabatClimate.ipynb cells 6, 10, 11, 14 (stored outputs of the reference)",
    inputs=dict(T_surf=280.0, P_i_bar=[260.0, 10.0, 1.0, 1.0e-10, 1.0e-10, 1.0e-10, 1.0e-10], P_top=1.0e-2, T_trop=150.0,
                RH=1.0, nz=100, planet_mass=5.972e27, planet_radius=6.371e8,
                note="species.yaml and Earth/settings.yaml of templates/AdiabatClimate (number-of-layers: 100)"),
    N_atmos=numbers(out(10)), N_surface=numbers(out(11)), P_surf_bar=numbers(out(14))[0],
    N_atmos_text=out(10).strip(), N_surface_text=out(11).strip(), P_surf_text=out(14).strip())
assert len(fix["N_atmos"]) == 7 and len(fix["N_surface"]) == 7
with open(os.path.join(HERE, "tutorial_adiabat_outputs.json"), "w") as f:
    json.dump(fix, f, indent=1)
print(fix["N_atmos"], fix["N_surface"][0], fix["P_surf_text"])
