"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/clima_rt.h declares, the ctypes mirror has the header's struct layout, and the product
path fails loudly (no CPU fallback) when no CUDA device is usable."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from cases import small_case
from clima_b200 import _abi, radtran
from conftest import HAS_GPU

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "clima_rt.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\bvoid\s+(clima_rt_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    names = _declared()
    for n in ("clima_rt_create", "clima_rt_destroy", "clima_rt_radiate_batch", "clima_rt_radiate_batch_device",
              "clima_rt_set_params", "clima_rt_set_custom_optical_properties"):
        assert n in names
    assert sorted(radtran.EXPORTS) == names, "clima_b200/radtran.py EXPORTS must mirror include/clima_rt.h"


def test_library_loads_and_exports_every_declared_symbol():
    lib = radtran.load_library()
    for n in _declared():
        assert hasattr(lib, n), n
    assert radtran.version().startswith("clima_rt_b200")


def test_no_torch_or_cxx_types_in_the_boundary():
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    assert "torch" not in src and "std::" not in src and "cudaStream_t" not in src
    out = subprocess.run(["nm", "-D", "--defined-only", radtran._LIB_PATH], capture_output=True, text=True).stdout
    exported = [ln.split()[-1] for ln in out.splitlines() if " T " in ln]
    assert set(_declared()) <= set(exported)


def test_ctypes_struct_layout_matches_the_header(tmp_path):
    """sizeof/offsetof from a C compile of the header == the ctypes mirror."""
    fields = {"clima_rt_tables": ["nw", "wavl", "nk", "k_log10k", "ncia", "ray_xs", "cont_temp", "part_gt",
                                  "ir_ind_start", "photons_sol"],
              "clima_rt_params": ["diurnal_fac", "has_hard_surface", "nzenith", "zenith_u", "surface_emissivity"],
              "clima_rt_out": ["fup_n_ir", "OLR", "fup_a_ir", "tau_band_sol", "status"]}
    prog = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(){"]
    for st, fl in fields.items():
        prog.append(f'printf("{st} %zu\\n", sizeof({st}));')
        for f in fl:
            prog.append(f'printf("{st}.{f} %zu\\n", offsetof({st}, {f}));')
    prog.append("return 0;}")
    c = tmp_path / "layout.c"
    c.write_text("\n".join(prog))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", str(c), "-o", str(exe)], check=True)
    got = dict(ln.split() for ln in subprocess.run([str(exe)], capture_output=True, text=True).stdout.splitlines())
    mirror = {"clima_rt_tables": _abi.CTables, "clima_rt_params": _abi.CParams, "clima_rt_out": _abi.COut}
    for st, fl in fields.items():
        assert int(got[st]) == C.sizeof(mirror[st]), st
        for f in fl:
            assert int(got[f"{st}.{f}"]) == getattr(mirror[st], f).offset, (st, f)


@pytest.mark.skipif(HAS_GPU, reason="only meaningful without a CUDA device")
def test_product_path_fails_loudly_without_a_gpu():
    tb, pr, cols = small_case(G=8, nw=6, nz=4, ncol=1)
    with pytest.raises(radtran.ClimaException, match="no usable CUDA device"):
        radtran.Radtran(tb, pr)


def test_error_string_convention():
    """err[1025], "" on success, message otherwise (clima/fortran/clima_c_api.f90:5,57-69)."""
    lib = radtran.load_library()
    err = C.create_string_buffer(b"x" * 10, _abi.ERR_LEN + 1)
    n = C.c_int32(1)
    lib.clima_rt_radiate_batch(None, C.byref(n), C.byref(n), C.byref(n), C.byref(n), None, None, None, None,
                               None, None, None, C.byref(n), C.byref(n), None, err)
    assert b"NULL handle" in err.value
    assert _abi.ERR_LEN == 1024


def test_zenith_quadrature_matches_reference_definition():
    """src/clima_eqns.f90:26-41: Gauss-Legendre on mu in [0,1], weights w/2 (sum = 1/2 * 2 = 1)."""
    for n in (1, 4, 8):
        u, w = _abi.zenith_angles_and_weights(n)
        assert abs(w.sum() - 1.0) < 1e-14 and np.all((u > 0) & (u < 1))
        assert abs((u * w).sum() - 0.5) < 1e-14  # integral of mu over [0,1]


def test_c_driver_compiles_against_the_header(tmp_path):
    """tests/c_driver.c (a C99 caller of the ABI, run on the GPU by tests/test_gpu_parity.py) compiles
    against include/clima_rt.h with gcc: the header is valid C, not only C++."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    assert gcc, "gcc is part of the image"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(root, "include"), "-c",
                        os.path.join(root, "tests", "c_driver.c"), "-o", str(tmp_path / "c_driver.o")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
