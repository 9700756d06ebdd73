"""`clima_rt_make_profile_batch_noalloc` (include/clima_rt.h): make_profile with its scratch in a static device array,
for callers that may not allocate.  No counterpart in the reference;
checked against the existing entry point `clima_rt_make_profile_batch`, through the C ABI, on two calls in a row.
The file name keeps it last in pytest's collection order."""
import numpy as np
import pytest

from test_adiabat import gpu_model, sweep


def _device_buffers(torch, nz, ng, ncol):
    f64 = dict(P=(ncol, nz), T=(ncol, nz), z=(ncol, nz), dz=(ncol, nz), f_i=(ncol, ng, nz), f_i_surf=(ncol, ng),
               N_surface=(ncol, ng), P_surf=(ncol,), P_trop=(ncol,))   # C order of the Fortran (nz, ..., ncol)
    bufs = {k: torch.full(s, float("nan"), dtype=torch.float64, device="cuda:0") for k, s in f64.items()}
    bufs["status"] = torch.full((ncol,), -1, dtype=torch.int32, device="cuda:0")
    bufs["nfcn"] = torch.full((ncol,), -1, dtype=torch.int32, device="cuda:0")
    return bufs


def _host(b):
    """device buffer (C order, column slowest) -> the Fortran-ordered array of the host entry point"""
    a = b.cpu().numpy()
    return np.transpose(a, tuple(range(a.ndim))[::-1])


@pytest.mark.gpu
def test_no_alloc_make_profile_matches_the_allocating_entry_point():
    import torch
    from clima_b200.radtran import ClimaException
    nz, ng = 20, 7
    _, _, ac = gpu_model(nz=nz)
    ac.T_trop = 150.0
    keys = ("P", "T", "z", "dz", "f_i", "f_i_surf", "N_surface", "P_surf", "P_trop", "status", "nfcn")
    # two calls in a row on the SAME output buffers and the same static scratch, different columns and batch sizes
    # (the second is smaller: the scratch and the tail of the buffers keep the first call's contents)
    nmax = 300
    bufs = _device_buffers(torch, nz, ng, nmax)
    for call, (ncol, seed) in enumerate(((300, 21), (37, 22))):
        T, P = sweep(ncol, seed=seed)
        T[1] = 100.0                                     # one failing column: reported in status, not in err
        want = ac.make_profile_batch(T, P, check=False)  # the existing entry point (host pointers, allocates)
        dT = torch.from_numpy(np.ascontiguousarray(T)).to("cuda:0")
        dP = torch.from_numpy(np.ascontiguousarray(P.T)).to("cuda:0")   # (ncol, ng) C order = (ng, ncol) Fortran
        before = torch.cuda.memory_allocated(0)
        ac.make_profile_batch_noalloc(ncol, dT.data_ptr(), dP.data_ptr(), {k: v.data_ptr() for k, v in bufs.items()})
        assert torch.cuda.memory_allocated(0) == before
        ok = want["status"] == 0
        assert np.array_equal(_host(bufs["status"])[:ncol], want["status"]) and want["status"][1] == 1 and ok.sum() == ncol - 1
        for k in keys[:9] + ("nfcn",):
            got = _host(bufs[k])[..., :ncol]
            # same kernel, same inputs: bit-identical (the tests of this path allow 1e-9 on T, 5e-7 on altitudes)
            assert np.array_equal(got[..., ok], want[k][..., ok]), (k, call)
    # one column more than the static array holds: a clear error, nothing launched
    too_many = ac.NOALLOC_MAX_WORK_DOUBLES // (7 * nz + 5) + 1
    with pytest.raises(ClimaException, match="exceed the static maximum"):
        ac.make_profile_batch_noalloc(too_many, dT.data_ptr(), dP.data_ptr(), {k: v.data_ptr() for k, v in bufs.items()})
    with pytest.raises(ClimaException, match="every member of `profile`"):
        ac.make_profile_batch_noalloc(4, dT.data_ptr(), dP.data_ptr(), {k: v.data_ptr() for k, v in bufs.items() if k != "z"})
