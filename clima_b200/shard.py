"""Column sharding of the RT hot path over the GPUs of one box.

Columns are independent units: rank r computes a contiguous block of columns with its own handle
(tables replicated), there is NO exchange during compute, and ONE all-gather returns every rank's
flux profiles `[fup_n_ir, fdn_n_ir, fup_n_sol, fdn_n_sol, f_total](nz+1, ncol) + ISR, OLR(ncol)` to
every rank -- `(nz+1)*5 + 2` doubles per column.  The collective goes through `torch.distributed`
(NCCL over NVLink on the GPU box, gloo on CPU in the tests); torch is plumbing only, the compute
callable is injected (`Radtran.radiate_batch` in production).

Two gathers: `radiate_sharded_nccl` is the product path -- `clima_rt_radiate_batch_sharded` of the C
ABI (ncclAllGather of the device-resident result block on the handle's stream); `radiate_sharded` is
the same partition logic over any `torch.distributed`-like object with a host-staged gather, kept for
the world_size-2 gloo tests on CPU (tests/test_shard_gloo.py), where no GPU and no NCCL exist.

Host logic only: nothing here touches the oracle or offers a CPU compute path.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Tuple

import numpy as np

RESULT_PROFILES = ["fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"]  # (nz+1, ncol)
RESULT_SCALARS = ["ISR", "OLR"]                                                  # (ncol)
_COL_ARRAYS = ["T_surface", "T", "P", "densities", "dz", "pdensities", "radii"]


def partition(ncol: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous blocks [begin, end) of columns per rank; the first `ncol % world` ranks hold one
    column more.  Ranks beyond `ncol` get empty blocks."""
    if world < 1:
        raise ValueError("world size must be >= 1")
    base, rem = divmod(int(ncol), world)
    out, b = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((b, b + n))
        b += n
    return out


def _reshape(cols: Dict, key: str) -> np.ndarray:
    """View of a (possibly flattened, Fortran-ordered) column array with the column index last."""
    a = np.asarray(cols[key])
    ncol = cols["ncol"]
    if a.ndim == 1 and a.size != ncol:
        return a.reshape((-1, ncol), order="F")
    return a


def slice_columns(cols: Dict, begin: int, end: int) -> Dict:
    """The sub-batch [begin, end) of a column dict (arrays keep the column index last, Fortran order)."""
    out = {k: v for k, v in cols.items() if k not in _COL_ARRAYS}
    out["ncol"] = end - begin
    for k in _COL_ARRAYS:
        if k in cols:
            a = _reshape(cols, k)
            out[k] = np.asfortranarray(a[..., begin:end])
    return out


def pack_results(res: Dict, nz: int, ncol: int) -> np.ndarray:
    """One contiguous message per rank: 5 profiles (nz+1, ncol) then ISR, OLR (ncol)."""
    L = nz + 1
    buf = np.empty(ncol * (5 * L + 2), dtype=np.float64)
    o = 0
    for k in RESULT_PROFILES:
        buf[o:o + L * ncol] = np.asarray(res[k], dtype=np.float64).reshape(-1, order="F")
        o += L * ncol
    for k in RESULT_SCALARS:
        buf[o:o + ncol] = np.asarray(res[k], dtype=np.float64).reshape(-1)
        o += ncol
    return buf


def unpack_results(buf: np.ndarray, nz: int, ncol: int) -> Dict:
    L = nz + 1
    out, o = {}, 0
    for k in RESULT_PROFILES:
        out[k] = buf[o:o + L * ncol].reshape((L, ncol), order="F")
        o += L * ncol
    for k in RESULT_SCALARS:
        out[k] = buf[o:o + ncol].copy()
        o += ncol
    return out


def gather_results(local: Dict, nz: int, blocks: List[Tuple[int, int]], dist=None, device="cpu") -> Dict:
    """All-gather of the per-rank result blocks (ragged blocks are padded to the largest one)."""
    import torch
    world = len(blocks)
    sizes = [e - b for b, e in blocks]
    per_col = 5 * (nz + 1) + 2
    nmax = max(sizes)
    rank = dist.get_rank() if (dist is not None and world > 1) else 0
    mine = torch.zeros(nmax * per_col, dtype=torch.float64, device=device)
    if sizes[rank] > 0:
        mine[:sizes[rank] * per_col] = torch.from_numpy(pack_results(local, nz, sizes[rank])).to(device)
    if world > 1:
        allb = torch.zeros(world * nmax * per_col, dtype=torch.float64, device=device)
        dist.all_gather_into_tensor(allb, mine)
        allb = allb.cpu().numpy()
    else:
        allb = mine.cpu().numpy()
    parts = [unpack_results(allb[r * nmax * per_col:r * nmax * per_col + sizes[r] * per_col], nz, sizes[r])
             for r in range(world) if sizes[r] > 0]
    out = {}
    for k in RESULT_PROFILES:
        out[k] = np.asfortranarray(np.concatenate([p[k] for p in parts], axis=1))
    for k in RESULT_SCALARS:
        out[k] = np.concatenate([p[k] for p in parts])
    return out


def radiate_sharded(compute: Callable[[Dict], Dict], cols: Dict, dist=None, device="cpu") -> Dict:
    """Every rank holds the full column batch description, computes its own block with `compute`
    (e.g. `Radtran(...).radiate_batch`) and receives all flux profiles.  With `dist=None` or world
    size 1 this is just `compute(cols)` restricted to the gathered keys."""
    world = dist.get_world_size() if dist is not None else 1
    rank = dist.get_rank() if dist is not None else 0
    blocks = partition(cols["ncol"], world)
    b, e = blocks[rank]
    local = compute(slice_columns(cols, b, e)) if e > b else {}
    return gather_results(local, cols["nz"], blocks, dist if world > 1 else None, device)


def radiate_sharded_nccl(rt, cols: Dict, rank: int, world: int, **kw) -> Dict:
    """Product path: this rank's block through `Radtran.radiate_batch_sharded` (C ABI: compute, then
    ncclAllGather on the handle's stream; the handle must have a communicator, `Radtran.comm_init`).
    Returns the results of ALL columns in their original order."""
    ncol = cols["ncol"]
    parts = partition(ncol, world)
    b, e = parts[rank]
    ncol_max = max(pe - pb for pb, pe in parts)
    mine = slice_columns(cols, b, e)
    out = rt.radiate_batch_sharded(mine, max(ncol_max, 1), **kw)
    res = {}
    for k in RESULT_PROFILES:
        res[k] = np.concatenate([out[k][:, :pe - pb, r] for r, (pb, pe) in enumerate(parts)], axis=1)
    for k in RESULT_SCALARS:
        res[k] = np.concatenate([out[k][:pe - pb, r] for r, (pb, pe) in enumerate(parts)])
    return res
