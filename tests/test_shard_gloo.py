"""CPU tests of the N>1 host logic (clima_b200/shard.py): column partition, ragged blocks, message
packing and the all-gather of flux profiles, run as 2 real processes over the `gloo` backend.
The compute callable injected here is the oracle (test infrastructure); on the GPU box it is
`Radtran.radiate_batch` (bench.py, tests/test_gpu_parity.py::test_sharded_matches_single)."""
import os
import socket

import numpy as np
import pytest

from cases import small_case
from clima_b200 import shard


def test_partition_covers_all_columns_contiguously():
    for ncol in (0, 1, 2, 5, 8, 4096, 65536 + 3):
        for world in (1, 2, 3, 4, 8):
            blocks = shard.partition(ncol, world)
            assert len(blocks) == world
            assert blocks[0][0] == 0 and blocks[-1][1] == ncol
            for (b0, e0), (b1, e1) in zip(blocks[:-1], blocks[1:]):
                assert e0 == b1
            sizes = [e - b for b, e in blocks]
            assert max(sizes) - min(sizes) <= 1 and sorted(sizes, reverse=True) == sizes
    with pytest.raises(ValueError):
        shard.partition(4, 0)


def test_slice_columns_keeps_fortran_layout():
    tb, pr, cols = small_case(G=8, nw=6, nz=6, ncol=5)
    sub = shard.slice_columns(cols, 1, 4)
    assert sub["ncol"] == 3 and sub["nz"] == cols["nz"]
    T = np.asarray(cols["T"]).reshape((cols["nz"], 5), order="F")
    assert np.array_equal(np.asarray(sub["T"]).reshape((cols["nz"], 3), order="F"), T[:, 1:4])
    d = np.asarray(cols["densities"]).reshape((cols["nz"], cols["ng"], 5), order="F")
    assert np.array_equal(np.asarray(sub["densities"]).reshape((cols["nz"], cols["ng"], 3), order="F"),
                          d[:, :, 1:4])
    assert np.array_equal(np.asarray(sub["T_surface"]).ravel(), np.asarray(cols["T_surface"]).ravel()[1:4])


def test_pack_unpack_roundtrip():
    rng = np.random.default_rng(0)
    nz, ncol = 7, 3
    res = {k: np.asfortranarray(rng.standard_normal((nz + 1, ncol))) for k in shard.RESULT_PROFILES}
    res.update({k: rng.standard_normal(ncol) for k in shard.RESULT_SCALARS})
    buf = shard.pack_results(res, nz, ncol)
    assert buf.size == ncol * (5 * (nz + 1) + 2)
    back = shard.unpack_results(buf, nz, ncol)
    for k in res:
        assert np.array_equal(back[k], res[k])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ncol, q):
    import torch.distributed as dist
    from oracle_lib import OracleRadtran
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tb, pr, cols = small_case(G=8, nw=6, nz=6, ncol=ncol)
        o = OracleRadtran(tb, pr, kind="parity")
        out = shard.radiate_sharded(lambda c: o.radiate_batch(c, nthreads=1), cols, dist=dist)
        q.put((rank, {k: np.array(v) for k, v in out.items()}))
    finally:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("ncol", [4, 5, 1])
def test_two_rank_gloo_gather_equals_single_process(ncol):
    """world_size 2: even split, ragged split (3+2) and a rank with an empty block (1+0)."""
    import torch.multiprocessing as mp
    from oracle_lib import OracleRadtran
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ncol, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tb, pr, cols = small_case(G=8, nw=6, nz=6, ncol=ncol)
    ref = OracleRadtran(tb, pr, kind="parity").radiate_batch(cols, nthreads=1)
    for rank in (0, 1):  # every rank receives every column's profiles
        for k in shard.RESULT_PROFILES + shard.RESULT_SCALARS:
            a = np.asarray(got[rank][k])
            b = np.asarray(ref[k]).reshape(a.shape, order="F")
            assert np.array_equal(a, b), (rank, k)


def test_world_size_one_is_plain_compute():
    from oracle_lib import OracleRadtran
    tb, pr, cols = small_case(G=8, nw=6, nz=6, ncol=3)
    o = OracleRadtran(tb, pr, kind="parity")
    out = shard.radiate_sharded(lambda c: o.radiate_batch(c, nthreads=1), cols, dist=None)
    ref = o.radiate_batch(cols, nthreads=1)
    for k in shard.RESULT_PROFILES + shard.RESULT_SCALARS:
        assert np.array_equal(out[k], np.asarray(ref[k]).reshape(out[k].shape, order="F")), k
