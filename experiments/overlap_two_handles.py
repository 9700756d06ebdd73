#!/usr/bin/env python
"""Experiment: do the issue-bound RORR stage and the latency/HBM-bound two-stream stage overlap when two
handles (two streams) work on half batches concurrently?  Knobs (environment): CLIMA_RT_RORR_BLOCKS_PER_SM
(cap of the staircase kernel's resident blocks), CLIMA_RT_TS_BLOCKS_PER_SM (persistent two-stream grid).
usage: overlap_two_handles.py [ncol_total] [nhandles] [stagger fraction of a half step]"""
import sys, os, time, threading, json
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran

ntot = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nth = int(sys.argv[2]) if len(sys.argv) > 2 else 2
stag = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
tb = synth.make_tables(); pr = synth.default_params(tb, nzenith=4, albedo=0.25)
allc = synth.make_columns(ntot, nz=100)


def part(i, n):
    return {k: (np.asfortranarray(v[..., i * n:(i + 1) * n]) if isinstance(v, np.ndarray) and v.shape[-1] == ntot else v) for k, v in allc.items()} | {"ncol": n}


def run(nhandles, reps=4):
    n = ntot // nhandles
    cols = [part(i, n) for i in range(nhandles)]
    hs = [Radtran(tb, pr) for _ in range(nhandles)]
    for h, c in zip(hs, cols): h.radiate_batch(c)
    def work(i):
        for _ in range(reps): hs[i].radiate_batch(cols[i])
    ths = [threading.Thread(target=work, args=(i,)) for i in range(nhandles)]
    t0 = time.perf_counter()
    for i, t in enumerate(ths):
        t.start()
        if i + 1 < nhandles: time.sleep(stag * 0.12 * n / 4096)
    for t in ths: t.join()
    dt = time.perf_counter() - t0
    return reps * ntot / dt
print(json.dumps(dict(rorr_blocks=os.environ.get("CLIMA_RT_RORR_BLOCKS_PER_SM", "-"), ts_blocks=os.environ.get("CLIMA_RT_TS_BLOCKS_PER_SM", "-"),
                      one_handle=round(run(1)), n_handles=nth, cols_per_s=round(run(nth)))))
