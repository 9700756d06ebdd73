#!/usr/bin/env python
"""Experiment: do the ALU-bound RORR stage and the latency/HBM-bound two-stream stage overlap when two
handles (two streams) work on half batches concurrently?  usage: overlap_two_handles.py [ncol_total] [nthreads]"""
import sys, os, time, threading
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran

ntot = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nth = int(sys.argv[2]) if len(sys.argv) > 2 else 2
tb = synth.make_tables(); pr = synth.default_params(tb, nzenith=4, albedo=0.25)
def run(nhandles, reps=4, stagger=True):
    n = ntot // nhandles
    cols = [synth.make_columns(n, nz=100, seed=synth.SEED + i) for i in range(nhandles)]
    hs = [Radtran(tb, pr) for _ in range(nhandles)]
    for h, c in zip(hs, cols): h.radiate_batch(c)
    def work(i):
        for _ in range(reps): hs[i].radiate_batch(cols[i])
    ths = [threading.Thread(target=work, args=(i,)) for i in range(nhandles)]
    t0 = time.perf_counter()
    for i, t in enumerate(ths):
        t.start()
        if stagger and i + 1 < nhandles: time.sleep(0.5 * 0.19 * n / 4096)
    for t in ths: t.join()
    dt = time.perf_counter() - t0
    return reps * ntot / dt
print("1 handle :", run(1))
print(f"{nth} handles:", run(nth))
