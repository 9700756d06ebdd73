#!/usr/bin/env python
"""Experiment: can a two-stream-only stream (frozen opacity: IR + solar kernels) run concurrently with a
stream doing full radiate calls (RORR-dominated)?  Handle A (low priority) loops full calls; handle B
(high priority, CLIMA_RT_TS_BLOCKS_PER_SM blocks) loops frozen-opacity calls.  Prints the rates alone and
together.  usage: overlap_rorr_ts.py [ncol] [ts_blocks_per_sm]"""
import sys, os, time, threading
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from clima_b200 import synth
from clima_b200.radtran import Radtran
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
tsb = sys.argv[2] if len(sys.argv) > 2 else "2"
tb = synth.make_tables(); pr = synth.default_params(tb, nzenith=4, albedo=0.25)
cols = synth.make_columns(n, nz=100)
os.environ["CLIMA_RT_STREAM_PRIORITY"] = "low"
A = Radtran(tb, pr); A.radiate_batch(cols)
os.environ["CLIMA_RT_STREAM_PRIORITY"] = "high"; os.environ["CLIMA_RT_TS_BLOCKS_PER_SM"] = tsb
B = Radtran(tb, pr); B.radiate_batch(cols)
def loopA(reps, out): 
    t0 = time.perf_counter()
    for _ in range(reps): A.radiate_batch(cols)
    out["A"] = reps * n / (time.perf_counter() - t0)
def loopB(reps, out):
    t0 = time.perf_counter()
    for _ in range(reps): B.radiate_batch(cols, compute_opacity=False)
    out["B"] = reps * n / (time.perf_counter() - t0)
o = {}; loopA(4, o); loopB(12, o); print("alone   : full calls %.0f col/s, two-stream-only calls %.0f col/s" % (o["A"], o["B"]))
o = {}
ta = threading.Thread(target=loopA, args=(6, o)); tbb = threading.Thread(target=loopB, args=(14, o))
t0 = time.perf_counter(); ta.start(); tbb.start(); ta.join(); tbb.join(); dt = time.perf_counter() - t0
print("together: full calls %.0f col/s, two-stream-only calls %.0f col/s  (wall %.3f s; serial estimate %.3f s)" % (o["A"], o["B"], dt, 0))
