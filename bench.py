#!/usr/bin/env python
"""bench.py -- column RT solves/s (IR + solar, RORR opacity recomputed per column, FP64).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

A "step" = one pass of the RT hot path (prep -> optics+RORR -> Planck -> two-stream IR ->
two-stream solar -> frequency integration) over one batch of synthetic columns per GPU:
BASELINE.json configs[1], the AdiabatClimate OLR-vs-T_surf sweep, 4096 columns, nz=100 doubled to
nz_r=202 radiative layers, 4 zenith angles, S-canon synthetic tables (nw=100, G=16, K=5;
SURVEY.md §8d -- the real photochem_clima_data tables are not on this box).

  value : columns/s with inputs resident in HBM (clima_rt_radiate_batch_device), CUDA events on the
          launching stream, max over ranks; N>1 adds the NCCL all-gather of the flux profiles.
  e2e   : the same through the host-pointer C-ABI call (clima_rt_radiate_batch): pinned host
          inputs -> H2D -> kernels -> D2H flux profiles, every step, wall clock around the
          (blocking) call, max over ranks.
  roofline : dominant kernel (optics+RORR), algorithmic FP64 ops / CUDA-event duration against
          the FP64 peak measured live by a DFMA micro-kernel (MEASURED_PEAKS.json has no FP64
          entry); the HBM view of the same kernel is in roofline["hbm"].
  cpu_baseline : oracle (C++ restatement of the reference algorithm, "port") on the host cores,
          bounded sample.  The Fortran reference cannot be built on this box.

--impl reference times that same CPU port (all host threads, OpenMP over columns) as the
reference arm: there is no compilable reference here (oracle/_ref does not exist).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

METRIC = "column RT solves/sec (IR+solar, RORR)"
UNIT = "columns/s"
NCOL_PER_GPU = 4096
TABLES = "nz=100 doubled (nz_r=202), Z=4, S-canon synthetic tables nw=100 G=16 K=5 (IR 64 bins, solar 64 bins)"
# BASELINE.json configs[1], [2], [4]; the column generators are synth.make_columns
CONFIGS = {
    "cfg2": dict(name="cfg2: AdiabatClimate OLR-vs-T_surf sweep 200-400 K", ncol=4096, kw={}, albedo=0.25, photon_scale=1.0),
    "cfg3": dict(name="cfg3: Early Mars CO2-H2O-N2 with CIA, pCO2 x T_surf grid, Mars gravity", ncol=4096,
                 kw=dict(pco2_range=(1.0e-2, 10 ** 0.5), T_surf_range=(200.0, 320.0), gravity=371.0), albedo=0.2,
                 photon_scale=0.4286),
    "cfg5": dict(name="cfg5: cfg2 generator with randomised (log-uniform) composition", ncol=8192,
                 kw=dict(randomize=True), albedo=0.25, photon_scale=1.0),
}


def workload_name(cfg, ncol):
    return f"{CONFIGS[cfg]['name']}, {ncol} columns per GPU, {TABLES}"


def host_cores():
    """Host threads the CPU arm may use: the affinity mask, NOT omp_get_max_threads() -- torchrun
    exports OMP_NUM_THREADS=1, which must not shrink the reference arm."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def make_workload(ncol, rank=0, cfg="cfg2"):
    from clima_b200 import synth
    c = CONFIGS[cfg]
    tb = synth.make_tables()
    pr = synth.default_params(tb, nzenith=4, albedo=c["albedo"])
    pr.photon_scale_factor = c["photon_scale"]
    kw = dict(c["kw"])
    if cfg == "cfg3":  # one slice of the 128 x 128 (pCO2, T_surf) grid per rank
        kw["seed"] = synth.SEED + rank
    else:
        kw["seed"] = synth.SEED + rank
    cols = synth.make_columns(ncol, nz=100, doubled=True, **kw)
    return tb, pr, cols


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons),
                "power_w_max": float(max(pw)), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
def cpu_port_rate(tb, pr, target_s=12.0, rank_seed=0):
    """Times the oracle's fast build (OpenMP over columns, all host threads) on a bounded sample of
    the same workload.  Returns (columns/s, cores, sample description)."""
    from clima_b200 import synth
    from oracle_lib import OracleRadtran, load
    load("fast")
    cores = host_cores()
    o = OracleRadtran(tb, pr, kind="fast")
    c1 = synth.make_columns(cores, nz=100, seed=synth.SEED + rank_seed)
    t = time.perf_counter()
    o.radiate_batch(c1, nthreads=cores)
    t1 = time.perf_counter() - t
    n = int(max(cores, min(64 * cores, cores * round(target_s / max(t1, 1e-3)))))
    cs = synth.make_columns(n, nz=100, seed=synth.SEED + rank_seed)
    t = time.perf_counter()
    o.radiate_batch(cs, nthreads=cores)
    dt = time.perf_counter() - t
    return n / dt, cores, f"{n} columns of the cfg2 sweep, {dt:.1f} s, OpenMP over columns"


def run_reference(args):
    """Reference arm: the reference's own CPU algorithm on the host cores.  The Fortran reference is
    not compilable here, so this is the C++ port (oracle/), all host threads, OpenMP over columns."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from clima_b200 import synth
    from oracle_lib import OracleRadtran, load
    tb, pr, _ = make_workload(1, 0, args.config)
    load("fast")
    cores = host_cores()
    if (os.cpu_count() or 1) > 1 and cores < 2:
        print("bench.py: reference arm confined to one host core by the affinity mask", file=sys.stderr)
    o = OracleRadtran(tb, pr, kind="fast")
    n = 16 * cores  # bounded sample per step (~2.5 s of the 16-thread port: long enough that thread start-up
    #                 and per-call allocation do not understate the CPU arm; r1r: 59 columns/s at 2 per thread,
    #                 100 columns/s at 40 per thread)
    _, _, cs = make_workload(n, 0, args.config)
    for _ in range(args.warmup):
        o.radiate_batch(cs, nthreads=cores)
    t = time.perf_counter()
    for _ in range(args.steps):
        o.radiate_batch(cs, nthreads=cores)
    dt = time.perf_counter() - t
    v = n * args.steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": workload_name(args.config, args.ncol if args.ncol > 0 else CONFIGS[args.config]["ncol"]),
                   "sample_columns_per_step": n},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{n} columns of the {args.config} workload per step, OpenMP over columns ({cores} threads); C++ "
                                   "restatement of the reference algorithm (the Fortran binary cannot be "
                                   "built on this box)"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
class DeviceLeg:
    """Inputs resident in HBM, results into one device block; the step adds the all-gather when world > 1."""

    def __init__(self, rt, cols, tb, dev, world):
        import torch
        self.rt, self.cols, self.world = rt, cols, world
        self.ncol = cols["T"].shape[-1]
        self.nz, self.ng = cols["nz"], cols["ng"]
        L = self.nz + 1
        self.names = ["T_surface", "T", "P", "densities", "dz"]
        self.host_in = {k: np.ascontiguousarray(cols[k].ravel(order="F")) for k in self.names}
        self.d_in = {k: torch.from_numpy(self.host_in[k]).to(dev) for k in self.names}
        self.nres = (5 * L + 2) * self.ncol
        self.d_res = torch.zeros(self.nres, dtype=torch.float64, device=dev)
        self.offs, o = {}, 0
        for k in ["fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total"]:
            self.offs[k] = o; o += L * self.ncol
        for k in ["ISR", "OLR"]:
            self.offs[k] = o; o += self.ncol
        self.outp = {k: self.d_res.data_ptr() + 8 * v for k, v in self.offs.items()}
        self.inp = {k: self.d_in[k].data_ptr() for k in self.names}
        self.d_all = torch.zeros(self.nres * world, dtype=torch.float64, device=dev) if world > 1 else None
        self.ext = torch.cuda.ExternalStream(rt.stream(), device=dev)

    def step(self):
        self.rt.radiate_batch_device(self.ncol, self.nz, self.ng, 0, self.inp, self.outp)
        if self.world > 1:
            # return every rank's flux profiles to every rank: ncclAllGather through the C ABI
            # (clima_rt_allgather_device), on the handle's stream -- the same path
            # clima_rt_radiate_batch_sharded takes; torch only launched the processes
            self.rt.allgather_device(self.d_res.data_ptr(), self.d_all.data_ptr(), self.nres)

    def timed(self, steps, warmup, barrier, dev, sampler=None):
        """(ms per step: CUDA events on the launching stream, max over ranks; stage ms; launches; wall s)"""
        import torch
        import torch.distributed as dist
        for _ in range(warmup):
            self.step()
        barrier()
        if sampler:
            sampler.start()
        n0 = self.rt.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stage = np.zeros(6)
        t_wall = time.perf_counter()
        e0.record(self.ext)
        for _ in range(steps):
            self.step()
            stage += np.array(self.rt.stage_ms())
        e1.record(self.ext)
        barrier()
        t_wall = time.perf_counter() - t_wall
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / steps, stage / steps, self.rt.launch_count() - n0, t_wall, clocks


def run_ours(args):
    import torch
    import torch.distributed as dist
    from clima_b200 import roofline
    from clima_b200.radtran import Radtran, measure_fp64_peak

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def init_comm(handle):
        """The handle's own NCCL communicator: rank 0 draws the id through the C ABI, torch.distributed
        (the launcher's plumbing) carries the 128 bytes to the other ranks."""
        if world == 1:
            return
        idt = torch.zeros(128, dtype=torch.uint8, device=dev)
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(Radtran.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        handle.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))

    cfg = args.config
    ncol = args.ncol if args.ncol > 0 else CONFIGS[cfg]["ncol"]
    tb, pr, cols = make_workload(ncol, rank, cfg)
    nz = cols["nz"]
    L = nz + 1
    rt = Radtran(tb, pr, device=local)
    init_comm(rt)
    leg = DeviceLeg(rt, cols, tb, dev, world)
    names, host_in, offs, nres, d_res = leg.names, leg.host_in, leg.offs, leg.nres, leg.d_res
    ms_per_step, stage, launches, t_wall, clocks = leg.timed(
        args.steps, args.warmup, barrier, dev, ClockSampler(local) if rank == 0 else None)
    value = world * ncol / (ms_per_step * 1e-3)
    rorr_stats = rt.rorr_stats()
    cut_stats = rt.rorr_cut_stats()

    # ---- e2e leg: host pointers through the C ABI, pinned buffers, H2D + D2H inside the timing
    from clima_b200._abi import alloc_out
    pin_in = {}
    for k in names:
        t = torch.empty(host_in[k].size, dtype=torch.float64).pin_memory()
        t.numpy()[:] = host_in[k]
        pin_in[k] = t
    pcols = dict(cols)
    for k in names:
        pcols[k] = pin_in[k].numpy()
    pin_out = torch.empty(nres, dtype=torch.float64).pin_memory()
    o_struct, o_arrs = alloc_out(nz, ncol, tb.nw_ir, tb.nw_sol, False)
    pout = pin_out.numpy()
    for k, v in offs.items():
        n = (L * ncol) if k not in ("ISR", "OLR") else ncol
        setattr(o_struct, k, pout[v:v + n].ctypes.data_as(type(o_struct.ISR)))
    e2e_steps = max(1, args.steps)
    rt.radiate_batch(pcols, out=(o_struct, o_arrs))
    barrier()
    t = time.perf_counter()
    for _ in range(e2e_steps):
        rt.radiate_batch(pcols, out=(o_struct, o_arrs))
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = world * ncol * e2e_steps / float(dt.item())
    h2d = int(sum(host_in[k].nbytes for k in names))
    d2h = int(nres * 8 + 4 * ncol)
    # sanity: host-path result equals the device-path result
    same = bool(np.array_equal(pout, d_res.cpu().numpy()))

    # ---- N > 1: the config north_star names for 8 GPUs (cfg5, 8192 randomised columns per GPU) and a
    # strong-scaling line (the N=1 batch split over the ranks), same timing rules, fewer steps
    extra = {}
    if world > 1 and not args.no_extra:
        xs, xw = max(2, min(args.steps, 3)), 2
        del leg
        tb5, pr5, cols5 = make_workload(CONFIGS["cfg5"]["ncol"], rank, "cfg5")
        rt5 = Radtran(tb5, pr5, device=local)
        init_comm(rt5)
        leg5 = DeviceLeg(rt5, cols5, tb5, dev, world)
        ms5, st5, _, _, _ = leg5.timed(xs, xw, barrier, dev)
        extra["cfg5_weak"] = {"workload": workload_name("cfg5", CONFIGS["cfg5"]["ncol"]), "scaling": "weak",
                              "value": world * CONFIGS["cfg5"]["ncol"] / (ms5 * 1e-3), "unit": UNIT,
                              "ms_per_step": ms5, "steps": xs, "warmup": xw,
                              "stage_ms_per_step": {"optics_rorr": st5[0], "kernels_total": st5[4]}}
        del leg5, rt5, cols5
        tot = CONFIGS[cfg]["ncol"]
        per = (tot + world - 1) // world
        lo, hi = min(rank * per, tot), min((rank + 1) * per, tot)
        _, _, call = make_workload(tot, 0, cfg)  # the SAME batch on every rank; this rank takes its block
        sl = {k: (np.asfortranarray(v[..., lo:hi]) if isinstance(v, np.ndarray) and v.shape[-1] == tot else v)
              for k, v in call.items()}
        if world * per == tot:
            legs = DeviceLeg(rt, sl, tb, dev, world)
            mss, sts, _, _, _ = legs.timed(xs, xw, barrier, dev)
            extra["strong"] = {"workload": workload_name(cfg, tot) + f" in total, {per} per GPU", "scaling": "strong",
                               "value": tot / (mss * 1e-3), "unit": UNIT, "ms_per_step": mss, "steps": xs, "warmup": xw,
                               "stage_ms_per_step": {"optics_rorr": sts[0], "kernels_total": sts[4], "prep": sts[5]}}

    # ---- N = 1: row N4, the batched AdiabatClimate%TOA_fluxes (make_profile on the device + the RT) on an
    # OLR-versus-T_surf sweep at fixed composition (the loop of examples/EarlyMars.ipynb cell 7), host inputs and
    # host results, after the headline legs so that it cannot disturb them
    if world == 1 and not args.no_extra and cfg == "cfg2":
        try:
            extra["adiabat_toa_fluxes"] = adiabat_leg(rt, ncol)
        except Exception as e:  # reported, never fatal for the headline line
            extra["adiabat_toa_fluxes"] = {"error": str(e)[:300]}

    if rank == 0:
        d = roofline.dims_of(tb, pr, cols)
        ops = roofline.stage_ops(d)
        byt = roofline.column_bytes(d)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        hbm_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        fp64_tf = measure_fp64_peak(local)
        k_ops = (ops["k_interp"] + ops["continua"] + ops["rorr"]) * ncol
        k_ms = stage[0]
        achieved = k_ops / (k_ms * 1e-3) / 1e12
        whole_tf = ops["total"] * ncol / (stage[4] * 1e-3) / 1e12
        # HBM view of the dominant kernel: reads the column inputs, writes tau/w0/g/tau_band of the
        # unique layers (those writes are an implementation intermediate, not algorithmic output)
        k_bytes = byt["in"] * ncol
        staircase = cut_stats["last_kernel"] in (0, 1)
        kernel_name = ("rorr_cut::rorr_cut_kernel<16> (sort-free staircase RORR, column chains)" if staircase
                       else "rorr_fast::optics_rorr_kernel<16> (sorting RORR)")
        dram_per_col = roofline.RORR_CUT_DRAM_BYTES_PER_COLUMN if staircase else roofline.RORR_FAST_DRAM_BYTES_PER_COLUMN
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(cfg, ncol), "columns_per_gpu": ncol, "nz_r": nz, "unique_layers": d.U,
                       "nw": d.nw, "ngauss": d.G, "n_kgases": d.K, "n_zenith": d.Z,
                       "l2": "per-step working set (tau/w0 of all columns, ~10 GB) is far larger than the 126 MB L2",
                       "sharding": f"columns block-partitioned over {world} rank(s), tables replicated"
                                   + (", ncclAllGather of flux profiles inside the step (clima_rt_allgather_device, "
                                      "the handle's own communicator and stream)" if world > 1 else "")},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "matches_device_path": same},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "stage_ms_per_step": {"optics_rorr": stage[0], "planck_ir": stage[1], "solar": stage[2],
                                  "integrate": stage[3], "kernels_total": stage[4], "prep": stage[5]},
            "wall_ms_per_step": 1e3 * t_wall / args.steps,
            "rorr": dict(zip(("units", "exact_kernel_units", "worklist_overflows"), rorr_stats),
                         passes_per_merge=cut_stats["warp_passes"] / max(cut_stats["warp_merges"], 1),
                         closed_form_share=cut_stats["closed_form_warp_merges"] / max(cut_stats["warp_merges"] + cut_stats["closed_form_warp_merges"], 1),
                         kernel=("staircase/slot chains", "staircase/column chains", "sorting", "exact sorting")[cut_stats["last_kernel"]]),
            "roofline": {
                "kernel": kernel_name, "bound": "fp64", "achieved": achieved,
                "peak": fp64_tf, "unit": "TFLOP/s", "frac": achieved / fp64_tf,
                "frac_of_issue_roof": achieved / (fp64_tf / 2.0),
                "peak_source": "DFMA micro-kernel measured live in this run (FMA = 2 flops); "
                               "issue roof = peak/2 because the work is not FMA-shaped",
                "algorithmic_ops_per_column": ops,
                # dram__bytes_read.sum + dram__bytes_write.sum of the fast RORR kernel from the
                # ncu --set full capture at 256 columns (profiles/r1p_ncu_full_summary.txt: 73.9 MB read,
                # 622.8 MB written = the tau/w0 lines of every unit), scaled to this launch's columns
                "traffic": dram_per_col * ncol,
                "traffic_over_algorithmic_bytes": dram_per_col / byt["in"],
                "traffic_source": roofline.RORR_DRAM_SOURCE["cut" if staircase else "fast"],
                "hbm": {"bound": "hbm", "achieved": k_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak,
                        "unit": "GB/s", "frac": k_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak,
                        "peak_source": hbm_src, "algorithmic_bytes_per_column": byt},
                "whole_step_tflops": whole_tf, "whole_step_frac": whole_tf / fp64_tf,
            },
        }
        if extra:
            line["extra"] = extra
        if world == 1 and not args.no_cpu:
            v, cores, sample = cpu_port_rate(tb, pr)
            line["cpu_baseline"] = {
                "value": v, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": sample + "; C++ restatement of the reference algorithm (the Fortran binary "
                                   "cannot be built on this box)"}
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _emit(line: dict):
    """The ONE JSON line goes to the real stdout; everything else any library prints on fd 1 during
    the run (e.g. NCCL's version banner) has been routed to stderr by main()."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def adiabat_leg(rt, ncol, nz=50, reps=2):
    """Row N4 on the bench handle: `ncol` surface temperatures 250..400 K over a 270-bar ocean inventory, 400 ppm
    CO2, 1 bar N2 (tests/test_adiabat.f90:31), the gases of the reference's species.yaml."""
    import torch
    from clima_b200.adiabat import AdiabatClimate
    with open(os.path.join(ROOT, "tests", "golden", "adiabat_species.json")) as f:
        sp = json.load(f)
    ac = AdiabatClimate(rt, sp, nz, 5.972e27, 6.371e8)
    T = np.linspace(250.0, 400.0, ncol)
    P = np.asfortranarray(np.repeat((np.array([270.0, 400e-6, 1.0, 1e-10, 1e-10, 1e-10, 1e-10]) * 1e6)[:, None], ncol, 1))
    ac.TOA_fluxes_batch(T[:64], P[:, :64])
    best, kern, stage = 1e30, 0.0, None
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        isr, olr = ac.TOA_fluxes_batch(T, P)
        dt = time.perf_counter() - t0
        if dt < best:
            best, kern, stage = dt, ac.kernel_ms(), rt.stage_ms()
    return {"workload": f"{ncol} columns, T_surf 250..400 K at fixed surface inventory, nz={nz} (nz_r={2 * nz + 2}), 7 gases",
            "value": ncol / best, "unit": UNIT, "ms_per_call": best * 1e3, "make_profile_kernel_ms": kern,
            "rt_kernels_ms": stage[4], "rorr_kernel": rt.rorr_cut_stats()["last_kernel"],
            "olr_min_max": [float(olr.min()), float(olr.max())],
            "timing": "host clock around the public call, host inputs and host results (an e2e number)"}


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ncol", type=int, default=0, help="columns per GPU (default: the config's)")
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS), help="workload (BASELINE.json configs)")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra legs (N>1: cfg5 / strong scaling; N=1: the batched AdiabatClimate TOA_fluxes)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
