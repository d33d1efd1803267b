#!/usr/bin/env python
"""bench.py -- the headline measurement (BASELINE.json): spectra/s of the HAT-P-11b-shape
transmission spectrum (config C2: 31 334 wavenumbers, 100 layers, 10^6 lines, 4 line-list
molecules, 2 CIA tables, line-by-line extinction) on N B200s, with the line-sum
G(line-samples)/s and the roofline of the dominant kernel beside it.

    python bench.py --gpus N --steps K --warmup W            (our arm; torchrun for N>1)
    python bench.py --impl reference ...                     (the unmodified reference binary
                                                               on the host cores, oracle/_ref)

A step = one do_transit (transit/src/transit.c:125-214): CIA interpolation + line-by-line
extinction of every layer + tau + modulation for one atmosphere.  `value` is timed with the
static inputs (line index, Voigt table, CIA tables) resident in HBM; `e2e` goes through the
host mirror of run_transit with HOST buffers: per step the atmosphere is copied host->device
and the spectrum device->host inside the timed region.  L2 is flushed between timed steps.
The oracle / reference are used here only as the timed CPU baseline, never as the product.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CLOCK_Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
           "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
           "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """SM clock / power / throttle reasons sampled DURING the timed region (B200_PROFILING.md): NVML
    polled every few ms from a thread (the timed region is tens of ms), nvidia-smi as the fallback."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int):
        self.index, self.rows, self.stop_flag, self.th, self.how = index, [], False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and vis.split(",")[0].isdigit() else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nv = pynvml
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.how = "nvml"
            self.th = threading.Thread(target=self._poll_nvml, daemon=True)
            self.th.start()
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={CLOCK_Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.how = "nvidia-smi"
            self.th = threading.Thread(target=self._read_smi, daemon=True)
            self.th.start()
        except OSError:
            self.how = None

    def _poll_nvml(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except AttributeError:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((sm, self.mx, pw, {n for bit, n in self.REASONS.items() if rs & bit}))
            except Exception:
                pass
            time.sleep(0.004)

    def _read_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                rs = {nm for k, nm in enumerate(names) if len(r) > 4 + k and r[4 + k].lower().startswith("active")}
                self.rows.append((float(r[1]), float(r[2]), float(r[3]), rs))
            except (ValueError, IndexError):
                continue

    def stop(self):
        if not self.how:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML / nvidia-smi"]}
        self.stop_flag = True
        if self.how == "nvidia-smi":
            self.proc.terminate()
        self.th.join(timeout=2)
        sm = [r[0] for r in self.rows]; mx = [r[1] for r in self.rows]; pw = [r[2] for r in self.rows]
        reasons = set().union(*[r[3] for r in self.rows]) if self.rows else set()
        # "under load" = samples drawing more than half of the highest power seen
        load = [s_ for s_, p_ in zip(sm, pw) if pw and p_ >= 0.5 * max(pw)] or sm
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "source": self.how,
                "reasons": sorted(reasons)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ------------------------------------------------------------------ reference arm (CPU)

REF_SO = os.path.join(ROOT, "oracle", "_ref", "libtransit_ref.so")
REF_SEC_PER_SPECTRUM = 27.0      # measured: one do_transit of full C2 on one 2025-class x86 core


def _ref_worker(cfg_path, conn):
    """One process = one instance of the UNMODIFIED reference (global state, non re-entrant):
    transit_init once (transit.c:25-74), then one do_transit (transit.c:125-214) per request."""
    import ctypes as C
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)                                   # the reference prints its banner per run
    L = C.CDLL(REF_SO)
    argv = [b"transit", b"-c", cfg_path.encode()]
    arr = (C.c_char_p * len(argv))(*argv)
    L.transit_init(len(argv), arr)
    L.get_no_samples.restype = C.c_int
    out = np.zeros(L.get_no_samples())
    conn.send(("ready", out.size))
    while True:
        msg = conn.recv()
        if msg == "stop":
            break
        t = time.perf_counter()
        L.do_transit(out.ctypes.data_as(C.POINTER(C.c_double)))
        conn.send((time.perf_counter() - t, float(out.sum())))


class ReferencePool:
    """`nproc` independent reference processes -- the only parallelism the reference supports
    (N independent processes over disjoint atmospheres)."""

    def __init__(self, cfg_paths):
        import multiprocessing as mp
        ctx = mp.get_context("fork")
        self.pipes, self.procs = [], []
        for cfg_path in cfg_paths:
            a, b = ctx.Pipe()
            p = ctx.Process(target=_ref_worker, args=(cfg_path, b), daemon=True)
            p.start()
            self.pipes.append(a); self.procs.append(p)
        for a in self.pipes:
            if not a.poll(600):
                raise RuntimeError("reference worker did not initialise (transit_init exits on bad input)")
            a.recv()

    def step(self):
        t = time.perf_counter()
        for a in self.pipes:
            a.send("go")
        res = [a.recv() for a in self.pipes]
        return time.perf_counter() - t, res

    def close(self):
        for a in self.pipes:
            a.send("stop")
        for p in self.procs:
            p.join(timeout=5)


def reference_inputs(workload, nbands, nproc, tag):
    """The workload as reference input files, split into `nbands` contiguous wavenumber bands (the
    reference's own way to spread one spectrum over processes); process i gets
    band i mod nbands.  Every per-spectrum cost of the reference (line sum, CIA, tau, modulation)
    is per wavenumber, so a band is a proportional sample of the spectrum."""
    from transit_b200 import workloads
    prob = workloads.make_problem(workload)
    d = os.path.join("/tmp", f"transit_b200_ref_{tag}_{os.getpid()}")
    edges = np.linspace(0, prob.nwn - 1, nbands + 1).astype(int)
    cfgs = []
    for i in range(nproc):
        b = i % nbands
        band = None if nbands == 1 else (int(edges[b]), int(edges[b + 1]))
        cfg, nlines = workloads.write_reference_inputs(prob, d, band=band, tag=f"w{i}")
        cfgs.append(cfg)
    return prob, cfgs, nlines


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    if not os.path.exists(REF_SO):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libtransit_ref.so not built"}))
        return 0
    cores = max(1, min(os.cpu_count() or 1, args.ref_procs or 32))
    # bounded sample: each process computes 1/nbands of the spectrum per step so that (K+W) steps
    # end within a few minutes
    nbands = int(np.ceil((args.steps + args.warmup) * REF_SEC_PER_SPECTRUM / args.ref_budget_s))
    nbands = max(1, min(nbands, 64))
    prob, cfgs, nlines = reference_inputs(args.workload, nbands, cores, "arm")
    pool = ReferencePool(cfgs)
    for _ in range(args.warmup):
        pool.step()
    t_steps = [pool.step()[0] for _ in range(args.steps)]
    pool.close()
    sec = float(np.sum(t_steps))
    value = (cores / nbands) * args.steps / sec            # whole spectra per second
    sample = (f"{cores} independent reference processes (libtransit_ref.so: transit_init once, do_transit per step), "
              f"each computing one of {nbands} contiguous wavenumber bands of the {args.workload} spectrum per step "
              f"(process i -> band i mod {nbands}; full {nlines}-line list on disk); value = processes / bands / step time")
    line = {"impl": "reference", "metric": "spectra/s", "value": value, "unit": "spectra/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(prob, args.workload, args.gpus),
            "cpu_baseline": {"value": value, "unit": "spectra/s", "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "spectra/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def workload_config(prob, name, world=1):
    return {"sharding": f"atmospheres x{world} (one spectrum per rank per step, NCCL all-gather of the spectra)"
            if world > 1 else "none",
            "workload": f"{name}: HAT-P-11b-shape transmission 0.3-5 um, line-by-line" if name == "C2" else name,
            "nwn": int(prob.nwn), "wnosamp": int(prob.wnosamp), "layers": int(prob.nrad), "lines": int(len(prob.wl)),
            "isotopologues": int(prob.niso), "line_molecules": len(prob.cfg["molecules"]), "cia_tables": len(prob.cia),
            "geometry": "eclipse" if prob.eclipse else "transit", "voigt_table": "60x60", "l2": "flushed between steps"}


# ------------------------------------------------------------------ our arm (B200)

def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md, 6.65 TB/s)"


def ncu_traffic(workload, kernel):
    """DRAM bytes per launch of `kernel` from an ncu --set full capture of THIS workload (profiles/ncu_traffic.json,
    keyed by workload), else None: a figure captured on another configuration is not reported."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(workload, {}).get(kernel, {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        return None


def rooflines(workload, k_ms, nsamp, nline_layers, nwn, nlayers, mb, hbm_peak, peak_src):
    """SURVEY.md 8d work models, one entry per timed kernel group, each against the peak measured LIVE in
    this run (tb_microbench: DFMA, exp, copy) or MEASURED_PEAKS.json (HBM).
      line sum:  2 flop per line-sample + (12 + 2 c_exp) flop per (line, layer); c_exp = DFMA flop/s over exp/s
      strength:  the (12 + 2 c_exp) N_line.layer part;  accumulate: 2 N_samples (+ 4 B per sample, 8 B per output)
      tau:       24 B per (layer, wavenumber) against HBM"""
    dfma = mb["dfma_tflops"]
    c_exp = dfma * 1e12 / (mb["dexp_gops"] * 1e9)               # flop-equivalents of one exp
    pair_flop = (12.0 + 2.0 * c_exp) * nline_layers
    samp_flop = 2.0 * nsamp
    src_f = "tb_microbench DFMA, measured in this run"
    out = {}

    def fp64(name, kernel, flop, note=None):
        if name in k_ms and k_ms[name] > 0:
            a = flop / (k_ms[name] * 1e-3) / 1e12
            out[name] = {"kernel": kernel, "bound": "fp64", "achieved": a, "peak": dfma, "unit": "TFLOP/s", "frac": a / dfma,
                         "ms": k_ms[name], "traffic": ncu_traffic(workload, name), "peak_source": src_f,
                         "algorithmic_flop": flop, "c_exp_flop": c_exp}
            if note:
                out[name]["note"] = note

    def hbm(name, kernel, nbytes):
        if name in k_ms and k_ms[name] > 0:
            a = nbytes / (k_ms[name] * 1e-3) / 1e9
            out[name] = {"kernel": kernel, "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak,
                         "ms": k_ms[name], "traffic": ncu_traffic(workload, name), "peak_source": peak_src,
                         "algorithmic_bytes": nbytes}

    fp64("molext", "line sum, all kernels (k_kmax + k_linesum_fused_w + k_strength + k_accumulate)", samp_flop + pair_flop)
    fp64("strength", "k_strength", pair_flop * k_ms.get("_unfused_frac", 1.0),
         "reference work of the layers this kernel evaluates (2 exp + 12 flop per (line, layer)); the log-domain "
         "pretest executes far fewer instructions, so this is throughput in reference work, not pipe use")
    fp64("fused", "k_linesum_fused_w (strength + gather of the narrow layers in one kernel, a warp per 32 bins x 4 states)",
         pair_flop * (1.0 - k_ms.get("_unfused_frac", 1.0)) + 2.0 * k_ms.get("_fused_samples", 0.0))
    fp64("accumulate", "k_accumulate (wide + narrow instances)", 2.0 * k_ms.get("_unfused_samples", nsamp))
    if "accumulate" in out:
        b = 4.0 * k_ms.get("_unfused_samples", nsamp) + 8.0 * nwn * nlayers * k_ms.get("_unfused_frac", 1.0)
        out["accumulate"]["hbm_gbs"] = b / (k_ms["accumulate"] * 1e-3) / 1e9
        out["accumulate"]["hbm_frac"] = out["accumulate"]["hbm_gbs"] / hbm_peak
    hbm("tau", "k_tau_transit", 24.0 * nwn * nlayers)
    return out


def ms_stats(ms_steps):
    a = np.asarray(ms_steps, float)
    return {"min": float(a.min()), "median": float(np.median(a)), "max": float(a.max()), "mean": float(a.mean())}


HOST_SO = os.path.join(ROOT, "host", "libtransit_host.so")


class HostLibrary:
    """The reference's library face as a user (BART) binds it -- transit_init(argv) once, then
    run_transit(in = [T | q_mol0 | q_mol1 ...], out[nwn]) per atmosphere (transit.i:11-19, transit.c:118-122) --
    on host/libtransit_host.so: cfg, atmosphere file, TLI and CS files are read from disk, reloadatm / radpress /
    makeradsample run per call, and the spectrum comes back in the caller's HOST buffer."""

    def __init__(self, prob, device, tag):
        import ctypes as C
        from transit_b200 import workloads
        if not os.path.exists(HOST_SO):
            raise SystemExit(f"{HOST_SO} is missing: run __graft_entry__.build()")
        d = os.path.join("/tmp", f"transit_b200_host_{tag}_{os.getpid()}")
        cfg, _ = workloads.write_reference_inputs(prob, d)
        k = prob.nrad // 3
        with open(cfg, "a") as f:        # run_transit needs the hydrostatic reference level (readatm.c:760-767)
            f.write(f"refpress {prob.p[k]:.6e}\nrefradius {prob.rad[k]:.6f}\ngsurf 1500.0\n")
        os.environ["TRANSIT_B200_DEVICE"] = str(device)
        self.C, self.L = C, C.CDLL(HOST_SO)
        argv = [b"transit", b"-c", cfg.encode()]
        arr = (C.c_char_p * len(argv))(*argv)
        so = os.dup(1)
        os.dup2(os.open(os.devnull, os.O_WRONLY), 1)     # keep stdout to the JSON line
        try:
            self.L.transit_init(len(argv), arr)
        finally:
            os.dup2(so, 1)
        self.L.get_no_samples.restype = C.c_int
        self.n = self.L.get_no_samples()
        self.inp = np.ascontiguousarray(np.concatenate([prob.t] + [prob.mol_q[j] for j in range(prob.nmol)]))
        self.out = np.zeros(self.n)
        self.P = C.POINTER(C.c_double)

    def run_transit(self):
        self.L.run_transit(self.inp.ctypes.data_as(self.P), self.inp.size, self.out.ctypes.data_as(self.P), self.n)
        return self.out

    def close(self):
        self.L.free_memory()


def cpu_baseline_leg(args):
    """The unmodified reference on a bounded sample of the same workload: `nb` processes, one
    wavenumber band each (together one whole spectrum)."""
    if not os.path.exists(REF_SO):
        return {"value": None, "unit": "spectra/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    nb = max(1, min(os.cpu_count() or 1, int(np.ceil(REF_SEC_PER_SPECTRUM / args.cpu_baseline_s))))
    prob, cfgs, nlines = reference_inputs(args.workload, nb, nb, "cpu")
    pool = ReferencePool(cfgs)
    sec, _ = pool.step()
    pool.close()
    return {"value": 1.0 / sec, "unit": "spectra/s", "cores": nb, "kind": "reference",
            "sample": f"one whole {args.workload} spectrum by the unmodified reference (oracle/_ref/libtransit_ref.so, "
                      f"do_transit), split into {nb} wavenumber bands on {nb} processes/cores: {sec:.1f} s"}


def bands_leg(args, prob, ref_out, one_gpu_ms, local_rank, rank, world, stream, flush):
    """Strong scaling of ONE spectrum: the wavenumber range of the same C2 problem cut into `world` bands, one per
    rank (north_star's first sharding).  Exchange inside the library: one ncclAllGather of the band slices (the exact
    strength maxima and the carried Doppler index are found by every rank itself); every rank ends each step with the whole spectrum.  Timed like the headline
    (CUDA events on the launching stream, L2 flushed between steps, max over ranks); no per-step barrier -- the
    collectives are what couples the ranks."""
    import torch
    import torch.distributed as dist
    from transit_b200 import sharding
    tb = sharding.BandShardedTransit(local_rank, rank, world)
    tb.ctx.set_stream(stream.cuda_stream)
    tb.transit_init(prob)
    first_cut = list(tb.bands)
    tb.rebalance(rounds=2)                    # measured rebalancing: cost per line group differs from band to band
    for _ in range(max(3, args.warmup)):
        g = tb.do_transit_gathered()
    same = bool(np.array_equal(g.cpu().numpy(), ref_out)) if rank == 0 else True
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms, kms = [], {}
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        ev0.record(stream)
        tb.do_transit_gathered()
        ev1.record(stream)
        torch.cuda.synchronize()
        ms.append(ev0.elapsed_time(ev1))
        for nm in ("kmax", "strength", "fused", "accumulate", "molext", "cia", "tau", "modulation", "gather"):
            try:
                kms.setdefault(nm, []).append(tb.ctx.kernel_ms(nm))
            except Exception:
                pass
    torch.cuda.synchronize(); dist.barrier()
    t = torch.tensor([float(np.sum(ms)), 0.0 if same else 1.0], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    names = ("molext", "kmax", "strength", "fused", "accumulate", "tau", "cia", "gather")
    mine = torch.tensor([float(np.median(kms.get(nm, [0.0]))) for nm in names] + [float(np.median(ms))],
                        dtype=torch.float64, device="cuda")
    allr = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allr, mine)
    per_rank = {nm: [round(float(a[i]), 4) for a in allr] for i, nm in enumerate(names + ("step",))}
    tb.free_memory()
    step_ms = float(t[0]) / args.steps
    return {"what": f"one {args.workload} spectrum per step, wavenumber range in {world} bands (strong scaling)",
            "ms_per_step": step_ms, "spectra_per_s": 1e3 / step_ms, "one_gpu_ms_per_step": one_gpu_ms,
            "speedup_vs_one_gpu": one_gpu_ms / step_ms, "efficiency": one_gpu_ms / step_ms / world,
            "bit_identical_to_single_gpu": bool(float(t[1]) == 0.0), "bands": tb.bands,
            "bands_before_rebalancing": first_cut,
            "exchange": "in-library NCCL: 1 x ncclAllGather per spectrum on the context stream (the strength maxima and the carried Doppler index are found by every rank itself: no all-reduce)",
            "rank0_kernel_ms": {k: float(np.mean(v)) for k, v in kms.items()}, "ms_step_stats": ms_stats(ms),
            "per_rank_median_ms": per_rank}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from transit_b200 import api, workloads, sharding

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL_DEBUG is left as the caller set it (the driver reads NCCL's own init lines to check the
        # communicator); the JSON line is printed LAST and is the only line of stdout that parses as JSON
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    prob = workloads.make_problem(args.workload, scale=args.scale)
    # N > 1: batched atmospheres sharded over the ranks (BART's use of the library: many T-P/abundance
    # profiles through one static problem) -- rank r computes the spectra of its own atmosphere,
    # no data-path collective; the spectra are gathered with NCCL.  Per-GPU work is fixed: weak scaling.
    bands_mode = args.sharding == "bands"
    # --sharding bands: ONE spectrum per step, its wavenumber range split over the ranks (each rank
    # evaluates the lines that reach its band, two exact all-reduce MAX complete the global
    # threshold / carried Doppler index, the spectrum slices are all-gathered): strong scaling.
    my_prob = prob if (rank == 0 or bands_mode) else workloads.perturbed(prob, seed=1000 + rank)
    t0 = time.perf_counter()
    tr = (sharding.BandShardedTransit if bands_mode else sharding.ShardedTransit)(local_rank, rank, world)
    stream = sharding.private_stream()         # kernels, copies and NCCL collectives share this stream
    tr.ctx.set_stream(stream.cuda_stream)              # kernels, copies and the NCCL gather share one stream
    tr.transit_init(my_prob)
    tr.ctx.sync()
    init_s = time.perf_counter() - t0
    ctx = tr.ctx
    if args.fuse_rmax:
        ctx.set_fused_reach(args.fuse_rmax, True)
    if args.narrow_r >= 0:
        ctx.set_narrow_reach(args.narrow_r)
    if args.fuse_variant >= 0:
        ctx.set_fused_variant(args.fuse_variant)
    out = np.zeros(prob.nwn)
    gathered = [None]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        if bands_mode:
            gathered[0] = tr.do_transit_gathered()
            return
        tr.do_transit(out)
        if world > 1:        # natm = world: one atmosphere per rank per step
            gathered[0] = sharding.gather_spectra(tr.device_spectrum_tensor().view(1, -1), world, rank, world)

    for _ in range(args.warmup):
        step()
    if bands_mode:
        out[:] = gathered[0].cpu().numpy()
    ref_out = out.copy()
    verified = None
    if bands_mode and args.verify and rank == 0:
        one = api.Transit(local_rank)
        one.transit_init(prob)
        verified = bool(np.array_equal(one.do_transit(None), ref_out))
        one.free_memory()
        assert verified, "band-sharded spectrum differs from the single-GPU spectrum"

    # ---- device-timed steps (static inputs resident; L2 flushed between steps)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    l0 = ctx.launch_count()
    ms_steps, k_ms = [], {}
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(args.steps):
        flush.zero_()
        barrier()
        ev0.record(stream)
        step()
        ev1.record(stream)
        torch.cuda.synchronize()
        ms_steps.append(ev0.elapsed_time(ev1))
        for nm in ("kmax", "strength", "fused", "accumulate", "molext", "cia", "tau", "modulation"):
            try:
                k_ms.setdefault(nm, []).append(ctx.kernel_ms(nm))
            except api.TransitError:
                pass
    barrier()
    launches = ctx.launch_count() - l0
    nsamp = int(tr.sample_counts().sum())                 # this rank's executed profile updates per step
    clocks = sampler.stop()
    if bands_mode:
        out[:] = gathered[0].cpu().numpy()
    assert np.array_equal(out, ref_out), "spectrum changed between steps: the path is not deterministic"
    if world > 1 and not bands_mode:
        assert np.array_equal(gathered[0][rank].cpu().numpy(), out), "gathered slice differs from the local spectrum"

    # ---- N > 1: the other sharding of north_star beside the headline -- one spectrum over N wavenumber bands
    bands = None
    if world > 1 and not bands_mode and not args.no_bands_leg:
        ref0 = [ref_out if rank == 0 else None]           # rank 0 ran the unperturbed problem: the single-GPU spectrum
        bands = bands_leg(args, prob, ref0[0], float(np.mean(ms_steps)), local_rank, rank, world, stream, flush)

    # ---- end to end with HOST buffers, every step: atmosphere in (H2D inside the call), spectrum out (D2H)
    e2e_api = args.e2e_api if not bands_mode else "mirror"
    if e2e_api == "host":
        # the reference's plugin call for this path: run_transit of the C host (reloadatm + radpress +
        # makeradsample on the host, then tb_set_atmosphere + tb_spectrum), one library instance per rank
        host = HostLibrary(my_prob, local_rank, f"r{rank}")
        assert host.n == prob.nwn
        for _ in range(max(2, args.warmup)):
            host.run_transit()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            host.run_transit()
        barrier()
        e2e_s = time.perf_counter() - t0
        assert np.all(np.isfinite(host.out)) and host.out.max() > 0
        host.close()
        e2e_api = "host/libtransit_host.so: run_transit(in=[T|q...], out[nwn]) per step, one instance per rank"
    else:
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            tr.set_atmosphere(my_prob)
            step()
        barrier()
        e2e_s = time.perf_counter() - t0
        e2e_api = "python mirror: tb_set_atmosphere + tb_spectrum (+ NCCL gather)"
    h2d = 8 * prob.nrad * (3 + prob.nmol + prob.niso)
    d2h = 8 * prob.nwn

    dev_ms = float(np.sum(ms_steps))
    t = torch.tensor([dev_ms, e2e_s, float(nsamp), float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms, e2e_s, nsamp, launches = float(tmax[0]), float(tmax[1]), int(tsum[2]), int(tsum[3])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    nspec = 1 if bands_mode else world           # spectra finished per step by the whole job
    hbm_peak, peak_src = measured_peaks()
    kms = {k: float(np.mean(v)) for k, v in k_ms.items() if np.isfinite(np.mean(v))}
    mol_ms = kms["molext"]
    ns, nwn = prob.nrad, tr.band_nwn()
    my_samp = int(tr.sample_counts().sum())
    # peaks measured in THIS run on this GPU (milliseconds of device time, outside the timed region)
    mb = ctx.microbench()
    frac = getattr(tr, "line_fraction", 1.0)               # band shard: the part of the line list this rank evaluates
    model = dict(kms)
    fs = ctx.fused_stats() if hasattr(ctx, "fused_stats") else None
    if fs:
        model["_unfused_frac"] = 1.0 - fs["fused_states"] / max(1, ns)
        model["_fused_samples"] = float(fs["fused_samples"])
        model["_unfused_samples"] = float(my_samp - fs["fused_samples"])
    rl = rooflines(args.workload, model, my_samp, len(prob.wl) * ns * frac, nwn, ns, mb, hbm_peak, peak_src)
    # the line's `roofline` is the kernel with the most device time in a step
    single = {k: v for k, v in rl.items() if k != "molext"}
    dominant = max(single, key=lambda k: single[k]["ms"]) if single else "molext"
    line = {
        "metric": "spectra/s", "value": nspec * args.steps / (dev_ms * 1e-3), "unit": "spectra/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if bands_mode else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": dict(workload_config(prob, args.workload, world),
                       **({"sharding": f"wavenumber bands x{world} (one spectrum per step)", "bands": tr.bands}
                          if bands_mode else {})),
        "e2e": {"value": nspec * args.steps / e2e_s, "unit": "spectra/s", "h2d_bytes_per_step": h2d * world,
                "d2h_bytes_per_step": d2h * nspec, "api": e2e_api},
        "gpu_launches": launches,
        "ms_step_stats": ms_stats(ms_steps),
        "line_sum": {"gsamples_per_s": nsamp / (mol_ms * 1e-3) / 1e9, "samples_per_step": nsamp,
                     "line_layers_per_s": len(prob.wl) * prob.nrad / (mol_ms * 1e-3), "molext_ms": mol_ms},
        "kernel_ms": kms,
        "roofline": dict(rl[dominant], dominant_of="kernel_ms (largest device time of a step)"),
        "roofline_line_sum": rl.get("molext"),
        "rooflines": rl,
        "microbench": mb,
        "clocks": clocks, "init_s": init_s,
    }
    if verified is not None:
        line["bit_identical_to_single_gpu"] = verified
    if bands is not None:
        line["bands"] = bands
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_leg(args)
    if world > 1:
        dist.destroy_process_group()
    sys.stdout.flush()
    print(json.dumps(line), flush=True)                   # last line of stdout
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--natm", type=int, default=1024, help="C5: atmospheres in the retrieval batch")
    ap.add_argument("--batch", type=int, default=32, help="C5: atmospheres per rank per step")
    ap.add_argument("--ctx-per-gpu", type=int, default=4, help="C5: contexts (streams + host threads) sharing one GPU")
    ap.add_argument("--c5-api", default="device", choices=["device", "batch", "single"],
                    help="C5: tb_spectrum_batch per context per step, or one tb_set_atmosphere + tb_spectrum per atmosphere")
    ap.add_argument("--block-layers", type=int, default=2, help="C3: layers (x 26 temperatures) per step")
    ap.add_argument("--verify", action="store_true",
                    help="--sharding bands: rank 0 also computes the spectrum unsharded and requires bit-identity")
    ap.add_argument("--sharding", default="atmospheres", choices=["atmospheres", "bands"],
                    help="N>1: one atmosphere per rank (weak scaling, default) or one spectrum split into "
                         "wavenumber bands (strong scaling)")
    ap.add_argument("--scale", type=float, default=1.0, help="line-count multiplier (debugging only; 1.0 = the named config)")
    ap.add_argument("--fuse-variant", type=int, default=-1, help="tuning: 0 block-wide fused kernel, 1 warp-autonomous")
    ap.add_argument("--narrow-r", type=int, default=-1, help="tuning: widest reach (coarse bins) of the one-thread-per-bin gather")
    ap.add_argument("--fuse-rmax", type=int, default=0, help="tuning: widest reach (coarse bins) of a fused state")
    ap.add_argument("--no-bands-leg", action="store_true", help="N>1: skip the wavenumber-band strong-scaling leg")
    ap.add_argument("--e2e-api", default="host", choices=["host", "mirror"],
                    help="e2e leg: run_transit of the C host (host/libtransit_host.so, the reference's plugin call, "
                         "a16 host work included) or the Python mirror (tb_set_atmosphere + tb_spectrum)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-baseline-s", type=float, default=14.0)
    ap.add_argument("--ref-procs", type=int, default=0)
    ap.add_argument("--ref-budget-s", type=float, default=150.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "b200" and args.workload in ("C3", "C4", "C5"):
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_configs
        return bench_configs.run(sys.modules[__name__], args)
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
