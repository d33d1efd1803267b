"""bench.py --workload C3 | C4 | C5: the other BASELINE.json configurations.

The headline (and the only line the round-end driver reads) is C2 in bench.py itself; these are
run by hand, their JSON lines are kept under profiles/.  Same timing rules: CUDA events on the
launching stream, barrier + synchronize on both sides, max over ranks, L2 flushed between steps.

  C3  opacity-grid build (calcopacity, opacity.c:281-427): an ExoMol-scale synthetic line list
      (10^8 lines at --scale 1) x 26 temperatures x 100 layers.  A step = the grid rows of
      --block-layers layers (x 26 T = states of the line sum, permol), delivered to HOST memory in
      file order.  N > 1: wavenumber bands (tb_set_shard), strong scaling.
  C4  high-resolution eclipse: 10^6 wavenumber samples, extinction interpolated from a resident
      grid (interpolmolext), CIA, eclipse tau, 5-angle intensity, flux.  A step = one do_transit.
  C5  BART-style retrieval: --natm perturbed T-P/abundance profiles through grid interpolation +
      CIA + tau + modulation on the C2 grid (built here by the line-by-line path).  A step = one
      batch of --batch atmospheres per rank; N > 1 shards the atmospheres (weak scaling).
"""
from __future__ import annotations

import json
import os
import time

import numpy as np

GRID_TEMPS = np.arange(500.0, 3000.1, 100.0)          # tlow 500 thigh 3000 tempdelt 100: 26 T


def _ziso(prob, temps):
    from transit_b200 import workloads
    iso_mol, iso_k = [], []
    for db in prob.dbs:
        for kk in range(len(db.iso_names)):
            iso_mol.append(db.molecule); iso_k.append(kk)
    return np.stack([workloads.partition_function(iso_mol[i], iso_k[i], temps) for i in range(prob.niso)])


def _mean_mass(prob):
    return (prob.mol_q * prob.mol_mass[:, None]).sum(0)


def _dist(bench, args):
    import torch
    import torch.distributed as dist
    rank, local_rank, world = bench.dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    return torch, dist, rank, local_rank, world


class _Timer:
    """K timed steps: L2 flush, barrier + synchronize, CUDA events on the launching stream."""

    def __init__(self, torch, dist, world, stream):
        self.torch, self.dist, self.world, self.stream = torch, dist, world, stream
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def run(self, step, steps, warmup):
        for i in range(warmup):
            step(i)
        ms = []
        for i in range(steps):
            self.flush.zero_()
            self.barrier()
            self.e0.record(self.stream)
            step(warmup + i)
            self.e1.record(self.stream)
            self.torch.cuda.synchronize()
            ms.append(self.e0.elapsed_time(self.e1))
        self.barrier()
        return float(np.sum(ms))

    def reduce(self, vals):
        t = self.torch.tensor(vals, dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            mx = t.clone(); self.dist.all_reduce(mx, op=self.dist.ReduceOp.MAX)
            sm = t.clone(); self.dist.all_reduce(sm, op=self.dist.ReduceOp.SUM)
            return mx.tolist(), sm.tolist()
        return t.tolist(), t.tolist()


def run_c5(bench, args):
    """1 024 atmospheres (default) through opacity interpolation + CIA + tau + modulation.

    One spectrum from the grid is a chain of small latency-bound kernels (CIA splines, tau, modulation:
    ~0.4 ms, a few blocks each); --ctx-per-gpu contexts, each on its own CUDA stream and driven by its own
    host thread, share a GPU so that those chains overlap (the library has no global state; every
    context keeps its own resident copy of the 2.6 GB grid)."""
    from concurrent.futures import ThreadPoolExecutor
    from transit_b200 import api, workloads, sharding
    torch, dist, rank, local_rank, world = _dist(bench, args)
    prob = workloads.make_problem("C2", scale=args.scale)
    K = 1 if getattr(args, "c5_api", "single") == "device" else max(1, args.ctx_per_gpu)
    stream = sharding.private_stream()         # timing events and the L2 flush live here
    t0 = time.perf_counter()
    first = sharding.ShardedTransit(local_rank, rank, world)
    first.ctx.set_stream(stream.cuda_stream)
    first.transit_init(prob)                             # line-by-line set-up, used once: to build the grid
    grid = first.ctx.opacity_build(GRID_TEMPS, _ziso(prob, GRID_TEMPS), prob.mol_q, _mean_mass(prob), 0, prob.ethresh)
    first.ctx.sync()
    build_s = time.perf_counter() - t0
    grid_bytes = grid.nbytes
    molID = first.ctx.line_molIDs()
    trs, wstreams = [first], [stream]
    for k in range(1, K):
        ws = torch.cuda.Stream()
        t = sharding.ShardedTransit(local_rank, rank, world)
        t.ctx.set_stream(ws.cuda_stream)
        t.transit_init(prob, opacity=dict(molID=molID, temp=GRID_TEMPS, press=prob.p * prob.p_fct, grid=grid))
        trs.append(t); wstreams.append(ws)
    if getattr(args, "c5_api", "single") != "device":
        del grid                                         # the resident copies serve every atmosphere
    mine = list(sharding.atmosphere_shard(args.natm, rank, world))
    batch = min(args.batch, max(1, len(mine)))
    atms = [workloads.perturbed(prob, seed=5000 + i, dT=150.0, dex=1.0) for i in mine[:batch * 2]]
    out = np.zeros((batch, prob.nwn))
    kms = {}
    pool = ThreadPoolExecutor(K)

    # --c5-api batch: each context takes a contiguous share of the step's atmospheres through ONE
    # tb_spectrum_batch call (no host synchronisation between its spectra); "single": one
    # tb_set_atmosphere + tb_spectrum per atmosphere, as BART calls run_transit
    use_batch = getattr(args, "c5_api", "single") == "batch"
    if getattr(args, "c5_api", "single") == "device":
        return run_c5_device(bench, args, torch, dist, rank, local_rank, world, prob, first, stream, build_s, grid_bytes, mine, grid, molID)
    share = [list(range(k * batch // K, (k + 1) * batch // K)) for k in range(K)]
    stacked = {}

    def stack(i, k):
        key = (i * batch) % len(atms), k
        if key not in stacked:
            sel = [atms[(i * batch + b) % len(atms)] for b in share[k]]
            stacked[key] = tuple(np.stack([np.asarray(getattr(x, f), np.float64) for x in sel])
                                 for f in ("rad", "p", "t", "mol_d", "iso_z"))
        return stacked[key]

    def work(k, i):
        tr = trs[k]
        if use_batch:
            if share[k]:
                rad, p, t, d, z = stack(i, k)               # host arrays; H2D inside the call, D2H of the spectra at its end
                tr.ctx.spectrum_batch(rad, prob.rad_fct, p, prob.p_fct, t, prob.t_fct, d, z, api.GEOM_TRANSIT,
                                      prob.toomuch, prob.ethresh, prob.starrad, prob.transp, None,
                                      out[share[k][0]:share[k][-1] + 1])
            return
        for b in range(k, batch, K):
            a = atms[(i * batch + b) % len(atms)]
            tr.set_atmosphere(a)                          # H2D: T, p, radii, densities, partition functions
            tr.do_transit(out[b])                         # D2H: the spectrum

    def step(i):
        if K > 1:                                         # worker streams start after, and end before, the timed events
            es = torch.cuda.Event(); es.record(stream)
            for ws in wstreams[1:]:
                ws.wait_event(es)
        list(pool.map(lambda k: work(k, i), range(K)))
        for ws in wstreams[1:]:
            ee = torch.cuda.Event(); ee.record(ws); stream.wait_event(ee)
        for nm in ("interp", "cia", "tau", "modulation"):
            try:
                kms.setdefault(nm, []).append(first.ctx.kernel_ms(nm))
            except api.TransitError:
                pass

    tm = _Timer(torch, dist, world, stream)
    sampler = bench.ClockSampler(local_rank)
    sampler.start()
    l0 = sum(t.ctx.launch_count() for t in trs)
    dev_ms = tm.run(step, args.steps, args.warmup)
    launches = sum(t.ctx.launch_count() for t in trs) - l0
    clocks = sampler.stop()
    assert np.isfinite(out).all() and (out > 0).all()
    # the same batch through ONE context must give the same spectra bit for bit
    chk = np.zeros(prob.nwn)
    first.set_atmosphere(atms[((args.warmup + args.steps - 1) * batch + batch - 1) % len(atms)])
    first.do_transit(chk)
    assert np.array_equal(chk, out[batch - 1]), "concurrent contexts changed a spectrum"
    tm.barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    tm.barrier()
    e2e_s = time.perf_counter() - t0
    (dev_ms, e2e_s), (_, _) = tm.reduce([dev_ms, e2e_s])
    if rank == 0:
        hbm_peak, src = bench.measured_peaks()
        nspec = world * batch * args.steps
        interp_ms = float(np.mean(kms["interp"]))
        nm = len(prob.cfg["molecules"])
        ibytes = (2 * nm * 8 + 8.0) * prob.nrad * prob.nwn
        print(json.dumps({
            "metric": "spectra/s", "value": nspec / (dev_ms * 1e-3), "unit": "spectra/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C5: BART-style retrieval, perturbed T-P/abundance profiles through opacity-grid "
                                   "interpolation + CIA + tau + modulation (C2 grid)", "natm": args.natm,
                       "batch_per_rank_per_step": batch, "contexts_per_gpu": K,
                       "api": "tb_spectrum_batch (one call per context per step)" if use_batch else "tb_set_atmosphere + tb_spectrum per atmosphere",
                       "sharding": f"atmospheres x{world}", "nwn": int(prob.nwn),
                       "layers": int(prob.nrad), "grid": [int(prob.nrad), len(GRID_TEMPS), nm, int(prob.nwn)],
                       "grid_gb": grid_bytes / 1e9, "grid_build_s": build_s, "l2": "flushed between steps"},
            "e2e": {"value": nspec / e2e_s, "unit": "spectra/s",
                    "h2d_bytes_per_step": 8 * prob.nrad * (3 + prob.nmol + prob.niso) * batch * world,
                    "d2h_bytes_per_step": 8 * prob.nwn * batch * world},
            "gpu_launches": launches * world,
            "kernel_ms": {k: float(np.mean(v)) for k, v in kms.items()},
            "roofline": {"kernel": "k_grid_interp", "bound": "hbm", "achieved": ibytes / (interp_ms * 1e-3) / 1e9,
                         "peak": hbm_peak, "unit": "GB/s", "frac": ibytes / (interp_ms * 1e-3) / 1e9 / hbm_peak,
                         "traffic": None, "peak_source": src, "ms": interp_ms, "algorithmic_bytes": ibytes,
                         "note": "kernel timed while the other contexts' kernels share the GPU"},
            "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()
    return 0


def retrieval_inputs(prob, idx, dT=150.0, dex=1.0):
    """run_transit inputs [T | q_mol0 | q_mol1 ...] (readatm.c:735-744) of seeded perturbed profiles; the bulk gas takes
    up the slack so that the abundances add up to one."""
    from transit_b200 import workloads
    rows = []
    for i in idx:
        a = workloads.perturbed(prob, seed=5000 + i, dT=dT, dex=dex)
        q = a.mol_q.copy()
        q[0] = 1.0 - q[1:].sum(0)
        rows.append(np.concatenate([a.t] + [q[j] for j in range(prob.nmol)]))
    return np.ascontiguousarray(np.stack(rows))


def run_c5_device(bench, args, torch, dist, rank, local_rank, world, prob, tr, stream, build_s, grid_bytes, mine, grid, molID):
    """C5 through tb_run_transit_batch: ONE context; a step = one call with --batch atmospheres as BART hands them over
    ([T | q...] per atmosphere, host memory), per-atmosphere set-up (densities, hydrostatic radii) on the device, every
    stage launched once per 32 atmospheres, the spectra back in host memory at the end of the call."""
    from transit_b200 import api
    ctx = tr.ctx
    k = prob.nrad // 3
    ctx.set_atmosphere_model(prob.p, prob.p_fct, prob.t_fct, prob.rad_fct, 0, prob.p[k], prob.rad[k], 1500.0)
    batch = min(args.batch, max(1, len(mine)))
    sets = [retrieval_inputs(prob, mine[:batch]), retrieval_inputs(prob, [i + 100000 for i in mine[:batch]])]
    out = np.zeros((batch, prob.nwn))
    kms = {}

    def step(i):
        ctx.run_transit_batch(sets[i % 2], prob.toomuch, prob.starrad, prob.transp, out)
        for nm in ("batch", "interp"):
            try:
                kms.setdefault(nm, []).append(ctx.kernel_ms(nm))
            except api.TransitError:
                pass

    tm = _Timer(torch, dist, world, stream)
    sampler = bench.ClockSampler(local_rank)
    sampler.start()
    l0 = ctx.launch_count()
    dev_ms = tm.run(step, args.steps, args.warmup)
    launches = ctx.launch_count() - l0
    clocks = sampler.stop()
    assert np.isfinite(out).all() and (out > 0).all()
    tm.barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    tm.barrier()
    e2e_s = time.perf_counter() - t0
    (dev_ms, e2e_s), (_, _) = tm.reduce([dev_ms, e2e_s])
    if rank == 0:
        hbm_peak, src = bench.measured_peaks()
        nspec = world * batch * args.steps
        nm = len(prob.cfg["molecules"])
        # interpolation 2*Nmol*8 B read + 8 B written, tau 24 B per (layer, wavenumber) and atmosphere
        ab = ((2 * nm * 8 + 8.0) + 24.0) * prob.nrad * prob.nwn * batch
        step_ms = dev_ms / args.steps
        line = {
            "metric": "spectra/s", "value": nspec / (dev_ms * 1e-3), "unit": "spectra/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C5: BART-style retrieval, perturbed T-P/abundance profiles through opacity-grid "
                                   "interpolation + CIA + tau + modulation (C2 grid)", "natm": args.natm,
                       "batch_per_rank_per_step": batch, "contexts_per_gpu": 1,
                       "api": "tb_run_transit_batch: inputs [T | q...] per atmosphere, set-up on the device, one launch per stage "
                              "and 32 atmospheres", "sharding": f"atmospheres x{world}", "nwn": int(prob.nwn),
                       "layers": int(prob.nrad), "grid": [int(prob.nrad), len(GRID_TEMPS), nm, int(prob.nwn)],
                       "grid_gb": grid_bytes / 1e9, "grid_build_s": build_s, "l2": "flushed between steps"},
            "e2e": {"value": nspec / e2e_s, "unit": "spectra/s",
                    "h2d_bytes_per_step": 8 * prob.nrad * (1 + prob.nmol) * batch * world,
                    "d2h_bytes_per_step": 8 * prob.nwn * batch * world,
                    "api": "the same call with host buffers in and out (what host/libtransit_host.so:run_transit_batch wraps)"},
            "gpu_launches": launches * world,
            "kernel_ms": {k2: float(np.mean(v)) for k2, v in kms.items()},
            "roofline": {"kernel": "whole batch step (k_atm_setup, k_grid_interp, k_cia_*, k_ray_geometry, k_tau_transit, "
                                   "k_modulation: all HBM streams)", "bound": "hbm", "achieved": ab / (step_ms * 1e-3) / 1e9,
                         "peak": hbm_peak, "unit": "GB/s", "frac": ab / (step_ms * 1e-3) / 1e9 / hbm_peak, "traffic": None,
                         "peak_source": src, "algorithmic_bytes": ab},
            "clocks": clocks}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = c5_cpu_baseline(bench, args, prob, grid, molID)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


REF_BIN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "transit")
REF_SO = os.path.join(os.path.dirname(REF_BIN), "libtransit_ref.so")


def write_opacity_file(path, molID, temps, press_cgs, wn, grid):
    """opacity.c:405-421: 4 longs, molID, temperatures, pressures, wavenumbers, o[r][t][m][w]."""
    Nl, Nt, Nm, Nw = grid.shape
    with open(path, "wb") as f:
        np.array([Nm, Nt, Nl, Nw], "<i8").tofile(f)
        np.asarray(molID, "<i4").tofile(f); np.asarray(temps, "<f8").tofile(f)
        np.asarray(press_cgs, "<f8").tofile(f); np.asarray(wn, "<f8").tofile(f)
        np.ascontiguousarray(grid, "<f8").tofile(f)


def _ref_retrieval_worker(cfg, inputs, conn):
    """One process = one instance of the unmodified reference: transit_init once, then run_transit per input."""
    import ctypes as C
    os.dup2(os.open(os.devnull, os.O_WRONLY), 1)
    L = C.CDLL(REF_SO)
    argv = [b"transit", b"-c", cfg.encode()]
    arr = (C.c_char_p * len(argv))(*argv)
    t0 = time.perf_counter()
    L.transit_init(len(argv), arr)
    L.get_no_samples.restype = C.c_int
    n = L.get_no_samples()
    init_s = time.perf_counter() - t0
    out = np.zeros(n)
    P = C.POINTER(C.c_double)
    conn.send(("ready", init_s))
    conn.recv()
    t0 = time.perf_counter()
    for x in inputs:
        x = np.ascontiguousarray(x)
        L.run_transit(x.ctypes.data_as(P), x.size, out.ctypes.data_as(P), n)
    conn.send((time.perf_counter() - t0, float(out.sum())))


def c5_cpu_baseline(bench, args, prob, grid, molID):
    """the unmodified reference serving run_transit calls from the SAME full-size grid (written
    here as an opacity file; the reference only reads and interpolates it), one process per core, each process its own
    atmospheres -- the only parallelism the reference has."""
    import multiprocessing as mp
    if not os.path.exists(REF_SO):
        return {"value": None, "unit": "spectra/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    from transit_b200 import workloads
    d = os.path.join("/tmp", f"transit_b200_c5ref_{os.getpid()}")
    cfg, _ = workloads.write_reference_inputs(prob, d)
    k = prob.nrad // 3
    opa = os.path.join(d, "grid.opa")
    write_opacity_file(opa, molID, GRID_TEMPS, prob.p * prob.p_fct, prob.wn_i + np.arange(prob.nwn) * prob.wn_d, grid)
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[k]:.6e}\nrefradius {prob.rad[k]:.6f}\ngsurf 1500.0\nopacityfile {opa}\n")
    nproc = max(1, min(os.cpu_count() or 1, 8))           # 8 x 2.6 GB of grid in host memory
    per = 3
    ctx = mp.get_context("fork")
    pipes, procs = [], []
    for i in range(nproc):
        a, b = ctx.Pipe()
        inputs = retrieval_inputs(prob, [900000 + i * per + j for j in range(per)])
        p_ = ctx.Process(target=_ref_retrieval_worker, args=(cfg, inputs, b), daemon=True)
        p_.start(); pipes.append(a); procs.append(p_)
    inits = []
    for a in pipes:
        if not a.poll(900):
            raise RuntimeError("reference worker did not initialise")
        inits.append(a.recv()[1])
    t0 = time.perf_counter()
    for a in pipes:
        a.send("go")
    res = [a.recv() for a in pipes]
    wall = time.perf_counter() - t0
    for p_ in procs:
        p_.join(timeout=5)
    try:
        os.remove(opa)
    except OSError:
        pass
    return {"value": nproc * per / wall, "unit": "spectra/s", "cores": nproc, "kind": "reference",
            "sample": f"{nproc} processes of the unmodified reference (libtransit_ref.so), each transit_init from the full-size C2 grid "
                      f"file ({grid.nbytes / 1e9:.1f} GB, {np.mean(inits):.1f} s, not counted) and then {per} run_transit calls with "
                      f"perturbed profiles: {wall:.1f} s for {nproc * per} spectra; one core alone: {np.mean([r[0] for r in res]) / per:.2f} s per spectrum"}


def c4_cpu_baseline(bench, args):
    """The unmodified reference binary on the C4 shape at a TENTH of the wavenumbers (1e5 samples, 0.96 GB grid file);
    every per-spectrum cost of the reference is per wavenumber, so the full size is 10x (stated in `sample`)."""
    import subprocess
    if not os.path.exists(REF_BIN):
        return {"value": None, "unit": "spectra/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    from transit_b200 import workloads
    prob = c4_problem(nwn=100001)
    temps = np.array([800.0, 1000.0, 1200.0, 1400.0, 1600.0, 1800.0])
    d = os.path.join("/tmp", f"transit_b200_c4ref_{os.getpid()}")
    cfg, _ = workloads.write_reference_inputs(prob, d)
    molID = np.array([prob.mol_ID[prob.species_names.index(s)] for s in ("H2O", "CH4")], np.int32)
    opa = os.path.join(d, "grid.opa")
    write_opacity_file(opa, molID, temps, prob.p * prob.p_fct, prob.wn_i + np.arange(prob.nwn) * prob.wn_d, c4_grid(prob, temps))
    with open(cfg, "a") as f:
        f.write(f"opacityfile {opa}\n")
    t0 = time.perf_counter()
    r = subprocess.run([REF_BIN, "-c", cfg], capture_output=True, text=True, cwd=d)
    wall = time.perf_counter() - t0
    os.remove(opa)
    if r.returncode != 0:
        return {"value": None, "unit": "spectra/s", "cores": 1, "kind": "reference", "sample": "reference run failed: " + r.stdout[-200:]}
    return {"value": 1.0 / (10.0 * wall), "unit": "spectra/s", "cores": 1, "kind": "reference",
            "sample": f"the unmodified reference binary, one process, C4 shape at 100 001 wavenumbers (a tenth): {wall:.1f} s wall "
                      f"including transit_init and reading the 0.96 GB grid file; value = 1 / (10 x that), a linear extrapolation in the "
                      f"wavenumber count (labelled: not a measurement at 1e6)"}


def c3_cpu_baseline(bench, args, ctx_factory):
    """the reference's --justOpacity build on a 1e5-line subsample of the C3 line list, 10 layers,
    at 2 and at 6 grid temperatures; the DIFFERENCE of the two runs is the cost of 40 more (layer, T) states without
    transit_init (TLI read, Voigt table), and the rate in line-states/s extrapolates linearly in lines x states."""
    import subprocess
    if not os.path.exists(REF_BIN):
        return {"value": None, "unit": "G line-samples/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    from transit_b200 import workloads
    c3_problem(1.0)                                        # registers CONFIGS["C3"]
    prob = workloads.make_problem("C3", scale=1e-3, nlayers=10)
    d = os.path.join("/tmp", f"transit_b200_c3ref_{os.getpid()}")
    cfg, nlines = workloads.write_reference_inputs(prob, d)
    walls = {}
    for nt, thigh in ((2, 900), (6, 1300)):
        opa = os.path.join(d, f"g{nt}.opa")
        t0 = time.perf_counter()
        r = subprocess.run([REF_BIN, "-c", cfg, "--opacityfile", opa, "--justOpacity", "--tlow", "800", "--thigh", str(thigh),
                            "--tempdelt", "100"], capture_output=True, text=True, cwd=d)
        walls[nt] = time.perf_counter() - t0
        if r.returncode != 0:
            return {"value": None, "unit": "G line-samples/s", "cores": 1, "kind": "reference", "sample": "reference run failed: " + r.stdout[-200:]}
        os.remove(opa)
    dstates = 10 * (6 - 2)
    dt = max(walls[6] - walls[2], 1e-3)
    # the same 40 states through the library, for the number of profile updates they execute (the metric's unit)
    temps = np.arange(1000.0, 1300.1, 100.0)
    ziso = _ziso(prob, temps)
    tr = ctx_factory(prob)
    T, D, Z = [], [], []
    for r_ in range(prob.nrad):
        for t, temp in enumerate(temps):
            T.append(temp)
            D.append(1.66053886e-24 * prob.mol_q[:, r_] * (prob.p[r_] * prob.p_fct) / 1.380658e-16 / temp * prob.mol_mass)
            Z.append(ziso[:, t])
    tr.ctx.molext_states(np.array(T), np.array(D), np.array(Z), prob.ethresh, permol=1)
    nsamp = int(tr.ctx.sample_counts(len(T)).sum())
    tr.free_memory()
    return {"value": nsamp / dt / 1e9, "unit": "G line-samples/s", "cores": 1, "kind": "reference",
            "line_states_per_s": nlines * dstates / dt,
            "sample": f"the unmodified reference binary, one process, --justOpacity on a {nlines}-line subsample (scale 1e-3 of C3's "
                      f"list), 10 layers: {walls[2]:.1f} s with 2 grid temperatures, {walls[6]:.1f} s with 6; the difference is {dstates} "
                      f"(layer, T) states = {nlines * dstates:.2e} line-states in {dt:.1f} s; the full C3 grid (1e8 lines x 2600 states) "
                      f"extrapolates to {1e8 * 2600 / (nlines * dstates / dt) / 3600:.0f} core-hours (linear in lines x states, labelled)"}


def c4_problem(nwn=1_000_000, nlayers=100):
    """C4 static problem: the C1/eclipse atmosphere shape on a 10^6-sample grid; extinction comes from a
    grid, so there is no line list."""
    from transit_b200 import workloads
    cfg = dict(workloads.CONFIGS["C1"])
    cfg.update(wn_lo=2000.0, wn_hi=2000.0 + (nwn - 1) * 0.01, wndelt=0.01, wnosamp=120, nlayers=nlayers, nlines=0,
               molecules={"H2O": (1, 0.5), "CH4": (1, 0.5)},
               species={"H2": 0.85, "He": 0.149, "H2O": 5e-4, "CH4": 4e-4}, cia=[("H2", "H2"), ("H2", "He")],
               seed=1004, toomuch=10.0)
    workloads.CONFIGS["C4"] = cfg
    return workloads.make_problem("C4")


def c4_grid(prob, temps, nm=2):
    """Synthetic opacity grid o[layer][T][molecule][wn] (cm2/g): smooth band structure, rising with T."""
    w = prob.wn_i + np.arange(prob.nwn) * prob.wn_d
    base = [1e-1 * np.exp(-((w - 3700.0) / 600.0) ** 2) * (1.2 + np.sin(w * 7.0)) + 1e-4,
            4e-2 * np.exp(-((w - 3000.0) / 400.0) ** 2) * (1.2 + np.cos(w * 11.0)) + 5e-5][:nm]
    grid = np.empty((prob.nrad, len(temps), nm, prob.nwn))
    for r in range(prob.nrad):
        for t, T in enumerate(temps):
            for m in range(nm):
                grid[r, t, m] = base[m] * (T / 1000.0) ** 1.5 * (1.0 + 0.002 * r)
    return grid


def run_c4(bench, args):
    from transit_b200 import api, sharding
    torch, dist, rank, local_rank, world = _dist(bench, args)
    assert world == 1, "C4 is a single-GPU line here (N independent replicas otherwise)"
    nwn = max(1000, int(1_000_000 * args.scale))
    prob = c4_problem(nwn)
    temps = np.array([800.0, 1000.0, 1200.0, 1400.0, 1600.0, 1800.0])
    grid = c4_grid(prob, temps)
    tr = sharding.ShardedTransit(local_rank, rank, world)
    stream = sharding.private_stream()         # kernels, copies and NCCL collectives share this stream
    tr.ctx.set_stream(stream.cuda_stream)
    molID = np.array([prob.mol_ID[prob.species_names.index(s)] for s in ("H2O", "CH4")], np.int32)
    tr.transit_init(prob, opacity=dict(molID=molID, temp=temps, press=prob.p * prob.p_fct, grid=grid))
    grid_gb = grid.nbytes / 1e9
    del grid
    out = np.zeros(prob.nwn)
    kms = {}

    def step(i):
        tr.do_transit(out)
        for nm in ("interp", "cia", "tau", "intensity"):
            try:
                kms.setdefault(nm, []).append(tr.ctx.kernel_ms(nm))
            except api.TransitError:
                pass

    tm = _Timer(torch, dist, world, stream)
    sampler = bench.ClockSampler(local_rank)
    sampler.start()
    l0 = tr.ctx.launch_count()
    dev_ms = tm.run(step, args.steps, args.warmup)
    launches = tr.ctx.launch_count() - l0
    clocks = sampler.stop()
    assert np.isfinite(out).all() and (out > 0).all()
    t0 = time.perf_counter()
    for i in range(args.steps):
        tr.set_atmosphere(prob)
        step(i)
    tm.barrier()
    e2e_s = time.perf_counter() - t0
    hbm_peak, src = bench.measured_peaks()
    nang, nm = len(prob.angles), 2
    # algorithmic bytes of one step: interpolation (2 planes per molecule in, e out), tau (e and e_cs in,
    # t out), intensity (t in once per angle, one intensity plane out per angle), CIA (e_cs out)
    ab = prob.nrad * prob.nwn * 8.0 * ((2 * nm + 1) + 3 + 1 + nang) + prob.nwn * 8.0 * (nang + 1)
    line = {
        "metric": "spectra/s", "value": args.steps / (dev_ms * 1e-3), "unit": "spectra/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C4: high-resolution eclipse, extinction interpolated from a resident opacity grid + CIA, "
                               "eclipse tau, multi-angle intensity, flux", "nwn": int(prob.nwn), "layers": int(prob.nrad),
                   "angles": [float(a) for a in prob.angles], "grid": [int(prob.nrad), len(temps), nm, int(prob.nwn)],
                   "grid_gb": grid_gb, "l2": "flushed between steps"},
        "samples_per_s": prob.nwn * args.steps / (dev_ms * 1e-3),
        "e2e": {"value": args.steps / e2e_s, "unit": "spectra/s",
                "h2d_bytes_per_step": 8 * prob.nrad * (3 + prob.nmol + prob.niso), "d2h_bytes_per_step": 8 * prob.nwn},
        "gpu_launches": launches, "kernel_ms": {k: float(np.mean(v)) for k, v in kms.items()},
        "roofline": {"kernel": "whole step (interp + CIA + tau + intensity, all HBM streams)", "bound": "hbm",
                     "achieved": ab / (dev_ms / args.steps * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": ab / (dev_ms / args.steps * 1e-3) / 1e9 / hbm_peak, "traffic": None, "peak_source": src,
                     "algorithmic_bytes": ab},
        "clocks": clocks}
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = c4_cpu_baseline(bench, args)
    print(json.dumps(line), flush=True)
    return 0


def c3_problem(scale=1.0):
    """C3 static problem: the C2 species on an ExoMol-scale line list (10^8 lines at scale 1: flat in
    wavenumber, steep gf distribution)."""
    from transit_b200 import workloads
    cfg = dict(workloads.CONFIGS["C2"])
    cfg.update(nlines=100_000_000, seed=1003)
    workloads.CONFIGS["C3"] = cfg
    return workloads.make_problem("C3", scale=scale)


def run_c3(bench, args):
    """Opacity-grid build: states (layer, T) of the permol line sum, rows delivered to host memory."""
    from transit_b200 import api, sharding
    torch, dist, rank, local_rank, world = _dist(bench, args)
    t0 = time.perf_counter()
    prob = c3_problem(args.scale)
    gen_s = time.perf_counter() - t0
    tr = sharding.ShardedTransit(local_rank, rank, world)
    stream = sharding.private_stream()         # kernels, copies and NCCL collectives share this stream
    tr.ctx.set_stream(stream.cuda_stream)
    t0 = time.perf_counter()
    tr.transit_init(prob)
    tr.ctx.sync()
    init_s = time.perf_counter() - t0
    bands = sharding.wavenumber_bands(sharding.line_weights(prob), world)
    if world > 1:
        tr.ctx.set_shard(bands[rank][0], bands[rank][1], rank, world, sharding.nccl_exchange())
    ziso = _ziso(prob, GRID_TEMPS)
    nt, nl = len(GRID_TEMPS), args.block_layers
    ns = nt * nl

    def states(block):
        T, D, Z = [], [], []
        for r in [(block * nl + k) % prob.nrad for k in range(nl)]:
            for t, temp in enumerate(GRID_TEMPS):
                T.append(temp)
                rho = api.AMU * prob.mol_q[:, r] * (prob.p[r] * prob.p_fct) / api.KB / temp
                D.append(rho * prob.mol_mass)               # opacity.c:387-403, stateeqnford by number
                Z.append(ziso[:, t])
        return np.array(T), np.array(D), np.array(Z)

    kms, samples, rows = {}, [0], [None]

    def step(i):
        T, D, Z = states(i)
        rows[0] = tr.ctx.molext_states(T, D, Z, prob.ethresh, permol=1)     # [ns][mol][nwn] on the host
        samples[0] = int(tr.ctx.sample_counts(ns).sum())
        for nm in ("kmax", "strength", "accumulate", "molext"):
            kms.setdefault(nm, []).append(tr.ctx.kernel_ms(nm))

    tm = _Timer(torch, dist, world, stream)
    sampler = bench.ClockSampler(local_rank)
    sampler.start()
    l0 = tr.ctx.launch_count()
    dev_ms = tm.run(step, args.steps, args.warmup)
    launches = tr.ctx.launch_count() - l0
    clocks = sampler.stop()
    lo, hi = bands[rank]
    assert np.isfinite(rows[0]).all() and rows[0][..., lo:hi].max() > 0
    (dev_ms,), (_,) = tm.reduce([dev_ms])
    _, (nsamp, mol_ms, acc_ms) = tm.reduce([float(samples[0]), float(np.mean(kms["molext"])),
                                            float(np.mean(kms["accumulate"]))])
    if rank == 0:
        hbm_peak, src = bench.measured_peaks()
        nlines = len(prob.wl)
        step_s = dev_ms / args.steps * 1e-3
        line = {
            "metric": "line-sum Glines.samples/s", "value": nsamp / step_s / 1e9, "unit": "G line-samples/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3: opacity-grid build (calcopacity), ExoMol-scale synthetic line list x 26 temperatures "
                                   "x layers, permol rows delivered to host memory", "lines": int(nlines),
                       "states_per_step": ns, "block_layers": nl, "temperatures": nt, "nwn": int(prob.nwn),
                       "wnosamp": int(prob.wnosamp), "line_molecules": len(prob.cfg["molecules"]),
                       "sharding": f"wavenumber bands x{world}" if world > 1 else "none", "bands": bands,
                       "l2": "flushed between steps", "line_list_gen_s": gen_s, "init_s": init_s},
            "line_states_per_s": nlines * ns / step_s, "states_per_s": ns / step_s,
            "full_grid_s_extrapolated": step_s * prob.nrad / nl,
            "e2e": {"value": nsamp / step_s / 1e9, "unit": "G line-samples/s",
                    "h2d_bytes_per_step": 8 * ns * (1 + prob.nmol + prob.niso) * world,
                    "d2h_bytes_per_step": 8 * ns * len(prob.cfg["molecules"]) * int(prob.nwn) * world,
                    "note": "the step already ends with the grid rows in host memory (D2H inside the timed region)"},
            "gpu_launches": int(launches * world),
            "kernel_ms": {k: float(np.mean(v)) for k, v in kms.items()},
            "clocks": clocks}
        mb = tr.ctx.microbench()
        c_exp = mb["dfma_tflops"] * 1e12 / (mb["dexp_gops"] * 1e9)
        flop = 2.0 * nsamp + (12.0 + 2.0 * c_exp) * nlines * ns            # SURVEY 8d, whole line sum of a step
        ach = flop / (float(np.mean(kms["molext"])) * 1e-3) / 1e12
        line["roofline"] = {"kernel": "line sum, all kernels (k_kmax + k_linesum_fused_w + k_strength + k_accumulate)", "bound": "fp64",
                            "achieved": ach, "peak": mb["dfma_tflops"], "unit": "TFLOP/s", "frac": ach / mb["dfma_tflops"],
                            "traffic": None, "peak_source": "tb_microbench DFMA, measured in this run", "ms": float(np.mean(kms["molext"])),
                            "algorithmic_flop": flop, "c_exp_flop": c_exp}
        line["microbench"] = mb
        if world == 1 and not args.no_cpu_baseline:
            def factory(pr):
                t2 = sharding.ShardedTransit(local_rank, rank, world)
                t2.ctx.set_fused(2)
                t2.transit_init(pr)
                return t2
            line["cpu_baseline"] = c3_cpu_baseline(bench, args, factory)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def run(bench, args):
    return {"C3": run_c3, "C4": run_c4, "C5": run_c5}[args.workload](bench, args)
