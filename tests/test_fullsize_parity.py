"""BASELINE.json's configurations at their REAL sizes against the UNMODIFIED reference (oracle/_ref, compiled from
/root/reference by oracle/Makefile and shipped to the GPU box with the tree; never read at run time from
/root/reference): the same input files go through

  * oracle/_ref/ref_dump_idx  -- the reference's own transit_init / do_transit stage sequence with full-precision
    taps, plus the per-line index tap (a patched COPY of extinction.c built only into oracle/_ref/obj), and
  * host/libtransit_host.so   -- our C host (file readers, a16 set-up) on top of the CUDA library, taps through
    the C ABI (tests/hostrun.py),

and are compared at the tolerances north_star states: line->bin assignment and grid indexing bit-exact; last[] and
computed[] exact; extinction <= 1e-10 relative on the layers the reference evaluated; spectra <= 1e-8 relative.
The worst values per configuration are written to gpurun_out/parity_fullsize.json (copied into DESIGN.md).

C1 is transit_demo.cfg VERBATIM: its keys, the shipped HD209458b_demo.atm (10 species), the
shipped CIA_H2H2_400-7000K.dat and molecules.dat, and the seeded 2e5-line CH4 TLI.
"""
import json
import os
import subprocess
import sys
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden as MG  # noqa: E402
from transit_b200 import synth, workloads  # noqa: E402

REFDIR = os.path.join(ROOT, "oracle", "_ref")
REF_IDX = os.path.join(REFDIR, "ref_dump_idx")
INPUTS = os.path.join(REFDIR, "inputs")
HOSTRUN = os.path.join(ROOT, "tests", "hostrun.py")
REPORT = os.path.join(ROOT, "gpurun_out", "parity_fullsize.json")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(REF_IDX), reason="oracle/_ref/ref_dump_idx not built")]


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0


def report(name, vals):
    os.makedirs(os.path.dirname(REPORT), exist_ok=True)
    try:
        d = json.load(open(REPORT))
    except (OSError, ValueError):
        d = {}
    d[name] = vals
    json.dump(d, open(REPORT, "w"), indent=1)
    print(f"\n[parity {name}] " + json.dumps(vals))


def run_both(tmp, cfg, idx_layers, extra_ref=(), extra_host=(), rerun=None, timeout=900):
    refbin, ours = os.path.join(tmp, "ref.bin"), os.path.join(tmp, "ours.npz")
    idx = ",".join(str(r) for r in idx_layers)
    t0 = time.perf_counter()
    cmd = [REF_IDX, refbin] + (["--idx", idx] if idx_layers else []) + (["--rerun", rerun] if rerun else []) + \
        list(extra_ref) + ["--", "-c", cfg]
    r = subprocess.run(cmd, cwd=tmp, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    ref_s = time.perf_counter() - t0
    cmd = [sys.executable, HOSTRUN, cfg, ours] + (["--trace", idx] if idx_layers else []) + \
        (["--rerun", rerun] if rerun else []) + list(extra_host)
    r = subprocess.run(cmd, cwd=tmp, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    return MG.read_container(refbin), np.load(ours), ref_s


def compare_lbl(name, g, o, idx_layers, eclipse):
    """The assertions every line-by-line configuration shares; returns the worst values."""
    nrad, nwn = int(g["nrad"].item()), int(g["wns_n"].item())
    assert tuple(o["sizes"][[0, 3]]) == (nrad, nwn)
    assert int(o["sizes"][4]) == int(g["nlines"].item())                   # same per-isotope TLI range selection
    comp_ref = g["computed"].astype(np.int16)
    assert np.array_equal(o["computed"], comp_ref), "lazily-evaluated layer set differs (tau.c:238-271)"
    assert np.array_equal(o["last"], g["last"]), "tau cut-off ray differs somewhere"
    worst = dict(nwn=nwn, layers=nrad, lines=int(g["nlines"].item()), layers_computed=int(comp_ref.sum()))
    e_ref, e = g["ext"], o["ext"]
    rows = np.nonzero(comp_ref)[0]
    assert np.array_equal(e[rows] == 0, e_ref[rows] == 0), "zero pattern of the extinction differs"
    m = e_ref[rows] > 0
    worst["ext_rel"] = rel(e[rows][m], e_ref[rows][m])
    assert worst["ext_rel"] < 1e-10
    worst["cia_rel"] = rel(o["e_cs"][g["e_cs"] > 0], g["e_cs"][g["e_cs"] > 0])
    assert worst["cia_rel"] < 1e-12 and np.array_equal(o["e_cs"] == 0, g["e_cs"] == 0)
    t_ref, t = g["tau"], o["tau"]
    assert np.array_equal(t == 0, t_ref == 0)                              # zero tail beyond last[]
    worst["tau_rel"] = rel(t[t_ref != 0], t_ref[t_ref != 0])
    assert worst["tau_rel"] < 1e-9
    if eclipse:
        worst["intens_rel"] = rel(o["intens"], g["intens"])
        assert worst["intens_rel"] < 1e-8
    worst["spectrum_rel"] = rel(o["out"], g["out"])
    assert worst["spectrum_rel"] < 1e-8
    # ---- the index tap: per processed line, against the reference itself (not the oracle)
    ps = o["profsize"]
    assert np.array_equal(ps, g["profsize"])
    iso = g["line_isoid"].astype(np.int64)
    osamp = int(g["owns_o"].item())
    nline = dict(evaluated=0, skipped=0, coadded=0)
    for r in idx_layers:
        st_ref, st = g[f"idx{r}_status"], o[f"idx{r}_status"].astype(np.int32)
        # our trace: 0 out of range, 1 evaluated, 2 co-added, 3 below threshold (include/transit_b200.h)
        assert np.array_equal(st, st_ref), f"layer {r}: {int((st != st_ref).sum())} lines differ in evaluated/co-added/skipped status"
        ev, sk = st_ref == 1, st_ref == 3
        assert np.array_equal(o[f"idx{r}_iown"][ev | sk], g[f"idx{r}_iown"][ev | sk]), f"layer {r}: iown"
        assert np.array_equal(o[f"idx{r}_idwn"][ev], g[f"idx{r}_idwn"][ev]), f"layer {r}: idwn"
        assert np.array_equal(o[f"idx{r}_idop"][ev], g[f"idx{r}_idop"][ev]), f"layer {r}: idop"
        ilor = o[f"idx{r}_ilor"][iso]
        assert np.array_equal(ilor[ev], g[f"idx{r}_ilor"][ev]), f"layer {r}: ilor"
        # minj / maxj from OUR indices and table (extinction.c:486-496)
        io, idw = o[f"idx{r}_iown"][ev].astype(np.int64), o[f"idx{r}_idwn"][ev].astype(np.int64)
        p = ps[o[f"idx{r}_idop"][ev], ilor[ev]]
        subw = io - idw * osamp
        # C integer division truncates toward zero
        minj = idw - np.trunc((p - subw) / osamp).astype(np.int64)
        maxj = idw + np.trunc((p + subw) / osamp).astype(np.int64)
        minj, maxj = np.maximum(minj, 0), np.minimum(maxj, nwn - 1)
        assert np.array_equal(minj, g[f"idx{r}_minj"][ev]) and np.array_equal(maxj, g[f"idx{r}_maxj"][ev]), f"layer {r}: minj/maxj"
        nline["evaluated"] += int(ev.sum()); nline["skipped"] += int(sk.sum()); nline["coadded"] += int((st_ref == 2).sum())
        # and that layer's extinction as the reference's own computemolext call returns it
        x_ref, x = g[f"idx{r}_ext"], o["ext"][r]
        assert np.array_equal(x == 0, x_ref == 0)
        worst["ext_rel"] = max(worst["ext_rel"], rel(x[x_ref > 0], x_ref[x_ref > 0]))
        assert worst["ext_rel"] < 1e-10
    worst["index_layers"] = list(idx_layers)
    worst["index_lines_compared"] = nline
    return worst


def test_c1_demo_cfg_verbatim(tmp_path):
    """transit_demo.cfg with its own keys and the shipped atmosphere / CIA / molecules files; only the paths are
    repointed and the TLI is the seeded synthetic CH4 list of SURVEY.md 8d (2e5 lines, 2-4 um)."""
    tmp = str(tmp_path)
    prob = workloads.make_problem("C1")
    synth.write_tli(os.path.join(tmp, "CH4_HITRAN_2-4um.tli"), prob.dbs, 1e4 / 5050.0, 1e4 / 2450.0)
    lines = []
    for l in open(os.path.join(INPUTS, "transit_demo.cfg"), encoding="utf-8", errors="replace"):
        k = l.split()[0] if l.split() else ""
        if k == "atm": l = f"atm     {INPUTS}/HD209458b_demo.atm\n"
        elif k == "linedb": l = f"linedb  {tmp}/CH4_HITRAN_2-4um.tli\n"
        elif k == "csfile": l = f"csfile  {INPUTS}/CIA_H2H2_400-7000K.dat\n"
        elif k == "molfile": l = f"molfile {INPUTS}/molecules.dat\n"
        elif k == "verb": l = "verb 0\n"
        elif k in ("outtoomuch", "outsample", "outspec"): l = f"{k} {tmp}/{k}_REF.dat\n"
        lines.append(l)
    cfg = os.path.join(tmp, "demo.cfg")
    open(cfg, "w", encoding="utf-8").writelines(lines)
    idx = [0, 14, 15, 22, 50, 99]
    g, o, ref_s = run_both(tmp, cfg, idx)
    assert int(g["wns_n"].item()) == 2501 and int(g["owns_n"].item()) == 5400001 and int(g["nrad"].item()) == 100
    assert int(g["nDop"].item()) == 60 and int(g["nLor"].item()) == 60 and int(g["nmol"].item()) == 10
    w = compare_lbl("C1", g, o, idx, eclipse=True)
    w["reference_s"] = ref_s
    w["ours_run_s"] = float(o["run_s"])
    report("C1 transit_demo.cfg verbatim", w)


@pytest.mark.parametrize("mode", ["default", "fuse-all"])
def test_c2_full_size(tmp_path, mode):
    """C2 as BASELINE.json names it: 31 334 wavenumbers x 100 layers x 1e6 lines, 60x60 table at wnosamp 2160,
    ONE whole-spectrum reference process (~30 core-seconds).  Twice: with the production split (narrow layers fused,
    the rest through k_strength + k_accumulate) and with EVERY layer through the fused kernel, the carried Doppler
    indices of the deep layers found by k_lastq_scan -- the per-line index tap covers both."""
    tmp = str(tmp_path)
    prob = workloads.make_problem("C2")
    cfg, nlines = workloads.write_reference_inputs(prob, tmp)
    idx = [0, 6, 12, 16, 19, 20, 21, 30, 60, 99]
    g, o, ref_s = run_both(tmp, cfg, idx, timeout=1500, extra_host=["--fuse-all"] if mode == "fuse-all" else [])
    assert int(g["wns_n"].item()) == 31334 and int(g["nrad"].item()) == 100 and int(g["nlines"].item()) >= 990000
    w = compare_lbl("C2", g, o, idx, eclipse=False)
    w["reference_s"] = ref_s
    w["ours_run_s"] = float(o["run_s"])
    report("C2 full size" + ("" if mode == "default" else " (every layer fused, carried indices by k_lastq_scan)"), w)


def _rerun_inputs(prob, k, seed0=7000, dT=150.0, dex=1.0):
    """k run_transit inputs [T | q_mol...] (readatm.c:735-744): seeded perturbations of the base profile."""
    rows = []
    for i in range(k):
        a = workloads.perturbed(prob, seed=seed0 + i, dT=dT, dex=dex)
        q = a.mol_q.copy()
        q[0] = 1.0 - q[1:].sum(0)                 # the bulk gas takes up the slack, as a retrieval's abundances do
        rows.append(np.concatenate([a.t] + [q[j] for j in range(prob.nmol)]))
    return np.stack(rows)


def test_c5_shape_retrieval_from_c2_grid(tmp_path):
    """C5 shape at the C2 grid's full size: the host BUILDS the 2.6 GB opacity grid (100 layers x 26 temperatures x 4
    molecules x 31 334 wavenumbers) with --opacityfile, the UNMODIFIED reference reads that file and serves the base
    spectrum + 4 run_transit calls (perturbed T-P / abundance profiles, hydrostatic radii from refpress / refradius /
    gsurf); the host serves the same calls one by one AND as one run_transit_batch (set-up on the device)."""
    tmp = str(tmp_path)
    prob = workloads.make_problem("C2")
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    k = prob.nrad // 3
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[k]:.6e}\nrefradius {prob.rad[k]:.6f}\ngsurf 1500.0\nopacityfile {tmp}/grid.opa\n")
    inp = _rerun_inputs(prob, 4)
    inp.astype("<f8").tofile(os.path.join(tmp, "rerun.f64"))
    # the host first (it builds and writes the grid), then the reference on the host's file
    ours = os.path.join(tmp, "ours.npz")
    t0 = time.perf_counter()
    r = subprocess.run([sys.executable, HOSTRUN, cfg, ours, "--rerun", os.path.join(tmp, "rerun.f64"), "--batch", "--no-taps"],
                       cwd=tmp, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    build_s = time.perf_counter() - t0
    assert os.path.getsize(os.path.join(tmp, "grid.opa")) > 2.5e9
    refbin = os.path.join(tmp, "ref.bin")
    t0 = time.perf_counter()
    r = subprocess.run([os.path.join(REFDIR, "ref_dump"), refbin, "--rerun", os.path.join(tmp, "rerun.f64"), "--", "-c", cfg],
                       cwd=tmp, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    ref_s = time.perf_counter() - t0
    g, o = MG.read_container(refbin), np.load(ours)
    assert int(g["op_Ntemp"].item()) == 26 and int(g["op_Nmol"].item()) == 4 and int(g["op_Nwave"].item()) == 31334
    w = dict(nwn=31334, layers=100, grid_temperatures=26, atmospheres=4)
    w["spectrum_rel"] = rel(o["out"], g["out"])
    w["run_transit_rel"] = rel(o["rerun_out_all"], g["rerun_out_all"])
    w["batch_vs_loop_rel"] = rel(o["batch_out_all"], o["rerun_out_all"])
    w["batch_vs_reference_rel"] = rel(o["batch_out_all"], g["rerun_out_all"])
    assert w["spectrum_rel"] < 1e-8 and w["run_transit_rel"] < 1e-8 and w["batch_vs_reference_rel"] < 1e-8
    assert w["batch_vs_loop_rel"] < 1e-11
    assert float(o["batch_device_ms"]) > 0, "run_transit_batch fell back to the loop of run_transit"
    w["batch_device_ms"] = float(o["batch_device_ms"])
    w["reference_s"] = ref_s
    w["ours_total_s_including_grid_build_and_file"] = build_s
    w["ours_loop_s"], w["ours_batch_s"] = float(o["rerun_s"]), float(o["batch_s"])
    report("C5 shape: run_transit x4 from the full-size C2 grid", w)


def test_c4_shape_high_resolution_eclipse(tmp_path):
    """C4 shape at a tenth of its wavenumbers (1e5 samples at 0.01 cm-1, 100 layers, 6 grid temperatures x 2
    molecules: a 0.96 GB opacity file both sides read), eclipse geometry, 5 angles: the largest size the reference
    serves in under two minutes."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_configs as BC
    tmp = str(tmp_path)
    prob = BC.c4_problem(nwn=100001)
    temps = np.array([800.0, 1000.0, 1200.0, 1400.0, 1600.0, 1800.0])
    grid = BC.c4_grid(prob, temps)
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    molID = np.array([prob.mol_ID[prob.species_names.index(s)] for s in ("H2O", "CH4")], np.int32)
    with open(os.path.join(tmp, "grid.opa"), "wb") as f:          # opacity.c:405-421
        np.array([2, len(temps), prob.nrad, prob.nwn], "<i8").tofile(f)
        molID.astype("<i4").tofile(f); temps.astype("<f8").tofile(f)
        (prob.p * prob.p_fct).astype("<f8").tofile(f)
        (prob.wn_i + np.arange(prob.nwn) * prob.wn_d).astype("<f8").tofile(f)
        grid.astype("<f8").tofile(f)
    del grid
    with open(cfg, "a") as f:
        f.write(f"opacityfile {tmp}/grid.opa\n")
    g, o, ref_s = run_both(tmp, cfg, [], timeout=1500)
    assert int(g["wns_n"].item()) == 100001 and int(g["nrad"].item()) == 100
    assert np.array_equal(o["last"], g["last"]) and np.array_equal(o["computed"], g["computed"].astype(np.int16))
    comp = g["computed"].astype(bool)
    w = dict(nwn=100001, layers=100, angles=5)
    w["ext_rel"] = rel(o["ext"][comp], g["ext"][comp])
    w["cia_rel"] = rel(o["e_cs"][g["e_cs"] > 0], g["e_cs"][g["e_cs"] > 0])
    nz = g["tau"] != 0
    assert np.array_equal(o["tau"] == 0, g["tau"] == 0)
    w["tau_rel"] = rel(o["tau"][nz], g["tau"][nz])
    w["intens_rel"] = rel(o["intens"], g["intens"])
    w["spectrum_rel"] = rel(o["out"], g["out"])
    assert w["ext_rel"] < 1e-10 and w["cia_rel"] < 1e-12 and w["tau_rel"] < 1e-9
    assert w["intens_rel"] < 1e-8 and w["spectrum_rel"] < 1e-8
    w["reference_s"] = ref_s
    w["ours_run_s"] = float(o["run_s"])
    report("C4 shape: 1e5 wavenumbers, eclipse from a grid", w)
