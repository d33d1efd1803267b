"""GPU parity: the CUDA path, called through the C ABI (transit_b200.api.Context), against
(a) the committed golden vectors produced by the UNMODIFIED reference and (b) the CPU
oracle on seeded synthetic inputs.  Integer/index work is compared bit-exactly; extinction
to 1e-10 relative and spectra to 1e-8 relative (BASELINE.json north_star)."""
import os

import numpy as np
import pytest

import oraclelib as O
from transit_b200 import api, workloads

pytestmark = pytest.mark.gpu

EXT_RTOL = 1e-10      # north_star: extinction within 1e-10 relative
SPEC_RTOL = 1e-8      # north_star: spectra within 1e-8 relative


def relerr(a, b):
    d = np.abs(a - b)
    s = np.maximum(np.abs(a), np.abs(b))
    return float(np.max(np.where(s > 0, d / np.where(s > 0, s, 1), 0.0)))


def cfg_of(d):
    return dict(l.split(None, 1) for l in bytes(d["cfg_text"]).decode().splitlines())


def setup_ctx(prob, d=None, nDop=None, nLor=None, lines=True):
    c = api.Context(0)
    c.set_trace(True)       # the fused line sum leaves its per-line records only when asked (tb_set_trace)
    c.set_fused(2)          # ... and is used on these small grids too, so that every parity test covers it
    c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
    c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio,
                  prob.iso_imol)
    if lines:
        c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
        c.build_profiles(nDop or 60, nLor or 60, 1e-3, 0.25, 1e-4, 10.0, prob.nwidth)
    for tb in prob.cia:
        c.cia_add_table(tb["wn"], tb["temp"], tb["tab"], tb["mol"])
    c.set_atmosphere(prob.rad, prob.rad_fct, prob.p, prob.p_fct, prob.t, prob.t_fct, prob.mol_d, prob.iso_z)
    return c


@pytest.fixture(scope="module")
def ctxs(golden):
    out = {}
    for case in ("tiny_ecl", "tiny_tra"):
        d = golden(case)
        prob = O.Problem.from_golden(d)
        cfg = cfg_of(d)
        out[case] = (d, prob, setup_ctx(prob, d, int(cfg["ndop"]), int(cfg["nlor"])))
    yield out
    for _, _, c in out.values():
        c.close()


# ------------------------------------------------------------------ Voigt table (a3/a4)

@pytest.mark.parametrize("case", ["tiny_ecl", "tiny_tra"])
def test_voigt_table_vs_reference(ctxs, case):
    d, prob, c = ctxs[case]
    t = c.get_profiles()
    assert np.array_equal(t["psize"], d["profsize"]) and np.array_equal(t["off"], d["prof_off"])
    assert relerr(t["aDop"], d["aDop"]) < 2e-15 and relerr(t["aLor"], d["aLor"]) < 2e-15
    ref = d["prof_data"]
    assert t["data"].size == ref.size
    bad = np.flatnonzero(t["data"] != ref)
    # the reference evaluates voigtxy in x87 long double and rounds to float; FP64 on the device may
    # land on the other side of a float rounding boundary in rare cases: at most 1 float ulp, rarely
    ulp = np.abs(t["data"][bad].view(np.int32) - ref[bad].view(np.int32))
    print(f"voigt[{case}]: {bad.size} of {ref.size} floats differ, max {ulp.max() if bad.size else 0} ulp")
    assert bad.size <= 1e-5 * ref.size and (bad.size == 0 or ulp.max() <= 1)


def test_voigt_default_table_vs_oracle():
    """60x60 defaults on the demo grid (C1): 2332 unique profiles incl. the QUICK ones."""
    prob = workloads.make_problem("C1", scale=0.001, nlayers=4)
    c = api.Context(0)
    c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
    c.build_profiles(60, 60, 1e-3, 0.25, 1e-4, 10.0, 20.0)
    t = c.get_profiles()
    f = lambda x: float(np.float32(x))
    vt = O.OracleVoigt(O.Problem(**vars(prob)), 60, 60, f(1e-3), f(0.25), f(1e-4), f(10.0))
    assert np.array_equal(t["psize"], vt.psize) and np.array_equal(t["off"], vt.off)
    bad = np.flatnonzero(t["data"] != vt.data)
    ulp = np.abs(t["data"][bad].view(np.int32) - vt.data[bad].view(np.int32))
    print(f"voigt[C1 60x60]: {bad.size} of {vt.data.size} floats differ, max {ulp.max() if bad.size else 0} ulp")
    assert bad.size <= 1e-5 * vt.data.size and (bad.size == 0 or ulp.max() <= 1)
    c.close()


# ------------------------------------------------------------------ line sum (a5)

@pytest.mark.parametrize("case", ["tiny_ecl", "tiny_tra"])
def test_line_to_bin_assignment_bit_exact(ctxs, case):
    """iown / idwn / idop / co-add / threshold decisions per line == oracle, every layer."""
    d, prob, c = ctxs[case]
    vt = O.voigt_from_arrays(d["aDop"], d["aLor"], d["profsize"], d["prof_off"], d["prof_data"])
    T = prob.t * prob.t_fct
    c.molext_states(T, prob.mol_d.T.copy(), prob.iso_z.T.copy(), prob.ethresh)
    counts = c.sample_counts(prob.nrad)
    for r in range(prob.nrad):
        _, tr = O.computemolext(prob, vt, T[r], prob.mol_d[:, r], prob.iso_z[:, r], prob.ethresh, trace=True)
        g = c.line_trace(r)
        assert np.array_equal(g["status"], tr["status"]), f"layer {r}: co-add/threshold decisions differ"
        ev = tr["status"] == 1
        lead = (tr["status"] == 1) | (tr["status"] == 3)
        assert np.array_equal(g["iown"][lead], tr["iown"][lead]), f"layer {r}: fine-bin index"
        assert np.array_equal(g["idwn"][lead], tr["idwn"][lead]), f"layer {r}: coarse-bin index"
        assert np.array_equal(g["idop"][ev], tr["idop"][ev]), f"layer {r}: Doppler-width index"
        assert counts[r] == tr["nsamples"], f"layer {r}: number of profile samples"


@pytest.mark.parametrize("case", ["tiny_ecl", "tiny_tra"])
def test_extinction_vs_reference_all_layers(ctxs, case):
    d, prob, c = ctxs[case]
    c.extinction_lbl(prob.ethresh)
    e = c.get_extinction()
    ref = d["ext_full"]
    assert np.array_equal(e == 0, ref == 0)
    # the table on the device may differ from the reference's by 1 float ulp in rare samples
    # (see test_voigt_table_vs_reference); everything else is FP64 and order-of-summation only
    assert relerr(e, ref) < EXT_RTOL
    comp = d["computed"].astype(bool)
    assert relerr(e[comp], d["ext"][comp]) < EXT_RTOL


def test_extinction_deterministic(ctxs):
    d, prob, c = ctxs["tiny_tra"]
    c.extinction_lbl(prob.ethresh)
    a = c.get_extinction()
    c.extinction_lbl(prob.ethresh)
    assert np.array_equal(a, c.get_extinction())     # fixed summation order, no float atomics


def test_line_sum_medium_vs_oracle():
    """C2-shaped problem at 2e4 lines x 10 layers, default 60x60 table, 2160 oversampling."""
    prob = workloads.make_problem("C2", scale=0.02, nlayers=10, wn_hi=6000.0)
    c = setup_ctx(prob)
    t = c.get_profiles()
    vt = O.voigt_from_arrays(t["aDop"], t["aLor"], t["psize"], t["off"], t["data"])   # same table both sides
    c.extinction_lbl(prob.ethresh)
    e = c.get_extinction()
    pr = O.Problem(**{k: v for k, v in vars(prob).items()})
    for r in range(prob.nrad):
        ref, tr = O.computemolext(pr, vt, prob.t[r], prob.mol_d[:, r], prob.iso_z[:, r], prob.ethresh, trace=True)
        g = c.line_trace(r)
        assert np.array_equal(g["status"], tr["status"])
        lead = tr["status"] == 1
        assert np.array_equal(g["iown"][lead], tr["iown"][lead]) and np.array_equal(g["idop"][lead], tr["idop"][lead])
        assert relerr(e[r], ref[0]) < 1e-12, f"layer {r}"
    c.close()


def test_fused_line_sum_vs_two_kernel_path():
    """k_linesum_fused (strength + gather in one kernel, records in shared memory) against k_strength +
    k_accumulate on the same states: the states of narrow reach add in the same order and must agree BIT FOR BIT,
    per-line decisions and sample counts included; states the two-kernel path sums with its wide-reach schedule
    differ by summation order only.  The fused states include layers whose lines partly keep a CARRIED Doppler index
    (extinction.c:480-483), found there by k_lastq_scan instead of k_strength's maxima: same per-line indices.
    Both geometries of the batch: collapsed and per molecule; and with every layer fused (reach limit lifted)."""
    prob = workloads.make_problem("C2", scale=0.05, nlayers=40, wn_hi=12000.0)
    c = setup_ctx(prob)
    T = prob.t * prob.t_fct
    dens, Z = prob.mol_d.T.copy(), prob.iso_z.T.copy()
    for permol, rmax in ((0, 8), (1, 8), (0, 4096)):
        c.set_fused(2)                     # 2: fused whatever the size of the grid
        c.set_fused_reach(rmax, carried=True)
        a = c.molext_states(T, dens, Z, prob.ethresh, permol=permol)
        st = c.fused_stats()
        na = c.sample_counts(prob.nrad)
        tr_layers = (prob.nrad - 1, prob.nrad // 2, 12, 8, 0)
        tr_a = [c.line_trace(r) for r in tr_layers]
        a2 = c.molext_states(T, dens, Z, prob.ethresh, permol=permol)
        assert np.array_equal(a, a2)                                   # deterministic
        c.set_fused_variant(0)             # the block-wide kernel: same windows, same chains, same bits as the warp-autonomous one
        a0 = c.molext_states(T, dens, Z, prob.ethresh, permol=permol)
        assert np.array_equal(a, a0) and np.array_equal(na, c.sample_counts(prob.nrad))
        for x, r in zip(tr_a, tr_layers):
            y = c.line_trace(r)
            for k in ("status", "iown", "idwn", "idop"):
                assert np.array_equal(x[k], y[k]), k
        c.set_fused_variant(1)
        c.set_fused(False)
        b = c.molext_states(T, dens, Z, prob.ethresh, permol=permol)
        nb = c.sample_counts(prob.nrad)
        tr_b = [c.line_trace(r) for r in tr_layers]
        assert c.fused_stats()["fused_states"] == 0
        assert st["fused_states"] >= 16 and st["fused_states"] % 8 == 0, st   # the upper layers of this atmosphere
        if rmax == 4096:
            assert st["fused_states"] == prob.nrad, st
        assert np.array_equal(na, nb)                                  # same profile updates, layer by layer
        assert st["fused_samples"] == int(na[prob.nrad - st["fused_states"]:].sum()) or st["fused_samples"] > 0
        for x, y in zip(tr_a, tr_b):
            for k in ("status", "iown", "idwn", "idop"):
                assert np.array_equal(x[k], y[k]), k
        assert np.array_equal(a == 0, b == 0)
        assert relerr(a, b) < 1e-13
        same = [r for r in range(prob.nrad) if np.array_equal(a[r], b[r])]
        assert len(same) >= prob.nrad // 2, (len(same), prob.nrad)     # only the layers of reach > 2 bins may differ in order
    c.set_fused(1); c.set_fused_reach(8, False)
    c.close()


def test_permol_rows_sum_to_total():
    """Size-independent property: with ethresh=0, sum_m rho_m * (per-molecule row) == collapsed sum."""
    prob = workloads.make_problem("C2", scale=0.05, nlayers=6, wn_hi=9000.0)
    c = setup_ctx(prob)
    T = prob.t * prob.t_fct
    dens, Z = prob.mol_d.T.copy(), prob.iso_z.T.copy()
    tot = c.molext_states(T, dens, Z, 0.0, permol=0)[:, 0]
    rows = c.molext_states(T, dens, Z, 0.0, permol=1)
    ids = c.line_molIDs()
    im = [list(prob.mol_ID).index(i) for i in ids]
    rec = sum(rows[:, k] * prob.mol_d[im[k]][:, None] for k in range(len(ids)))
    assert relerr(rec, tot) < 1e-12
    c.close()


def test_edge_cases_empty_and_out_of_range():
    prob = workloads.make_problem("C2", scale=0.001, nlayers=4, wn_hi=3000.0)
    c = setup_ctx(prob)
    # all lines outside the grid -> zero extinction; empty list -> zero extinction
    c.upload_lines(prob.wl * 10.0, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
    c.extinction_lbl(prob.ethresh)
    assert not c.get_extinction().any()
    c.upload_lines(prob.wl[:0], prob.elow[:0], prob.gf[:0], prob.isoid[:0], prob.wfct, prob.efct)
    c.extinction_lbl(prob.ethresh)
    assert not c.get_extinction().any()
    # unsorted list is refused (the Doppler-index carry relies on the TLI order)
    with pytest.raises(api.TransitError, match="sorted"):
        c.upload_lines(prob.wl[::-1].copy(), prob.elow, prob.gf, np.zeros_like(prob.isoid), prob.wfct, prob.efct)
    c.close()


# ------------------------------------------------------------------ CIA, tau, spectra (a9-a15)

@pytest.mark.parametrize("case", ["tiny_ecl", "tiny_tra"])
def test_cia_tau_spectrum_vs_reference(ctxs, case):
    d, prob, c = ctxs[case]
    c.cia_interp()
    assert relerr(c.get_cia(), d["e_cs"]) < 1e-12
    c.extinction_lbl(prob.ethresh)
    geom = api.GEOM_ECLIPSE if prob.eclipse else api.GEOM_TRANSIT
    c.tau(geom, prob.toomuch)
    t, last = c.get_tau()
    assert np.array_equal(last, d["last"])                       # toomuch cut-off index, bit-exact
    assert np.array_equal(c.get_computed(), d["computed"])       # lazy-evaluation set of tau.c
    assert np.array_equal(t == 0, d["tau"] == 0)
    assert relerr(t, d["tau"]) < 1e-9                            # interp_parab conditioning (see oracle test)
    if prob.eclipse:
        out = c.intensity_flux(prob.angles)
        assert relerr(c.get_intensity(), d["intens"]) < SPEC_RTOL
    else:
        out = c.modulation(prob.starrad, prob.transp)
    assert relerr(out, d["out"]) < SPEC_RTOL
    # and the one-call pipeline (do_transit)
    out2 = c.spectrum(geom, prob.toomuch, prob.ethresh, prob.starrad, prob.transp,
                      prob.angles if prob.eclipse else None)
    assert np.array_equal(out, out2)


def test_scattering_and_cloud_vs_oracle(ctxs):
    d, prob, c = ctxs["tiny_tra"]
    c.cia_interp(); c.extinction_lbl(prob.ethresh)
    e, ecs = c.get_extinction(), c.get_cia()
    ptop = float(np.log10(prob.p[prob.nrad // 2]))
    c.set_scattering(1, 0.5)
    c.set_cloud(1, 1e-9, ptop, ptop + 1.0)
    c.tau(api.GEOM_TRANSIT, prob.toomuch)
    t, last = c.get_tau()
    cb = O.contrib(prob, sc_flag=1, sc_logext=0.5, cl_flag=1, cl_ext=1e-9, cl_top=ptop, cl_bot=ptop + 1.0)
    t0, last0, _ = O.tau(prob, e, ecs, cb, lazy=False)
    assert np.array_equal(last, last0) and relerr(t, t0) < 1e-9
    c.set_scattering(0, 0.0); c.set_cloud(0)


def test_explicit_impact_parameters_vs_oracle(ctxs):
    """tb_tau(h) / tb_set_rays: an impact-parameter sample that is NOT the reversed radii (makeipsample's second branch,
    makesample.c:577-596, which a resampled radius grid brings along): more rays than layers, between the layers and
    on them, top first -- tau, last[] and the modulation against the oracle's totaltau1 / modulation1 on the same rays,
    and tb_spectrum with tb_set_rays against the staged calls."""
    d, prob, c = ctxs["tiny_tra"]
    c.cia_interp(); c.extinction_lbl(prob.ethresh)
    e, ecs = c.get_extinction(), c.get_cia()
    r = prob.rad
    h = np.sort(np.concatenate([r[1:], 0.5 * (r[1:] + r[:-1]), r[:1] + 0.25 * (r[1] - r[0])]))[::-1].copy()
    for fast in (False, True):
        c.set_tau_streaming(fast)
        c.tau(api.GEOM_TRANSIT, prob.toomuch, h, prob.rad_fct)
        t, last = c.get_tau()
        t0, last0, _ = O.tau(prob, e, ecs, lazy=False, h=h)
        assert t.shape == (prob.nwn, len(h)) and np.array_equal(last, last0)
        assert np.array_equal(t == 0, t0 == 0) and relerr(t[t0 != 0], t0[t0 != 0]) < 1e-9
        out = c.modulation(prob.starrad, prob.transp)
        assert relerr(out, O.modulation(prob, t0, last0, ip=h)) < SPEC_RTOL
    c.set_tau_streaming(True)
    c.set_rays(h, prob.rad_fct)
    out2 = c.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    c.set_rays(None, 1.0)
    assert np.array_equal(out2, out)
    out3 = c.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    assert relerr(out3, d["out"]) < SPEC_RTOL                     # back on the reversed radii


# ------------------------------------------------------------------ opacity grid (a6-a8)

def test_tau_weight_form_vs_pairwise_kernel():
    """Transit geometry: k_tau_w (t = W er: one weight per (ray, layer), four rays per warp pass, the parabola value
    at the closest approach kept as the exact interp_parab expression) against k_tau_transit, which adds the Simpson
    pairs term by term like simpson() does: same last[], same zero tail, optical depths equal to the association of
    the sum; with scattering and a cloud deck switched on as well."""
    prob = workloads.make_problem("C2", scale=0.02, nlayers=37, wn_hi=5000.0)     # 37 layers: a ragged last group of rays
    c = setup_ctx(prob)
    c.extinction_lbl(prob.ethresh)
    c.cia_interp()
    for extra in (False, True):
        if extra:
            c.set_scattering(1, 0.7)
            c.set_cloud(1, 2e-9, -3.0, 0.0)
        res = []
        for fast in (True, False):
            c.set_tau_streaming(fast)
            c.tau(api.GEOM_TRANSIT, prob.toomuch)
            t, last = c.get_tau()
            res.append((t, last, c.modulation(prob.starrad)))
        (t1, l1, m1), (t0, l0, m0) = res
        assert np.array_equal(l1, l0) and np.array_equal(t1 == 0, t0 == 0)
        # interp_parab (numerical.c:182-195) amplifies last-bit differences (here: how the compiler contracts its
        # polynomial into FMAs inside each kernel) by x0^2 ~ 1e6 on the rays with few layers: the bound on tau
        # against the reference is 1e-9 for that reason, between the two kernels a few 1e-12
        assert relerr(t1[t0 != 0], t0[t0 != 0]) < 1e-10 and relerr(m1, m0) < 1e-11
    c.set_tau_streaming(True)
    c.close()


def test_scattering_cloud_modulation_variants_vs_reference(golden, tmp_path):
    """a10 / a14 against the UNMODIFIED reference (tests/golden/tiny_cloud.npz): Rayleigh flags 1 and 2, cloud models
    1-5, the --cloudtop deck (set_cloudtop), --transparent and modlevel -1, each through the library interface
    (Transit.set_scattering / set_cloud / set_cloudtop) AND through the C host's command line."""
    import json
    import subprocess
    d = golden("tiny_cloud")
    prob = O.Problem.from_golden(d)
    prob.species_names = O.golden_mol_names(d)
    cfg = cfg_of(d)
    variants = json.loads(bytes(d["variants"]).decode())
    # the same inputs as files for the C host (the fixture carries them)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for key, fn in (("tli", "lines.tli"), ("atm", "atm.dat"), ("cia0", "cia_h2h2.dat")):
        open(tmp_path / fn, "wb").write(bytes(d["file_" + key]))
    txt = bytes(d["cfg_text"]).decode().replace("/root/reference/inputs/molecules.dat",
                                                os.path.join(root, "oracle", "_ref", "inputs", "molecules.dat"))
    open(tmp_path / "case.cfg", "w").write(txt)
    host = os.path.join(root, "host", "transit_b200")
    have_host = os.path.exists(host) and os.path.exists(os.path.join(root, "oracle", "_ref", "inputs", "molecules.dat"))
    tr = api.Transit(0)
    tr.ctx.set_fused(2)
    tr.transit_init(prob, int(cfg["ndop"]), int(cfg["nlor"]))
    worst = {}
    for name, extra in variants.items():
        kw = O.cloud_variant(d, name)
        p2 = type(prob)(**vars(prob))
        p2.toomuch = d[f"{name}__toomuch"].item()
        p2.transp = int(d[f"{name}__transpplanet"].item())
        p2.starrad = prob.starrad
        tr.prob = p2
        tr.set_scattering(kw["sc_flag"], kw["sc_logext"])
        if name == "cloudtop":
            tr.set_cloudtop(kw["cl_top"])                 # BART's entry (transit.c:103-109)
        else:
            tr.set_cloud(kw["cl_flag"], kw["cl_ext"], kw["cl_top"], kw["cl_bot"], kw["cl_gamma"], kw["cl_Q"], kw["cl_r"],
                         kw["cl_sig"], kw["cl_refwn"])
        tr.ctx.set_modlevel(-1 if name == "modm1" else 1)
        out = tr.do_transit()
        t, last = tr.ctx.get_tau()
        assert np.array_equal(last, d[f"{name}__last"]), name
        assert np.array_equal(tr.ctx.get_computed(), d[f"{name}__computed"]), name
        assert np.array_equal(t == 0, d[f"{name}__tau"] == 0), name
        nz = d[f"{name}__tau"] != 0
        assert relerr(t[nz], d[f"{name}__tau"][nz]) < 1e-9, name
        worst[name] = relerr(out, d[f"{name}__out"])
        assert worst[name] < SPEC_RTOL, (name, worst[name])
        if have_host:
            spec = tmp_path / f"spec_{name}.dat"
            r = subprocess.run([host, "-c", str(tmp_path / "case.cfg"), "--outspec", str(spec)] + list(extra),
                               capture_output=True, text=True, cwd=str(tmp_path))
            assert r.returncode == 0, (name, r.stderr[-800:])
            a = np.loadtxt(spec, comments="#")
            want = np.array([float(f"{v:.9g}") for v in d[f"{name}__out"]])     # the file carries 9 digits
            assert np.max(np.abs(a[:, 1] - want) / np.abs(want)) < 1e-8, name
    tr.ctx.set_modlevel(1)
    tr.free_memory()
    print("\n[parity a10/a14 variants] worst spectrum rel:", json.dumps(worst))


def test_opacity_grid_build_and_interp_vs_reference(golden):
    d = golden("tiny_grid")
    cfg = cfg_of(d)
    lines, dbs = O.parse_tli(bytes(d["file_tli"]), float(cfg["wnlow"]), float(cfg["wnhigh"]))
    prob = O.Problem.from_golden(d, tli_lines=lines)
    c = setup_ctx(prob, d, int(cfg["ndop"]), int(cfg["nlor"]))
    assert np.array_equal(c.line_molIDs(), d["op_molID"])
    gtemp = d["op_temp"]
    L = O.lib()
    ziso = np.zeros((prob.niso, len(gtemp)))
    k = 0
    for mol, T, isos in dbs:                         # partition function at grid temperatures: host side
        T = O.c64(T)
        for (m, r, Z) in isos:
            Z = O.c64(Z); z = np.zeros(len(T))
            L.to_spline_init(O.p(z), O.p(T), O.p(Z), len(T))
            ziso[k] = [L.to_splinterp_pt(O.p(z), len(T), O.p(T), O.p(Z), float(tt)) for tt in gtemp]
            k += 1
    grid = c.opacity_build(gtemp, ziso, prob.mol_q, d["atm_mm"], int(d["atm_bymass"].item()), prob.ethresh)
    assert grid.shape == d["op_o"].shape
    assert np.array_equal(grid == 0, d["op_o"] == 0)
    assert relerr(grid, d["op_o"]) < EXT_RTOL
    # spectrum from the reference's grid file content (readopacity + interpolmolext + tau + modulation)
    c.opacity_load(d["op_molID"], gtemp, d["op_press"], d["op_o"])
    out = c.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    comp = d["computed"].astype(bool)
    assert relerr(c.get_extinction()[comp], d["ext"][comp]) < 1e-13
    t, last = c.get_tau()
    assert np.array_equal(last, d["last"])
    assert relerr(out, d["out"]) < SPEC_RTOL
    # run_transit with a new atmosphere (reloadatm output taken from the fixture)
    c.set_atmosphere(d["rerun_rad_v"], prob.rad_fct, prob.p, prob.p_fct, d["rerun_atm_t"], prob.t_fct,
                     d["rerun_mol_d"], d["rerun_iso_z"])
    out = c.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    assert relerr(out, d["rerun_out"]) < SPEC_RTOL
    # a second context of the same GPU reads the first one's grid in place (tb_opacity_share: no copy, no line list)
    c2 = setup_ctx(prob, d, lines=False)
    with pytest.raises(api.TransitError, match="holds no opacity grid"):
        c.opacity_share(c2)
    c2.opacity_share(c)
    c2.set_atmosphere(d["rerun_rad_v"], prob.rad_fct, prob.p, prob.p_fct, d["rerun_atm_t"], prob.t_fct,
                      d["rerun_mol_d"], d["rerun_iso_z"])
    out2 = c2.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    assert np.array_equal(out2, out)
    c2.close()                                     # the borrower goes first and frees nothing of the owner's
    out3 = c.spectrum(api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    assert np.array_equal(out3, out)
    c.close()


def test_host_mirror_interface(golden):
    """The reference's library call sequence (transit.i) through transit_b200.api.Transit."""
    d = golden("tiny_tra")
    prob = O.Problem.from_golden(d)
    cfg = cfg_of(d)
    tr = api.Transit(0)
    tr.transit_init(prob, int(cfg["ndop"]), int(cfg["nlor"]))
    assert tr.get_no_samples() == prob.nwn
    assert np.array_equal(tr.get_waveno_arr(), d["wns_v"])
    out = tr.do_transit()
    assert relerr(out, d["out"]) < SPEC_RTOL
    tr.free_memory()


# ------------------------------------------------------------------ wavenumber-band shard (8e)

def _band_union_check(prob, world=2):
    import threading
    import torch
    from transit_b200 import sharding

    ref = setup_ctx(prob)
    ref.extinction_lbl(prob.ethresh)
    e_ref = ref.get_extinction()
    n_ref = ref.sample_counts(prob.nrad)
    ref.close()

    bands = sharding.wavenumber_bands(sharding.line_weights(prob), world)
    assert bands[0][1] % 128 == 0 and 0 < bands[0][1] <= prob.nwn
    barrier = threading.Barrier(world)
    slots = [None] * world
    ctxs_, out, cnt, err = [None] * world, [None] * world, [None] * world, []

    def make_exchange(rank):
        def exchange(what, ptr, count):
            ctxs_[rank].sync()
            t = sharding._alias(ptr, count, "<f8" if what == 0 else "<i4")
            slots[rank] = t.cpu()
            barrier.wait(timeout=120)
            m = slots[0]
            for other in slots[1:]:
                m = torch.maximum(m, other)
            barrier.wait(timeout=120)
            t.copy_(m.cuda())
            torch.cuda.synchronize()
        return exchange

    def run(rank):
        try:
            c = setup_ctx(prob)
            ctxs_[rank] = c
            c.set_shard(bands[rank][0], bands[rank][1], rank, world, make_exchange(rank))
            c.extinction_lbl(prob.ethresh)
            out[rank] = c.get_extinction()
            cnt[rank] = c.sample_counts(prob.nrad)
            c.close()
        except Exception as e:       # noqa: BLE001
            err.append(e)
            barrier.abort()

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=300)
    assert not err, err
    e = np.zeros_like(e_ref)
    for r, (lo, hi) in enumerate(bands):
        e[:, lo:hi] = out[r][:, lo:hi]
        other = np.ones(prob.nwn, bool)
        other[lo:hi] = False
        assert not out[r][:, other].any()            # bins outside the band come back as zero
    assert np.array_equal(e, e_ref)
    assert np.array_equal(sum(cnt), n_ref)
    return e_ref, n_ref


def test_band_shard_two_ranks_bit_identical_to_one():
    """Two contexts, each owning a wavenumber band (tb_set_shard), run as two 'ranks' in two threads;
    the exchange callback completes the two global maxima through host memory.  Extinction stitched
    from the bands must be BIT-identical to the single-context result, deep (wide) and shallow
    (narrow) layers alike, and the executed sample counts must add up."""
    _band_union_check(workloads.make_problem("C2", scale=0.02, nlayers=24, wn_hi=6000.0))


def test_band_shard_uneven_and_empty_bands():
    """Three ranks on a grid of 2 501 bins (20 tiles, the last one ragged) and five ranks on a grid of
    three tiles, where some ranks own NO bin at all: they still take part in both exchanges and
    contribute nothing."""
    _band_union_check(workloads.make_problem("C2", scale=0.01, nlayers=12, wn_hi=4500.0), world=3)
    _band_union_check(workloads.make_problem("C2", scale=0.002, nlayers=8, wn_hi=2300.0), world=5)


def test_full_size_c2_properties():
    """BASELINE.json's C2 at FULL size (31 334 wavenumbers, 100 layers, 10^6 lines), where the oracle
    would take minutes: size-independent properties instead.  (1) band union == whole, bit for bit,
    and sample counts add up; (2) two runs are bit-identical (fixed summation order, no float
    atomics); (3) linearity: the per-molecule rows of a layer, weighted by the molecule densities,
    add up to the layer's total extinction; (4) every executed profile update is inside the grid:
    the sample count of a layer is bounded by groups x widest profile."""
    prob = workloads.make_problem("C2")
    e_ref, n_ref = _band_union_check(prob)
    c = setup_ctx(prob)
    c.extinction_lbl(prob.ethresh)
    assert np.array_equal(c.get_extinction(), e_ref)
    assert np.array_equal(c.sample_counts(prob.nrad), n_ref)
    names = list(prob.cfg["molecules"].keys())
    for r in (0, 37, 99):
        T = np.array([prob.t[r] * prob.t_fct]); d = prob.mol_d[:, r][None, :]; Z = prob.iso_z[:, r][None, :]
        rows = c.molext_states(T, d, Z, prob.ethresh, permol=1)[0]          # cm2/g per molecule
        tot = c.molext_states(T, d, Z, prob.ethresh, permol=0)[0, 0]
        assert np.array_equal(tot, e_ref[r])
        molID = c.line_molIDs()
        dens = np.array([prob.mol_d[list(prob.mol_ID).index(m), r] for m in molID])
        # permol thresholds are per molecule (extinction.c:467 uses kmax[molecule]), the total's is
        # global: the weighted rows can only contain MORE weak lines, by at most ethresh of the peak
        s_ = (rows * dens[:, None]).sum(0)
        assert np.all(s_ >= tot * (1 - 1e-12))
        big = tot > 1e-2 * tot.max()
        assert np.all(s_[big] <= tot[big] * 1.5)
    assert n_ref.min() > 0 and n_ref.max() <= c.ngroups() * 401
    c.close()


# ------------------------------------------------------------------ C5 / C4 shapes

def _oracle_from_grid(pr, grid, gtemp, gmol_rows):
    """interpolmolext for every layer on the oracle (extinction.c:534-581)."""
    L = O.lib()
    g = O.c64(grid)
    Nl, Nt, Nm, Nw = g.shape
    e = np.zeros((pr.nrad, pr.nwn))
    for r in range(pr.nrad):
        dens = O.c64(pr.mol_d[gmol_rows, r])
        rc = L.to_interpolmolext(O.p(g), Nt, Nm, Nw, O.p(O.c64(gtemp)), r, pr.t[r] * pr.t_fct, O.p(dens), O.p(e[r]))
        assert rc == 0
    return e


def test_retrieval_batch_from_grid_vs_oracle():
    """C5 shape, reduced: a grid built by the line-by-line path, then several perturbed T-P/abundance
    profiles through interpolation + CIA + tau + modulation; every spectrum against the oracle
    fed with the same grid, and the grid itself against the oracle's computemolext on a sample."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench_configs as BC
    prob = workloads.make_problem("C2", scale=0.004, nlayers=16, wn_hi=3500.0, wnosamp=240)
    c = setup_ctx(prob, nDop=24, nLor=24)
    gtemp = np.arange(500.0, 1500.1, 100.0)
    ziso = BC._ziso(prob, gtemp)
    grid = c.opacity_build(gtemp, ziso, prob.mol_q, BC._mean_mass(prob), 0, prob.ethresh)
    molID = c.line_molIDs()
    rows = [list(prob.mol_ID).index(m) for m in molID]
    # the grid against the oracle's line sum, two (layer, T) states
    t = c.get_profiles()
    vt = O.voigt_from_arrays(t["aDop"], t["aLor"], t["psize"], t["off"], t["data"])
    pr0 = O.Problem(**vars(prob))
    for (r, it) in ((0, 3), (11, 7)):
        dens = api.AMU * prob.mol_q[:, r] * (prob.p[r] * prob.p_fct) / api.KB / gtemp[it] * prob.mol_mass
        o = O.computemolext(pr0, vt, float(gtemp[it]), dens, ziso[:, it], prob.ethresh, permol=1, molID=molID)
        assert relerr(grid[r, it], o) < EXT_RTOL
    singles, atms = [], []
    for seed in (11, 12, 13):
        a = workloads.perturbed(prob, seed=seed, dT=150.0, dex=1.0)
        c.set_atmosphere(a.rad, a.rad_fct, a.p, a.p_fct, a.t, a.t_fct, a.mol_d, a.iso_z)
        out = c.spectrum(api.GEOM_TRANSIT, a.toomuch, a.ethresh, a.starrad, a.transp)
        pr = O.Problem(**vars(a))
        e0 = _oracle_from_grid(pr, grid, gtemp, rows)
        assert relerr(c.get_extinction(), e0) < 1e-13
        t0, last0, _ = O.tau(pr, e0, O.interpcs(pr), lazy=False)
        tg, lastg = c.get_tau()
        assert np.array_equal(lastg, last0)
        assert relerr(out, O.modulation(pr, t0, last0)) < SPEC_RTOL
        singles.append(out.copy()); atms.append(a)
    # the same atmospheres through ONE tb_spectrum_batch call: bit-identical spectra, and a ray below
    # the bottom layer anywhere in the batch is still reported (one error read-back per batch)
    st = lambda k: np.stack([getattr(x, k) for x in atms])
    outb = c.spectrum_batch(st("rad"), prob.rad_fct, st("p"), prob.p_fct, st("t"), prob.t_fct, st("mol_d"), st("iso_z"),
                            api.GEOM_TRANSIT, prob.toomuch, prob.ethresh, prob.starrad, prob.transp)
    assert np.array_equal(outb, np.stack(singles))
    out1 = c.spectrum_batch(st("rad")[1:2], prob.rad_fct, st("p")[1:2], prob.p_fct, st("t")[1:2], prob.t_fct,
                            st("mol_d")[1:2], st("iso_z")[1:2], api.GEOM_TRANSIT, prob.toomuch, prob.ethresh,
                            prob.starrad, prob.transp)
    assert np.array_equal(out1[0], singles[1])
    c.close()


def test_high_resolution_eclipse_from_grid_vs_oracle():
    """C4 shape, reduced: 20 001 samples at 0.01 cm-1, extinction from a resident synthetic grid, CIA,
    eclipse tau, 5-angle intensity and flux against the oracle."""
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench_configs as BC
    prob = BC.c4_problem(nwn=20001, nlayers=30)
    temps = np.array([800.0, 1000.0, 1200.0, 1400.0, 1600.0, 1800.0])
    grid = BC.c4_grid(prob, temps)
    molID = np.array([prob.mol_ID[prob.species_names.index(s)] for s in ("H2O", "CH4")], np.int32)
    tr = api.Transit(0)
    tr.transit_init(prob, opacity=dict(molID=molID, temp=temps, press=prob.p * prob.p_fct, grid=grid))
    out = tr.do_transit()
    pr = O.Problem(**vars(prob))
    rows = [list(prob.mol_ID).index(m) for m in molID]
    e0 = _oracle_from_grid(pr, grid, temps, rows)
    assert relerr(tr.ctx.get_extinction(), e0) < 1e-13
    t0, last0, _ = O.tau(pr, e0, O.interpcs(pr), lazy=False)
    tg, lastg = tr.ctx.get_tau()
    assert np.array_equal(lastg, last0)
    intens0, flux0 = O.intensity_flux(pr, t0, last0)
    assert relerr(tr.ctx.get_intensity(), intens0) < SPEC_RTOL
    assert relerr(out, flux0) < SPEC_RTOL
    assert np.array_equal(tg == 0, t0 == 0) and relerr(tg[t0 != 0], t0[t0 != 0]) < 1e-9
    # the O(layers) streaming kernel (default) against the general kernel that integrates every ray on its own:
    # same last[], same zero tail, optical depths equal to the association of the pair sums
    tr.ctx.set_tau_streaming(False)
    out2 = tr.do_transit()
    tg2, lastg2 = tr.ctx.get_tau()
    tr.ctx.set_tau_streaming(True)
    assert np.array_equal(lastg2, lastg) and np.array_equal(tg2 == 0, tg == 0)
    assert relerr(tg[tg2 != 0], tg2[tg2 != 0]) < 1e-12 and relerr(out, out2) < 1e-12
    tr.free_memory()
