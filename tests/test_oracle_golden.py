"""Pins oracle/transit_oracle.c against the UNMODIFIED reference (tests/golden/*.npz,
produced by tests/golden/make_golden.py from oracle/_ref/ref_dump).  CPU only."""
import numpy as np
import pytest

import oraclelib as O


def relerr(a, b):
    d = np.abs(a - b)
    s = np.maximum(np.abs(a), np.abs(b))
    return float(np.max(np.where(s > 0, d / np.where(s > 0, s, 1), 0.0)))


CASES = ["tiny_ecl", "tiny_tra"]


@pytest.mark.parametrize("case", CASES + ["tiny_grid"])
def test_voigt_table_bit_exact(golden, case):
    d = golden(case)
    prob = O.Problem.from_golden(d, tli_lines={})
    cfg = dict(l.split(None, 1) for l in bytes(d["cfg_text"]).decode().splitlines())
    # dmin..lmax are `float` in struct transithint (structures_tr.h:333)
    f = lambda x: float(np.float32(x))
    vt = O.OracleVoigt(prob, int(cfg["ndop"]), int(cfg["nlor"]), f(1e-3), f(0.25), f(1e-4), f(10.0))
    # the reference's -ffast-math turns pow(10,x) into exp10-style code: a few ulp apart
    assert relerr(vt.aDop, d["aDop"]) < 2e-15 and relerr(vt.aLor, d["aLor"]) < 2e-15
    assert np.array_equal(vt.psize, d["profsize"])
    assert np.array_equal(vt.off, d["prof_off"])
    # float table: bit-exact against the reference's (built with -ffast-math, long double)
    assert vt.data.size == d["prof_data"].size
    nbad = int(np.count_nonzero(vt.data != d["prof_data"]))
    assert nbad == 0, f"{nbad} of {vt.data.size} float samples differ"


@pytest.mark.parametrize("case", CASES)
def test_computemolext_all_layers(golden, case):
    d = golden(case)
    prob = O.Problem.from_golden(d)
    vt = O.voigt_from_arrays(d["aDop"], d["aLor"], d["profsize"], d["prof_off"], d["prof_data"])
    worst = 0.0
    for r in range(prob.nrad):
        e = O.computemolext(prob, vt, prob.t[r] * prob.t_fct, prob.mol_d[:, r], prob.iso_z[:, r],
                            prob.ethresh)
        ref = d["ext_full"][r]
        assert np.array_equal(e[0] == 0, ref == 0), f"layer {r}: support differs (bin assignment)"
        worst = max(worst, relerr(e[0], ref))
    assert worst < 1e-12, worst


@pytest.mark.parametrize("case", CASES + ["tiny_grid"])
def test_interpcs(golden, case):
    d = golden(case)
    prob = O.Problem.from_golden(d, tli_lines={})
    e_cs = O.interpcs(prob)
    assert relerr(e_cs, d["e_cs"]) < 1e-12


@pytest.mark.parametrize("case", CASES)
def test_tau_and_spectrum(golden, case):
    d = golden(case)
    prob = O.Problem.from_golden(d)
    t, last, comp = O.tau(prob, d["ext_full"], d["e_cs"], lazy=True)
    assert np.array_equal(comp, d["computed"])
    assert np.array_equal(last, d["last"])
    # interp_parab (numerical.c:182-195) evaluates a parabola with x0 = x[0]/dx ~ 1e3, which
    # amplifies last-bit (compiler-licence) differences to ~1e-11 relative on single rays
    assert relerr(t, d["tau"]) < 1e-9
    if prob.eclipse:
        intens, out = O.intensity_flux(prob, d["tau"], d["last"])
        assert relerr(intens, d["intens"]) < 1e-11
    else:
        out = O.modulation(prob, d["tau"], d["last"])
    assert relerr(out, d["out"]) < 1e-11


def test_opacity_grid_build(golden):
    """calcopacity (opacity.c:281-427): computemolext with permol=1 on the (layer, T) grid."""
    d = golden("tiny_grid")
    cfg = dict(l.split(None, 1) for l in bytes(d["cfg_text"]).decode().splitlines())
    lines, dbs = O.parse_tli(bytes(d["file_tli"]), float(cfg["wnlow"]), float(cfg["wnhigh"]))
    prob = O.Problem.from_golden(d, tli_lines=lines)
    vt = O.voigt_from_arrays(d["aDop"], d["aLor"], d["profsize"], d["prof_off"], d["prof_data"])
    L = O.lib()
    gtemp = d["op_temp"]; press = d["op_press"]; molID = d["op_molID"]
    assert np.array_equal(molID, prob.line_molIDs())
    # partition function splined to grid temperatures (opacity.c:325-339)
    ziso = np.zeros((prob.niso, len(gtemp)))
    k = 0
    for mol, T, isos in dbs:
        T = O.c64(T)
        for (m, r, Z) in isos:
            Z = O.c64(Z); z = np.zeros(len(T))
            L.to_spline_init(O.p(z), O.p(T), O.p(Z), len(T))
            for it, tt in enumerate(gtemp):
                ziso[k, it] = L.to_splinterp_pt(O.p(z), len(T), O.p(T), O.p(Z), float(tt))
            k += 1
    AMU, KB = 1.66053886e-24, 1.380658e-16
    worst = 0.0
    for r in range(0, prob.nrad, 3):
        for it in range(0, len(gtemp), 2):
            dens = AMU * prob.mol_q[:, r] * press[r] / KB / gtemp[it] * prob.mol_mass   # stateeqnford
            o = O.computemolext(prob, vt, float(gtemp[it]), dens, ziso[:, it], prob.ethresh,
                                permol=1, molID=molID)
            worst = max(worst, relerr(o, d["op_o"][r, it]))
    assert worst < 1e-12, worst


def test_interpolmolext_spectrum_from_grid(golden):
    d = golden("tiny_grid")
    prob = O.Problem.from_golden(d, tli_lines={})
    L = O.lib()
    grid = O.c64(d["op_o"]); gtemp = O.c64(d["op_temp"])
    Nl, Nt, Nm, Nw = grid.shape
    gm = [list(prob.mol_ID).index(m) for m in d["op_molID"]]
    e = np.zeros((prob.nrad, prob.nwn))
    for r in range(prob.nrad):
        dens = O.c64(prob.mol_d[gm, r])
        L.to_interpolmolext(O.p(grid), Nt, Nm, Nw, O.p(gtemp), r, prob.t[r] * prob.t_fct,
                            O.p(dens), O.p(e[r]))
    comp = d["computed"].astype(bool)
    assert relerr(e[comp], d["ext"][comp]) < 1e-13
    t, last, comp2 = O.tau(prob, e, d["e_cs"], lazy=True)
    assert np.array_equal(comp2, d["computed"]) and np.array_equal(last, d["last"])
    assert relerr(t, d["tau"]) < 1e-9
    assert relerr(O.modulation(prob, t, last), d["out"]) < 1e-11


def test_scattering_clouds_and_modulation_variants(golden):
    """a10 / a14 pinned to the UNMODIFIED reference (tests/golden/tiny_cloud.npz: one problem through the reference
    once per variant): computeextscat flags 1 and 2, computeextcloud models 1-5, the --cloudtop deck
    (set_cloudtop's fields), --transparent and modlevel -1 -- tau, last[], the lazily evaluated layer set and the
    spectrum of every variant."""
    import json
    d = golden("tiny_cloud")
    prob = O.Problem.from_golden(d)
    names = O.golden_mol_names(d)
    assert names[:2] == ["H2", "He"]
    nH, md = O.cloud_aux(prob, names)
    variants = json.loads(bytes(d["variants"]).decode())
    assert set(variants) >= {"sc1", "sc2", "cl1", "cl2", "cl3", "cl4", "cl5", "cloudtop", "transparent", "modm1"}
    base_out = d["none__out"]
    for name in variants:
        kw = O.cloud_variant(d, name)
        toomuch = d[f"{name}__toomuch"].item()
        cb = O.contrib(prob, nH=nH, md=md, **kw)
        t, last, comp = O.tau(prob, d["ext_full"], d["e_cs"], cb, lazy=True, toomuch=toomuch)
        assert np.array_equal(last, d[f"{name}__last"]), name
        assert np.array_equal(comp, d[f"{name}__computed"]), name
        assert relerr(t, d[f"{name}__tau"]) < 1e-9, (name, relerr(t, d[f"{name}__tau"]))
        out = O.modulation(prob, d[f"{name}__tau"], d[f"{name}__last"], transp=int(d[f"{name}__transpplanet"].item()),
                           modlevel=-1 if name == "modm1" else 1, toomuch=toomuch)
        assert relerr(out, d[f"{name}__out"]) < 1e-11, (name, relerr(out, d[f"{name}__out"]))
        if name not in ("none",):
            assert not np.array_equal(d[f"{name}__out"], base_out), f"variant {name} does not change the spectrum"
