"""The C host (host/): the reference's CLI / cfg syntax / file formats and library interface on
top of libtransit_b200.so.  CPU tests cover the option machinery and the loud failure without a
GPU; the GPU tests run the same cfg through `host/transit_b200` and through the UNMODIFIED
reference binary / library (oracle/_ref, built from /root/reference by oracle/Makefile) and
compare the spectrum files and run_transit outputs."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from transit_b200 import workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host", "transit_b200")
HOSTLIB = os.path.join(ROOT, "host", "libtransit_host.so")
# test builds with the dry-run tap compiled in (-DTH_TEST_TAPS, host/Makefile); the shipped host has none
HOST_TAPS = os.path.join(ROOT, "host", "transit_b200_taps")
HOSTLIB_TAPS = os.path.join(ROOT, "host", "libtransit_host_taps.so")
REF = os.path.join(ROOT, "oracle", "_ref", "transit")
REFLIB = os.path.join(ROOT, "oracle", "_ref", "libtransit_ref.so")


def _build():
    if not (os.path.exists(HOST) and os.path.exists(HOST_TAPS) and os.path.exists(HOSTLIB_TAPS)):
        from transit_b200 import api
        api.build_library()
        subprocess.run(["make", "-C", os.path.join(ROOT, "host")], check=True, stdout=subprocess.DEVNULL)


def _inputs(tmp_path, name, **kw):
    prob = workloads.make_problem(name, **kw)
    cfg, _ = workloads.write_reference_inputs(prob, str(tmp_path))
    return prob, cfg


def test_cli_version_and_unknown_option():
    _build()
    r = subprocess.run([HOST, "--version"], capture_output=True, text=True)
    assert r.returncode == 0 and "transit" in r.stdout
    r = subprocess.run([HOST, "--nosuchoption", "1"], capture_output=True, text=True)
    assert r.returncode != 0 and "option" in r.stderr


def test_cfg_prefix_match_and_loud_failure_without_gpu(tmp_path):
    """`ethresh` must select `ethreshold` (first prefix match in table order, procopt.c:665-675); a cfg
    with an unknown key is refused; with valid inputs and no GPU the run dies at tb_create with the
    library's message (no CPU fallback)."""
    _build()
    prob, cfg = _inputs(tmp_path, "C2", scale=0.002, nlayers=8, wn_hi=2300.0, wnosamp=40)
    bad = tmp_path / "bad.cfg"
    bad.write_text(open(cfg).read() + "nosuchkey 3\n")
    r = subprocess.run([HOST, "-c", str(bad)], capture_output=True, text=True)
    assert r.returncode != 0 and "nosuchkey" in r.stderr
    txt = open(cfg).read().replace("ethreshold", "ethresh")
    ok = tmp_path / "ok.cfg"
    ok.write_text(txt)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    r = subprocess.run([HOST, "-c", str(ok)], capture_output=True, text=True)
    assert r.returncode != 0 and ("CUDA" in r.stderr or "no CPU fallback" in r.stderr)


def test_shipped_host_has_no_dry_run_switch(tmp_path):
    """TRANSIT_B200_DRYRUN is a hook of the TEST build only: the shipped binary and library must not even
    contain the string, so no environment variable can turn their device calls into no-ops."""
    _build()
    for f in (HOST, HOSTLIB):
        assert b"TRANSIT_B200_DRYRUN" not in open(f, "rb").read(), f
    assert b"TRANSIT_B200_DRYRUN" in open(HOST_TAPS, "rb").read()


def _read_spec(path):
    return np.loadtxt(path, comments="#")


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", [("C2", dict(scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120)),
                                      ("C1", dict(scale=0.02, nlayers=24, wn_hi=2900.0, wnosamp=120))])
def test_cli_spectrum_file_matches_reference_binary(tmp_path, name, kw):
    _build()
    prob, cfg = _inputs(tmp_path, name, **kw)
    ours = tmp_path / "ours.dat"
    r = subprocess.run([HOST, "-c", cfg, "--outspec", str(ours)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([REF, "-c", cfg], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    ref = os.path.join(str(tmp_path), "spec_run.dat")
    a, b = _read_spec(str(ours)), _read_spec(ref)
    assert a.shape == b.shape and a.shape[0] == prob.nwn
    assert open(str(ours)).readline() == open(ref).readline()              # same header
    assert np.array_equal(a[:, 0], b[:, 0])                                  # wavelengths as printed
    assert np.max(np.abs(a[:, 1] - b[:, 1]) / np.abs(b[:, 1])) < 1e-8        # 9 printed digits


@pytest.mark.gpu
def test_library_interface_run_transit_matches_reference(tmp_path):
    """transit_init + get_no_samples + get_waveno_arr + run_transit (reloadatm path: new T and
    abundances, hydrostatic radii from refpress/refradius/gsurf) against libtransit_ref.so."""
    _build()
    prob = workloads.make_problem("C2", scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120)
    cfg, _ = workloads.write_reference_inputs(prob, str(tmp_path))
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[6]:.6e}\nrefradius {prob.rad[6]:.6f}\ngsurf 2200.0\n")
    argv = [b"transit", b"-c", cfg.encode()]
    inp = np.concatenate([prob.t * 1.01 + 3.0] + [prob.mol_q[j] for j in range(prob.nmol)])
    outs = []
    for lib in (HOSTLIB, REFLIB):
        code = (f"import ctypes as C, numpy as np, sys, os\n"
                f"os.dup2(os.open(os.devnull, os.O_WRONLY), 1)\n"
                f"L = C.CDLL({lib!r}); argv = {argv!r}; arr = (C.c_char_p * len(argv))(*argv)\n"
                f"L.transit_init(len(argv), arr); L.get_no_samples.restype = C.c_int; n = L.get_no_samples()\n"
                f"wn = np.zeros(n); out = np.zeros(n); inp = np.load({str(tmp_path / 'inp.npy')!r})\n"
                f"P = C.POINTER(C.c_double)\n"
                f"L.get_waveno_arr(wn.ctypes.data_as(P), n)\n"
                f"L.run_transit(inp.ctypes.data_as(P), len(inp), out.ctypes.data_as(P), n)\n"
                f"np.save({str(tmp_path)!r} + '/out_' + os.path.basename({lib!r}) + '.npy', np.stack([wn, out]))\n")
        np.save(tmp_path / "inp.npy", inp)
        r = subprocess.run(["python", "-c", code], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(str(tmp_path) + "/out_" + os.path.basename(lib) + ".npy"))
    assert np.array_equal(outs[0][0], outs[1][0])
    assert np.max(np.abs(outs[0][1] - outs[1][1]) / np.abs(outs[1][1])) < 1e-8


def _golden_files(tmp_path, case):
    """Input files a golden fixture carries + its cfg with the molecules file repointed to the shipped copy."""
    d = np.load(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    names = {"tli": "lines.tli", "atm": "atm.dat", "cia0": "cia_h2h2.dat", "cia1": "cia_h2he.dat", "opa": "ref.opa"}
    for k in d.files:
        if k.startswith("file_"):
            open(tmp_path / names[k[5:]], "wb").write(bytes(d[k]))
    txt = bytes(d["cfg_text"]).decode().replace("/root/reference/inputs/", os.path.join(ROOT, "oracle", "_ref", "inputs") + "/")
    return d, txt


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "inputs", "molecules.dat")),
                    reason="oracle/_ref/inputs not built")
def test_opacity_file_face_across_implementations(tmp_path):
    """a6 / a7 as FILES (opacity.c:405-423 writer, :432-503 reader): (1) the host builds a grid with --opacityfile
    and its file has the reference file's header byte for byte and its grid within 1e-10; (2) the host reads the file
    the REFERENCE wrote (tests/golden/tiny_grid.npz:file_opa) and reproduces the reference's spectrum; (3) the
    unmodified reference binary reads the file the HOST wrote and prints the spectrum it prints from its own file."""
    _build()
    d, txt = _golden_files(tmp_path, "tiny_grid")
    base = "".join(l + "\n" for l in txt.splitlines() if not l.startswith(("opacityfile", "outspec")))
    def run(binary, opa, spec, extra=()):
        cfg = tmp_path / f"c_{os.path.basename(spec)}.cfg"
        cfg.write_text(base + f"opacityfile {opa}\noutspec {spec}\n")
        r = subprocess.run([binary, "-c", str(cfg)] + list(extra), capture_output=True, text=True, cwd=str(tmp_path))
        assert r.returncode == 0, (binary, r.stdout[-1500:], r.stderr[-1500:])
    # (1) build with the host
    run(HOST, tmp_path / "ours.opa", tmp_path / "s_build.dat")
    ref_b, our_b = open(tmp_path / "ref.opa", "rb").read(), open(tmp_path / "ours.opa", "rb").read()
    assert len(ref_b) == len(our_b)
    Nmol, Nt, Nl, Nw = np.frombuffer(ref_b[:32], "<i8")
    hdr = 32 + 4 * Nmol + 8 * (Nt + Nl + Nw)
    assert ref_b[:hdr] == our_b[:hdr], "opacity file header (sizes, molID, temperatures, pressures, wavenumbers) differs"
    g_ref, g_our = np.frombuffer(ref_b[hdr:], "<f8"), np.frombuffer(our_b[hdr:], "<f8")
    assert g_ref.size == Nmol * Nt * Nl * Nw and np.array_equal(g_ref == 0, g_our == 0)
    nz = g_ref != 0
    assert np.max(np.abs(g_our[nz] - g_ref[nz]) / np.abs(g_ref[nz])) < 1e-10
    # (2) the host on the reference's file
    run(HOST, tmp_path / "ref.opa", tmp_path / "s_ours_on_ref.dat")
    a = _read_spec(str(tmp_path / "s_ours_on_ref.dat"))
    want = np.array([float(f"{v:.9g}") for v in d["out"]])
    assert a.shape[0] == want.size and np.max(np.abs(a[:, 1] - want) / np.abs(want)) < 1e-8
    # (3) the reference on the host's file, against the reference on its own file
    run(REF, tmp_path / "ours.opa", tmp_path / "s_ref_on_ours.dat")
    run(REF, tmp_path / "ref.opa", tmp_path / "s_ref_on_ref.dat")
    b, c = _read_spec(str(tmp_path / "s_ref_on_ours.dat")), _read_spec(str(tmp_path / "s_ref_on_ref.dat"))
    assert np.array_equal(b[:, 0], c[:, 0]) and np.max(np.abs(b[:, 1] - c[:, 1]) / np.abs(c[:, 1])) < 1e-8


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "inputs", "molecules.dat")),
                    reason="oracle/_ref/inputs not built")
@pytest.mark.parametrize("case", ["tiny_ecl", "tiny_tra"])
def test_side_files_match_reference(tmp_path, case):
    """a17: outtoomuch (printtoomuch, tau.c:611-640), outintens (printintens, eclipse.c:293-348), outsample with
    `savefiles yes` (makesample.c:598,743-772) and saveext (extinction.c:62-91) as the host writes them, against the
    files the unmodified reference writes from the same cfg: text files equal to the printed digits, the extinction
    savefile's magic and computed[] exact and its values within 1e-10."""
    _build()
    d, txt = _golden_files(tmp_path, case)
    base = "".join(l + "\n" for l in txt.splitlines() if not l.startswith("outspec"))
    res = {}
    for who, binary in (("ours", HOST), ("ref", REF)):
        sub = tmp_path / who
        sub.mkdir()
        cfg = tmp_path / f"{who}.cfg"
        cfg.write_text(base + f"outspec {sub}/spec.dat\nouttoomuch {sub}/toomuch.dat\noutsample {sub}/sample.dat\n"
                              f"outintens {sub}/intens.dat\nsaveext {sub}/ext.sav\nsavefiles yes\n")
        r = subprocess.run([binary, "-c", str(cfg)], capture_output=True, text=True, cwd=str(tmp_path))
        assert r.returncode == 0, (who, r.stdout[-1500:], r.stderr[-1500:])
        res[who] = sub
    def numbers(path):
        return [l.split() for l in open(path) if not l.startswith("#") and l.strip()]
    # sampling file: identical text
    assert open(res["ours"] / "sample.dat").read() == open(res["ref"] / "sample.dat").read()
    # toomuch: same header, same radius index and radius, optical depth to its 6 printed digits
    o, r = open(res["ours"] / "toomuch.dat").read().splitlines(), open(res["ref"] / "toomuch.dat").read().splitlines()
    assert o[:2] == r[:2] and len(o) == len(r)
    for lo, lr in zip(o[2:], r[2:]):
        fo, fr = lo.split(), lr.split()
        assert fo[0] == fr[0] and fo[2:] == fr[2:], (lo, lr)
        assert abs(float(fo[1]) - float(fr[1])) <= 2e-5 * abs(float(fr[1])), (lo, lr)
    # extinction savefile
    bo, br = open(res["ours"] / "ext.sav", "rb").read(), open(res["ref"] / "ext.sav", "rb").read()
    assert len(bo) == len(br) and bo[:5] == br[:5] == b"@E@S@"
    nrad, nwn = int(d["nrad"].item()), int(d["wns_n"].item())
    eo, er = np.frombuffer(bo[5:5 + 8 * nrad * nwn], "<f8"), np.frombuffer(br[5:5 + 8 * nrad * nwn], "<f8")
    assert bo[5 + 8 * nrad * nwn:] == br[5 + 8 * nrad * nwn:], "computed[] (shorts) differ"
    assert np.array_equal(eo == 0, er == 0)
    assert np.max(np.abs(eo[er != 0] - er[er != 0]) / np.abs(er[er != 0])) < 1e-10
    if case == "tiny_ecl":      # intensities per angle: header lines identical, values to the printed digits
        o, r = open(res["ours"] / "intens.dat").read().splitlines(), open(res["ref"] / "intens.dat").read().splitlines()
        assert o[:2] == r[:2] and len(o) == len(r)
        a, b = np.array([[float(x) for x in l.split()] for l in o[2:]]), np.array([[float(x) for x in l.split()] for l in r[2:]])
        assert np.array_equal(a[:, 0], b[:, 0]) and np.max(np.abs(a[:, 1:] - b[:, 1:]) / np.abs(b[:, 1:])) < 1e-8


def _read_dry(path):
    """The host's dry-run dump (th_api.c do_transit, TRANSIT_B200_DRYRUN): header of 8 doubles, then the
    sampled state in the order it is handed to the device."""
    a = np.fromfile(path, "<f8")
    nl, nm, ni, nwn, onwn, nlines = (int(x) for x in a[:6])
    d = dict(nlayer=nl, nmol=nm, niso=ni, nwn=nwn, onwn=onwn, nlines=nlines, wn_i=a[6], wn_d=a[7])
    pos = 8

    def take(n, shape=None):
        nonlocal pos
        v = a[pos:pos + n].copy(); pos += n
        return v if shape is None else v.reshape(shape)
    d["rad"], d["p"], d["t"], d["mm"] = take(nl), take(nl), take(nl), take(nl)
    d["dens"] = take(nm * nl, (nm, nl)); d["Z"] = take(ni * nl, (ni, nl))
    d["molmass"], d["molradius"] = take(nm), take(nm)
    d["isomass"], d["isoratio"] = take(ni), take(ni)
    if nlines:
        d["wl"], d["elow"], d["gf"] = take(nlines), take(nlines), take(nlines)
    assert pos == len(a)
    return d


@pytest.mark.parametrize("name,kw,win", [("C2", dict(scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120), (2300.0, 2700.5)),
                                          ("C1", dict(scale=0.02, nlayers=24, wn_hi=2900.0, wnosamp=120), (2600.0, 2800.0)),
                                          # three samples: most isotopologues have no line at all inside
                                          ("C2", dict(scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120), (2500.0, 2502.0))])
def test_host_inputs_without_gpu(tmp_path, name, kw, win):
    """Host logic on CPU through the dry-run tap (no device call is made): cfg + atmosphere file + TLI +
    molecules.dat in, and out come the wavenumber grid (makewnsample), the layers, densities
    (stateeqnford), per-layer partition functions (makeradsample splines) and the per-isotope
    wavelength-range selection of the TLI (readdatarng, readlineinfo.c:416-537).  The cfg asks for a
    window narrower than the database, so both ends of every isotope's block are cut."""
    _build()
    import oraclelib as O
    prob, cfg = _inputs(tmp_path, name, **kw)
    txt = "".join(l for l in open(cfg) if not l.startswith(("wnlow", "wnhigh")))
    cfg = str(tmp_path / "window.cfg")
    open(cfg, "w").write(txt + f"wnlow {win[0]}\nwnhigh {win[1]}\n")
    dump = tmp_path / "dry.f64"
    env = dict(os.environ, TRANSIT_B200_DRYRUN=str(dump))
    r = subprocess.run([HOST_TAPS, "-c", cfg], capture_output=True, text=True, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr[-800:]
    d = _read_dry(str(dump))
    nwn = int(O.lib().to_grid_count(win[0], win[1], prob.wn_d, 1))
    assert (d["nlayer"], d["nmol"], d["niso"]) == (prob.nrad, prob.nmol, prob.niso)
    assert (d["nwn"], d["onwn"]) == (nwn, (nwn - 1) * prob.wnosamp + 1)
    assert d["wn_i"] == win[0] and d["wn_d"] == prob.wn_d
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    assert rel(d["rad"], prob.rad) < 1e-6 and rel(d["p"], prob.p) < 1e-6 and rel(d["t"], prob.t) < 1e-6   # file digits
    assert np.array_equal(d["molmass"], prob.mol_mass) and rel(d["isomass"], prob.iso_mass) == 0
    assert rel(d["isoratio"], prob.iso_ratio) == 0
    # densities and partition functions from the values the HOST read (the files carry ~7 digits)
    AMU, KB = 1.66053886e-24, 1.380658e-16
    q = prob.mol_q
    dens = AMU * q * (d["p"] * prob.p_fct)[None, :] / (KB * d["t"][None, :] * prob.t_fct) * prob.mol_mass[:, None]
    assert rel(d["dens"], dens) < 2e-6
    assert rel(d["Z"], prob.iso_z) < 1e-5
    # TLI selection, in file order, bit for bit
    raw = open(os.path.join(str(tmp_path), "lines.tli"), "rb").read()
    allw, _ = O.parse_tli(raw, 1.0, 1e7)
    want, _ = O.parse_tli(raw, win[0], win[1])     # wns.i .. wns.f as hinted, not the last grid point (readlineinfo.c:302-303)
    assert 0 < len(want["wl"]) < len(allw["wl"])
    assert d["nlines"] == len(want["wl"])
    assert np.array_equal(d["wl"], want["wl"]) and np.array_equal(d["elow"], want["elow"])
    assert np.array_equal(d["gf"], want["gf"])
    # ... and the same inputs through the UNMODIFIED reference (oracle/_ref/ref_dump, here only): what the host
    # hands to the device is what the reference's own readers and samplers produce
    ref_dump = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
    if os.path.exists(ref_dump):
        import sys
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        import make_golden as MG
        r = subprocess.run([ref_dump, "ref.bin", "--", "-c", cfg], cwd=str(tmp_path), capture_output=True, text=True)
        assert r.returncode == 0, r.stdout[-500:]
        g = MG.read_container(os.path.join(str(tmp_path), "ref.bin"))
        assert int(g["wns_n"].item()) == d["nwn"] and int(g["owns_n"].item()) == d["onwn"]
        assert np.array_equal(g["line_wl"], d["wl"]) and np.array_equal(g["line_elow"], d["elow"])
        assert np.array_equal(g["line_gf"], d["gf"])
        assert np.array_equal(g["rad_v"], d["rad"]) and np.array_equal(g["atm_t"], d["t"])
        # the reference passes p, densities and abundances through splinterp at the layers' own radii
        # (makesample.c:508-530), compiled with -ffast-math: the file's values come back within an ulp or two
        assert rel(d["p"], g["atm_p"]) < 1e-14
        assert rel(d["dens"], g["mol_d"]) < 1e-14
        assert rel(d["Z"], g["iso_z"]) < 1e-14


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_dump")),
                    reason="oracle/_ref/ref_dump not built")
def test_host_run_transit_setup_without_gpu(tmp_path):
    """run_transit's per-atmosphere setup on CPU (dry-run tap of the test build libtransit_host_taps.so): reloadatm with new
    temperatures and abundances (readatm.c:722-784), densities, hydrostatic radii from
    refpress/refradius/gsurf (radpress, readatm.c:787-865) and the per-layer partition functions, against
    the same call on the UNMODIFIED reference (ref_dump --rerun)."""
    _build()
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as MG
    prob = workloads.make_problem("C2", scale=0.004, nlayers=16, wn_hi=2600.0, wnosamp=60)
    tmp = str(tmp_path)
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[5]:.6e}\nrefradius {prob.rad[5]:.6f}\ngsurf 2200.0\n")
    inp = np.concatenate([prob.t * 1.01 + 3.0] + [prob.mol_q[j] for j in range(prob.nmol)])
    inp.astype("<f8").tofile(os.path.join(tmp, "rerun.f64"))
    np.save(tmp_path / "inp.npy", inp)
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_dump"), "ref.bin", "--rerun", "rerun.f64", "--",
                        "-c", cfg], cwd=tmp, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-800:]
    g = MG.read_container(os.path.join(tmp, "ref.bin"))
    argv = [b"transit", b"-c", cfg.encode()]
    dump = os.path.join(tmp, "dry.f64")
    code = (f"import ctypes as C, numpy as np, os\n"
            f"os.dup2(os.open(os.devnull, os.O_WRONLY), 1)\n"
            f"L = C.CDLL({HOSTLIB_TAPS!r}); argv = {argv!r}; arr = (C.c_char_p * len(argv))(*argv)\n"
            f"L.transit_init(len(argv), arr); L.get_no_samples.restype = C.c_int; n = L.get_no_samples()\n"
            f"out = np.zeros(n); inp = np.load({str(tmp_path / 'inp.npy')!r}); P = C.POINTER(C.c_double)\n"
            f"L.run_transit(inp.ctypes.data_as(P), len(inp), out.ctypes.data_as(P), n)\n")
    r = subprocess.run(["python", "-c", code], capture_output=True, text=True,
                       env=dict(os.environ, TRANSIT_B200_DRYRUN=dump))
    assert r.returncode == 0, r.stderr[-2000:]
    d = _read_dry(dump)
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    assert np.array_equal(d["t"], g["rerun_atm_t"])
    assert rel(d["rad"], g["rerun_rad_v"]) < 1e-13          # log/pow of the -ffast-math build: a few ulp
    assert np.any(d["rad"] != prob.rad)                      # the radii really are the hydrostatic ones
    assert rel(d["dens"], g["rerun_mol_d"]) < 1e-13
    assert rel(d["Z"], g["rerun_iso_z"]) < 1e-13


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_dump")),
                    reason="oracle/_ref/ref_dump not built")
@pytest.mark.parametrize("extra", [["--raddelt", "500"],
                                   ["--raddelt", "300", "--radlow", "21000", "--radhigh", "26000"],
                                   ["--qmol", "CH4", "--qscale", "0.5"],
                                   ["--raddelt", "410", "--qmol", "CH4 H2", "--qscale", "-0.3 0.0"]])
def test_host_resampled_layers_and_qscale_without_gpu(tmp_path, extra):
    """makeradsample with raddelt != -1 (makesample.c:145-300,484-530: an equidistant radius sample, temperature,
    pressure, mean molecular mass, densities and abundances carried over by natural splines, partition functions at
    the new temperatures) and qmol/qscale (readatm.c:394-406,519-541), through the dry-run tap of the test build
    against the UNMODIFIED reference's samplers (ref_dump)."""
    _build()
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as MG
    prob = workloads.make_problem("C1", scale=0.02, nlayers=20)
    tmp = str(tmp_path)
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    dump = os.path.join(tmp, "dry.f64")
    r = subprocess.run([HOST_TAPS, "-c", cfg] + extra, capture_output=True, text=True, cwd=tmp,
                       env=dict(os.environ, TRANSIT_B200_DRYRUN=dump))
    assert r.returncode == 0, r.stderr[-800:]
    d = _read_dry(dump)
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_dump"), "ref.bin", "--", "-c", cfg] + extra,
                       cwd=tmp, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-500:]
    g = MG.read_container(os.path.join(tmp, "ref.bin"))
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    assert d["nlayer"] == len(g["rad_v"])
    if "--raddelt" in extra:
        assert d["nlayer"] != prob.nrad                      # really a new radius sample
    else:
        assert rel(d["dens"][2], prob.mol_d[2]) > 0.5        # really scaled: CH4 by 10^0.5
    assert np.array_equal(d["rad"], g["rad_v"])
    # the reference is a -ffast-math build: its splines come back within a few ulp
    assert rel(d["t"], g["atm_t"]) < 1e-14 and rel(d["p"], g["atm_p"]) < 1e-14
    assert rel(d["dens"], g["mol_d"]) < 1e-14 and rel(d["Z"], g["iso_z"]) < 1e-14


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw,nnew,more", [
    ("C2", dict(scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120), 26.3, []),
    ("C1", dict(scale=0.02, nlayers=24, wn_hi=2900.0, wnosamp=120), 17.6, []),
    ("C2", dict(scale=0.01, nlayers=20, wn_hi=3000.0, wnosamp=120), 15.2, ["--qmol", "H2O CH4", "--qscale", "0.4 -0.7"])])
def test_resampled_radius_grid_and_qscale_match_reference(tmp_path, name, kw, nnew, more):
    """raddelt != -1 end to end (transit: the impact-parameter sample of makeipsample's second branch goes to the
    device through tb_set_rays; eclipse: the reversed resampled radii), with and without qmol/qscale, full
    precision through transit_init + do_transit of both libraries.

    Tolerances.  Eclipse: 1e-8 like every spectrum.  Transit: makeipsample builds the impact parameters from the top
    radius DOWN (i + k*(-d)) and the radii from the bottom UP, so a third of them land one ulp below a layer radius
    (checked on the reference's dump).  totaltau1 then integrates from r0 to that layer over a first interval of
    ~40 cm in path length beside intervals of ~1e7 cm: the unequal-interval Simpson weights (numerical.c:454-525)
    cancel to (h1/h0)^2 ~ 1e10 rounding errors, and the reference's own -ffast-math arithmetic is one realisation of
    that noise (the CPU oracle, fed the reference's exact samples, differs from it by 3e-6 as well).  1e-4 here.

    run_transit on a resampled grid is exercised on OUR side only: the reference's interpcs zeroes
    wns.n*rads.n doubles of an array whose rows keep the stride of the FIRST call's layer count
    (crosssec.c:38-40,284), so with a changed layer count it adds the new CIA extinction on top of stale values
    (or writes past the array when the count grows); the per-atmosphere state it builds (radii, T, densities, Z) is
    compared on CPU in test_host_resampled_layers_and_qscale_without_gpu / ..._rerun."""
    _build()
    prob, cfg = _inputs(tmp_path, name, **kw)
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[6]:.6e}\nrefradius {prob.rad[6]:.6f}\ngsurf 2200.0\n")
    delt = (prob.rad[-1] - prob.rad[0]) / nnew
    argv = [b"transit", b"-c", cfg.encode(), b"--raddelt", f"{delt:.4f}".encode()] + [m.encode() for m in more]
    # ... and one run_transit on top: reloadatm moves the hydrostatic radii, so the sample is rebuilt (another layer count)
    np.save(tmp_path / "inp.npy", np.concatenate([prob.t * 1.02 + 5.0] + [prob.mol_q[j] for j in range(prob.nmol)]))
    outs = []
    for lib in (HOSTLIB, REFLIB):
        dst = str(tmp_path) + "/out_" + os.path.basename(lib) + ".npy"
        rerun = (f"inp = np.load({str(tmp_path / 'inp.npy')!r})\n"
                 f"for k in (1, 2): L.run_transit(inp.ctypes.data_as(P), len(inp), out[k].ctypes.data_as(P), n)\n") if lib == HOSTLIB else ""
        code = (f"import ctypes as C, numpy as np, os\n"
                f"os.dup2(os.open(os.devnull, os.O_WRONLY), 1)\n"
                f"L = C.CDLL({lib!r}); argv = {argv!r}; arr = (C.c_char_p * len(argv))(*argv)\n"
                f"L.transit_init(len(argv), arr); L.get_no_samples.restype = C.c_int; n = L.get_no_samples()\n"
                f"P = C.POINTER(C.c_double); out = np.zeros((3, n)); L.do_transit(out[0].ctypes.data_as(P))\n"
                + rerun + f"np.save({dst!r}, out)\n")
        r = subprocess.run(["python", "-c", code], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(np.load(dst))
    ours, ref = outs
    assert ours.shape == (3, prob.nwn) and np.all(ref[0] != 0)
    assert np.max(np.abs(ours[0] - ref[0]) / np.abs(ref[0])) < (1e-8 if prob.eclipse else 1e-4)
    assert np.all(np.isfinite(ours[1])) and np.all(ours[1] > 0) and np.array_equal(ours[1], ours[2])
    assert np.max(np.abs(ours[1] - ours[0]) / np.abs(ours[0])) > 1e-4              # the second atmosphere is another one


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_dump")),
                    reason="oracle/_ref/ref_dump not built")
def test_host_resampled_layers_rerun_without_gpu(tmp_path):
    """run_transit with raddelt != -1: reloadatm updates the atmosphere FILE's layers (new T and abundances,
    hydrostatic radii), makeradsample then rebuilds the equidistant sample from the moved radii -- another layer
    count than at initialisation -- against the state the UNMODIFIED reference holds after the same call."""
    _build()
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as MG
    prob = workloads.make_problem("C2", scale=0.004, nlayers=23, wn_hi=2600.0, wnosamp=60)
    tmp = str(tmp_path)
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[6]:.6e}\nrefradius {prob.rad[6]:.6f}\ngsurf 2200.0\n")
    delt = f"{(prob.rad[-1] - prob.rad[0]) / 26.3:.4f}"
    inp = np.concatenate([prob.t * 1.02 + 5.0] + [prob.mol_q[j] for j in range(prob.nmol)])
    inp.astype("<f8").tofile(os.path.join(tmp, "rerun.f64"))
    np.save(tmp_path / "inp.npy", inp)
    r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "ref_dump"), "ref.bin", "--rerun", "rerun.f64", "--",
                        "-c", cfg, "--raddelt", delt], cwd=tmp, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-800:]
    g = MG.read_container(os.path.join(tmp, "ref.bin"))
    argv = [b"transit", b"-c", cfg.encode(), b"--raddelt", delt.encode()]
    dump = os.path.join(tmp, "dry.f64")
    code = (f"import ctypes as C, numpy as np, os\n"
            f"os.dup2(os.open(os.devnull, os.O_WRONLY), 1)\n"
            f"L = C.CDLL({HOSTLIB_TAPS!r}); argv = {argv!r}; arr = (C.c_char_p * len(argv))(*argv)\n"
            f"L.transit_init(len(argv), arr); L.get_no_samples.restype = C.c_int; n = L.get_no_samples()\n"
            f"out = np.zeros(n); inp = np.load({str(tmp_path / 'inp.npy')!r}); P = C.POINTER(C.c_double)\n"
            f"L.run_transit(inp.ctypes.data_as(P), len(inp), out.ctypes.data_as(P), n)\n")
    r = subprocess.run(["python", "-c", code], capture_output=True, text=True,
                       env=dict(os.environ, TRANSIT_B200_DRYRUN=dump))
    assert r.returncode == 0, r.stderr[-2000:]
    d = _read_dry(dump)
    rel = lambda a, b: float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300)))
    assert d["nlayer"] == len(g["rerun_rad_v"]) and d["nlayer"] != len(g["rad_v"]) and d["nlayer"] != prob.nrad
    assert rel(d["rad"], g["rerun_rad_v"]) < 1e-13 and rel(d["t"], g["rerun_atm_t"]) < 1e-13
    assert rel(d["dens"], g["rerun_mol_d"]) < 1e-12 and rel(d["Z"], g["rerun_iso_z"]) < 1e-12
