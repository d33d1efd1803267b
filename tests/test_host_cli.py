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
REF = os.path.join(ROOT, "oracle", "_ref", "transit")
REFLIB = os.path.join(ROOT, "oracle", "_ref", "libtransit_ref.so")


def _build():
    if not os.path.exists(HOST):
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
