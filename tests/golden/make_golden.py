#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/ref_dump).

Run in the dev container only (needs /root/reference compiled by `make -C oracle ref`):

    python tests/golden/make_golden.py            # all cases
    python tests/golden/make_golden.py tiny_ecl   # one case

Every case is a small, seeded, synthetic problem in the reference's own file formats
(transit_b200/synth.py); the reference reads them through its stock CLI path and
oracle/ref_dump.c taps each stage at full double precision.  The committed .npz files
are what pins oracle/transit_oracle.c (tests/test_oracle_golden.py) and, through it, the
CUDA path.  Fixtures store float64 arrays exactly (np.savez_compressed is lossless).
"""
from __future__ import annotations

import json
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from transit_b200 import synth  # noqa: E402

REF_DUMP = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
REF_INPUTS = "/root/reference/inputs"
DT = {0: "<f8", 1: "<i4", 2: "<i8", 3: "<f4", 4: "<i2"}


def read_container(path):
    out = {}
    with open(path, "rb") as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        (nl,) = struct.unpack_from("<I", data, pos); pos += 4
        name = data[pos:pos + nl].decode(); pos += nl
        dtype, ndim = struct.unpack_from("<II", data, pos); pos += 8
        dims = struct.unpack_from(f"<{ndim}Q", data, pos); pos += 8 * ndim
        n = int(np.prod(dims)) if ndim else 1
        arr = np.frombuffer(data, DT[dtype], n, pos).reshape(dims).copy()
        pos += arr.nbytes
        out[name] = arr
    return out


def run_case(name, build, keep=None):
    with tempfile.TemporaryDirectory() as tmp:
        cfg, extra = build(tmp)
        cfgp = os.path.join(tmp, "case.cfg")
        synth.write_cfg(cfgp, **cfg)
        outp = os.path.join(tmp, "dump.bin")
        cmd = [REF_DUMP, outp] + extra.get("dump_args", []) + ["--", "-c", cfgp]
        pre = {}
        for args, keys in extra.get("pre_dumps", []):  # e.g. build an opacity file first
            prep = os.path.join(tmp, "pre.bin")
            subprocess.run([REF_DUMP, prep, "--"] + args, cwd=tmp, check=True,
                           stdout=subprocess.DEVNULL)
            pd = read_container(prep)
            pre.update({k: pd[k] for k in keys})
        subprocess.run(cmd, cwd=tmp, check=True, stdout=subprocess.DEVNULL)
        d = read_container(outp)
        d.update(pre)
        # ship the exact input files too, so the host-side readers can be tested on them
        for key, fn in extra.get("files", {}).items():
            with open(os.path.join(tmp, fn), "rb") as f:
                d["file_" + key] = np.frombuffer(f.read(), np.uint8)
        d["cfg_text"] = np.frombuffer(open(cfgp, "rb").read(), np.uint8)
        if keep:
            d = {k: v for k, v in d.items() if keep(k)}
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **d)
    print(f"{name}: {len(d)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


def _common(tmp, seed, wn_lo, wn_hi, nlines, nlayers, molecules, species, **atm_kw):
    rng = np.random.default_rng(seed)
    dbs = []
    for mol, (niso, frac) in molecules.items():
        db = synth.make_database(mol, niso)
        synth.make_lines(db, int(nlines * frac), wn_lo + 0.3, wn_hi - 0.3, rng, bands=3)
        dbs.append(db)
    synth.write_tli(os.path.join(tmp, "lines.tli"), dbs, 1e4 / (wn_hi + 50), 1e4 / max(wn_lo - 50, 1.0))
    atm = synth.make_atmosphere(nlayers, species, rng=rng, **atm_kw)
    synth.write_atmosphere(os.path.join(tmp, "atm.dat"), atm)
    return atm


def case_tiny_ecl(tmp):
    """C1-shaped: one CH4 database, eclipse, shipped H2-H2 CIA table."""
    _common(tmp, 1001, 2900.0, 3100.0, 3000, 24, {"CH4": (4, 1.0)},
            {"H2": 0.85, "He": 0.1496, "CH4": 4e-4}, t_top=900.0, t_bot=1600.0)
    cfg = dict(atm="atm.dat", linedb="lines.tli", molfile=f"{REF_INPUTS}/molecules.dat",
               csfile=f"{REF_INPUTS}/CIA_H2H2_400-7000K.dat",
               wnlow=2900.0, wnhigh=3100.0, wndelt=1.0, wnosamp=120, wnfct=1.0, wlfct=1e-4,
               solution="eclipse", raygrid="0 20 40 60 80", toomuch=10, ethreshold=1e-50,
               nwidth=20, ndop=14, nlor=14, verb=0, outspec="spec.dat")
    return cfg, dict(dump_args=["--full-ext"],
                     files=dict(tli="lines.tli", atm="atm.dat"))


def case_tiny_tra(tmp):
    """C2-shaped: four molecules, transit geometry, two synthetic CIA tables, ethresh active."""
    _common(tmp, 1002, 4000.0, 4300.0, 6000, 30,
            {"H2O": (4, 0.5), "CH4": (3, 0.2), "CO": (3, 0.05), "CO2": (4, 0.25)},
            {"H2": 0.85, "He": 0.1485, "H2O": 5e-4, "CH4": 3e-4, "CO": 4e-4, "CO2": 3e-4})
    wn = np.arange(20.0, 6000.1, 20.0)
    temps = np.array([200., 400., 600., 800., 1000., 1500., 2000., 3000.])
    synth.write_cia(os.path.join(tmp, "cia_h2h2.dat"), "H2", "H2", wn, temps, 1e-7, 1)
    synth.write_cia(os.path.join(tmp, "cia_h2he.dat"), "H2", "He", wn, temps, 4e-8, 2)
    cfg = dict(atm="atm.dat", linedb="lines.tli", molfile=f"{REF_INPUTS}/molecules.dat",
               csfile="cia_h2h2.dat,cia_h2he.dat",
               wnlow=4000.0, wnhigh=4300.0, wndelt=1.0, wnosamp=240, wnfct=1.0, wlfct=1e-4,
               solution="transit", toomuch=10, ethreshold=1e-6, starrad=0.75,
               nwidth=20, ndop=12, nlor=16, verb=0, outspec="spec.dat")
    return cfg, dict(dump_args=["--full-ext"],
                     files=dict(tli="lines.tli", atm="atm.dat", cia0="cia_h2h2.dat",
                                cia1="cia_h2he.dat"))


def case_tiny_grid(tmp):
    """C3/C5-shaped: opacity-grid build (permol=1), then a spectrum interpolated from it,
    then one run_transit() re-evaluation with a perturbed atmosphere (reloadatm)."""
    atm = _common(tmp, 1003, 4000.0, 4120.0, 2500, 16,
                  {"H2O": (3, 0.6), "CO2": (3, 0.4)},
                  {"H2": 0.85, "He": 0.149, "H2O": 6e-4, "CO2": 4e-4},
                  p_top_bar=1e-5, p_bot_bar=1e1)
    wn = np.arange(20.0, 6000.1, 20.0)
    temps = np.array([200., 400., 600., 800., 1000., 1500., 2000., 3000.])
    synth.write_cia(os.path.join(tmp, "cia_h2h2.dat"), "H2", "H2", wn, temps, 1e-7, 1)
    cfg = dict(atm="atm.dat", linedb="lines.tli", molfile=f"{REF_INPUTS}/molecules.dat",
               csfile="cia_h2h2.dat", opacityfile="grid.opa",
               tlow=600, thigh=1300, tempdelt=100,
               wnlow=4000.0, wnhigh=4120.0, wndelt=1.0, wnosamp=120, wnfct=1.0, wlfct=1e-4,
               solution="transit", toomuch=10, ethreshold=1e-8, starrad=0.75,
               refpress=0.1, refradius=float(np.interp(np.log(0.1), np.log(atm["p_bar"][::-1]),
                                                       atm["radius_km"][::-1])),
               gsurf=1500.0, nwidth=20, ndop=10, nlor=12, verb=0, outspec="spec.dat")
    # perturbed atmosphere for run_transit: [T | q_mol0 | q_mol1 ...] (readatm.c:735-744)
    n = len(atm["T"])
    rng = np.random.default_rng(77)
    T2 = atm["T"] + 60.0 * np.sin(np.linspace(0, 3, n)) + rng.uniform(-3, 3, n)
    q2 = [atm["q"][s] * (10 ** rng.uniform(-0.5, 0.5) if s in ("H2O", "CO2") else 1.0)
          for s in atm["q"]]
    np.concatenate([T2] + q2).astype("<f8").tofile(os.path.join(tmp, "rerun.f64"))
    cfgp = os.path.join(tmp, "build.cfg")
    synth.write_cfg(cfgp, **cfg)
    pre = (["-c", cfgp, "--justOpacity"],
           ["nDop", "nLor", "aDop", "aLor", "profsize", "prof_off", "prof_data", "timesalpha"])
    return cfg, dict(dump_args=["--rerun", "rerun.f64"], pre_dumps=[pre],
                     files=dict(tli="lines.tli", atm="atm.dat", cia0="cia_h2h2.dat",
                                opa="grid.opa"))


CLOUD_VARIANTS = {
    # name -> extra reference CLI arguments (argum.c:636-715: --cloud type,ext,top,bot[,gamma[,Q,r | sig,refwn]];
    # --scattering <log ext> | polar; --cloudtop <log bar>)
    "none": [],
    "sc1": ["--scattering", "0.5"],
    "sc2": ["--scattering", "polar"],
    "cl1": ["--cloud", "ext,3e-9,-4.0,-1.0"],
    "cl2": ["--cloud", "opa,2e-3,-4.0,-1.0"],
    "cl3": ["--cloud", "B17,4e-19,-4.0,-1.0,4.0"],
    "cl4": ["--cloud", "F18,2e-3,-4.0,-1.0,4.0,40.0,1e-4"],
    "cl5": ["--cloud", "P19,1e5,-4.0,-1.0,4.0,1e-27,2857.0"],
    "cloudtop": ["--cloudtop", "-2.5"],
    "sc1_cl2": ["--scattering", "0.5", "--cloud", "opa,2e-3,-4.0,-1.0"],
    "transparent": ["--transparent"],
    "modm1": ["--modlevel", "-1", "--toomuch", "1.0"],
}


def make_tiny_cloud():
    """a10 / a14 fixtures: ONE small transit problem through the unmodified reference once per scattering / cloud /
    modulation variant (computeextscat, computeextcloud, set_cloudtop's fields, --transparent, modlevel -1);
    the static inputs are kept once, tau / last / spectrum per variant."""
    with tempfile.TemporaryDirectory() as tmp:
        _common(tmp, 1010, 4000.0, 4080.0, 1500, 24,
                {"H2O": (3, 0.6), "CO2": (3, 0.4)},
                {"H2": 0.85, "He": 0.1485, "H2O": 9e-4, "CO2": 6e-4}, p_top_bar=1e-6, p_bot_bar=1e1)
        wn = np.arange(20.0, 6000.1, 20.0)
        temps = np.array([200., 400., 600., 800., 1000., 1500., 2000., 3000.])
        synth.write_cia(os.path.join(tmp, "cia_h2h2.dat"), "H2", "H2", wn, temps, 1e-7, 1)
        cfg = dict(atm="atm.dat", linedb="lines.tli", molfile=f"{REF_INPUTS}/molecules.dat", csfile="cia_h2h2.dat",
                   wnlow=4000.0, wnhigh=4080.0, wndelt=1.0, wnosamp=120, wnfct=1.0, wlfct=1e-4,
                   solution="transit", toomuch=10, ethreshold=1e-8, starrad=0.75,
                   nwidth=20, ndop=10, nlor=12, verb=0, outspec="spec.dat")
        cfgp = os.path.join(tmp, "case.cfg")
        synth.write_cfg(cfgp, **cfg)
        out = {}
        for name, extra in CLOUD_VARIANTS.items():
            outp = os.path.join(tmp, f"dump_{name}.bin")
            subprocess.run([REF_DUMP, outp] + (["--full-ext"] if name == "none" else []) + ["--", "-c", cfgp] + extra,
                           cwd=tmp, check=True, stdout=subprocess.DEVNULL)
            d = read_container(outp)
            if name == "none":
                out.update({k: v for k, v in d.items() if not k.startswith(("prof_", "profsize"))})
            for k in ("tau", "last", "out", "computed"):
                out[f"{name}__{k}"] = d[k]
            for k in ("sc_flag", "sc_logext", "cl_flag", "cl_ext", "cl_top", "cl_bot", "cl_gamma", "cl_Q", "cl_r", "cl_sig",
                      "cl_refwn", "transpplanet", "toomuch"):
                out[f"{name}__{k}"] = d[k]
        out["cfg_text"] = np.frombuffer(open(cfgp, "rb").read(), np.uint8)
        out["variants"] = np.frombuffer(json.dumps(CLOUD_VARIANTS).encode(), np.uint8)
        for key, fn in dict(tli="lines.tli", atm="atm.dat", cia0="cia_h2h2.dat").items():
            out["file_" + key] = np.frombuffer(open(os.path.join(tmp, fn), "rb").read(), np.uint8)
    path = os.path.join(HERE, "tiny_cloud.npz")
    np.savez_compressed(path, **out)
    print(f"tiny_cloud: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


CASES = dict(tiny_ecl=case_tiny_ecl, tiny_tra=case_tiny_tra, tiny_grid=case_tiny_grid)

if __name__ == "__main__":
    if not os.path.exists(REF_DUMP):
        sys.exit("oracle/_ref/ref_dump missing: run `make -C oracle ref` first")
    for nm in (sys.argv[1:] or list(CASES) + ["tiny_cloud"]):
        if nm == "tiny_cloud":
            make_tiny_cloud()
        else:
            run_case(nm, CASES[nm])
