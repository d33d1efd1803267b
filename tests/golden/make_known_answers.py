"""Collect the known answers the reference's OWN test material holds for the path into one small
fixture, tests/golden/ref_known_answers.npz (run here, where /root/reference exists; the fixture
travels, the reference does not):

* transit/test/slantpath.080604.dat -- archived printout of transit/test/test_slantpath.c
  (test_tau :522-539, test_mod :491-517): for every parameter combination the analytic expectation
  (6 digits) and what the 2004 build observed (10 digits) with its relative error;
  `RefIdx constant technique` = totaltau1, the only ray solution on the path.
* transit/test/voigt.080904.dat -- archived printout of pu/test/test_voigt.c: the Voigt function on
  10 001 points over +-75 cm-1, Lorentz width 1.5, Doppler width 1, 4 significant digits; the
  centred column `v(w+0)` is kept.

    python tests/golden/make_known_answers.py [/root/reference]
"""
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MOD_KINDS = {"Constant": 0, "Increasing": 1, "Decreasing": 2}


def parse_slantpath(path):
    tau, mod = [], []
    mode = None
    alpha = prm = kind = rm = ip = exp = nrad = star = first_r = None
    for ln in open(path).read().splitlines():
        m = re.match(r"Testing OPTICAL DEPTH.*alpha=(\S+)", ln)
        if m:
            mode, alpha = "tau", float(m.group(1)); continue
        m = re.match(r"Testing MODULATION for (\w+) Tau; parameter (\S+)", ln)
        if m:
            mode, kind, prm = "mod", MOD_KINDS[m.group(1)], float(m.group(2)); continue
        m = re.match(r"\s+For planets of radius (\S+)", ln)
        if m:
            rm = float(m.group(1)); continue
        m = re.match(r"\s+For a star of radius (\S+)", ln)
        if m:
            star = float(m.group(1)); continue
        m = re.match(r"\s+For a planet of radius (\S+)", ln)
        if m:
            rm = float(m.group(1)); continue
        m = re.match(r"\s+For rays crossing at (\S+), expected value is (\S+)", ln)
        if m:
            ip, exp = float(m.group(1)), float(m.group(2)); continue
        m = re.match(r"\s+Atmosphere starting at (\S+)\. Expected result is (\S+)", ln)
        if m:
            first_r, exp = float(m.group(1)), float(m.group(2)); continue
        m = re.match(r"\s+Using (\d+) (radial )?layers", ln)
        if m:
            nrad = int(m.group(1)); continue
        m = re.match(r"\s+RefIdx constant technique: observed\s+(\S+) \(error (\S+)\)", ln)
        if m and mode == "tau":
            tau.append((alpha, rm, ip, nrad, exp, float(m.group(1)), float(m.group(2)))); continue
        m = re.match(r"\s+Full eclipse, no LD\s+: observed\s+(\S+) \(error (\S+)\)", ln)
        if m and mode == "mod":
            mod.append((kind, prm, star, rm, first_r / rm, nrad, exp, float(m.group(1)), float(m.group(2))))
    return np.array(tau), np.array(mod)


def parse_voigt(path):
    with open(path) as f:
        head = f.readline()
        cols = f.readline()
        m = re.match(r"#Total range:(\S+), Lorentz width:(\S+), Doppler Width (\S+)", head)
        rng, lor, dop = float(m.group(1)), float(m.group(2)), float(m.group(3))
        names = re.findall(r"v\(w([+-])\s*([0-9.e-]+)\)", cols)
        centre = [i for i, (s, v) in enumerate(names) if float(v) == 0.0][0]
        a = np.loadtxt(f)
    return dict(voigt_range=rng, voigt_lor=lor, voigt_dop=dop, voigt_x=a[:, 0],
                voigt_v0=a[:, 1 + centre].astype(np.float32))


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    tdir = os.path.join(ref, "transit", "test")
    tau, mod = parse_slantpath(os.path.join(tdir, "slantpath.080604.dat"))
    v = parse_voigt(os.path.join(tdir, "voigt.080904.dat"))
    out = os.path.join(HERE, "ref_known_answers.npz")
    np.savez_compressed(out, tau_cases=tau, mod_cases=mod, **v)
    print(f"{out}: {len(tau)} optical-depth cases, {len(mod)} modulation cases, "
          f"{len(v['voigt_x'])} Voigt points; {os.path.getsize(out)} bytes")


if __name__ == "__main__":
    main()
