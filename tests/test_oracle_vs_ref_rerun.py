"""The oracle against the UNMODIFIED reference on the run_transit path (reloadatm: new temperatures
and abundances, hydrostatic radii from refpress/refradius/gsurf).  Hydrostatic radii are not the
round numbers of an atmosphere file, and that is where the reference's -ffast-math build shows:
tau.c:274 `h*hfct/rfct` is compiled as (h*hfct)*(1/rfct); with an IEEE division a few impact
parameters land one ulp under their own layer radius and tau changes by ~20 % for those rays.
Needs oracle/_ref/ref_dump (built from /root/reference by oracle/Makefile); skipped without it."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oraclelib as O
from transit_b200 import workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DUMP = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


@pytest.mark.skipif(not os.path.exists(REF_DUMP), reason="oracle/_ref/ref_dump not built")
def test_oracle_matches_reference_on_hydrostatic_radii(tmp_path):
    import make_golden as MG
    prob = workloads.make_problem("C2", scale=0.004, nlayers=16, wn_hi=2600.0, wnosamp=60)
    tmp = str(tmp_path)
    cfg, _ = workloads.write_reference_inputs(prob, tmp)
    with open(cfg, "a") as f:
        f.write(f"refpress {prob.p[5]:.6e}\nrefradius {prob.rad[5]:.6f}\ngsurf 2200.0\n")
    inp = np.concatenate([prob.t * 1.01 + 3.0] + [prob.mol_q[j] for j in range(prob.nmol)])
    inp.astype("<f8").tofile(os.path.join(tmp, "rerun.f64"))
    r = subprocess.run([REF_DUMP, "ref.bin", "--full-ext", "--rerun", "rerun.f64", "--rerun-taps", "--", "-c", cfg],
                       cwd=tmp, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1000:]
    d = MG.read_container(os.path.join(tmp, "ref.bin"))
    rad, t, dd, Z = d["rerun_rad_v"], d["rerun_atm_t"], d["rerun_mol_d"], d["rerun_iso_z"]
    # hydrostatic radii: impact parameters that do NOT survive (h*1e5)/1e5 exist in this case
    ip = rad[::-1]
    assert np.any((ip * 1e5) / 1e5 != ip)
    pr = O.Problem(**vars(prob)); pr.rad = rad.copy(); pr.t = t.copy(); pr.mol_d = dd; pr.iso_z = Z
    f32 = lambda x: float(np.float32(x))        # dmin..lmax are `float` in struct transithint
    vt = O.OracleVoigt(O.Problem(**vars(prob)), 60, 60, f32(1e-3), f32(0.25), f32(1e-4), f32(10.0))
    e0 = np.stack([O.computemolext(pr, vt, t[k], dd[:, k].copy(), Z[:, k].copy(), prob.ethresh)[0]
                   for k in range(prob.nrad)])
    ef = d["rr_ext_full"]
    m = ef > 0
    assert np.max(np.abs(e0[m] - ef[m]) / ef[m]) < 1e-10
    ecs0 = O.interpcs(pr, temp=t, mol_d=dd)
    assert np.max(np.abs(ecs0 - d["rr_e_cs"]) / np.maximum(d["rr_e_cs"], 1e-300)) < 1e-12
    t0, last0, comp = O.tau(pr, e0, ecs0, lazy=True, rad=rad)
    assert np.array_equal(last0, d["rr_last"]) and np.array_equal(comp, d["rr_computed"])
    big = d["rr_tau"] > 1e-9                     # below that: pure cancellation in interp_parab
    assert np.max(np.abs(t0[big] - d["rr_tau"][big]) / d["rr_tau"][big]) < 1e-5
    out0 = O.modulation(pr, t0, last0, rad=rad)
    assert np.max(np.abs(out0 - d["rr_out"]) / d["rr_out"]) < 1e-8
    assert np.array_equal(d["rr_out"], d["rerun_out"])
