"""Census of (line group, layer) pairs on C2: how many are below threshold, evaluated, and of the
evaluated how many have a profile that reaches no grid point (from tb_get_line_trace + the profile
table).  usage: python tools/pair_census.py [layer,layer,...]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from transit_b200 import api, workloads

prob = workloads.make_problem("C2")
c = api.Context(0)
c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio, prob.iso_imol)
c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
c.build_profiles(60, 60, 1e-3, 0.25, 1e-4, 10.0, prob.nwidth)
t = c.get_profiles()
o = prob.wnosamp
layers = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 10, 16, 24, 40, 60, 80, 99]
print("layer  lines  evaluated  coadded  below  evaluated_noreach  samples")
for r in layers:
    T = np.array([prob.t[r] * prob.t_fct]); d = prob.mol_d[:, r][None, :].copy(); Z = prob.iso_z[:, r][None, :].copy()
    c.molext_states(T, d, Z, prob.ethresh, 0)
    tr = c.line_trace(0); iown, idwn, idop, status = tr["iown"], tr["idwn"], tr["idop"], tr["status"]
    ev = status == 1
    # Lorentz index per isotope is not in the trace: take the profile half-size of (idop, any ilor) as the
    # MAXIMUM over ilor -> a lower bound on the no-reach count
    psmax = t["psize"].max(axis=1)
    ps = psmax[np.clip(idop[ev], 0, len(psmax) - 1)]
    io = iown[ev].astype(np.int64)
    dist = np.abs(io - ((io + o // 2) // o) * o)
    nr = int(np.sum(dist > ps))
    print(f"{r:4d} {len(status):8d} {int(ev.sum()):9d} {int((status == 2).sum()):8d} {int((status == 3).sum()):8d} {nr:9d} {c.sample_counts(1)[0]:12d}")
