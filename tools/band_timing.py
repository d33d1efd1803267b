"""Single-GPU check of what a wavenumber band costs: one layer replicated NREP times, whole grid vs one band."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from transit_b200 import api, workloads
nrep = 20
prob = workloads.make_problem("C2")
c = api.Context(0)
c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio, prob.iso_imol)
c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
c.build_profiles(60, 60, 1e-3, 0.25, 1e-4, 10.0, prob.nwidth)
for r in (0, 10, 60):
    T = np.full(nrep, prob.t[r] * prob.t_fct); d = np.tile(prob.mol_d[:, r], (nrep, 1)); Z = np.tile(prob.iso_z[:, r], (nrep, 1))
    for band in ((0, 0), (0, 3968), (15744, 19584), (27520, 31334)):
        c.set_shard(band[0], band[1], 0, 1, None)
        for permol in (0, 1):
            for _ in range(2):
                c.molext_states(T, d, Z, prob.ethresh, permol)
            print(f"layer {r} band {band} permol {permol}: " + " ".join(f"{k} {c.kernel_ms(k)*1e3/nrep:7.1f}us" for k in ("kmax", "strength", "accumulate")),
                  "samples", int(c.sample_counts(nrep)[0]))
