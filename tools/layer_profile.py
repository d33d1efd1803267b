"""Per-layer cost of the line sum on C2: each layer's state replicated NREP times in one batch,
device time of k_strength / k_accumulate / k_kmax per replica, samples per state.
usage: python tools/layer_profile.py [NREP] [step] [layer,layer,...|-] [config]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from transit_b200 import api, workloads

nrep = int(sys.argv[1]) if len(sys.argv) > 1 else 20
step = int(sys.argv[2]) if len(sys.argv) > 2 else 5
prob = workloads.make_problem(sys.argv[4] if len(sys.argv) > 4 else "C2")
c = api.Context(0)
c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio, prob.iso_imol)
c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
c.build_profiles(60, 60, 1e-3, 0.25, 1e-4, 10.0, prob.nwidth)
print("layer  p[bar]      T   samples/state   strength_us accumulate_us kmax_us fused_us  (per state)")
tot = {"strength": 0.0, "accumulate": 0.0, "kmax": 0.0, "fused": 0.0}
if os.environ.get("LP_FUSE_RMAX"):
    c.set_fused(2); c.set_fused_reach(int(os.environ["LP_FUSE_RMAX"]), True)
layers = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 and sys.argv[3] != "-" else range(0, prob.nrad, step)
for r in layers:
    T = np.full(nrep, prob.t[r] * prob.t_fct)
    d = np.tile(prob.mol_d[:, r], (nrep, 1))
    Z = np.tile(prob.iso_z[:, r], (nrep, 1))
    for _ in range(2):
        c.molext_states(T, d, Z, prob.ethresh, 0)
    ns = c.sample_counts(nrep)
    ms = {}
    for k in tot:
        try: ms[k] = c.kernel_ms(k) * 1e3 / nrep
        except api.TransitError: ms[k] = 0.0
    for k in tot:
        tot[k] += ms[k] * step
    print(f"{r:4d} {prob.p[r]*prob.p_fct/1e6:9.2e} {T[0]:7.1f} {ns[0]:12d}   {ms['strength']:9.2f} {ms['accumulate']:9.2f} {ms['kmax']:9.2f} {ms['fused']:9.2f}")
print("sum over layers (us):", {k: round(v, 1) for k, v in tot.items()})
