"""tb_tau_transit_static -- the transit optical depth with its scratch in a static device array, for callers that
cannot call cudaMalloc -- against tb_tau on the same inputs, through the C ABI, twice in a row (the array keeps the
last call's contents: the second call, on OTHER rays and a smaller problem, must not see the first one's).  The file
name makes pytest collect it last: a failure here hides nothing under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

from transit_b200 import api, workloads

pytestmark = pytest.mark.gpu


def relerr(a, b):
    d = np.abs(a - b)
    s = np.maximum(np.abs(a), np.abs(b))
    return float(np.max(np.where(s > 0, d / np.where(s > 0, s, 1), 0.0)))


def test_static_scratch_tau_matches_tb_tau_twice_in_a_row():
    import torch
    dev = torch.device("cuda:0")
    prob = workloads.make_problem("C2", scale=0.01, nlayers=37, wn_hi=3500.0, wnosamp=120)   # 37 rays: a ragged group of four
    c = api.Context(0)
    c.set_wn_grid(prob.wn_i, prob.wn_d, prob.wnosamp, prob.nwn, prob.onwn, prob.wn_fct)
    c.set_species(prob.mol_mass, prob.mol_radius, prob.mol_pol, prob.mol_ID, prob.iso_mass, prob.iso_ratio, prob.iso_imol)
    c.upload_lines(prob.wl, prob.elow, prob.gf, prob.isoid, prob.wfct, prob.efct)
    c.build_profiles(20, 20, 1e-3, 0.25, 1e-4, 10.0, prob.nwidth)
    for tb in prob.cia:
        c.cia_add_table(tb["wn"], tb["temp"], tb["tab"], tb["mol"])
    c.set_atmosphere(prob.rad, prob.rad_fct, prob.p, prob.p_fct, prob.t, prob.t_fct, prob.mol_d, prob.iso_z)
    c.extinction_lbl(prob.ethresh)
    c.cia_interp()
    e = c.get_extinction()                       # [nrad][nwn]
    ecs = np.ascontiguousarray(c.get_cia().T)    # tb_get_cia hands [nwn][nrad]; the device layout is [nrad][nwn]
    assert e.shape == (prob.nrad, prob.nwn) and ecs.shape == e.shape
    r = prob.rad
    rays = [np.ascontiguousarray(r[::-1]),       # the reversed radii (makeipsample, raddelt -1)
            np.sort(np.concatenate([r[2:30:3], 0.5 * (r[3:20] + r[2:19])]))[::-1].copy()]   # fewer, other rays
    d_e, d_ecs = torch.from_numpy(e).to(dev), torch.from_numpy(ecs).to(dev)
    d_rad = torch.from_numpy(np.ascontiguousarray(r)).to(dev)
    for k, h in enumerate(rays):
        nh = len(h)
        # existing entry point: tb_tau with explicit rays (context buffers from cudaMalloc)
        c.tau(api.GEOM_TRANSIT, prob.toomuch, h, prob.rad_fct)
        t0, last0 = c.get_tau()                  # [nwn][nh], [nwn]
        # the new one: caller-owned device buffers, filled with junk to prove they are written
        d_h = torch.from_numpy(h).to(dev)
        d_tau = torch.full((nh, prob.nwn), 7.0, dtype=torch.float64, device=dev)
        d_last = torch.full((prob.nwn,), -5, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]      # cudaMemGetInfo: whatever allocates on the device shows here
        c.tau_transit_static(prob.nrad, d_rad.data_ptr(), prob.rad_fct, nh, d_h.data_ptr(), prob.rad_fct, prob.nwn,
                             d_e.data_ptr(), d_ecs.data_ptr(), prob.toomuch, d_tau.data_ptr(), d_last.data_ptr(), True)
        torch.cuda.synchronize()
        if k == 1:                               # (the first call may still load the kernels' module lazily)
            assert torch.cuda.mem_get_info()[0] >= free_before, "the call allocated device memory"
        t1 = d_tau.cpu().numpy().T
        last1 = d_last.cpu().numpy().astype(np.int64)
        assert np.array_equal(last1, last0), f"call {k}"
        assert np.array_equal(t1 == 0, t0 == 0), f"call {k}"
        assert relerr(t1, t0) < 1e-10, f"call {k}"              # the tolerance of tau between two launches of these kernels
        assert np.array_equal(t1, t0), f"call {k}"              # ... and in fact the same kernels on the same numbers
    # sizes beyond the static array are refused, loudly, before anything is launched
    with pytest.raises(api.TransitError, match="static scratch holds"):
        c.tau_transit_static(129, d_rad.data_ptr(), prob.rad_fct, 4, d_h.data_ptr(), prob.rad_fct, prob.nwn,
                             d_e.data_ptr(), d_ecs.data_ptr(), prob.toomuch, d_tau.data_ptr(), d_last.data_ptr(), True)
    with pytest.raises(api.TransitError, match="static scratch holds"):
        c.tau_transit_static(prob.nrad, d_rad.data_ptr(), prob.rad_fct, 129, d_h.data_ptr(), prob.rad_fct, prob.nwn,
                             d_e.data_ptr(), d_ecs.data_ptr(), prob.toomuch, d_tau.data_ptr(), d_last.data_ptr(), True)
    c.close()
