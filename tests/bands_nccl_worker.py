"""Worker of tests/test_multigpu_bands.py, launched with torchrun (one rank per GPU): one spectrum over `world`
wavenumber bands with the exchange inside the library (tb_comm_init / tb_set_bands) must come back WHOLE on every
rank and bit-identical to the single-GPU spectrum; the per-column stages must really be restricted to the band."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from transit_b200 import api, sharding, workloads  # noqa: E402


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")            # plumbing for the 128-byte id only: the data path is the library's NCCL
    cases = [("C2", dict(scale=0.05, nlayers=30, wn_hi=9000.0)), ("C1", dict(scale=0.2, nlayers=30)),
             ("C2", dict())]
    for name, kw in cases:
        prob = workloads.make_problem(name, **kw)
        one = api.Transit(local)
        one.transit_init(prob)
        ref = one.do_transit(None).copy()
        one.free_memory()
        tb = sharding.BandShardedTransit(local, rank, world)
        tb.transit_init(prob)
        out = np.zeros(prob.nwn)
        for _ in range(3):
            tb.do_transit(out)
        assert np.array_equal(out, ref), f"{name} {kw}: rank {rank} spectrum differs from the single-GPU one"
        lo, hi = tb.bands[rank]
        t, last = tb.ctx.get_tau()
        assert (hi == lo) or np.any(t[lo:hi] != 0)
        assert not np.any(t[:lo] != 0) and not np.any(t[hi:] != 0), "tau was computed outside the band"
        if rank == 0:
            print(f"bands ok: {name} {kw or 'full size'} world {world} bands {tb.bands}", flush=True)
        tb.free_memory()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
