"""Multi-GPU plumbing: one process per GPU (torchrun), torch.distributed for the exchange.

The path shards over *atmospheres* (BART's use of the library: many
T-P/abundance profiles through one static problem).  Every rank holds the static problem (line
index, Voigt table, CIA tables or opacity grid) and computes the spectra of its own atmospheres;
there is no collective on the data path, only one all-gather of the finished spectra.  The kernels,
the copies and the gather share one CUDA stream (tb_set_stream / tb_device_spectrum).

`atmosphere_shard` and `gather_spectra` are backend-agnostic (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import numpy as np

from . import api


def atmosphere_shard(natm: int, rank: int, world: int):
    """Contiguous, balanced block of atmosphere indices owned by `rank` (the first natm % world
    ranks get one more)."""
    base, extra = divmod(natm, world)
    lo = rank * base + min(rank, extra)
    return range(lo, lo + base + (1 if rank < extra else 0))


def gather_spectra(local, natm: int, rank: int, world: int, group=None):
    """All-gather per-rank spectra [n_local, nwn] into [natm, nwn] in atmosphere order on every rank.
    Blocks may differ by one row between ranks: they are padded to the largest for the collective."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local
    nmax = -(-natm // world)
    nwn = local.shape[1]
    buf = torch.zeros(nmax, nwn, dtype=local.dtype, device=local.device)
    buf[:local.shape[0]] = local
    out = torch.empty(world * nmax, nwn, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    rows = []
    for r in range(world):
        n = len(atmosphere_shard(natm, r, world))
        rows.append(out[r * nmax:r * nmax + n])
    return torch.cat(rows, 0)


class ShardedTransit(api.Transit):
    """api.Transit plus what a rank of a multi-GPU run needs."""

    def __init__(self, device=0, rank=0, world=1):
        super().__init__(device)
        self.rank, self.world = rank, world

    def device_spectrum_tensor(self):
        """The last spectrum as a torch CUDA tensor aliasing the context's device buffer."""
        import torch
        ptr, n = self.ctx.device_spectrum()

        class _Alias:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}
        return torch.as_tensor(_Alias(), device="cuda")

    def set_atmosphere(self, prob):
        self.ctx.set_atmosphere(prob.rad, prob.rad_fct, prob.p, prob.p_fct, prob.t, prob.t_fct, prob.mol_d, prob.iso_z)

    def sample_counts(self):
        return self.ctx.sample_counts(self.prob.nrad)

    def ngroups(self):
        return self.ctx.ngroups()

    def band_nwn(self):
        return self.ctx.nwn
