"""Wavenumber-band sharding of one spectrum over the GPUs of a box.

One process per GPU.  Every rank holds the whole (layer-independent) line index but owns one
contiguous band of the coarse wavenumber grid; the only exchanges are the per-layer strength
maxima / carried Doppler indices (exact integer-ordered maxima, so results are bit-identical to
the single-GPU run) and the final gather of the spectrum slices.
"""
from __future__ import annotations

import numpy as np

from . import api


class ShardedTransit(api.Transit):
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
