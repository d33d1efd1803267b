"""The N>1 host logic on CPU: two gloo ranks shard a batch of atmospheres, each "computes" its own
spectra (a deterministic stand-in: the CUDA path needs a GPU) and the gather must return every
spectrum, once, in atmosphere order, on every rank -- also when the batch does not divide evenly."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transit_b200 import sharding


def _fake_spectrum(i, nwn):
    return torch.arange(nwn, dtype=torch.float64) * 1e-3 + i


def _worker(rank, world, port, natm, nwn, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = sharding.atmosphere_shard(natm, rank, world)
    local = torch.stack([_fake_spectrum(i, nwn) for i in mine]) if len(mine) else torch.zeros(0, nwn, dtype=torch.float64)
    full = sharding.gather_spectra(local, natm, rank, world)
    q.put((rank, list(mine), full.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("natm", [4, 5, 1])
def test_two_ranks_gather_in_atmosphere_order(natm):
    world, nwn = 2, 17
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, natm, nwn, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    owned = sorted(i for _, mine, _ in res for i in mine)
    assert owned == list(range(natm))                       # every atmosphere exactly once
    want = np.stack([_fake_spectrum(i, nwn).numpy() for i in range(natm)])
    for _, _, full in res:
        assert full.shape == (natm, nwn) and np.array_equal(full, want)


def test_shard_is_balanced_and_contiguous():
    for natm in (1, 7, 8, 1024):
        for world in (1, 2, 4, 8):
            blocks = [sharding.atmosphere_shard(natm, r, world) for r in range(world)]
            sizes = [len(b) for b in blocks]
            assert sum(sizes) == natm and max(sizes) - min(sizes) <= 1
            flat = [i for b in blocks for i in b]
            assert flat == list(range(natm))
