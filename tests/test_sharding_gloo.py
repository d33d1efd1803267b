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


# ------------------------------------------------------------------ wavenumber bands

@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_wavenumber_bands_cover_aligned_balanced(world):
    rng = np.random.default_rng(3)
    nwn = 31334
    w = rng.random(nwn) * (1 + 5 * np.exp(-((np.arange(nwn) - 9000) / 3000.0) ** 2))   # a dense band
    bands = sharding.wavenumber_bands(w, world)
    assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == nwn
    for r in range(world - 1):
        assert bands[r][1] == bands[r + 1][0] and bands[r][1] % sharding.BAND_ALIGN == 0
    share = np.array([w[a:b].sum() for a, b in bands]) / w.sum()
    assert share.max() - share.min() < 0.03, share       # balanced by weight, not by width


def test_wavenumber_bands_more_ranks_than_tiles():
    bands = sharding.wavenumber_bands(np.ones(300), 8)
    assert bands[0][0] == 0 and bands[-1][1] == 300
    assert all(a <= b for a, b in bands) and sum(b - a for a, b in bands) == 300


def _band_worker(rank, world, port, nwn, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bands = sharding.wavenumber_bands(np.ones(nwn), world)
    truth = torch.arange(2 * nwn, dtype=torch.float64).view(2, nwn) * 0.5
    local = torch.full((2, nwn), -7.0, dtype=torch.float64)      # garbage outside the own band
    lo, hi = bands[rank]
    local[:, lo:hi] = truth[:, lo:hi]
    full = sharding.gather_bands(local, bands, rank, world)
    # the two exact maxima tb_set_shard asks for: float64 strengths and int32 line indices
    k = torch.tensor([1.0 + rank * 2.0 ** -52, 3.0 - rank], dtype=torch.float64)
    i = torch.tensor([rank * 7 - 1, 5 - rank], dtype=torch.int32)
    sharding.allreduce_max_(k)
    sharding.allreduce_max_(i)
    q.put((rank, full.numpy(), truth.numpy(), k.numpy(), i.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gather_bands_and_exact_maxima():
    world, nwn = 2, 1000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_band_worker, args=(r, world, port, nwn, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    for rank, full, truth, k, i in res:
        assert np.array_equal(full, truth)
        assert k[0] == 1.0 + 2.0 ** -52 and k[1] == 3.0      # last-bit differences survive the MAX
        assert list(i) == [6, 5]
