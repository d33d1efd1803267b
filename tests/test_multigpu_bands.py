"""Wavenumber-band sharding with the exchange inside the library (tb_comm.cu: NCCL bound at run time), on REAL
GPUs: needs at least two (gpurun --gpus 2); skipped on a single-GPU box.  (The host-callback exchange is covered
on one GPU by tests/test_gpu_parity.py with two contexts, and the plumbing on CPU by tests/test_sharding_gloo.py.)"""
import os
import subprocess
import sys

import numpy as np
import pytest

from transit_b200 import workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host", "transit_b200")


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


pytestmark = [pytest.mark.gpu, pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")]


def test_library_exchange_bit_identical_to_one_gpu():
    n = min(_ngpu(), 8)
    for world in sorted({2, n}):
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                            "--master-addr", "127.0.0.1", "--master-port", "29631",
                            os.path.join(ROOT, "tests", "bands_nccl_worker.py")], capture_output=True, text=True, timeout=900)
        assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
        assert r.stdout.count("bands ok") == 3, r.stdout[-2000:]


def test_c_host_on_two_gpus_writes_the_same_files(tmp_path):
    """TRANSIT_B200_GPUS=2 host/transit_b200 -c cfg: the spectrum file is byte-identical to the single-GPU one, and
    an --opacityfile build writes the same grid file (every rank pwrites its own column slices)."""
    prob = workloads.make_problem("C2", scale=0.05, nlayers=24, wn_hi=9000.0)
    cfg, _ = workloads.write_reference_inputs(prob, str(tmp_path))
    outs = {}
    for g in (1, 2):
        spec = tmp_path / f"spec{g}.dat"
        env = dict(os.environ, TRANSIT_B200_GPUS=str(g))
        r = subprocess.run([HOST, "-c", cfg, "--outspec", str(spec)], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[g] = open(spec, "rb").read()
    assert outs[1] == outs[2] and len(outs[1]) > 1000
    grids = {}
    for g in (1, 2):
        opa = tmp_path / f"grid{g}.opa"
        env = dict(os.environ, TRANSIT_B200_GPUS=str(g))
        r = subprocess.run([HOST, "-c", cfg, "--opacityfile", str(opa), "--tlow", "600", "--thigh", "1200", "--tempdelt",
                            "100", "--outspec", str(tmp_path / f"gs{g}.dat")], capture_output=True, text=True, env=env, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        grids[g] = open(opa, "rb").read()
        outs[g] = open(tmp_path / f"gs{g}.dat", "rb").read()
    assert grids[1] == grids[2] and len(grids[1]) > 10000
    assert outs[1] == outs[2]
