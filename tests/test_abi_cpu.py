"""CPU-side checks of the boundary: the library loads, exports every symbol the header
declares, and refuses to run without a GPU (no silent CPU fallback)."""
import os
import re

import numpy as np
import pytest

from transit_b200 import api, workloads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "transit_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(tb_[a-z_A-Z0-9]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = api.load_library()
    syms = header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/transit_b200.h but not exported"
        assert s in api.ABI, f"{s} has no ctypes signature in transit_b200/api.py"
    assert sorted(api.ABI) == syms


def test_grid_count_matches_makesample1():
    lib = api.load_library()
    # transit_demo.cfg: 2500-5000 cm-1 at 1 cm-1, oversampling 2160 (C1)
    assert lib.tb_grid_count(2500.0, 5000.0, 1.0, 1) == 2501
    assert lib.tb_grid_count(2500.0, 5000.0, 1.0, 2160) == 5400001
    assert workloads.grid_count(2000.0, 33333.0, 1.0, 1) == 31334


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.TransitError, match="no CPU fallback"):
        api.Context(0)


def test_product_does_not_import_oracle():
    """The product path must never route through oracle/ (test infrastructure only)."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "transit_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".c")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oraclelib" not in txt and "liboracle" not in txt and "transit_oracle" not in txt, f


def test_workload_shapes():
    p = workloads.make_problem("C2", scale=0.002, nlayers=12)
    assert p.nwn == 31334 and p.onwn == (31334 - 1) * 2160 + 1
    assert p.mol_d.shape == (6, 12) and p.iso_z.shape == (26, 12)
    # per isotope ascending wavelength (TLI order)
    for i in range(p.niso):
        w = p.wl[p.isoid == i]
        assert np.all(np.diff(w) >= 0)
