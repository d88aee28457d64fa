"""CPU: the C-ABI library loads and exports every symbol declared in include/rnnwf.h; host-only logic
(parameter layout / counts pinned by the reference's notebooks); loud failure without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

from rnnwavefunctions_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "rnnwf.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(rnnwf_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 18
    lib = C.CDLL(_lib.LIB_PATH)
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} is declared in include/rnnwf.h but not exported by librnnwf.so"
    assert lib.rnnwf_version() >= 1


def test_parameter_counts_pinned_by_the_reference_notebooks():
    """422 (pRNN) and 444 (cRNN) parameters: Tutorial_1DTFIM.ipynb cell 15, Tutorial_1DJ1J2.ipynb cell 15."""
    count = lambda **kw: _lib.lib.rnnwf_count_params(C.byref(_lib.Config(**kw)))
    assert count(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=10, num_layers=1) == 422
    assert count(cell=_lib.CELL_GRU1D, complex_head=1, u1=1, nx=10, ny=1, num_units=10, num_layers=1) == 444
    # W1 8102 / 38502, W8 5352, W10 180604, W16 801288
    assert count(cell=_lib.CELL_GRU1D, nx=20, ny=1, num_units=50, num_layers=1) == 8102
    assert count(cell=_lib.CELL_GRU1D, nx=20, ny=1, num_units=50, num_layers=3) == 38502
    assert count(cell=_lib.CELL_VANILLA2D, nx=12, ny=12, num_units=50) == 5352
    assert count(cell=_lib.CELL_TENSGRU2D, complex_head=1, u1=1, nx=10, ny=10, num_units=100) == 180604
    assert count(cell=_lib.CELL_GRU4, nx=8, ny=8, num_units=256) == 801288
    assert count(cell=99, nx=4, ny=4, num_units=8) < 0


def test_philox_twin_is_a_valid_uniform_stream():
    from helpers import philox_uniforms
    u = philox_uniforms(111, 2000, 20)
    assert u.dtype == np.float32 and u.min() >= 0.0 and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 0.01 and abs(u.var() - 1 / 12) < 0.005
    # counter-based: a shard offset reproduces the rows of the full stream (multi-GPU sharding rule)
    np.testing.assert_array_equal(philox_uniforms(111, 50, 20, row_offset=700), u[700:750])
    # Known-answer vector of Philox4x32-10 (Random123 kat_vectors): counter 0, key 0 -> 6627e8d5 ...
    from helpers import philox_uniforms as p
    assert abs(float(p(0, 1, 1)[0, 0]) - (0x6627e8d5 >> 8) / 16777216.0) < 1e-12


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback():
    h = C.c_void_p()
    cfg = _lib.Config(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=10, num_layers=1)
    rc = _lib.lib.rnnwf_create(C.byref(cfg), C.byref(h))
    assert rc == -2 and not h.value
    from rnnwavefunctions_b200.engine import Engine
    with pytest.raises(RuntimeError):
        Engine(_lib.CELL_GRU1D, 10, 1, num_units=10)
