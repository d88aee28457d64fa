"""rnnwf_logpsi_static: the log-amplitude entry point for hosts that cannot allocate device scratch
(a process whose runtime already owns the whole GPU, or a bare ctypes script holding only its input and
output arrays).  It must give what rnnwf_logpsi gives -- same kernel, the history ring merely lives in the
16 MiB array built into the library -- and keep doing so on back-to-back calls, because the array is shared
by every call and is never cleared between them.

This file sorts last on purpose: a failure here cannot hide the result of any other test under `-x`.
Called straight through the C-ABI (ctypes), not through the Engine wrapper.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine

RTOL = 1e-5  # the log-amplitude tolerance of tests/test_gpu_forward.py


def _static_logpsi(eng, samples):
    s = samples.to(device=eng.device, dtype=torch.int8).contiguous()
    rows = s.shape[0]
    logp = torch.full((rows,), float("nan"), dtype=torch.float64, device=eng.device)
    phase = torch.full((rows,), float("nan"), dtype=torch.float64, device=eng.device)
    rc = _lib.lib.rnnwf_logpsi_static(eng._h, rows, C.c_void_p(s.data_ptr()), C.c_void_p(logp.data_ptr()),
                                      C.c_void_p(phase.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, logp, phase


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [
    dict(cell=_lib.CELL_TENSGRU2D, nx=6, ny=6, num_units=100, complex_head=True, u1=True),
    dict(cell=_lib.CELL_GRU1D, nx=20, ny=1, num_units=50, num_layers=2),
    dict(cell=_lib.CELL_GRU4, nx=3, ny=3, num_units=40),
])
def test_static_scratch_matches_workspace_entry_point_twice(cuda, cfg):
    eng = Engine(seed=41, **cfg)
    g = torch.Generator().manual_seed(42)
    for call, rows in enumerate((3000, 517)):   # second call: other rows, other tile shape, stale ring contents
        u = torch.rand(rows, eng.nsteps, generator=g, dtype=torch.float32)
        samples, _, _ = eng.sample(rows, uniforms=u)
        want_lp, want_ph = eng.logpsi(samples)
        rc, lp, ph = _static_logpsi(eng, samples)
        assert rc == 0, _lib.lib.rnnwf_last_error().decode()
        torch.cuda.synchronize()
        np.testing.assert_allclose(lp.cpu().numpy(), want_lp.cpu().numpy(), rtol=RTOL, atol=1e-6, err_msg=f"call {call}")
        np.testing.assert_allclose(ph.cpu().numpy(), want_ph.cpu().numpy(), rtol=RTOL, atol=1e-5, err_msg=f"call {call}")


@pytest.mark.gpu
def test_static_scratch_refuses_models_that_do_not_fit(cuda):
    # 12x12 periodic cells x 1024 units: one block's ring is 144*32*1024*4 B = 18.9 MB > 16 MiB
    eng = Engine(cell=_lib.CELL_GRU4, nx=12, ny=12, num_units=1024, seed=1)
    samples = torch.zeros(4, 12, 12, 3, dtype=torch.int8)
    rc, _, _ = _static_logpsi(eng, samples)
    assert rc == -3
    msg = _lib.lib.rnnwf_last_error().decode()
    assert "RNNWF_STATIC_SCRATCH_BYTES" in msg and "rnnwf_logpsi" in msg
    # the regular entry point still serves the same model
    lp, _ = eng.logpsi(samples)
    assert torch.isfinite(lp).all()


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [
    dict(cell=_lib.CELL_TENSGRU2D, nx=6, ny=6, num_units=48, complex_head=True, u1=True),
    dict(cell=_lib.CELL_GRU4, nx=3, ny=3, num_units=72),
])
def test_static_scratch_on_a_tensor_core_handle_runs_the_fp32_sweep(cuda, cfg):
    """A handle created in the reduced-precision regime sizes its workspace ring for the tensor-core kernels; the
    entry point without a workspace must not inherit that shape: it always runs the FP32 sweep on the same weights."""
    tc = Engine(seed=43, precision="tc", **cfg)
    f32 = Engine(seed=43, **cfg)
    f32.set_weights(tc.params)
    u = torch.rand(700, tc.nsteps, generator=torch.Generator().manual_seed(44), dtype=torch.float32)
    samples, _, _ = f32.sample(700, uniforms=u)
    want_lp, want_ph = f32.logpsi(samples)
    for call in range(2):
        rc, lp, ph = _static_logpsi(tc, samples)
        assert rc == 0, _lib.lib.rnnwf_last_error().decode()
        torch.cuda.synchronize()
        np.testing.assert_allclose(lp.cpu().numpy(), want_lp.cpu().numpy(), rtol=RTOL, atol=1e-6, err_msg=f"call {call}")
        np.testing.assert_allclose(ph.cpu().numpy(), want_ph.cpu().numpy(), rtol=RTOL, atol=1e-5, err_msg=f"call {call}")
