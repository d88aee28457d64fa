"""Degenerate lattices (single site, single row, single column, 2x2) through both precision regimes of the two-neighbour
2D cells: the FP32 regime against the float64 oracle (1e-5), the tensor-core regime against the FP32 regime (1e-3 on
log psi / E_loc, 1e-2 per tensor on the gradient)."""
import numpy as np
import pytest
import torch

from oracle import costs
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run

LATTICES = [(1, 1), (1, 4), (4, 1), (2, 2), (1, 2)]


def _pair(cell, nx, ny, nh, **kw):
    f32 = Engine(cell, nx, ny, num_units=nh, seed=7, **kw)
    g = torch.Generator().manual_seed(8)
    p = f32.params.cpu()
    f32.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))
    tc = Engine(cell, nx, ny, num_units=nh, seed=7, precision="tc", **kw)
    tc.set_weights(f32.params)
    return f32, tc


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny", LATTICES)
@pytest.mark.parametrize("cell", ["vanilla", "tensgru"])
def test_degenerate_lattices_both_regimes(cuda, cell, nx, ny):
    if cell == "vanilla":
        f32, tc = _pair(_lib.CELL_VANILLA2D, nx, ny, 20)
        ham, hp = _lib.HAM_TFIM2D, [1.3] + [1.0] * (nx * ny)
    else:
        f32, tc = _pair(_lib.CELL_TENSGRU2D, nx, ny, 20, complex_head=True)
        ham, hp = _lib.HAM_HEIS2D, [0.7]
    ns = 150
    u = torch.rand(ns, f32.nsteps, generator=torch.Generator().manual_seed(9), dtype=torch.float32)
    sv = f32.new_saved(ns)
    s, lp, ph = f32.sample(ns, uniforms=u, saved=sv)
    # FP32 regime against the oracle
    with torch.no_grad():
        ref = oracle_run(f32, oracle_params(f32), uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-6).numpy()
    assert (s.cpu().long().numpy()[ok] == ref["samples"].numpy()[ok]).all()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].numpy()[ok], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(ph.cpu().numpy()[ok], ref["phase"].numpy()[ok], rtol=1e-5, atol=1e-5)
    # tensor-core regime against the FP32 regime, on the FP32 regime's samples
    sv2 = tc.new_saved(ns)
    lp2, ph2 = tc.logpsi(s, saved=sv2)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-3, atol=1e-3)
    np.testing.assert_allclose(ph2.cpu().numpy(), ph.cpu().numpy(), rtol=1e-3, atol=1e-2)
    e = f32.local_energies(ham, hp, s, lp, ph, saved=sv).cpu().numpy()
    e2 = tc.local_energies(ham, hp, s, lp2, ph2, saved=sv2).cpu().numpy()
    np.testing.assert_allclose(e2, e, rtol=2e-3, atol=2e-3 * max(1e-30, np.abs(e).max()))
    gen = torch.Generator().manual_seed(10)
    sl = costs.seeds_real(torch.randn(ns, generator=gen, dtype=torch.float64))
    sp = costs.seeds_real(torch.randn(ns, generator=gen, dtype=torch.float64)) if f32.complex_head else None
    ga, gb = tc.backward(s, sv2, sl, sp).cpu(), f32.backward(s, sv, sl, sp).cpu()
    for (n, a), (_, b) in zip(tc.views(ga).items(), f32.views(gb).items()):
        assert float((a - b).abs().max()) <= 1e-2 * float(b.abs().max()) + 1e-9, f"gradient of {n} differs between the regimes"
