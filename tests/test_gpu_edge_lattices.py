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


@pytest.mark.gpu
@pytest.mark.parametrize("n,layers,complex_head", [(1, 1, False), (2, 1, False), (2, 2, False), (3, 1, True), (2, 1, True), (1, 3, False)])
def test_short_chains_1d_gru(cuda, n, layers, complex_head):
    """Chains of one to three sites through the 1D GRU (sampling, log psi, TFIM / J1-J2 local energies, gradient with the
    weight gradients batched over all sites) against the oracle."""
    from oracle import local_energies as ole
    from helpers import oracle_logpsi_fn
    eng = Engine(_lib.CELL_GRU1D, n, 1, num_units=12, num_layers=layers, complex_head=complex_head, seed=3)
    g = torch.Generator().manual_seed(4)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))
    ns = 130
    u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    P = oracle_params(eng, requires_grad=True)
    ref = oracle_run(eng, P, uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-6).numpy()
    assert (s.cpu().long().numpy()[ok] == ref["samples"].detach().numpy()[ok]).all()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].detach().numpy()[ok], rtol=1e-5, atol=1e-5)
    if complex_head:
        hp = [0.0, 0.0] + [1.0] * n + [0.2] * n + [0.0] * n
        e = eng.local_energies(_lib.HAM_J1J2, hp, s, lp, ph, saved=saved).cpu().numpy()
    else:
        e = eng.local_energies(_lib.HAM_TFIM1D, [0.8] + [1.0] * max(1, n - 1), s, lp, ph, saved=saved).cpu().numpy()
    assert np.isfinite(e).all()
    Er = torch.randn(ns, generator=g, dtype=torch.float64)
    Ei = torch.randn(ns, generator=g, dtype=torch.float64)
    ref2 = oracle_run(eng, P, samples=s.cpu().numpy())
    if complex_head:
        cost = costs.cost_complex(ref2["logp"], ref2["phase"], torch.complex(Er, Ei))
        sl, sp = costs.seeds_complex(lp.cpu(), torch.complex(Er, Ei))
    else:
        cost = costs.cost_real(ref2["logp"], Er)
        sl, sp = costs.seeds_real(Er), None
    cost.backward()
    grad = eng.backward(s, saved, sl, sp).cpu()
    for name, v in eng.views(grad).items():
        gr = P[name].grad
        gr = torch.zeros_like(P[name]) if gr is None else gr
        assert float((v.double() - gr).abs().max()) <= 1e-4 * float(gr.abs().max()) + 1e-9, f"gradient of {name} differs"


@pytest.mark.gpu
@pytest.mark.parametrize("ns", [1, 127, 128, 129, 257])
def test_tile_boundaries_tensor_core_regimes(cuda, ns):
    """Batch sizes around the 128-row tile of the tensor-core sweeps (one row; one short of, exactly, one past a tile)."""
    for cell, kw in ((_lib.CELL_TENSGRU2D, dict(complex_head=True, u1=True)), (_lib.CELL_GRU4, dict())):
        f32, tc = _pair(cell, 3, 2, 24, **kw)
        u = torch.rand(ns, f32.nsteps, generator=torch.Generator().manual_seed(ns), dtype=torch.float32)
        s, lp, ph = f32.sample(ns, uniforms=u)
        sv = tc.new_saved(ns)
        lp2, ph2 = tc.logpsi(s, saved=sv)
        np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-3, atol=1e-3)
        ham, hp = (_lib.HAM_HEIS2D, [1.0]) if cell == _lib.CELL_TENSGRU2D else (_lib.HAM_RYDBERG, [1.7 ** 6, 3.3, 1.0, 2.001])
        e = f32.local_energies(ham, hp, s, lp, ph).cpu().numpy()
        e2 = tc.local_energies(ham, hp, s, lp2, ph2, saved=sv).cpu().numpy()
        np.testing.assert_allclose(e2, e, rtol=2e-3, atol=2e-3 * max(1e-30, np.abs(e).max()))
        grad = tc.backward(s, sv, costs.seeds_real(torch.randn(ns, generator=torch.Generator().manual_seed(1), dtype=torch.float64)) if ns > 1
                           else torch.ones(1, dtype=torch.float64), None)
        assert bool(torch.isfinite(grad).all())
