"""GPU parity of the hand-written backward against the oracle's autograd (torch float64 on the CPU) of the
reference's surrogate costs, through the C-ABI.  Tolerance: 1e-4 of the largest gradient entry of each
tensor (FP32 kernels with atomically accumulated weight gradients vs a float64 reference)."""
import numpy as np
import pytest
import torch

from oracle import costs
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run

MODELS = {
    "gru1d": dict(cell=_lib.CELL_GRU1D, nx=12, ny=1, num_units=20, num_layers=1),
    "gru1d_3layer": dict(cell=_lib.CELL_GRU1D, nx=8, ny=1, num_units=14, num_layers=3),
    "cgru1d_u1": dict(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=17, complex_head=True, u1=True),
    "vanilla2d": dict(cell=_lib.CELL_VANILLA2D, nx=4, ny=4, num_units=18),
    "tens2d_u1": dict(cell=_lib.CELL_TENS2D, nx=4, ny=4, num_units=11, complex_head=True, u1=True),
    "tensgru2d_u1": dict(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=20, complex_head=True, u1=True),
    "tensgru2d_free": dict(cell=_lib.CELL_TENSGRU2D, nx=3, ny=4, num_units=9, complex_head=True),
    "rydberg": dict(cell=_lib.CELL_GRU4, nx=3, ny=3, num_units=24),
    # periodic lattices two cells wide: a cell's successor is its predecessor, two neighbour slots name the same cell and
    # both push their input gradient into it (a non-atomic scatter lost one of the two)
    "rydberg_2x2": dict(cell=_lib.CELL_GRU4, nx=2, ny=2, num_units=24),
    "rydberg_1x2": dict(cell=_lib.CELL_GRU4, nx=1, ny=2, num_units=16),
}


def _randomize(eng, seed):
    g = torch.Generator().manual_seed(seed)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("from_sampling", [True, False])
def test_gradient_matches_oracle_autograd(cuda, name, from_sampling):
    ns = 257
    eng = Engine(seed=31, **MODELS[name])
    _randomize(eng, 32)
    g = torch.Generator().manual_seed(33)
    saved = eng.new_saved(ns)
    u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
    if from_sampling:
        s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    else:
        s, _, _ = eng.sample(ns, uniforms=u)
        lp, ph = eng.logpsi(s, saved=saved)
    # arbitrary (complex) "local energies": the seeds of the reference's surrogate costs
    Er = torch.randn(ns, generator=g, dtype=torch.float64)
    Ei = torch.randn(ns, generator=g, dtype=torch.float64)
    P = oracle_params(eng, requires_grad=True)
    ref = oracle_run(eng, P, samples=s.cpu().numpy())
    if eng.complex_head:
        T = 0.3
        cost = costs.cost_complex(ref["logp"], ref["phase"], torch.complex(Er, Ei), T=T)
        sl, sp = costs.seeds_complex(lp.cpu(), torch.complex(Er, Ei), T=T)
    else:
        cost = costs.cost_real(ref["logp"], Er)
        sl, sp = costs.seeds_real(Er), None
    cost.backward()
    grad = eng.backward(s, saved, sl, sp).cpu()
    bad = []
    for n, v in eng.views(grad).items():
        gr = P[n].grad
        scale = float(gr.abs().max()) + 1e-30
        err = float((v.double() - gr).abs().max()) / scale
        if not err < 1e-4 + 1e-6 / scale:      # (absolute floor 1e-6: float atomics on a bias gradient that nearly cancels)
            bad.append((n, err, scale))
    assert not bad, f"gradient mismatch (tensor, rel err, scale): {bad}"


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp32", "tc"])
def test_free_energy_cost_gradient(cuda, precision):
    """G3: the Rydberg scripts minimise the free energy, F_loc = E_loc + T log P in place of E_loc
    (RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-159).  The product forms the seeds in vmc.seeds(free_energy=True); the
    oracle differentiates oracle/costs.py:cost_free_energy."""
    from rnnwavefunctions_b200 import vmc
    ns, T = 300, 0.7
    eng = Engine(seed=41, precision=precision, cell=_lib.CELL_GRU4, nx=3, ny=3, num_units=40)
    _randomize(eng, 42)
    g = torch.Generator().manual_seed(43)
    saved = eng.new_saved(ns)
    s, lp, _ = eng.sample(ns, uniforms=torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32), saved=saved)
    E = torch.randn(ns, generator=g, dtype=torch.float64)
    Ec = torch.complex(E, torch.zeros_like(E)).cuda()
    n, mean_e, var, mean_lp = vmc.energy_moments(Ec, lp)
    sl, sp = vmc.seeds(Ec, mean_e, n, False, logp=lp, mean_logp=mean_lp, T=T, free_energy=True)
    assert sp is None
    grad = eng.backward(s, saved, sl).cpu()
    P = oracle_params(eng, requires_grad=True)
    costs.cost_free_energy(oracle_run(eng, P, samples=s.cpu().numpy())["logp"], E, T).backward()
    tol = 1e-4 if precision == "fp32" else 1e-2
    for name, v in eng.views(grad).items():
        gr = P[name].grad
        assert float((v.double() - gr).abs().max()) < tol * (float(gr.abs().max()) + 1e-30), name
