"""GPU parity: sampling and log-amplitudes of every wavefunction family against the float64 oracle,
through the C-ABI, on identical weights / samples / uniforms.

Tolerances (BASELINE.json north_star): samples bit-exact except rows where a draw falls within 1e-6 of
a CDF threshold; log-amplitudes within 1e-5 relative (FP32 mode)."""
import numpy as np
import pytest
import torch

from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run, philox_uniforms

RTOL = 1e-5

MODELS = {
    "gru1d": dict(cell=_lib.CELL_GRU1D, nx=20, ny=1, num_units=50, num_layers=1),
    "gru1d_3layer": dict(cell=_lib.CELL_GRU1D, nx=12, ny=1, num_units=18, num_layers=3),
    "cgru1d_u1": dict(cell=_lib.CELL_GRU1D, nx=20, ny=1, num_units=50, num_layers=1, complex_head=True, u1=True),
    "vanilla2d": dict(cell=_lib.CELL_VANILLA2D, nx=6, ny=6, num_units=50),
    "tens2d_u1": dict(cell=_lib.CELL_TENS2D, nx=4, ny=4, num_units=21, complex_head=True, u1=True),
    "tensgru2d_u1": dict(cell=_lib.CELL_TENSGRU2D, nx=6, ny=6, num_units=100, complex_head=True, u1=True),
    "tensgru2d_free": dict(cell=_lib.CELL_TENSGRU2D, nx=3, ny=5, num_units=13, complex_head=True, u1=False),
    "rydberg": dict(cell=_lib.CELL_GRU4, nx=3, ny=3, num_units=40),
    "rydberg_256": dict(cell=_lib.CELL_GRU4, nx=2, ny=2, num_units=256),
}


def _randomize(eng, seed):
    """Break the symmetric default initialisation (biases 0/1) so every parameter matters."""
    g = torch.Generator().manual_seed(seed)
    p = eng.params.cpu()
    p = p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5)
    eng.set_weights(p)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("ns", [37, 700])
def test_sample_and_logpsi_match_oracle(cuda, name, ns):
    eng = Engine(seed=11, **MODELS[name])
    _randomize(eng, 5)
    g = torch.Generator().manual_seed(ns)
    u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u)
    P = oracle_params(eng)
    with torch.no_grad():
        ref = oracle_run(eng, P, uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-6).numpy()
    assert ok.mean() > 0.9
    s_cpu = s.cpu().long().numpy()
    assert (s_cpu[ok] == ref["samples"].numpy()[ok]).all()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].numpy()[ok], rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(ph.cpu().numpy()[ok], ref["phase"].numpy()[ok], rtol=RTOL, atol=1e-5)
    if eng.u1:
        assert (s_cpu.reshape(ns, -1).sum(1) == eng.nsteps // 2).all()
    # teacher-forced evaluation of the GPU's own samples reproduces the sampling pass and the oracle
    lp2, ph2 = eng.logpsi(s)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    with torch.no_grad():
        ref2 = oracle_run(eng, P, samples=s_cpu)
    np.testing.assert_allclose(lp2.cpu().numpy(), ref2["logp"].numpy(), rtol=RTOL, atol=1e-6)
    np.testing.assert_allclose(ph2.cpu().numpy(), ref2["phase"].numpy(), rtol=RTOL, atol=1e-5)


@pytest.mark.gpu
def test_philox_sampling_matches_host_twin(cuda):
    eng = Engine(seed=3, **MODELS["gru1d"])
    ns = 300
    s1, lp1, _ = eng.sample(ns, seed=111, row_offset=1000)
    u = torch.from_numpy(philox_uniforms(111, ns, eng.nsteps, row_offset=1000))
    s2, lp2, _ = eng.sample(ns, uniforms=u)
    assert torch.equal(s1, s2)
    assert torch.equal(lp1, lp2)


@pytest.mark.gpu
def test_normalisation_brute_force(cuda):
    """sum over all 2^N configurations of P(sigma) = 1 (N = 10 chain; 3x3 lattice)."""
    for cfg, shape in ((dict(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=16), (10,)),
                       (dict(cell=_lib.CELL_VANILLA2D, nx=3, ny=3, num_units=10), (3, 3)),
                       (dict(cell=_lib.CELL_TENSGRU2D, nx=3, ny=3, num_units=10, complex_head=True), (3, 3))):
        eng = Engine(seed=2, **cfg)
        _randomize(eng, 9)
        n = int(np.prod(shape))
        allcfg = ((np.arange(2 ** n)[:, None] >> np.arange(n)[None, :]) & 1).astype(np.int8).reshape((-1,) + shape)
        lp, _ = eng.logpsi(torch.from_numpy(allcfg))
        assert abs(float(torch.exp(lp).sum()) - 1.0) < 1e-5


@pytest.mark.gpu
def test_errors_are_loud(cuda):
    eng = Engine(**MODELS["gru1d"])
    with pytest.raises(ValueError):
        eng.logpsi(torch.zeros(4, 7, dtype=torch.int8))
    with pytest.raises(ValueError):
        Engine(cell=_lib.CELL_TENSGRU2D, nx=8, ny=2, num_units=4, u1=True, magnetization=0, complex_head=False, precision="tc3", num_layers=1).local_energies(
            _lib.HAM_TFIM1D, [1.0] * 8, torch.zeros(2, 8, 2, dtype=torch.int8))       # TFIM1D needs a chain
    chain2d = Engine(cell=_lib.CELL_GRU1D, nx=4, ny=2, num_units=4)                    # the 2DTFIM_1DRNN twin: flat [B, 8] rows
    assert chain2d.cfg_shape == (8,)
    with pytest.raises(ValueError):
        chain2d.local_energies(_lib.HAM_TFIM1D, [1.0] * 8, torch.zeros(2, 8, dtype=torch.int8))
