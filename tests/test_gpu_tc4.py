"""GPU parity of the tensor-core (tcgen05, FP16 operands) regime of the four-neighbour 2D GRU of the Rydberg Kagome
wavefunction (csrc/sweep_tc4.cuh) against the float64 oracle.  Tolerance: 1e-3 relative on log-probabilities and E_loc
(the reduced-precision bound of BASELINE.json's north_star); samples must agree with the oracle wherever no draw falls
within 1e-3 of a CDF threshold; the gradient (FFMA backward fed with the saved state this sweep wrote) within 1e-2 per
tensor of the oracle's autograd."""
import numpy as np
import pytest
import torch

from oracle import local_energies as ole
from oracle import costs
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run, oracle_logpsi_fn, assert_rel

RTOL = 1e-3
HAM = [1.7 ** 6, 3.3, 1.0, 2.001]


def _mk(nx, ny, nh, seed=71, precision="tc"):
    eng = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=seed, precision=precision)
    g = torch.Generator().manual_seed(seed + 1)
    p = eng.params.cpu()
    eng.set_weights(p + 0.1 * (torch.rand(p.shape, generator=g) - 0.5))
    return eng


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny,nh,ns", [(3, 3, 40, 120), (2, 3, 64, 37), (4, 4, 256, 300), (5, 3, 100, 200), (4, 5, 130, 20000)])
def test_tc4_sample_and_logpsi(cuda, nx, ny, nh, ns):
    eng = _mk(nx, ny, nh)
    assert "tcgen05" in eng.describe()
    nref = min(ns, 300)                      # the oracle checks the first rows; the rest through the FP32 regime below
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(ns), dtype=torch.float32)
    s, lp, _ = eng.sample(ns, uniforms=u)
    P = oracle_params(eng)
    with torch.no_grad():
        ref = oracle_run(eng, P, uniforms=u[:nref], numsamples=nref)
    ok = (ref["margin"].min(dim=1).values > 1e-3).numpy()
    assert ok.mean() > 0.5
    s_cpu = s.cpu().long().numpy()
    assert (s_cpu[:nref][ok] == ref["samples"].numpy()[ok]).all()
    scale = np.abs(ref["logp"].numpy()).max()
    np.testing.assert_allclose(lp.cpu().numpy()[:nref][ok], ref["logp"].numpy()[ok], rtol=RTOL, atol=RTOL * scale * 0.1)
    # teacher forcing reproduces the sampling pass bit for bit, and agrees with the oracle on the drawn configurations
    lp2, _ = eng.logpsi(s)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    with torch.no_grad():
        ref2 = oracle_run(eng, P, samples=s_cpu[:nref])
    np.testing.assert_allclose(lp2.cpu().numpy()[:nref], ref2["logp"].numpy(), rtol=RTOL, atol=RTOL * scale * 0.1)
    if ns > nref:                            # more tiles than SMs: every row against the FP32 regime on the same weights
        f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=71)
        f32.set_weights(eng.params)
        lp3, _ = f32.logpsi(s)
        np.testing.assert_allclose(lp2.cpu().numpy(), lp3.cpu().numpy(), rtol=RTOL, atol=RTOL * scale * 0.1)


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
@pytest.mark.parametrize("nx,ny,nh", [(3, 3, 40), (4, 3, 256)])
def test_tc4_rydberg_local_energies(cuda, reuse, nx, ny, nh):
    eng = _mk(nx, ny, nh, seed=81)
    ns = 150
    saved = eng.new_saved(ns) if reuse else None
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(5), dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    e = eng.local_energies(_lib.HAM_RYDBERG, HAM, s, lp, ph, saved=saved).cpu().numpy()
    pos = ole.generate_kagome(nx, ny)
    nb, dist = ole.neighbouring_atoms(pos, HAM[3], nx, ny)
    ref = ole.rydberg_local_energies(HAM[0], HAM[1], HAM[2], nb, dist, s.cpu().numpy().astype(np.int64),
                                     oracle_logpsi_fn(eng, oracle_params(eng)))
    assert_rel(e.real, ref, RTOL, "E_loc tc")


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny,nh,ns", [(3, 3, 72, 200), (2, 2, 40, 17), (4, 3, 256, 300), (3, 4, 130, 19500)])
def test_tc4_backward(cuda, nx, ny, nh, ns):
    """Tensor-core backward (bwd_tc4.cuh: persistent reverse sweep + weight gradients) fed with the hidden states and gates
    the tensor-core sweep saved, against the oracle's autograd; the last case has more tiles than SMs."""
    eng = _mk(nx, ny, nh, seed=91)
    saved = eng.new_saved(ns)
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(9), dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    E = torch.randn(ns, generator=torch.Generator().manual_seed(10), dtype=torch.float64)
    Pg = oracle_params(eng, requires_grad=True)
    nref = min(ns, 400)          # the oracle differentiates the cost of the first rows: zero seeds for the rest
    E[nref:] = E[:nref].mean()
    r = oracle_run(eng, Pg, samples=s.cpu().long().numpy()[:nref])
    (costs.cost_real(r["logp"], E[:nref]) * (nref / ns)).backward()
    grad = eng.backward(s, saved, costs.seeds_real(E)).cpu()
    for n, v in eng.views(grad).items():
        gr = Pg[n].grad
        assert float((v.double() - gr).abs().max()) <= 1e-2 * float(gr.abs().max()) + 1e-30, f"gradient of {n} differs"
    if ns > nref:
        # every tile carries weight now (more tiles than persistent CTAs): against the FP32 regime on the same samples
        f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=91)
        f32.set_weights(eng.params)
        sv2 = f32.new_saved(ns)
        f32.logpsi(s, saved=sv2)
        seeds = costs.seeds_real(torch.randn(ns, generator=torch.Generator().manual_seed(11), dtype=torch.float64))
        ga, gb = eng.backward(s, saved, seeds).cpu(), f32.backward(s, sv2, seeds).cpu()
        for (n, a), (_, b) in zip(eng.views(ga).items(), f32.views(gb).items()):
            assert float((a - b).abs().max()) <= 1e-2 * float(b.abs().max()) + 1e-30, f"gradient of {n} differs from the FP32 regime"


@pytest.mark.gpu
def test_tc4_large_lattice_matches_fp32_regime(cuda):
    """8x8 unit cells (192 atoms, the primary C5 size): log P and the prefix-reuse local energies of the tensor-core regime
    against the FP32 regime on the same weights and samples (the float64 oracle is too slow at this size); the
    periodic look-back spans the whole lattice here (ring of 64 cells, row 0 feeds the last row)."""
    nx = ny = 8
    nh, ns = 96, 300
    eng = _mk(nx, ny, nh, seed=101)
    f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=101)
    f32.set_weights(eng.params)
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, seed=7, row_offset=0, saved=saved)
    lp32, ph32 = f32.logpsi(s)
    scale = float(lp32.abs().max())
    np.testing.assert_allclose(lp.cpu().numpy(), lp32.cpu().numpy(), rtol=RTOL, atol=RTOL * scale * 0.1)
    e = eng.local_energies(_lib.HAM_RYDBERG, HAM, s, lp, ph, saved=saved).cpu().numpy().real
    e_noreuse = eng.local_energies(_lib.HAM_RYDBERG, HAM, s, lp, ph).cpu().numpy().real
    e32 = f32.local_energies(_lib.HAM_RYDBERG, HAM, s, lp32, ph32).cpu().numpy().real
    np.testing.assert_allclose(e, e32, rtol=2e-3, atol=2e-3 * np.abs(e32).max())
    np.testing.assert_allclose(e_noreuse, e, rtol=1e-4, atol=1e-4 * np.abs(e).max())


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny", [(1, 1), (1, 2), (2, 1), (1, 5)])
def test_tc4_degenerate_lattices(cuda, nx, ny):
    """Single cells and chains of cells: sites with no neighbour at all (only the input / bias job runs) and lattices whose
    periodic successor is the predecessor; sampling, teacher forcing, prefix-reuse local energies and the gradient from a
    teacher-forced saved state, against the FP32 regime on the same weights."""
    nh, ns = 40, 200
    eng = _mk(nx, ny, nh, seed=111)
    f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=111)
    f32.set_weights(eng.params)
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(3), dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u)
    s32, lp32, ph32 = f32.sample(ns, uniforms=u)
    same = (s == s32).reshape(ns, -1).all(dim=1).cpu().numpy()
    assert same.mean() > 0.9
    np.testing.assert_allclose(lp.cpu().numpy()[same], lp32.cpu().numpy()[same], rtol=RTOL, atol=1e-3)
    saved = eng.new_saved(ns)
    lp2, ph2 = eng.logpsi(s, saved=saved)           # teacher forcing writes the saved state the backward consumes
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    e = eng.local_energies(_lib.HAM_RYDBERG, HAM, s, lp2, ph2, saved=saved).cpu().numpy().real
    sv32 = f32.new_saved(ns)
    lp3, ph3 = f32.logpsi(s, saved=sv32)
    e32 = f32.local_energies(_lib.HAM_RYDBERG, HAM, s, lp3, ph3, saved=sv32).cpu().numpy().real
    np.testing.assert_allclose(e, e32, rtol=2e-3, atol=2e-3 * np.abs(e32).max())
    seeds = costs.seeds_real(torch.randn(ns, generator=torch.Generator().manual_seed(4), dtype=torch.float64))
    ga, gb = eng.backward(s, saved, seeds).cpu(), f32.backward(s, sv32, seeds).cpu()
    for (n, a), (_, b) in zip(eng.views(ga).items(), f32.views(gb).items()):
        assert float((a - b).abs().max()) <= 1e-2 * float(b.abs().max()) + 1e-30, f"gradient of {n} differs from the FP32 regime"


@pytest.mark.gpu
def test_tc4_local_energies_more_tiles_than_sms(cuda):
    """Every persistent CTA walks several tiles of connected configurations (317 tiles of 128 rows), each restarting from
    its parents' saved history at a different cell: against the FP32 regime on the same samples."""
    nx, ny, nh, ns = 3, 3, 64, 1500
    eng = _mk(nx, ny, nh, seed=121)
    f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=121)
    f32.set_weights(eng.params)
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, seed=11, row_offset=0, saved=saved)
    e = eng.local_energies(_lib.HAM_RYDBERG, HAM, s, lp, ph, saved=saved).cpu().numpy().real
    sv32 = f32.new_saved(ns)
    lp32, ph32 = f32.logpsi(s, saved=sv32)
    e32 = f32.local_energies(_lib.HAM_RYDBERG, HAM, s, lp32, ph32, saved=sv32).cpu().numpy().real
    np.testing.assert_allclose(e, e32, rtol=2e-3, atol=2e-3 * np.abs(e32).max())
    assert abs(e.mean() - e32.mean()) < 2e-4 * abs(e32.mean())


@pytest.mark.gpu
def test_tc4_training_tracks_the_fp32_regime(cuda):
    """End to end (north_star: 'variational energies agree within statistical error after equal training steps'): the same
    VMC run -- 3x3 Kagome cells, same Philox seeds, TF-Adam -- in both precision regimes.  The first step sees identical
    weights (energies agree to 1e-3); afterwards both trajectories descend and stay statistically equivalent."""
    from rnnwavefunctions_b200 import vmc
    out = {}
    for prec in ("fp32", "tc"):
        eng = Engine(_lib.CELL_GRU4, 3, 3, num_units=32, seed=5, precision=prec)
        eng.init_params(seed=5)
        v = vmc.VMC(eng, _lib.HAM_RYDBERG, HAM, 512, lr=5e-3, seed=21)
        es = []
        for it in range(60):
            mean_e, var = v.step()
            es.append(float(mean_e.real))
        out[prec] = np.array(es)
    a, b = out["fp32"], out["tc"]
    assert abs(a[0] - b[0]) < 1e-3 * abs(a[0])
    assert np.isfinite(b).all()
    assert a[-10:].mean() < a[:3].mean() and b[-10:].mean() < b[:3].mean()
    spread = max(a[-10:].std(), b[-10:].std(), 1e-3 * abs(a[-10:].mean()))
    assert abs(a[-10:].mean() - b[-10:].mean()) < 4 * spread + 0.02 * abs(a[-10:].mean())
