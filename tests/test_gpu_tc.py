"""GPU parity of the tensor-core (tcgen05, FP16 operands / FP32 accumulate) regime of the tensorized 2D GRU sweep against the
float64 oracle.  Tolerance: 1e-3 relative on log-amplitudes and E_loc (the reduced-precision bound of BASELINE.json's north_star);
samples must agree with the oracle wherever no draw falls within 1e-3 of a CDF threshold."""
import numpy as np
import pytest
import torch

from oracle import local_energies as ole
from oracle import costs
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run, oracle_logpsi_fn, assert_rel

RTOL = 1e-3


def _mk(nx, ny, nh, seed=51, u1=True):
    eng = Engine(_lib.CELL_TENSGRU2D, nx, ny, num_units=nh, complex_head=True, u1=u1, seed=seed, precision="tc")
    g = torch.Generator().manual_seed(seed + 1)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))
    return eng


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny,nh,ns", [(4, 4, 24, 100), (6, 6, 100, 700), (3, 5, 13, 37), (10, 10, 100, 300), (16, 16, 20, 150)])
def test_tc_sample_and_logpsi(cuda, nx, ny, nh, ns):
    u1 = (nx * ny) % 2 == 0     # zero magnetisation needs an even number of sites
    eng = _mk(nx, ny, nh, u1=u1)
    assert "tcgen05" in eng.describe()
    g = torch.Generator().manual_seed(ns)
    u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u)
    P = oracle_params(eng)
    with torch.no_grad():
        ref = oracle_run(eng, P, uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-3).numpy()
    assert ok.mean() > 0.4     # (256 draws per row on the largest lattice: more rows come within 1e-3 of a threshold)
    s_cpu = s.cpu().long().numpy()
    assert (s_cpu[ok] == ref["samples"].numpy()[ok]).all()
    if u1:
        assert (s_cpu.reshape(ns, -1).sum(1) == eng.nsteps // 2).all()
    scale = np.abs(ref["logp"].numpy()).max()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].numpy()[ok], rtol=RTOL, atol=RTOL * scale * 0.1)
    np.testing.assert_allclose(ph.cpu().numpy()[ok], ref["phase"].numpy()[ok], rtol=RTOL, atol=RTOL * 0.5 * scale)
    lp2, ph2 = eng.logpsi(s)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    with torch.no_grad():
        ref2 = oracle_run(eng, P, samples=s_cpu)
    np.testing.assert_allclose(lp2.cpu().numpy(), ref2["logp"].numpy(), rtol=RTOL, atol=RTOL * scale * 0.1)


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
@pytest.mark.parametrize("J2", [0.0, 1.0])
def test_tc_local_energies(cuda, reuse, J2):
    eng = _mk(6, 6, 48, seed=61)
    ns = 200
    saved = eng.new_saved(ns) if reuse else None
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(3), dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    e = eng.local_energies(_lib.HAM_HEIS2D, [J2], s, lp, ph, saved=saved).cpu().numpy()
    ref, _ = ole.heis2d_local_energies(J2, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    assert_rel(e, ref, RTOL, "E_loc tc")


@pytest.mark.gpu
def test_tc_matches_fp32_regime_and_feeds_backward(cuda):
    """Same weights in both regimes: log psi within 1e-3, and the backward (which consumes the saved states the
    tensor-core sweep wrote) within 1e-2 of the oracle's autograd per tensor (the saved hidden states carry the
    reduced-precision regime's perturbation)."""
    kw = dict(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=20, complex_head=True, u1=True, seed=71)
    e32, e16 = Engine(precision="fp32", **kw), Engine(precision="tc", **kw)
    g = torch.Generator().manual_seed(72)
    p = e32.params.cpu() + 0.15 * (torch.rand(e32.params.shape, generator=g) - 0.5)
    e32.set_weights(p)
    e16.set_weights(p)
    ns = 257
    u = torch.rand(ns, 16, generator=g, dtype=torch.float32)
    s, lp32, ph32 = e32.sample(ns, uniforms=u)
    saved = e16.new_saved(ns)
    lp16, ph16 = e16.logpsi(s, saved=saved)
    np.testing.assert_allclose(lp16.cpu().numpy(), lp32.cpu().numpy(), rtol=RTOL, atol=1e-3)
    np.testing.assert_allclose(ph16.cpu().numpy(), ph32.cpu().numpy(), rtol=RTOL, atol=1e-2)
    Er = torch.randn(ns, generator=g, dtype=torch.float64)
    Ei = torch.randn(ns, generator=g, dtype=torch.float64)
    P = oracle_params(e16, requires_grad=True)
    ref = oracle_run(e16, P, samples=s.cpu().numpy())
    costs.cost_complex(ref["logp"], ref["phase"], torch.complex(Er, Ei)).backward()
    sl, sp = costs.seeds_complex(lp16.cpu(), torch.complex(Er, Ei))
    grad = e16.backward(s, saved, sl, sp).cpu()
    for n, v in e16.views(grad).items():
        gr = P[n].grad
        scale = float(gr.abs().max()) + 1e-30
        assert float((v.double() - gr).abs().max()) / scale < 1e-2, n


@pytest.mark.gpu
def test_tc_persistent_tiles_are_row_independent(cuda):
    """More 128-row tiles than SMs (every CTA walks several tiles, the last one ragged): each row's result must not
    depend on the batch it was evaluated in, and the backward of a big batch must equal the sum over its halves."""
    eng = _mk(4, 4, 20, seed=91)
    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    ns = nsm * 128 * 2 + 77
    s, lp, ph = eng.sample(ns, seed=5)
    saved = eng.new_saved(ns)
    lp2, ph2 = eng.logpsi(s, saved=saved)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    cut = 128 * 37 + 5                                   # a split that is not tile-aligned
    for lo, hi in ((0, cut), (cut, ns)):
        l3, p3 = eng.logpsi(s[lo:hi])
        assert torch.equal(l3, lp2[lo:hi]) and torch.equal(p3, ph2[lo:hi])
    g = torch.Generator().manual_seed(3)
    sl = (torch.randn(ns, generator=g, dtype=torch.float64) / ns).cuda()
    sp = (torch.randn(ns, generator=g, dtype=torch.float64) / ns).cuda()
    grad = eng.backward(s, saved, sl, sp).double()
    parts = torch.zeros_like(grad)
    for lo, hi in ((0, cut), (cut, ns)):
        sv = eng.new_saved(hi - lo)
        eng.logpsi(s[lo:hi], saved=sv)
        parts += eng.backward(s[lo:hi], sv, sl[lo:hi], sp[lo:hi]).double()
    scale = float(grad.abs().max())
    assert float((grad - parts).abs().max()) < 2e-4 * scale    # same terms, different (atomic / split) summation order


@pytest.mark.gpu
@pytest.mark.parametrize("complex_head,ns", [(False, 130), (True, 1), (True, 3)])
def test_tc_real_head_and_tiny_batches(cuda, complex_head, ns):
    """Real-valued heads (two logits per unit in the packed layout) and batches smaller than one tile, against the
    FP32 regime on the same weights: sample with shared uniforms, log psi, local energies, gradient."""
    kw = dict(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=20, complex_head=complex_head, u1=True, seed=17)
    e32, e16 = Engine(precision="fp32", **kw), Engine(precision="tc", **kw)
    g = torch.Generator().manual_seed(18)
    p = e32.params.cpu() + 0.15 * (torch.rand(e32.params.shape, generator=g) - 0.5)
    e32.set_weights(p)
    e16.set_weights(p)
    u = torch.rand(ns, 16, generator=g, dtype=torch.float32)
    s, lp32, ph32 = e32.sample(ns, uniforms=u)
    saved = e16.new_saved(ns)
    lp16, ph16 = e16.logpsi(s, saved=saved)
    np.testing.assert_allclose(lp16.cpu().numpy(), lp32.cpu().numpy(), rtol=RTOL, atol=1e-3)
    np.testing.assert_allclose(ph16.cpu().numpy(), ph32.cpu().numpy(), rtol=RTOL, atol=1e-2)
    ea = e32.local_energies(_lib.HAM_HEIS2D, [1.0], s, lp32, ph32).cpu().numpy()
    eb = e16.local_energies(_lib.HAM_HEIS2D, [1.0], s, lp16, ph16, saved=saved).cpu().numpy()
    np.testing.assert_allclose(eb, ea, rtol=5 * RTOL, atol=5 * RTOL * np.abs(ea).max())
    sl = torch.randn(ns, generator=g, dtype=torch.float64) / ns
    sp = torch.randn(ns, generator=g, dtype=torch.float64) / ns if complex_head else None
    sv32 = e32.new_saved(ns)
    e32.logpsi(s, saved=sv32)
    ga = e32.backward(s, sv32, sl, sp).cpu().double()
    gb = e16.backward(s, saved, sl, sp).cpu().double()
    for n, v in e16.views(gb).items():
        ref = e32.views(ga)[n]
        assert float((v - ref).abs().max()) <= 1e-2 * (float(ref.abs().max()) + 1e-30), n


@pytest.mark.gpu
def test_tc_training_tracks_the_fp32_regime(cuda):
    """End to end (north_star: 'variational energies agree within statistical error after equal training steps'):
    the same VMC run -- 4x4 Heisenberg, Sz = 0, same Philox seeds, TF-Adam -- in both precision regimes.  The first
    step sees identical weights (energies agree to 1e-3); afterwards the trajectories are statistically equivalent
    and both move towards the exact ground state (-9.1892 for the 4x4 open square lattice)."""
    from rnnwavefunctions_b200 import vmc
    out = {}
    for prec in ("fp32", "tc"):
        eng = Engine(_lib.CELL_TENSGRU2D, 4, 4, num_units=20, complex_head=True, u1=True, seed=5, precision=prec)
        eng.init_params(seed=5)
        v = vmc.VMC(eng, _lib.HAM_HEIS2D, [0.0], 512, lr=5e-3, seed=21)
        es = []
        for it in range(80):
            mean_e, var = v.step()
            es.append(float(mean_e.real))
        out[prec] = np.array(es)
    a, b = out["fp32"], out["tc"]
    assert abs(a[0] - b[0]) < 1e-3 * abs(a[0])
    assert np.isfinite(b).all()
    assert a[-10:].mean() < a[:3].mean() - 1.0 and b[-10:].mean() < b[:3].mean() - 1.0
    assert abs(a[-10:].mean() - b[-10:].mean()) < 0.25
    assert -9.2 < b[-10:].mean() < -5.0


def _mk_vanilla(nx, ny, nh, seed=31, precision="tc"):
    eng = Engine(_lib.CELL_VANILLA2D, nx, ny, num_units=nh, seed=seed, precision=precision)
    g = torch.Generator().manual_seed(seed + 1)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))
    return eng


@pytest.mark.gpu
@pytest.mark.parametrize("nx,ny,nh,ns", [(5, 4, 50, 300), (12, 12, 50, 130), (3, 3, 10, 40), (4, 4, 64, 129)])
def test_tc_vanilla_sample_and_logpsi(cuda, nx, ny, nh, ns):
    """2DTFIM_2DRNN vanilla cell in the tensor-core regime against the float64 oracle (same bounds as the tGRU)."""
    eng = _mk_vanilla(nx, ny, nh)
    assert "tcgen05" in eng.describe()
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(ns), dtype=torch.float32)
    s, lp, ph = eng.sample(ns, uniforms=u)
    P = oracle_params(eng)
    with torch.no_grad():
        ref = oracle_run(eng, P, uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-3).numpy()
    assert ok.mean() > 0.4
    s_cpu = s.cpu().long().numpy()
    assert (s_cpu[ok] == ref["samples"].numpy()[ok]).all()
    scale = np.abs(ref["logp"].numpy()).max()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].numpy()[ok], rtol=RTOL, atol=RTOL * scale * 0.1)
    lp2, _ = eng.logpsi(s)
    np.testing.assert_allclose(lp2.cpu().numpy(), lp.cpu().numpy(), rtol=1e-12, atol=1e-12)
    with torch.no_grad():
        ref2 = oracle_run(eng, P, samples=s_cpu)
    np.testing.assert_allclose(lp2.cpu().numpy(), ref2["logp"].numpy(), rtol=RTOL, atol=RTOL * scale * 0.1)


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
def test_tc_vanilla_tfim2d_local_energies_and_gradient(cuda, reuse):
    Nx, Ny, ns = 5, 4, 300
    eng = _mk_vanilla(Nx, Ny, 50, seed=41)
    e32 = Engine(_lib.CELL_VANILLA2D, Nx, Ny, num_units=50, seed=41, precision="fp32")
    e32.set_weights(eng.params)
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, seed=6, saved=saved)
    Jz = np.random.RandomState(0).uniform(0.5, 1.5, (Nx, Ny))
    hp = np.concatenate([[3.0], Jz.reshape(-1)])
    e = eng.local_energies(_lib.HAM_TFIM2D, hp, s, lp, ph, saved=saved if reuse else None).cpu().numpy()
    ref = ole.ising2d_local_energies(Jz, 3.0, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    # The un-gated elu network amplifies the FP16 operand rounding of this regime: 1.1e-3 here, ~1e-2 at the 12x12 lattice
    # of C3 (tools/cmp_regimes.py) -- outside north_star's 1e-3.  The regime that serves the vanilla cell on the tensor
    # cores is the split-precision one (precision="tc3", tests/test_gpu_tc3.py); this one stays available, bound stated here.
    assert_rel(e.real, ref, 2e-3, "E_loc tc (vanilla cell)")
    # gradient: generic backward over the history the tensor-core sweep saved, against the FP32 regime
    g = torch.Generator().manual_seed(3)
    sl = torch.randn(ns, generator=g, dtype=torch.float64) / ns
    sv32 = e32.new_saved(ns)
    e32.logpsi(s, saved=sv32)
    ga = e32.backward(s, sv32, sl).cpu().double()
    gb = eng.backward(s, saved, sl).cpu().double()
    for n, v in eng.views(gb).items():
        r = e32.views(ga)[n]
        assert float((v - r).abs().max()) <= 1e-2 * (float(r.abs().max()) + 1e-30), n


@pytest.mark.gpu
@pytest.mark.parametrize("nh", [20, 24])
def test_tc_ungated_tensorized_cell(cuda, nh):
    """2DHeisenberg/MDTensorizedcell.py (tanh(T:W + b)) in the tensor-core regime: samples / log psi / E_loc against the
    oracle, gradient against the oracle's autograd (both the bias-in-MMA and the plain variants: nh = 20 / 24)."""
    eng = Engine(_lib.CELL_TENS2D, 4, 4, num_units=nh, complex_head=True, u1=True, seed=77, precision="tc")
    g = torch.Generator().manual_seed(78)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))
    assert "tcgen05" in eng.describe()
    ns = 257
    u = torch.rand(ns, 16, generator=g, dtype=torch.float32)
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
    P = oracle_params(eng)
    with torch.no_grad():
        ref = oracle_run(eng, P, uniforms=u, numsamples=ns)
    ok = (ref["margin"].min(dim=1).values > 1e-3).numpy()
    s_cpu = s.cpu().long().numpy()
    assert ok.mean() > 0.4 and (s_cpu[ok] == ref["samples"].numpy()[ok]).all()
    scale = np.abs(ref["logp"].numpy()).max()
    np.testing.assert_allclose(lp.cpu().numpy()[ok], ref["logp"].numpy()[ok], rtol=RTOL, atol=RTOL * scale * 0.1)
    np.testing.assert_allclose(ph.cpu().numpy()[ok], ref["phase"].numpy()[ok], rtol=RTOL, atol=RTOL * 0.5 * scale)
    e = eng.local_energies(_lib.HAM_HEIS2D, [1.0], s, lp, ph, saved=saved).cpu().numpy()
    eref, _ = ole.heis2d_local_energies(1.0, s_cpu, oracle_logpsi_fn(eng, P))
    np.testing.assert_allclose(e, eref, rtol=RTOL, atol=RTOL * np.abs(eref).max())
    Er = torch.randn(ns, generator=g, dtype=torch.float64)
    Ei = torch.randn(ns, generator=g, dtype=torch.float64)
    Pg = oracle_params(eng, requires_grad=True)
    r = oracle_run(eng, Pg, samples=s_cpu)
    costs.cost_complex(r["logp"], r["phase"], torch.complex(Er, Ei)).backward()
    sl, sp = costs.seeds_complex(lp.cpu(), torch.complex(Er, Ei))
    grad = eng.backward(s, saved, sl, sp).cpu()
    for n, v in eng.views(grad).items():
        gr = Pg[n].grad
        assert float((v.double() - gr).abs().max()) / (float(gr.abs().max()) + 1e-30) < 1e-2, n


@pytest.mark.gpu
@pytest.mark.parametrize("cell,nx,ny,nh,ns", [(_lib.CELL_TENSGRU2D, 5, 4, 20, 700), (_lib.CELL_TENSGRU2D, 6, 6, 48, 200),
                                              (_lib.CELL_VANILLA2D, 5, 4, 50, 300), (_lib.CELL_TENS2D, 4, 4, 24, 130)])
def test_tc_row_spreading_and_on_sm_paths_match_their_fallbacks(cuda, monkeypatch, cell, nx, ny, nh, ns):
    """The sweeps spread batches of fewer tiles than SMs over all SMs (tile_rows live rows per 128-row tile), the reverse
    sweep keeps the horizontal half of d h on the SM and the biases' gradients come out of the weight-gradient MMAs: each of
    these has a plain fallback (full tiles; read-modify-write in global memory; a row-sum kernel) selected by an
    environment switch read per call -- samples and log psi must be identical, gradients equal to rounding."""
    cplx = cell != _lib.CELL_VANILLA2D
    eng = Engine(cell, nx, ny, num_units=nh, complex_head=cplx, u1=cplx, seed=5, precision="tc")
    p = eng.params.cpu()
    eng.set_weights(p + 0.1 * (torch.rand(p.shape, generator=torch.Generator().manual_seed(6)) - 0.5))
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(7), dtype=torch.float32)
    seeds = torch.randn(ns, generator=torch.Generator().manual_seed(8), dtype=torch.float64)

    def run():
        saved = eng.new_saved(ns)
        s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
        g = eng.backward(s, saved, seeds, seeds if cplx else None)
        sv2 = eng.new_saved(ns)
        lp2, ph2 = eng.logpsi(s, saved=sv2)
        g2 = eng.backward(s, sv2, seeds, seeds if cplx else None)
        return s.cpu(), lp.cpu(), ph.cpu(), lp2.cpu(), g.cpu().double(), g2.cpu().double()

    ref = run()
    assert torch.equal(ref[1], ref[3])
    # (the row-sum fallback adds the biases' gradients in FP32, the MMAs round d(pre-activations) to TF32 first: 10 mantissa bits)
    for var, tol in (("RNNWF_NO_VTILES", 2e-5), ("RNNWF_NO_HCHAIN", 2e-5), ("RNNWF_NO_BIAS_MMA", 3e-3)):
        monkeypatch.setenv(var, "1")
        alt = run()
        monkeypatch.delenv(var)
        assert torch.equal(ref[0], alt[0]) and torch.equal(ref[1], alt[1]) and torch.equal(ref[2], alt[2]), var
        for a, b in ((ref[4], alt[4]), (ref[5], alt[5])):
            err = float((a - b).abs().max()) / float(b.abs().max())
            print(f"[parity] {var}: gradient differs from the default path by {err:.2e} of its largest entry (bound {tol:.0e})")
            assert err <= tol, var


@pytest.mark.gpu
@pytest.mark.parametrize("cell,nx,ny,nh,prec", [(_lib.CELL_TENSGRU2D, 6, 5, 40, "tc"), (_lib.CELL_TENSGRU2D, 5, 6, 36, "tc3"),
                                               (_lib.CELL_VANILLA2D, 7, 4, 50, "tc"), (_lib.CELL_TENS2D, 3, 8, 24, "tc")])
def test_tc_ring_slots_by_live_range_change_nothing(cuda, monkeypatch, cell, nx, ny, nh, prec):
    """The history ring's slots are assigned by live range (one per lattice column on the zigzag path) instead of
    step mod (look-back + 1): same arithmetic on the same values, so samples, log psi and E_loc (prefix reuse on and off)
    must be bit-identical (E_loc: to the order of its FP64 atomic sums) to the modulo slots that RNNWF_TC_DEBUG bit 64
    restores (read per launch)."""
    cplx = cell != _lib.CELL_VANILLA2D
    eng = Engine(cell, nx, ny, num_units=nh, complex_head=cplx, u1=cplx and (nx * ny) % 2 == 0, seed=15, precision=prec)
    p = eng.params.cpu()
    eng.set_weights(p + 0.1 * (torch.rand(p.shape, generator=torch.Generator().manual_seed(16)) - 0.5))
    ns = 333
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(17), dtype=torch.float32)
    ham, hp = (_lib.HAM_TFIM2D, [2.5] + [1.0] * (nx * ny)) if cell == _lib.CELL_VANILLA2D else (_lib.HAM_HEIS2D, [1.0])

    def run():
        saved = eng.new_saved(ns)
        s, lp, ph = eng.sample(ns, uniforms=u, saved=saved)
        lp2, ph2 = eng.logpsi(s)
        e = eng.local_energies(ham, hp, s, lp, ph, saved=saved)
        e2 = eng.local_energies(ham, hp, s, lp, ph, saved=None)
        return [t.cpu() for t in (s, lp, ph, lp2, ph2, e, e2)]

    ref = run()
    monkeypatch.setenv("RNNWF_TC_DEBUG", "64")
    alt = run()
    monkeypatch.delenv("RNNWF_TC_DEBUG")
    for a, b, name in zip(ref, alt, ("samples", "log P", "phase", "log P (evaluation)", "phase (evaluation)", "E_loc", "E_loc without reuse")):
        if name.startswith("E_loc"):      # (summed over the children with FP64 atomics: order-dependent in the last bits)
            assert float((a - b).abs().max()) <= 1e-12 * float(a.abs().max()), name
        else:
            assert torch.equal(a, b), name
