"""SURVEY 8f "next" rows: schedules / gradient transforms (N1), TF-named state dict (N3), final-estimate pass (N4)."""
import math

import numpy as np
import pytest
import torch

from rnnwavefunctions_b200 import checkpoint, schedules


def test_lr_schedules_match_reference_formulas():
    assert schedules.lr_tfim2d_1drnn(0, 5e-3) == pytest.approx(5e-3)
    assert schedules.lr_tfim2d_1drnn(100, 5e-3) == pytest.approx(1.0 / (200.0 + 10.0))
    assert schedules.lr_tfim2d_2drnn(5000, 1e-3) == pytest.approx(5e-4)
    kw = dict(lr=5e-4, lrthreshold=1e-4, lrdecaytime=100.0, lrthreshold_conv=1e-4, lrdecaytime_conv=50.0,
              Nwarmup=10, Nannealing=4, Ntrain=5)   # annealing ends after step index 29
    assert schedules.lr_heisenberg(0, **kw) == pytest.approx(5e-4)
    assert schedules.lr_heisenberg(29, **kw) == pytest.approx(max(1e-4, 5e-4 / 1.29))
    assert schedules.lr_heisenberg(10 ** 6, lr=5e-4, lrthreshold=1e-4, lrdecaytime=100.0, lrthreshold_conv=1e-4,
                                   lrdecaytime_conv=50.0, Nwarmup=10 ** 7, Nannealing=0, Ntrain=5) == pytest.approx(1e-4)
    assert schedules.lr_heisenberg(30, **kw) == pytest.approx(1e-4)                  # first convergence step
    assert schedules.lr_heisenberg(80, **kw) == pytest.approx(1e-4 / 2.0)


def test_temperature_schedules():
    T = [schedules.temperature_heisenberg(it, 2.0, 10, 4, 5) for it in range(40)]
    assert T[:10] == [2.0] * 10                      # warm-up: it + 1 <= 10
    assert T[10] == 2.0 and T[13] == 2.0             # (it + 1 - 10) // 5 == 0
    assert T[14] == pytest.approx(2.0 * (1 - 1 / 4))
    assert T[29] == pytest.approx(0.0)               # (30 - 10) // 5 = 4 -> T0 (1 - 4/4)
    assert T[30] == 0.0
    assert schedules.temperature_heisenberg(3, 0.0, 10, 4, 5) == 0.0
    ryd = schedules.RydbergTemperature(1.0, numwarmup=10, Nannealing=2, Neq=5)
    seq = [ryd(it) for it in range(1, 25)]
    assert seq[:9] == [1.0] * 9                      # it = 1..9
    assert seq[9] == pytest.approx(1.0)              # it = 10: 1 - 0/10
    assert seq[13] == pytest.approx(1.0)             # it = 14 keeps the value set at it = 10
    assert seq[14] == pytest.approx(0.5)             # it = 15
    assert seq[19] == pytest.approx(0.5)             # it = 20 == end of annealing: neither branch fires
    assert seq[20] == 0.0                            # it = 21 > end
    assert schedules.lr_rydberg(19, 1e-3, 1e-4, 10, 2, 5) == 1e-3
    assert schedules.lr_rydberg(20, 1e-3, 1e-4, 10, 2, 5) == 1e-4


def test_gradient_transforms():
    g = torch.tensor([-3.0, -0.5, 0.0, 0.2, 7.0])
    assert torch.equal(schedules.clip_by_value_(g.clone(), 1.0), torch.tensor([-1.0, -0.5, 0.0, 0.2, 1.0]))
    gen = torch.Generator().manual_seed(5)
    n = schedules.normalized_gradient(g, gen)
    assert torch.equal(torch.sign(n), torch.sign(g)) and float(n.abs().max()) < 1.0
    gen2 = torch.Generator().manual_seed(5)
    assert torch.equal(n, schedules.normalized_gradient(g, gen2))   # same seed on every rank -> same update


def test_tf_variable_names():
    f = checkpoint.tf_variable_name
    assert f("gru0.Wg", False) == "RNNwavefunction/multi_rnn_cell/cell_0/cudnn_compatible_gru_cell/gates/kernel"
    assert f("gru2.bch", False) == "RNNwavefunction/multi_rnn_cell/cell_2/cudnn_compatible_gru_cell/candidate/hidden_projection/bias"
    assert f("gru1.Wci", True, scope="S") == "S/multi_rnn_cell/cell_1/cudnn_compatible_gru_cell/candidate/input_projection/kernel"
    assert f("dense.W", False) == "RNNwavefunction/wf_dense/kernel"
    assert f("dense.b", True) == "RNNwavefunction/wf_dense_ampl/bias"
    assert f("dense_phase.W", True) == "RNNwavefunction/wf_dense_phase/kernel"
    assert f("Wmerge", True) == "RNNwavefunction/Wmerge_rnn_0"
    assert f("Uh", False) == "RNNwavefunction/Uh_rnn_0"
    # the complex 2D Heisenberg model calls its amplitude head 'wf_dense' (2DHeisenberg/MDComplexRNN.py:63), only the 1D
    # cRNN has 'wf_dense_ampl' (J1J2/ComplexRNNwavefunction.py:42)
    assert f("dense.W", True, gru1d=False) == "RNNwavefunction/wf_dense/kernel"


def test_variable_names_match_the_executed_reference():
    """Variable-name set per reference package: the names the reference's own code registers when it is executed under
    tests/golden/tf_shim.py (recorded in the wf_*.npz fixtures: get_variable / Dense names are literals of the reference
    source) against what checkpoint.tf_variable_name produces for the engine's tensors of that model."""
    import os
    G = os.path.join(os.path.dirname(__file__), "golden")
    gru = lambda l: [f"gru{l}.{t}" for t in ("Wg", "bg", "Wci", "Wch", "bci", "bch")]
    cases = {
        "wf_gru1d": (gru(0) + ["dense.W", "dense.b"], False, True),
        "wf_gru1d_3layer": (gru(0) + gru(1) + gru(2) + ["dense.W", "dense.b"], False, True),
        "wf_cgru1d_u1": (gru(0) + ["dense.W", "dense.b", "dense_phase.W", "dense_phase.b"], True, True),
        "wf_vanilla2d": (["Wh", "Uh", "Wv", "Uv", "b", "dense.W", "dense.b"], False, False),
        "wf_tens2d_u1": (["W", "b", "dense.W", "dense.b", "dense_phase.W", "dense_phase.b"], True, False),
        "wf_tensgru2d_u1": (["W", "b", "Wg", "bg", "Wmerge", "dense.W", "dense.b", "dense_phase.W", "dense_phase.b"], True, False),
        "wf_rydberg_3x3": (["W", "b", "Wg", "bg", "Wmerge", "dense.W", "dense.b"], False, False),
    }
    for fixture, (names, cplx, gru1d) in cases.items():
        z = np.load(os.path.join(G, fixture + ".npz"))
        ref = {k[5:].replace("__", "/") for k in z.files if k.startswith("var__")}
        ours = {checkpoint.tf_variable_name(n, cplx, gru1d=gru1d) for n in names}
        assert ours == ref, (fixture, ours ^ ref)


@pytest.mark.gpu
def test_state_dict_roundtrip_and_warm_start(cuda, tmp_path):
    from rnnwavefunctions_b200 import _lib
    from rnnwavefunctions_b200.engine import Engine
    a = Engine(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=12, complex_head=True, u1=True, seed=3)
    a.init_params(seed=9)
    sd = checkpoint.state_dict(a)
    assert sd["RNNwavefunction/W_rnn_0"].shape == (12, 24, 4) and sd["RNNwavefunction/wf_dense_phase/kernel"].shape == (12, 2)
    # a larger lattice takes the same variables (Run_RNNWF_2DHeis.py:119-125 warm start)
    b = Engine(cell=_lib.CELL_TENSGRU2D, nx=6, ny=6, num_units=12, complex_head=True, u1=True, seed=4)
    b.init_params(seed=1)
    checkpoint.load_state_dict(b, {k + ":0": v for k, v in sd.items()})
    assert torch.equal(a.params.cpu(), b.params.cpu())
    path = str(tmp_path / "ckpt.npz")
    checkpoint.save_npz(a, path, meanEnergy=np.arange(3.0))
    c = Engine(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=12, complex_head=True, u1=True, seed=5)
    c.init_params(seed=2)
    extra = checkpoint.load_npz(c, path)
    assert torch.equal(a.params.cpu(), c.params.cpu()) and np.array_equal(extra["meanEnergy"], np.arange(3.0))
    s = torch.randint(0, 2, (8, 4, 4), dtype=torch.int8)
    s[:, :2] = 1 - s[:, 2:]                          # any configurations: same weights -> same amplitudes
    la, _ = a.logpsi(s)
    lc, _ = c.logpsi(s)
    assert torch.equal(la, lc)
    with pytest.raises(KeyError):
        checkpoint.load_state_dict(c, {"RNNwavefunction/W_rnn_0": sd["RNNwavefunction/W_rnn_0"]})


@pytest.mark.gpu
def test_final_estimate_pass(cuda):
    from rnnwavefunctions_b200 import _lib, vmc
    from rnnwavefunctions_b200.engine import Engine
    eng = Engine(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=12, complex_head=True, u1=True, seed=3)
    eng.init_params(seed=9)
    out = vmc.estimate_energy(eng, _lib.HAM_HEIS2D, [1.0], numsamples_estimation=600, numbatches=200, nsites=16,
                              seed=7, return_samples=True)
    e = out["eloc"].cpu().numpy()
    assert out["n"] == 600 and out["samples"].shape == (600, 4, 4)
    assert out["mean"] == pytest.approx(np.mean(e).real / 16, rel=1e-12)
    assert out["err"] == pytest.approx(math.sqrt(np.var(e) / 600) / 16, rel=1e-9)
    # batches draw disjoint Philox rows: the same 600 samples as one call
    s_all, lp, ph = eng.sample(600, seed=7)
    assert torch.equal(s_all, out["samples"])
    dens, dens_err, avg, avg_err = vmc.site_densities(out["samples"])
    x = out["samples"].cpu().numpy()
    np.testing.assert_allclose(dens.cpu().numpy(), x.mean(0), rtol=1e-12)
    assert avg == pytest.approx(x.mean()) and avg_err == pytest.approx(np.std(x.mean(axis=(1, 2))) / math.sqrt(600))


@pytest.mark.gpu
def test_fused_adam_step_matches_tf_adam_semantics(cuda):
    """rnnwf_adam_step (one fused kernel) against the TF-Adam formulas of SURVEY App. B written out in float64,
    several steps in a row, with and without clip-by-value; called through the C-ABI by vmc.TFAdam."""
    import math
    from rnnwavefunctions_b200 import vmc
    n = 180604
    g = torch.Generator().manual_seed(5)
    for clip in (0.0, 0.3):
        p0 = torch.randn(n, generator=g)
        p = p0.clone().to(cuda)
        opt = vmc.TFAdam(n, cuda)
        m = torch.zeros(n, dtype=torch.float64)
        v = torch.zeros(n, dtype=torch.float64)
        ref = p0.double()
        for t in range(1, 6):
            grad = torch.randn(n, generator=g)
            lr = 1e-2 / t
            opt.step(p, grad.to(cuda), lr, clip=clip)
            gd = grad.double().clamp(-clip, clip) if clip else grad.double()
            m = 0.9 * m + 0.1 * gd
            v = 0.999 * v + 0.001 * gd * gd
            ref = ref - lr * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * m / (v.sqrt() + 1e-8)
            torch.testing.assert_close(p.cpu().double(), ref, rtol=2e-6, atol=2e-6)


@pytest.mark.gpu
def test_sign_sgd_step_is_the_normalized_gradients_update(cuda):
    """rnnwf_sign_sgd_step against theta - lr * sign(g) * u (2DHeisenberg/Run_RNNWF_2DHeis.py:206-213) with the uniforms of the
    library's counter-based generator reproduced on the host (helpers.philox_uniforms); entries with g = 0 stay untouched."""
    from rnnwavefunctions_b200 import schedules
    from helpers import philox_uniforms
    n = 38502
    g = torch.Generator().manual_seed(7)
    p0 = torch.randn(n, generator=g)
    grad = torch.randn(n, generator=g)
    grad[::17] = 0.0
    p = p0.clone().to(cuda)
    for t in (0, 1, 5):
        before = p.cpu().clone()
        schedules.sign_sgd_step_(p, grad.to(cuda), 3e-3, seed=111, t=t)
        u = torch.from_numpy(philox_uniforms(111, n, t + 1)[:, t].copy())
        ref = before - (torch.tensor(3e-3, dtype=torch.float32) * torch.sign(grad)) * u
        torch.testing.assert_close(p.cpu(), ref, rtol=0, atol=5e-7)          # (the kernel may fuse the multiply-subtract)
        assert torch.equal(p.cpu()[::17], before[::17])                      # sign(0) = 0
        assert float(u.min()) >= 0.0 and float(u.max()) < 1.0 and abs(float(u.mean()) - 0.5) < 0.01


@pytest.mark.gpu
@pytest.mark.parametrize("G,with_phase", [(4, True), (16, True), (6, False), (1, True)])
def test_symmetry_epilogue_and_seeds_match_the_reference_formulas(cuda, G, with_phase):
    """rnnwf_symmetry_epilogue / rnnwf_symmetry_seeds against 2DHeisenberg/MDComplexRNN.py:23-25,290-316 written in torch
    float64 (regularised real part with the sign under stop-gradient) and its autograd."""
    from rnnwavefunctions_b200 import _lib
    from rnnwavefunctions_b200.engine import Engine
    eng = Engine(_lib.CELL_TENSGRU2D, 2, 2, num_units=4, complex_head=True, u1=True, seed=1)
    g = torch.Generator().manual_seed(G)
    rows = 1037
    lp = (-30 * torch.rand(G, rows, generator=g, dtype=torch.float64)).requires_grad_(True)
    ph = (6 * torch.rand(G, rows, generator=g, dtype=torch.float64) - 3).requires_grad_(True)
    chi = [1.0 if (k % 3) else -1.0 for k in range(G)]
    out_lp, out_ph = eng.symmetrize(lp.detach(), ph.detach() if with_phase else None, chi)
    ref_lp = torch.log(torch.exp(lp).sum(0) / G)
    torch.testing.assert_close(out_lp.cpu(), ref_lp.detach(), rtol=1e-12, atol=1e-12)
    s_lp = torch.randn(rows, generator=g, dtype=torch.float64)
    s_ph = torch.randn(rows, generator=g, dtype=torch.float64)
    cost = (ref_lp * s_lp).sum()
    if with_phase:
        c = torch.tensor(chi, dtype=torch.float64).view(G, 1)
        re, im = (torch.cos(ph) / c).sum(0), (torch.sin(ph) / c).sum(0)
        sign = torch.where(re >= 0, torch.ones_like(re), -torch.ones_like(re)).detach()
        ref_ph = torch.atan2(im, sign * torch.sqrt(re * re + 1e-8))
        torch.testing.assert_close(out_ph.cpu(), ref_ph.detach(), rtol=1e-11, atol=1e-11)
        cost = cost + (ref_ph * s_ph).sum()
    cost.backward()
    seed_lp, seed_ph = eng.symmetry_seeds(lp.detach(), ph.detach() if with_phase else None, chi, s_lp, s_ph if with_phase else None)
    torch.testing.assert_close(seed_lp.cpu(), lp.grad, rtol=1e-10, atol=1e-12)
    if with_phase:
        torch.testing.assert_close(seed_ph.cpu(), ph.grad, rtol=1e-8, atol=1e-10)
