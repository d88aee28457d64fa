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
