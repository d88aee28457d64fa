"""GPU: the reference-named Python surface (drop-in modules) end to end, against the oracle."""
import numpy as np
import pytest
import torch

from oracle import local_energies as ole
from oracle import wavefunctions as owf
from rnnwavefunctions_b200 import TFIM1D, J1J2, TFIM2D, Heisenberg2D, RydbergKagome
from rnnwavefunctions_b200.compat import placeholder, Session
from helpers import oracle_params, oracle_logpsi_fn, oracle_run


@pytest.mark.gpu
def test_tfim1d_reference_call_sequence(cuda):
    """The call sequence of 1DTFIM/TrainingRNN_1DTFIM.py:186-205 with the reference's argument order."""
    N, ns = 12, 64
    wf = TFIM1D.RNNwavefunction(N, units=[16], seed=111)
    samples_ = wf.sample(numsamples=ns, inputdim=2)
    assert samples_.dtype == torch.int64 and tuple(samples_.shape) == (ns, N)
    inputs = placeholder(dtype="int32", shape=[None, N])
    log_probs_tensor = wf.log_probability(inputs, inputdim=2)
    sess = Session()
    queue_samples = np.zeros((N + 1, ns, N), dtype=np.int32)
    log_probs = np.zeros((N + 1) * ns, dtype=np.float64)
    samples = samples_.cpu().numpy()
    Jz = np.ones(N)
    E = TFIM1D.Ising_local_energies(Jz, 1.0, samples, queue_samples, log_probs_tensor, inputs, log_probs, sess)
    assert E.dtype == np.float64 and E.shape == (ns,)
    lp = sess.run(log_probs_tensor, feed_dict={inputs: samples})
    P = oracle_params(wf.engine)
    ref = ole.ising1d_local_energies(Jz, 1.0, samples, oracle_logpsi_fn(wf.engine, P))
    np.testing.assert_allclose(E, ref, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(lp, oracle_run(wf.engine, P, samples=samples)["logp"].detach().numpy(), rtol=1e-5)
    # parity-symmetrised ansatz
    wfp = TFIM1D.RNNwavefunction_paritysym(N, units=[16], seed=111)
    lps = wfp.log_probability(samples_, 2).cpu().numpy()
    Pp = oracle_params(wfp.engine)
    want = owf.rnn1d_paritysym_logp(Pp, N, [16], torch.as_tensor(samples)).detach().numpy()
    np.testing.assert_allclose(lps, want, rtol=1e-5)


@pytest.mark.gpu
def test_j1j2_and_heisenberg_surface(cuda):
    N, ns = 10, 50
    wf = J1J2.RNNwavefunction(N, units=[12], seed=111)
    s = wf.sample(ns, 2)
    la = wf.log_amplitude(s, 2)
    assert la.dtype == torch.complex64
    E = J1J2.J1J2_local_energies(np.ones(N), 0.2 * np.ones(N), np.zeros(N), s, wf)
    ref, _ = ole.j1j2_local_energies(np.ones(N), 0.2 * np.ones(N), np.zeros(N), s.cpu().numpy(), oracle_logpsi_fn(wf.engine, oracle_params(wf.engine)))
    np.testing.assert_allclose(E, ref, rtol=2e-5, atol=2e-5)
    sig = np.zeros((2 * N, N), dtype=np.int32)
    me = np.zeros(2 * N)
    num = J1J2.J1J2MatrixElements(np.ones(N), 0.2 * np.ones(N), np.zeros(N), s[0].cpu().numpy(), sig, me)
    c, h = ole.j1j2_connected(np.ones(N), 0.2 * np.ones(N), np.zeros(N), s[0].cpu().numpy())
    assert num == len(h) and (sig[:num] == c).all() and np.allclose(me[:num], h)

    L = 4
    wf2 = Heisenberg2D.RNNwavefunction(L, L, cell=Heisenberg2D.MDRNNcell, units=[14], seed=111, mag_fixed=True, magnetization=0)
    s2, e2 = Heisenberg2D.Get_Samples_and_Elocs(1.0, L, L, np.ones((40, L, L), dtype=np.int32), log_amps=wf2)
    assert s2.dtype == np.int64 and e2.dtype == np.complex128
    P2 = oracle_params(wf2.engine)
    ref2, _ = ole.heis2d_local_energies(1.0, s2, oracle_logpsi_fn(wf2.engine, P2))
    np.testing.assert_allclose(e2, ref2, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(Heisenberg2D.Get_Elocs(1.0, L, L, s2, log_amps=wf2), ref2, rtol=1e-5, atol=1e-5)
    # symmetry-projected amplitude (c4v, A1, spin parity +1) against the oracle's epilogue
    st = torch.as_tensor(s2)
    got = wf2.log_amplitude_c4vsym_spinparity(st, 2, "A1", 1).cpu().numpy()
    ops = [lambda x: x, lambda x: torch.rot90(x, -1, (1, 2)), lambda x: torch.rot90(x, -2, (1, 2)), lambda x: torch.rot90(x, -3, (1, 2)),
           lambda x: torch.flip(x, [1]), lambda x: torch.flip(x, [2]), lambda x: x.transpose(1, 2),
           lambda x: torch.rot90(x, -2, (1, 2)).transpose(1, 2)]
    lst = [o(st) for o in ops] + [o(1 - st) for o in ops]
    lps, phs = [], []
    for c in lst:
        r = oracle_run(wf2.engine, P2, samples=c.numpy())
        lps.append(r["logp"].detach())
        phs.append(r["phase"].detach())
    want = owf.log_amplitudes_from_symmetry_group(torch.stack(lps), torch.stack(phs), [1] * 16).numpy()
    np.testing.assert_allclose(got.real, want.real, rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(np.exp(1j * got.imag), np.exp(1j * want.imag), atol=1e-4)


@pytest.mark.gpu
def test_tfim2d_and_rydberg_surface(cuda):
    wf = TFIM2D.RNNwavefunction(4, 3, cell=TFIM2D.MDRNNcell, units=[10], seed=111)
    s = wf.sample(30, 2)
    E = TFIM2D.Ising2D_local_energies(np.ones((4, 3)), 3.0, 4, 3, s, None, wf)
    ref = ole.ising2d_local_energies(np.ones((4, 3)), 3.0, s.cpu().numpy(), oracle_logpsi_fn(wf.engine, oracle_params(wf.engine)))
    np.testing.assert_allclose(E, ref, rtol=1e-5, atol=1e-5)

    Lx = Ly = 3
    wr = RydbergKagome.MDRNNWavefunction(Lx, Ly, cell=RydbergKagome.MDRNNcell, units=[12], seed=111)
    sr = wr.sample(25, 2)
    assert tuple(sr.shape) == (25, Lx, Ly, 3)
    pos = RydbergKagome.generate_kagome(Lx, Ly)
    nb, dist = RydbergKagome.neighbouring_atoms(pos, 2.001, Lx, Ly)
    E = RydbergKagome.Rydberg_local_energies(1.7 ** 6, 3.3, 1.0, nb, dist, sr, None, wr)
    P = oracle_params(wr.engine)
    ref = ole.rydberg_local_energies(1.7 ** 6, 3.3, 1.0, *ole.neighbouring_atoms(ole.generate_kagome(Lx, Ly), 2.001, Lx, Ly),
                                     sr.cpu().numpy(), oracle_logpsi_fn(wr.engine, P))
    np.testing.assert_allclose(E, ref, rtol=1e-5, atol=1e-5)
    # C6-symmetrised probability against the oracle
    got = wr.log_probability_symmetry(sr, 2, "c6sym").cpu().numpy()
    lst = [sr.cpu()]
    for _ in range(5):
        lst.append(owf.rotate60(lst[-1]))
    lps = torch.stack([oracle_run(wr.engine, P, samples=c.numpy())["logp"].detach() for c in lst])
    np.testing.assert_allclose(got, (torch.logsumexp(lps, 0) - np.log(6)).numpy(), rtol=1e-5)
    assert torch.equal(RydbergKagome.intarray_to_binary(RydbergKagome.binary_to_intarray(sr)), sr)


@pytest.mark.gpu
def test_short_training_run_lowers_the_energy(cuda):
    """run_1DTFIM for a few dozen steps on N=10 moves towards the ED energy -12.3815 (Tutorial notebook)."""
    import io, contextlib
    with contextlib.redirect_stdout(io.StringIO()):
        meanE, varE = TFIM1D.run_1DTFIM(numsteps=150, systemsize=10, num_units=10, Bx=1, num_layers=1, numsamples=200,
                                        learningrate=5e-3, seed=111)
    assert np.mean(meanE[:5]) > np.mean(meanE[-10:]) + 0.5
    assert -12.5 < np.mean(meanE[-10:]) < -11.0


@pytest.mark.gpu
@pytest.mark.parametrize("layers", [1, 2])
def test_2dtfim_1drnn_twin(cuda, layers):
    """2DTFIM_1DRNN/: the GRU chain raster-scanned over an Nx x Ny lattice with the 2D Ising bonds
    (RNNwavefunction.py:40-130, Training1DRNN_2DTFIM.py:13-81), reference call sequence, against the oracle."""
    from rnnwavefunctions_b200 import TFIM2D_1DRNN
    from helpers import assert_rel
    Nx, Ny, ns, nh = 4, 3, 60, 18
    wf = TFIM2D_1DRNN.RNNwavefunction(Nx, Ny, units=[nh] * layers, seed=5)
    g = torch.Generator().manual_seed(6)
    p = wf.engine.params.cpu()
    wf.engine.set_weights(p + 0.2 * (torch.rand(p.shape, generator=g) - 0.5))
    samples_ = wf.sample(ns, 2)
    assert tuple(samples_.shape) == (ns, Nx * Ny)
    inputs = placeholder(dtype="int32", shape=[None, Nx * Ny])
    lpt = wf.log_probability(inputs, 2)
    Jz = np.random.RandomState(1).uniform(0.5, 1.5, (Nx, Ny))
    samples = samples_.cpu().numpy()
    E = TFIM2D_1DRNN.Ising2D_local_energies(Jz, 2.0, Nx, Ny, samples, None, lpt, inputs, None, Session())
    P = oracle_params(wf.engine)
    f = oracle_logpsi_fn(wf.engine, P)
    ref = ole.ising2d_local_energies(Jz, 2.0, samples.reshape(ns, Nx, Ny), lambda c: f(c.reshape(c.shape[0], -1)))
    assert_rel(E, ref, 1e-5, f"2DTFIM_1DRNN E_loc, {layers} layer(s)")
    lp = Session().run(lpt, feed_dict={inputs: samples})
    assert_rel(lp, oracle_run(wf.engine, P, samples=samples)["logp"].detach().numpy(), 1e-5, "2DTFIM_1DRNN log P")


@pytest.mark.gpu
def test_training_twins_feed_the_reference_learning_rates(cuda, monkeypatch):
    """The learning rate each training twin hands to the optimiser step, iteration by iteration, is the reference script's:
    1DTFIM/TrainingRNN_1DTFIM.py:222 and J1J2/TrainingRNN_J1J2.py:306 constant, 2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:228
    lr (1 + it/5000)^-1, 2DTFIM_1DRNN/Training1DRNN_2DTFIM.py:229 1 / (1/lr + it/10)."""
    import io, contextlib
    from rnnwavefunctions_b200 import TFIM2D, TFIM2D_1DRNN, J1J2, vmc
    seen = []
    real_step = vmc.VMC.step

    def spy(self, uniforms=None, lr=None, update=True):
        seen.append(self.lr if lr is None else lr)
        return real_step(self, uniforms=uniforms, lr=lr, update=update)

    monkeypatch.setattr(vmc.VMC, "step", spy)
    its = np.arange(4)
    for fn, kw, want in (
            (TFIM1D.run_1DTFIM, dict(systemsize=6, num_units=6, numsamples=20, learningrate=5e-3), 5e-3 + 0 * its),
            (J1J2.run_J1J2, dict(systemsize=6, J2_=0.2, num_units=6, numsamples=20, learningrate=2.5e-4), 2.5e-4 + 0 * its),
            (TFIM2D.run_2DTFIM, dict(systemsize_x=3, systemsize_y=3, num_units=6, numsamples=20, learningrate=5e-3), 5e-3 / (1 + its / 5000)),
            (TFIM2D_1DRNN.run_2DTFIM, dict(systemsize_x=3, systemsize_y=3, num_units=6, numsamples=20, learningrate=1e-3), 1 / (1 / 1e-3 + its / 10))):
        seen.clear()
        with contextlib.redirect_stdout(io.StringIO()):
            fn(numsteps=3, **kw)
        np.testing.assert_allclose(np.array(seen, dtype=np.float64), want, rtol=1e-12)
