"""GPU parity of the symmetry-projected wavefunctions: the group copies are
formed on the device through index-remap tables (rnnwf_symmetry_create / _apply), and amplitudes, local energies
(rnnwf_local_energies_sym) and the gradient are held to the oracle, whose group lists and epilogue are pinned to the
unmodified reference by tests/test_oracle_wf_golden.py.  FP32 regime, 1e-5."""
import numpy as np
import pytest
import torch

from oracle import local_energies as ole
from oracle import wavefunctions as owf
from oracle import costs
from rnnwavefunctions_b200 import _lib, Heisenberg2D, RydbergKagome, TFIM1D
from rnnwavefunctions_b200.compat import placeholder, Session
from rnnwavefunctions_b200.vmc import VMC
from helpers import oracle_params, oracle_run, assert_rel

HEIS_CASES = [("rotsym", "A", None), ("rotsym", "B", -1), ("rotreducedsym", "B", None), ("rotreducedsym", "A", 1),
              ("c2sym", "B", None), ("c2sym", "A", -1), ("c2vsym", "A2", None), ("c2vsym", "B1", 1),
              ("c2dsym", "A1", None), ("c2dsym", "B2", -1), ("c4vsym", "A1", None), ("c4vsym", "B1", -1)]


def _jitter(eng, seed, amp=0.15):
    g = torch.Generator().manual_seed(seed)
    p = eng.params.cpu()
    eng.set_weights(p + amp * (torch.rand(p.shape, generator=g) - 0.5))


def _heis_oracle(wf, P, cfg, group, ch, sp, grad=False, cond=None):
    lst, signs = owf.heis_symmetry_copies(torch.as_tensor(np.asarray(cfg)).long(), group, ch, sp)
    ctx = torch.enable_grad() if grad else torch.no_grad()
    with ctx:
        r = oracle_run(wf.engine, P, samples=torch.cat(lst))
        la = owf.log_amplitudes_from_symmetry_group(r["logp"].view(len(lst), -1), r["phase"].view(len(lst), -1), signs)
    if cond is not None:      # conditioning of arg z per configuration: (|Re z|, |z|), z = sum_g e^{i phi_g} / chi_g
        chs = torch.as_tensor(signs, dtype=torch.float64).reshape(-1, 1)
        ph_g = r["phase"].detach().view(len(lst), -1)
        zre, zim = (torch.cos(ph_g) / chs).sum(0), (torch.sin(ph_g) / chs).sum(0)
        cond.append(torch.stack([zre.abs(), torch.sqrt(zre ** 2 + zim ** 2)], dim=1).numpy())
    return la


def _method(wf, group, sp):
    if sp is None:
        return getattr(wf, f"log_amplitude_{group}")
    return getattr(wf, "log_amplitude_c2v_spinparity" if group == "c2vsym" else f"log_amplitude_{group}_spinparity")


@pytest.mark.gpu
@pytest.mark.parametrize("group,ch,sp", HEIS_CASES)
def test_heisenberg_groups(cuda, group, ch, sp):
    """Every group of 2DHeisenberg/MDComplexRNN.py:320-577: copies (bit-exact), amplitudes, E_loc and gradient."""
    L, nh, ns = 4, 10, 48
    wf = Heisenberg2D.RNNwavefunction(L, L, cell=Heisenberg2D.MDRNNcell, units=[nh], seed=5)
    _jitter(wf.engine, 6)
    eng = wf.engine
    # large phase-head weights: with the small phases of a fresh network every non-trivial character projects to ~0
    p = eng.params.clone()
    eng.views(p)["dense_phase.W"].mul_(8.0)
    eng.set_weights(p)
    s = wf.sample(ns, 2)
    P = oracle_params(eng)
    args = (ch,) if sp is None else (ch, sp)
    # copies through the device tables == the reference's list_samples
    grp = wf.symmetry_group(group, ch, sp)
    lst, signs = owf.heis_symmetry_copies(s.cpu(), group, ch, sp)
    assert torch.equal(eng.symmetry_apply(grp, s).cpu().long(), torch.cat(lst))
    assert grp.characters == [float(c) for c in signs]
    # amplitudes
    la = _method(wf, group, sp)(s, 2, *args).cpu().numpy()
    ref = _heis_oracle(wf, P, s.cpu(), group, ch, sp).numpy()
    assert_rel(la.real, ref.real, 1e-5, f"{group} {ch} {sp} Re log psi")
    # phase = arg z, z = sum_g e^{i phi_g} / chi_g with Re regularised: the network's phases are FP32 (abs error ~1e-6 per
    # copy), so arg z is known to ~G * 1e-6 / |z|; where Re z cancels exactly its sign, hence the phase, is arbitrary
    with torch.no_grad():
        r = oracle_run(eng, P, samples=torch.cat(lst))
    chs = torch.as_tensor(signs, dtype=torch.float64).reshape(-1, 1)
    ph_g = r["phase"].view(len(lst), -1)
    zre, zim = (torch.cos(ph_g) / chs).sum(0).numpy(), (torch.sin(ph_g) / chs).sum(0).numpy()
    zmag = np.sqrt(zre ** 2 + zim ** 2)
    nondeg = (np.abs(zre) > 1e-3) & (zmag > 0.3)
    dphi = np.abs(np.exp(1j * (la.imag - ref.imag)) - 1)                  # compared modulo 2 pi
    assert (dphi[np.abs(zre) > 1e-3] < (1e-5 + 4e-6 * len(lst) / zmag[np.abs(zre) > 1e-3])).all()
    # local energies through the reference-named surface: the handle of the symmetric method selects the group
    inputs = placeholder(dtype="int32", shape=[None, L, L])
    log_amps = _method(wf, group, sp)(inputs, 2, *args)
    E = Heisenberg2D.Get_Elocs(1.0, L, L, s.cpu().numpy(), inputs=inputs, log_amps=log_amps, sess=Session(), RNN_symmetry=group)
    cond = []

    def logpsi(cfg):
        z = _heis_oracle(wf, P, cfg, group, ch, sp, cond=cond)
        return (2 * z.real).numpy(), z.imag.numpy()
    Eref, counts = ole.heis2d_local_energies(1.0, s.cpu().numpy(), logpsi)
    # a sample is compared when it and all of its connected configurations have a well-conditioned phase
    cond = np.concatenate(cond)
    ends = np.cumsum(counts)
    keep = np.array([(cond[e - c:e, 0] > 1e-3).all() and (cond[e - c:e, 1] > 0.3).all() for c, e in zip(counts, ends)])
    assert keep.sum() >= 4, keep.sum()
    assert_rel(E[keep], Eref[keep], 2e-5, f"{group} {ch} {sp} E_loc ({int(keep.sum())} of {ns} samples well-conditioned)")
    # gradient of the complex cost through the projection, on the rows whose phase sum does not cancel (where it does,
    # d phase / d theta ~ 1 / 1e-4 amplifies the FP32 rounding of the network's phases: ill-conditioned upstream too)
    sg = s[torch.as_tensor(nondeg)]
    nsg = sg.shape[0]
    assert nsg >= 4
    v = VMC(eng, _lib.HAM_HEIS2D, [1.0], nsg, symmetry=grp, T=0.2)
    g = torch.Generator().manual_seed(3)
    Er = torch.complex(torch.randn(nsg, generator=g, dtype=torch.float64), torch.randn(nsg, generator=g, dtype=torch.float64))
    lp_s, ph_s, _, _, _ = eng.logpsi_sym(grp, sg)
    grad, _, _ = v.gradient(sg, lp_s, ph_s, Er.cuda())
    Pg = oracle_params(eng, requires_grad=True)
    z = _heis_oracle(wf, Pg, sg.cpu(), group, ch, sp, grad=True)
    costs.cost_complex(2 * z.real, z.imag, Er, T=0.2).backward()
    for n, t in eng.views(grad.cpu()).items():
        gr = Pg[n].grad
        assert float((t.double() - gr).abs().max()) <= 2e-4 * (float(gr.abs().max()) + 1e-30), n


@pytest.mark.gpu
@pytest.mark.parametrize("symmetry,Lx,Ly", [("c6sym", 3, 3), ("translationsym", 3, 2), ("c6sym", 2, 2), ("translationsym", 4, 4)])
def test_rydberg_symmetry(cuda, symmetry, Lx, Ly):
    """RydbergKagome_2DGRU/MDRNNWavefunction.py:242-291 (c6sym, translationsym): copies, log P, E_loc (a symmetric handle
    must give the symmetric local energies, Run_RNNWF_Rydberg.py:155,262) and the free-energy gradient."""
    nh, ns = 12, 20
    wf = RydbergKagome.MDRNNWavefunction(Lx, Ly, cell=RydbergKagome.MDRNNcell, units=[nh], seed=8)
    _jitter(wf.engine, 9, 0.1)
    eng = wf.engine
    s = wf.sample(ns, 2)
    P = oracle_params(eng)
    grp = wf.symmetry_group(symmetry)
    lst = owf.rydberg_symmetry_copies(s.cpu(), symmetry)
    assert torch.equal(eng.symmetry_apply(grp, s).cpu().long(), torch.cat(lst))

    def logp_sym(cfg, grad=False):
        c = owf.rydberg_symmetry_copies(torch.as_tensor(np.asarray(cfg)).long(), symmetry)
        with (torch.enable_grad() if grad else torch.no_grad()):
            lp = oracle_run(eng, P if not grad else Pg, samples=torch.cat(c))["logp"].view(len(c), -1)
            return owf.log_probability_symmetry(lp)
    lp = wf.log_probability_symmetry(s, 2, symmetry).cpu().numpy()
    assert_rel(lp, logp_sym(s.cpu()).numpy(), 1e-5, f"Rydberg {symmetry} log P")
    V, delta, Omega, Rc = 1.7 ** 6, 3.3, 1.0, 2.001
    nb, dist = RydbergKagome.neighbouring_atoms(RydbergKagome.generate_kagome(Lx, Ly), Rc, Lx, Ly)
    inputs = placeholder(dtype="int64", shape=[None, Lx, Ly, 3])
    handle = wf.log_probability_symmetry(inputs, 2, symmetry)
    E = RydbergKagome.Rydberg_local_energies(V, delta, Omega, nb, dist, s.cpu().numpy(), None, handle, inputs, None, Session(), Rcutoff=Rc)
    onb, odist = ole.neighbouring_atoms(ole.generate_kagome(Lx, Ly), Rc, Lx, Ly)
    Eref = ole.rydberg_local_energies(V, delta, Omega, onb, odist, s.cpu().numpy(), lambda c: (logp_sym(c).numpy(), 0.0))
    assert_rel(E, Eref, 1e-5, f"Rydberg {symmetry} E_loc")
    # the plain handle still gives the plain local energies (and they differ)
    Eplain = RydbergKagome.Rydberg_local_energies(V, delta, Omega, nb, dist, s.cpu().numpy(), None, wf, inputs, None, Session(), Rcutoff=Rc)
    assert np.abs(Eplain - E).max() > 1e-4
    # free-energy cost (G3) through the projection
    T = 0.4
    v = VMC(eng, _lib.HAM_RYDBERG, [V, delta, Omega, Rc], ns, symmetry=grp, T=T, free_energy=True)
    lp_s = eng.logpsi_sym(grp, s)[0]
    Et = torch.as_tensor(Eref)
    grad, _, _ = v.gradient(s, lp_s, None, torch.complex(Et, torch.zeros_like(Et)).cuda())
    Pg = oracle_params(eng, requires_grad=True)
    costs.cost_free_energy(logp_sym(s.cpu(), grad=True), Et, T).backward()
    for n, t in eng.views(grad.cpu()).items():
        gr = Pg[n].grad
        assert float((t.double() - gr).abs().max()) <= 2e-4 * (float(gr.abs().max()) + 1e-30), n


@pytest.mark.gpu
def test_parity_symmetric_chain(cuda):
    """1DTFIM/RNNwavefunction_paritysym.py:122-145 end to end: log P, Ising_local_energies, gradient."""
    N, ns = 12, 40
    wf = TFIM1D.RNNwavefunction_paritysym(N, units=[14, 14], seed=3)
    _jitter(wf.engine, 4)
    eng = wf.engine
    s = wf.sample(ns, 2)
    P = oracle_params(eng)
    units = [14, 14]
    lp = wf.log_probability(s, 2).cpu().numpy()
    assert_rel(lp, owf.rnn1d_paritysym_logp(P, N, units, s.cpu()).detach().numpy(), 1e-5, "parity-sym log P")
    Jz = np.linspace(0.8, 1.2, N)
    inputs = placeholder(dtype="int32", shape=[None, N])
    E = TFIM1D.Ising_local_energies(Jz, 0.9, s.cpu().numpy(), None, wf.log_probability(inputs, 2), inputs, None, Session())
    Eref = ole.ising1d_local_energies(Jz, 0.9, s.cpu().numpy(),
                                      lambda c: (owf.rnn1d_paritysym_logp(P, N, units, torch.as_tensor(c).long()).detach().numpy(), 0.0))
    assert_rel(E, Eref, 1e-5, "parity-sym E_loc")
    v = VMC(eng, _lib.HAM_TFIM1D, np.concatenate([[0.9], Jz[:N - 1]]), ns, symmetry=wf.symmetry_group())
    Et = torch.as_tensor(Eref)
    grad, _, _ = v.gradient(s, torch.as_tensor(lp).cuda(), None, torch.complex(Et, torch.zeros_like(Et)).cuda())
    Pg = oracle_params(eng, requires_grad=True)
    costs.cost_real(owf.rnn1d_paritysym_logp(Pg, N, units, s.cpu()), Et).backward()
    for n, t in eng.views(grad.cpu()).items():
        gr = Pg[n].grad
        assert float((t.double() - gr).abs().max()) <= 2e-4 * (float(gr.abs().max()) + 1e-30), n


@pytest.mark.gpu
def test_symmetric_vmc_step_runs(cuda):
    """One full symmetric VMC step (sample -> E_loc_sym -> gradient -> Adam) on the device."""
    wf = Heisenberg2D.RNNwavefunction(4, 4, cell=Heisenberg2D.MDRNNcell, units=[8], seed=2)
    v = VMC(wf.engine, _lib.HAM_HEIS2D, [1.0], 64, symmetry=wf.symmetry_group("c4vsym", "A1", 1), lr=1e-3)
    m0, var0 = v.step()
    m1, var1 = v.step()
    assert np.isfinite(float(m0.real)) and np.isfinite(float(m1.real)) and float(var1) >= 0
