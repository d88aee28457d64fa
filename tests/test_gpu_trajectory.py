"""GPU, whole pipeline: a VMC training trajectory (sample -> connected configurations -> E_loc -> cost seeds -> backward ->
TF-Adam -> repack, step after step) against the same loop written with the float64 oracle (oracle sampling on the same
uniforms, oracle local energies, torch autograd of the reference's cost, TF-Adam formulas of SURVEY App. B).  The two are
chaotic systems fed with identical random numbers: they must agree sample for sample and to ~1e-5 in the energy over the
first dozens of steps; the comparison stops at the first step where a draw falls within 1e-6 of a CDF threshold (the FP32
and FP64 cumulative sums may then legitimately pick different symbols)."""
import math

import numpy as np
import pytest
import torch

from oracle import costs
from oracle import local_energies as ole
from rnnwavefunctions_b200 import _lib, vmc
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_run, oracle_logpsi_fn


def _oracle_adam(P, grads, state, t, lr, b1=0.9, b2=0.999, eps=1e-8):
    lr_t = lr * math.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    for k in P:
        g = grads[k]
        state["m"][k] = b1 * state["m"][k] + (1 - b1) * g
        state["v"][k] = b2 * state["v"][k] + (1 - b2) * g * g
        P[k] = (P[k].detach() - lr_t * state["m"][k] / (state["v"][k].sqrt() + eps)).requires_grad_(True)


CASES = {
    # C1-like: positive 1D GRU, TFIM chain (1DTFIM/TrainingRNN_1DTFIM.py:157-222)
    "tfim1d": dict(eng=dict(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=10), ham=_lib.HAM_TFIM1D, hp=[1.0] + [1.0] * 9, complex=False),
    # C2-like: complex 1D GRU with U(1), J1-J2 chain without Marshall sign (J1J2/TrainingRNN_J1J2.py:197-301)
    "j1j2": dict(eng=dict(cell=_lib.CELL_GRU1D, nx=10, ny=1, num_units=10, complex_head=True, u1=True), ham=_lib.HAM_J1J2,
                 hp=[0.0, 0.0] + [1.0] * 10 + [0.2] * 10 + [0.0] * 10, complex=True),
    # C4-like: tensorized complex 2D GRU with U(1), triangular Heisenberg, FP32 regime (2DHeisenberg/Run_RNNWF_2DHeis.py:199-349)
    "heis2d": dict(eng=dict(cell=_lib.CELL_TENSGRU2D, nx=4, ny=4, num_units=12, complex_head=True, u1=True), ham=_lib.HAM_HEIS2D,
                   hp=[1.0], complex=True),
}


def _oracle_eloc(name, eng, P, s):
    f = oracle_logpsi_fn(eng, P)
    if name == "tfim1d":
        return torch.from_numpy(np.asarray(ole.ising1d_local_energies(np.ones(9), 1.0, s, f))).to(torch.complex128)
    if name == "j1j2":
        e, _ = ole.j1j2_local_energies(np.ones(10), 0.2 * np.ones(10), np.zeros(10), s, f, periodic=False, Marshall_sign=False)
        return torch.from_numpy(np.asarray(e)).to(torch.complex128)
    e, _ = ole.heis2d_local_energies(1.0, s, f)
    return torch.from_numpy(np.asarray(e)).to(torch.complex128)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_training_trajectory_matches_the_oracle_loop(cuda, name):
    c = CASES[name]
    ns, steps, lr = 120, 30, 5e-3
    eng = Engine(seed=111, **c["eng"])
    P = oracle_params(eng, requires_grad=True)
    state = dict(m={k: torch.zeros_like(v) for k, v in P.items()}, v={k: torch.zeros_like(v) for k, v in P.items()})
    V = vmc.VMC(eng, c["ham"], c["hp"], ns, lr=lr, seed=1)
    g = torch.Generator().manual_seed(5)
    compared, worst = 0, 0.0
    for t in range(1, steps + 1):
        u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
        # ---- oracle step
        with torch.no_grad():
            r0 = oracle_run(eng, P, uniforms=u, numsamples=ns)
        if float(r0["margin"].min()) <= 1e-6:
            break                                   # a draw on a CDF threshold: the two loops may part ways here
        s = r0["samples"].numpy().astype(np.int64)
        E = _oracle_eloc(name, eng, P, s.reshape((ns,) + eng.cfg_shape))
        r = oracle_run(eng, P, samples=s)
        cost = costs.cost_complex(r["logp"], r["phase"], E) if c["complex"] else costs.cost_real(r["logp"], E.real)
        grads = dict(zip(P.keys(), torch.autograd.grad(cost, list(P.values()), allow_unused=True)))
        grads = {k: (torch.zeros_like(P[k]) if v is None else v) for k, v in grads.items()}
        # ---- GPU step on the same uniforms
        mean_e, var = V.step(uniforms=u.to(cuda))
        assert torch.equal(V.last["samples"].cpu().long().reshape(ns, -1), torch.from_numpy(s).reshape(ns, -1)), f"step {t}: samples differ"
        ref_mean = complex(E.mean())
        err = abs(complex(mean_e) - ref_mean) / abs(ref_mean)
        worst = max(worst, err)
        assert err < 2e-4, f"step {t}: <E> {complex(mean_e)} vs oracle {ref_mean}"
        _oracle_adam(P, grads, state, t, lr)
        compared += 1
    print(f"[parity] {name}: {compared} training steps compared, worst relative error of <E> {worst:.2e}")
    assert compared >= 10
    # the parameters after the compared steps (float32 on the device, float64 in the oracle loop)
    for k, v in eng.views().items():
        d = float((v.detach().double().cpu() - P[k].detach()).abs().max())
        assert d < 2e-3 * max(1.0, float(P[k].detach().abs().max())), f"{k} drifted by {d}"
