"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- the reference's surrogate costs (SURVEY 8a
rows G1-G3) in torch float64; their autograd gradient is the oracle for the backward kernels."""
import torch


def cost_real(logp, Eloc):
    """G1: 1DTFIM/TrainingRNN_1DTFIM.py:157, 2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:164."""
    Eloc = Eloc.detach()
    return (logp * Eloc).mean() - Eloc.mean() * logp.mean()


def cost_complex(logp, phase, Eloc, T=0.0):
    """G2: J1J2/TrainingRNN_J1J2.py:197; with the annealing term of 2DHeisenberg/Run_RNNWF_2DHeis.py:199.
    log psi = 0.5*logp + i*phase, Eloc complex."""
    logpsi = torch.complex(0.5 * logp, phase)
    Eloc = Eloc.detach()
    c = 2 * torch.real((torch.conj(logpsi) * Eloc).mean() - torch.conj(logpsi).mean() * Eloc.mean())
    if T != 0.0:
        re = torch.real(logpsi)
        c = c + 4 * T * ((re * re.detach()).mean() - re.mean() * re.detach().mean())
    return c


def cost_free_energy(logp, Eloc, T):
    """G3: RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-159."""
    Floc = (Eloc + T * logp).detach()
    return (logp * Floc).mean() - logp.mean() * Floc.mean()


def seeds_real(Eloc):
    """d cost / d logp_b for G1."""
    return (Eloc - Eloc.mean()) / Eloc.numel()


def seeds_complex(logp, Eloc, T=0.0):
    """(d cost/d logp_b, d cost/d phase_b) for G2."""
    ns = Eloc.numel()
    dE = Eloc - Eloc.mean()
    s_logp = torch.real(dE) / ns + T * (logp - logp.mean()) / ns
    s_phase = 2 * torch.imag(dE) / ns
    return s_logp, s_phase
