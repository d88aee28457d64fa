"""Drop-in for the reference package ``J1J2/``.
Reference: J1J2/ComplexRNNwavefunction.py, TrainingRNN_J1J2.py, run_j1j2.py."""
import numpy as np
import torch

from . import _lib
from .compat import Handle, Placeholder, Session, find_wf, to_numpy
from .engine import Engine
from .vmc import VMC


class RNNwavefunction:
    """J1J2/ComplexRNNwavefunction.py:15-43: GRU stack, sqrt-softmax amplitude head, pi*softsign phase head,
    zero-magnetisation projection."""

    def __init__(self, systemsize, cell=None, units=[10, 10], scope='RNNwavefunction', seed=111):
        if len(set(units)) != 1:
            raise ValueError("all layers must have the same number of units")
        self.N, self.scope, self.units, self.seed = systemsize, scope, list(units), seed
        self.engine = Engine(_lib.CELL_GRU1D, systemsize, 1, num_units=units[0], num_layers=len(units),
                             complex_head=True, u1=True, seed=seed)
        self._step = 0

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """J1J2/ComplexRNNwavefunction.py:45-103 -> int64 [numsamples, N], every row at zero magnetisation."""
        assert inputdim == 2
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, ph = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        return self.samples

    def log_amplitude(self, samples, inputdim):
        """J1J2/ComplexRNNwavefunction.py:105-169 -> complex64 [B]  (log psi = sum_n log(a_n e^{i phi_n}))."""
        assert inputdim == 2
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.log_amplitude(s, inputdim), samples)
        lp, ph = self.engine.logpsi(torch.as_tensor(samples))
        self.log_amplitudes = torch.complex(0.5 * lp, ph).to(torch.complex64)
        return self.log_amplitudes


def J1J2MatrixElements(J1, J2, Bz, sigmap, sigmaH, matrixelements, periodic=False, Marshall_sign=False):
    """J1J2/TrainingRNN_J1J2.py:12-93: connected configurations of ONE configuration into the caller's arrays;
    returns their number.  Compatibility helper (host arrays in, host arrays out); the training path uses
    :func:`J1J2_local_energies`, which enumerates on the device."""
    s = to_numpy(sigmap).astype(np.int64)
    N = len(Bz)
    pairs = [(i, (i + 1) % N, J1[i], True) for i in range(N if periodic else N - 1)]
    pairs += [(i, (i + 2) % N, J2[i], False) for i in range(N if periodic else N - 2)]
    diag = float(np.dot(s - 0.5, Bz))
    for i, j, J, _ in pairs:
        diag += (0.25 if s[i] == s[j] else -0.25) * J
    sigmaH[0] = s
    matrixelements[0] = diag
    num = 1
    for i, j, J, nearest in pairs:
        if J != 0.0 and s[i] != s[j]:
            c = s.copy()
            c[i], c[j] = s[j], s[i]
            sigmaH[num] = c
            matrixelements[num] = (-J / 2) if (nearest and Marshall_sign) else (J / 2)
            num += 1
    return num


def J1J2Slices(J1, J2, Bz, sigmasp, sigmas, H, sigmaH, matrixelements, Marshall_sign):
    """J1J2/TrainingRNN_J1J2.py:95-127 -> (slices, total).  ``Marshall_sign`` has its intended meaning here
    (upstream passes it positionally into ``periodic``)."""
    slices, length = [], 0
    for n in range(sigmasp.shape[0]):
        num = J1J2MatrixElements(J1, J2, Bz, sigmasp[n], sigmaH, matrixelements, False, Marshall_sign)
        slices.append(slice(length, length + num))
        H[slices[n]] = matrixelements[:num]
        sigmas[slices[n]] = sigmaH[:num]
        length += num
    return slices, length


def J1J2_local_energies(J1, J2, Bz, samples, wf, Marshall_sign=False, periodic=False):
    """The block J1J2/TrainingRNN_J1J2.py:256-279 (J1J2Slices + chunked log_amplitude + reduce) as one fused
    device call -> complex64 ndarray [numsamples]."""
    wf = find_wf(wf)
    ham = np.concatenate([[float(periodic), float(Marshall_sign)], np.asarray(J1, dtype=np.float64),
                          np.asarray(J2, dtype=np.float64), np.asarray(Bz, dtype=np.float64)])
    e = wf.engine.local_energies(_lib.HAM_J1J2, ham, torch.as_tensor(to_numpy(samples)))
    return e.cpu().numpy().astype(np.complex64)


def run_J1J2(numsteps=10 ** 5, systemsize=20, J1_=1.0, J2_=0.0, Marshall_sign=False, num_units=50, num_layers=1,
             numsamples=500, learningrate=2.5e-4, seed=111):
    """J1J2/TrainingRNN_J1J2.py:131-308 -> (meanEnergy, varEnergy).  Constant learning rate (:305-306: "without learning
    decay"; the Tutorials copy of the script decays it as 1/(1/lr + it/10) instead)."""
    N = systemsize
    wf = RNNwavefunction(N, units=[num_units] * num_layers, seed=seed)
    ham = np.concatenate([[0.0, float(Marshall_sign)], J1_ * np.ones(N), J2_ * np.ones(N), np.zeros(N)])
    v = VMC(wf.engine, _lib.HAM_J1J2, ham, numsamples, lr=learningrate, seed=seed)
    meanEnergy, varEnergy = [], []
    for it in range(numsteps + 1):
        m, var = v.step(lr=learningrate)
        e = v.last["eloc"]
        meanEnergy.append(complex(m))
        varEnergy.append(float(e.real.var(unbiased=False)))   # the reference takes the variance of the real part (:282)
        if it % 10 == 0:
            print('mean(E): {0}, var(E): {1}, #samples {2}, #Step {3} \n\n'.format(meanEnergy[-1], varEnergy[-1], numsamples, it))
    return meanEnergy, varEnergy
