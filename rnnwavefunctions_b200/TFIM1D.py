"""Drop-in for the reference package ``1DTFIM/`` (the twin ``2DTFIM_1DRNN/`` -- the same GRU raster-scanned over an
Nx x Ny lattice with 2D Ising bonds -- is ``TFIM2D_1DRNN.py``).

Reference: 1DTFIM/RNNwavefunction.py, RNNwavefunction_paritysym.py, TrainingRNN_1DTFIM.py, run_1dTFIM.py."""
import math

import numpy as np
import torch

from . import _lib
from .compat import Handle, Placeholder, Session, find_wf, to_numpy
from .engine import Engine
from .symmetry import parity_tables
from .vmc import VMC


class RNNwavefunction:
    """1DTFIM/RNNwavefunction.py:7-33.  ``cell`` is accepted for signature compatibility (the stacked
    CudnnCompatibleGRUCell is the only cell the reference's 1D scripts use)."""

    def __init__(self, systemsize, cell=None, units=[10], scope='RNNwavefunction', seed=111):
        if len(set(units)) != 1:
            raise ValueError("all layers must have the same number of units")
        self.N, self.scope, self.units = systemsize, scope, list(units)
        self.engine = Engine(_lib.CELL_GRU1D, systemsize, 1, num_units=units[0], num_layers=len(units), seed=seed)
        self._step = 0
        self.seed = seed

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """1DTFIM/RNNwavefunction.py:35-74 -> int64 [numsamples, N].  ``uniforms`` [numsamples, N] float32 or
        ``philox=(seed, row_offset)`` select the random stream (default: Philox keyed by the constructor seed)."""
        assert inputdim == 2
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, ph = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        self.log_probs = lp
        return self.samples

    def log_probability(self, samples, inputdim):
        """1DTFIM/RNNwavefunction.py:76-118 -> float64 [B]."""
        assert inputdim == 2
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.engine.logpsi(s)[0], samples)
        self.log_probs = self.engine.logpsi(torch.as_tensor(samples))[0]
        return self.log_probs


class RNNwavefunction_paritysym(RNNwavefunction):
    """1DTFIM/RNNwavefunction_paritysym.py:80-145: log(0.5*(P(s) + P(reversed s))).  The mirror image is formed on the
    device through a two-element remap table; both copies go through one sweep and the fused epilogue."""

    def symmetry_group(self):
        if getattr(self, "_parity", None) is None:
            self._parity = self.engine.symmetry_group(parity_tables(self.N)[0])
        return self._parity

    def log_probability(self, samples, inputdim):
        grp = self.symmetry_group()
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.log_probability(s, inputdim), samples, symmetry=grp)
        return self.engine.logpsi_sym(grp, torch.as_tensor(samples))[0]


def Ising_local_energies(Jz, Bx, samples, queue_samples=None, log_probs_tensor=None, samples_placeholder=None,
                         log_probs=None, sess=None):
    """1DTFIM/TrainingRNN_1DTFIM.py:13-75 -> float64 ndarray [numsamples].  Pass the wavefunction (or the handle
    returned by ``wf.log_probability(placeholder, 2)``) as ``log_probs_tensor``; the queue / log_probs scratch arrays
    of the reference are not needed (the connected configurations are enumerated and reduced on the device)."""
    wf = find_wf(log_probs_tensor, sess, samples_placeholder)
    N = wf.N
    ham = np.concatenate([[float(Bx)], np.asarray(Jz, dtype=np.float64).reshape(-1)[:max(1, N - 1)]])
    if isinstance(wf, RNNwavefunction_paritysym):
        e = wf.engine.local_energies_sym(wf.symmetry_group(), _lib.HAM_TFIM1D, ham, torch.as_tensor(to_numpy(samples)))[0]
    else:
        e = wf.engine.local_energies(_lib.HAM_TFIM1D, ham, torch.as_tensor(to_numpy(samples)))
    return e.real.cpu().numpy()


def run_1DTFIM(numsteps=10 ** 4, systemsize=20, num_units=50, Bx=1, num_layers=1, numsamples=500, learningrate=5e-3, seed=111):
    """1DTFIM/TrainingRNN_1DTFIM.py:79-230 -> (meanEnergy, varEnergy); checkpoint I/O is out of scope."""
    N = systemsize
    wf = RNNwavefunction(N, units=[num_units] * num_layers, seed=seed)
    ham = np.concatenate([[float(Bx)], np.ones(max(1, N - 1))])
    v = VMC(wf.engine, _lib.HAM_TFIM1D, ham, numsamples, lr=learningrate, seed=seed)
    meanEnergy, varEnergy = [], []
    for it in range(numsteps + 1):
        m, var = v.step()
        meanEnergy.append(float(m.real))
        varEnergy.append(float(var))
        if it % 10 == 0:
            print('mean(E): {0}, var(E): {1}, #samples {2}, #Step {3} \n\n'.format(meanEnergy[-1], varEnergy[-1], numsamples, it))
    return meanEnergy, varEnergy
