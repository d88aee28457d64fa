"""Drop-in for the reference package ``2DTFIM_2DRNN/``.
Reference: 2DTFIM_2DRNN/MDRNNcell.py, RNNwavefunction.py, Training2DRNN_2DTFIM.py, run_2dTFIM.py."""
import numpy as np
import torch

from . import _lib, schedules
from .compat import Handle, Placeholder, Session, find_wf, to_numpy
from .engine import Engine
from .vmc import VMC


class MDRNNcell:
    """Marker for the vanilla 2D cell elu(x_h Uh + h_h Wh + x_v Uv + h_v Wv + b) (2DTFIM_2DRNN/MDRNNcell.py:6-66);
    the arithmetic lives in the sweep kernel."""
    cell_id = _lib.CELL_VANILLA2D

    def __init__(self, num_units=None, num_in=None, name=None, dtype=None, reuse=None):
        self.num_units, self.num_in, self.name = num_units, num_in, name


class RNNwavefunction:
    """2DTFIM_2DRNN/RNNwavefunction.py:5-33 (zig-zag sweep, positive wavefunction)."""

    def __init__(self, systemsize_x, systemsize_y, cell=None, units=[10], scope='RNNwavefunction', seed=111):
        self.Nx, self.Ny, self.N, self.scope, self.seed = systemsize_x, systemsize_y, systemsize_x * systemsize_y, scope, seed
        self.engine = Engine(getattr(cell, "cell_id", _lib.CELL_VANILLA2D), systemsize_x, systemsize_y, num_units=units[0], seed=seed)
        self._step = 0

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """2DTFIM_2DRNN/RNNwavefunction.py:35-118 -> int64 [numsamples, Nx, Ny]."""
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, ph = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        return self.samples

    def log_probability(self, samples, inputdim):
        """2DTFIM_2DRNN/RNNwavefunction.py:120-200 -> float64 [B]."""
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.engine.logpsi(s)[0], samples)
        self.log_probs = self.engine.logpsi(torch.as_tensor(samples))[0]
        return self.log_probs


def Ising2D_local_energies(Jz, Bx, Nx, Ny, samples, queue_samples=None, log_probs_tensor=None, samples_placeholder=None,
                           log_probs=None, sess=None):
    """2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:13-83 -> float64 ndarray [numsamples]."""
    wf = find_wf(log_probs_tensor, sess, samples_placeholder)
    ham = np.concatenate([[float(Bx)], np.asarray(Jz, dtype=np.float64).reshape(-1)])
    e = wf.engine.local_energies(_lib.HAM_TFIM2D, ham, torch.as_tensor(to_numpy(samples)))
    return e.real.cpu().numpy()


def run_2DTFIM(numsteps=2 * 10 ** 4, systemsize_x=5, systemsize_y=5, Bx=+2, num_units=50, numsamples=500, learningrate=5e-3, seed=111):
    """2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:88-231 -> (meanEnergy, varEnergy).  lr schedule lr*(1+it/5000)^-1 (:228)."""
    Nx, Ny = systemsize_x, systemsize_y
    wf = RNNwavefunction(Nx, Ny, cell=MDRNNcell, units=[num_units], seed=seed)
    ham = np.concatenate([[float(Bx)], np.ones(Nx * Ny)])
    v = VMC(wf.engine, _lib.HAM_TFIM2D, ham, numsamples, lr=learningrate, seed=seed)
    meanEnergy, varEnergy = [], []
    for it in range(numsteps + 1):
        m, var = v.step(lr=schedules.lr_tfim2d_2drnn(it, learningrate))
        meanEnergy.append(float(m.real))
        varEnergy.append(float(var))
        if it % 10 == 0:
            print('mean(E): {0}, var(E): {1}, #samples {2}, #Step {3} \n\n'.format(meanEnergy[-1], varEnergy[-1], numsamples, it))
    return meanEnergy, varEnergy
