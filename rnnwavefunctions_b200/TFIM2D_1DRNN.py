"""Drop-in for the reference package ``2DTFIM_1DRNN/``: a stacked 1D GRU raster-scanned over an Nx x Ny lattice
(site (i, j) at flat index i * Ny + j) with the 2D transverse-field Ising bonds.
Reference: 2DTFIM_1DRNN/RNNwavefunction.py, Training1DRNN_2DTFIM.py, run_2dTFIM.py."""
import numpy as np
import torch

from . import _lib, schedules
from .compat import Handle, Placeholder, Session, find_wf, to_numpy
from .engine import Engine
from .vmc import VMC


class RNNwavefunction:
    """2DTFIM_1DRNN/RNNwavefunction.py:8-38.  ``cell`` / ``activation`` are accepted for signature compatibility (the
    training script passes CudnnCompatibleGRUCell, Training1DRNN_2DTFIM.py:105)."""

    def __init__(self, systemsize_x, systemsize_y, cell=None, activation=None, units=[10], scope='RNNwavefunction', seed=111):
        if len(set(units)) != 1:
            raise ValueError("all layers must have the same number of units")
        self.Nx, self.Ny, self.N, self.scope, self.seed = systemsize_x, systemsize_y, systemsize_x * systemsize_y, scope, seed
        self.engine = Engine(_lib.CELL_GRU1D, systemsize_x, systemsize_y, num_units=units[0], num_layers=len(units), seed=seed)
        self._step = 0

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """2DTFIM_1DRNN/RNNwavefunction.py:40-84 -> int64 [numsamples, Nx*Ny]."""
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, _ = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        return self.samples

    def log_probability(self, samples, inputdim):
        """2DTFIM_1DRNN/RNNwavefunction.py:86-130 -> float64 [B]; samples [B, Nx*Ny]."""
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.engine.logpsi(s)[0], samples)
        self.log_probs = self.engine.logpsi(torch.as_tensor(samples).reshape(-1, self.N))[0]
        return self.log_probs


def Ising2D_local_energies(Jz, Bx, Nx, Ny, samples, queue_samples=None, log_probs_tensor=None, samples_placeholder=None,
                           log_probs=None, sess=None):
    """2DTFIM_1DRNN/Training1DRNN_2DTFIM.py:13-81 -> float64 ndarray [numsamples]; samples [numsamples, Nx*Ny], Jz [Nx, Ny]
    (bond (i,j)-(i+1,j) weighs Jz[i, j], bond (i,j)-(i,j+1) weighs Jz[i, j], as upstream's Jz[i,:] / Jz[:,i] rows)."""
    wf = find_wf(log_probs_tensor, sess, samples_placeholder)
    if wf.engine.cell != _lib.CELL_GRU1D or (wf.engine.nx, wf.engine.ny) != (Nx, Ny):
        raise ValueError("Ising2D_local_energies (2DTFIM_1DRNN) needs the 1D-RNN wavefunction of the same Nx x Ny lattice")
    ham = np.concatenate([[float(Bx)], np.asarray(Jz, dtype=np.float64).reshape(-1)])
    e = wf.engine.local_energies(_lib.HAM_TFIM2D, ham, torch.as_tensor(to_numpy(samples)).reshape(-1, Nx * Ny))
    return e.real.cpu().numpy()


def run_2DTFIM(numsteps=2 * 10 ** 4, systemsize_x=5, systemsize_y=5, Bx=+2, num_units=50, num_layers=1, numsamples=500,
               learningrate=1e-3, seed=333):
    """2DTFIM_1DRNN/Training1DRNN_2DTFIM.py:85-233 -> (meanEnergy, varEnergy).  lr schedule 1/(1/lr + it/10) (:229)."""
    Nx, Ny = systemsize_x, systemsize_y
    wf = RNNwavefunction(Nx, Ny, units=[num_units] * num_layers, seed=seed)
    ham = np.concatenate([[float(Bx)], np.ones(Nx * Ny)])
    v = VMC(wf.engine, _lib.HAM_TFIM2D, ham, numsamples, lr=learningrate, seed=seed)
    meanEnergy, varEnergy = [], []
    for it in range(numsteps + 1):
        m, var = v.step(lr=schedules.lr_tfim2d_1drnn(it, learningrate))
        meanEnergy.append(float(m.real))
        varEnergy.append(float(var))
        if it % 10 == 0:
            print('mean(E): {0}, var(E): {1}, #samples {2}, #Step {3} \n\n'.format(meanEnergy[-1], varEnergy[-1], numsamples, it))
    return meanEnergy, varEnergy
