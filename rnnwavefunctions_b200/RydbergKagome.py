"""Drop-in for the reference package ``RydbergKagome_2DGRU/``.
Reference: RydbergKagome_2DGRU/MDGRUcell.py, MDRNNWavefunction.py, Helping_Functions.py, Run_RNNWF_Rydberg.py."""
import math

import numpy as np
import torch

from . import _lib
from .compat import Handle, Placeholder, Session, find_wf, find_symmetry, to_numpy
from .engine import Engine
from .symmetry import rydberg_tables


class MDRNNcell:
    """Periodic 2D GRU cell with four neighbours (RydbergKagome_2DGRU/MDGRUcell.py:6-59)."""
    cell_id = _lib.CELL_GRU4


def binary_to_intarray(binary_samples):
    """MDRNNWavefunction.py:12-18: bits (b0,b1,b2) on the last axis -> class b0 + 2 b1 + 4 b2."""
    b = torch.as_tensor(binary_samples).long()
    return b[..., 0] + 2 * b[..., 1] + 4 * b[..., 2]


def intarray_to_binary(samples, bin_length=3):
    """MDRNNWavefunction.py:20-34."""
    s = torch.as_tensor(samples).long()
    return torch.stack([(s >> k) & 1 for k in range(bin_length)], dim=-1)


def rotate60(samples):
    """MDRNNWavefunction.py:281-291 (counter-clockwise 60 degree rotation of the periodic Kagome lattice)."""
    s = torch.as_tensor(samples)
    _, Lx, Ly, _ = s.shape
    out = torch.empty_like(s)
    for i in range(Lx):
        for j in range(Ly):
            out[:, i, j, 0] = s[:, (i + j + 1) % Lx, (Ly - i) % Ly, 1]
            out[:, i, j, 1] = s[:, (i + j) % Lx, (Ly - i) % Ly, 2]
            out[:, i, j, 2] = s[:, (i + j + 1) % Lx, (Ly - (i + 1)) % Ly, 0]
    return out


class MDRNNWavefunction:
    """MDRNNWavefunction.py:37-68."""

    def __init__(self, systemsize_x, systemsize_y, cell=None, activation=None, units=[10], inputdim=2, scope='RNNwavefunction', seed=111):
        self.Nx, self.Ny, self.N, self.scope, self.seed = systemsize_x, systemsize_y, systemsize_x * systemsize_y, scope, seed
        self.engine = Engine(_lib.CELL_GRU4, systemsize_x, systemsize_y, num_units=units[0], seed=seed)
        self._step = 0
        self._groups = {}

    def symmetry_group(self, symmetry):
        """Device remap tables of the six rotate60 copies / all translations (MDRNNWavefunction.py:246-263), cached."""
        if symmetry not in self._groups:
            perm, _, _ = rydberg_tables(self.Nx, self.Ny, symmetry)
            self._groups[symmetry] = self.engine.symmetry_group(perm)
        return self._groups[symmetry]

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """MDRNNWavefunction.py:73-153 -> int64 [numsamples, Nx, Ny, 3]."""
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, _ = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        return self.samples

    def log_probability(self, samples, inputdim):
        """MDRNNWavefunction.py:157-238 -> float64 [B]."""
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.engine.logpsi(s)[0], samples)
        self.log_probs = self.engine.logpsi(torch.as_tensor(samples))[0]
        return self.log_probs

    def log_probability_symmetry(self, samples, inputdim=2, symmetry="nosym"):
        """MDRNNWavefunction.py:242-279: log(mean_g P_g) over the C6 rotations or all translations; ``"nosym"``
        (missing upstream) is the plain log-probability."""
        if symmetry not in ("nosym", "c6sym", "translationsym"):
            raise ValueError(symmetry)
        grp = None if symmetry == "nosym" else self.symmetry_group(symmetry)
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.log_probability_symmetry(s, inputdim, symmetry), samples, symmetry=grp)
        s = torch.as_tensor(samples).to(self.engine.device)
        if grp is None:
            return self.log_probability(s, inputdim)
        return self.engine.logpsi_sym(grp, s)[0]     # copies through the remap tables, one sweep, fused log(mean_g P_g)


def generate_kagome(Nx, Ny):
    """Helping_Functions.py:33-48: atom positions, index (i*Ny+j)*3+k, k: 0 up, 1 left, 2 right."""
    pos = np.zeros((Nx * Ny * 3, 2))
    for i in range(Nx):
        for j in range(Ny):
            left = np.array([2 * (i + j / 2), math.sqrt(3) * j])
            pos[(i * Ny + j) * 3 + 1] = left
            pos[(i * Ny + j) * 3 + 2] = left + [1.0, 0.0]
            pos[(i * Ny + j) * 3 + 0] = left + [0.5, math.sqrt(3) / 2]
    return pos


def neighbouring_atoms(positions, cutoff, Nx, Ny):
    """Helping_Functions.py:50-96 (host helper; the device path rebuilds the same table from Rcutoff)."""
    a1, a2 = np.array([2.0, 0.0]), np.array([1.0, math.sqrt(3)])
    images = [positions, positions - Nx * a1, positions - Ny * a2, positions - Nx * a1 - Ny * a2]
    nb, dist = [[], [], []], [[], [], []]
    for i in range(Nx):
        for j in range(Ny):
            for k in range(3):
                for kc in range(3):
                    for img in images:
                        d = float(np.linalg.norm(img[(i * Ny + j) * 3 + k] - positions[kc]))
                        if 0 < d <= cutoff:
                            nb[kc].append([i, j, k])
                            dist[kc].append(d)
    return nb, dist


def Rydberg_local_energies(V, delta, Omega, neighbours, distances, binary_samples, queue_samples=None, log_probs_tensor=None,
                           samples_placeholder=None, log_probs=None, sess=None, Rcutoff=None):
    """Helping_Functions.py:100-153 -> float64 ndarray [numsamples].  The neighbour table is rebuilt on the library
    side from the cutoff (largest listed distance unless ``Rcutoff`` is given)."""
    wf = find_wf(log_probs_tensor, sess, samples_placeholder)
    if Rcutoff is None:
        Rcutoff = max(max(d) for d in distances if len(d)) + 1e-9
    ham = [float(V), float(delta), float(Omega), float(Rcutoff)]
    smp = torch.as_tensor(to_numpy(binary_samples))
    grp = find_symmetry(log_probs_tensor)       # handle of wf.log_probability_symmetry(..., symmetry=RNN_symmetry): Run_RNNWF_Rydberg.py:155,262
    if grp is not None:
        e = wf.engine.local_energies_sym(grp, _lib.HAM_RYDBERG, ham, smp)[0]
    else:
        e = wf.engine.local_energies(_lib.HAM_RYDBERG, ham, smp)
    return e.real.cpu().numpy()
