"""Drop-in for the reference package ``2DHeisenberg/``.
Reference: 2DHeisenberg/MDTensorizedcell.py, MDTensorizedGRUcell.py, MDComplexRNN.py, Helper_Functions.py,
Run_RNNWF_2DHeis.py."""
import numpy as np
import torch

from . import _lib
from .compat import Handle, Placeholder, Session, find_wf, find_symmetry, to_numpy
from .engine import Engine
from .symmetry import HEIS_GROUPS, heisenberg_tables


class MDTensorizedcell:
    """tanh(sum_ik [h_h;h_v]_i [x_h;x_v]_k W_jik + b)  (2DHeisenberg/MDTensorizedcell.py:7-49)."""
    cell_id = _lib.CELL_TENS2D


class MDTensorizedGRUcell:
    """Tensorized 2D GRU cell (2DHeisenberg/MDTensorizedGRUcell.py:7-61) -- the reference's ``MDRNNcell`` when
    imported from MDTensorizedGRUcell, as Run_RNNWF_2DHeis.py does."""
    cell_id = _lib.CELL_TENSGRU2D


MDRNNcell = MDTensorizedGRUcell

_GROUPS = HEIS_GROUPS      # group elements / character tables in the reference's order (symmetry.py)


def regularized_identity(x, epsilon=1e-4):
    """2DHeisenberg/MDComplexRNN.py:23-25."""
    sign = torch.where(x >= 0, torch.ones_like(x), -torch.ones_like(x))
    return sign * torch.sqrt(x * x + epsilon * epsilon)


class RNNwavefunction:
    """2DHeisenberg/MDComplexRNN.py:27-65."""

    def __init__(self, systemsize_x, systemsize_y, cell=None, activation=None, units=[10], scope='RNNwavefunction', seed=111,
                 mag_fixed=True, magnetization=0):
        self.Nx, self.Ny, self.N = systemsize_x, systemsize_y, systemsize_x * systemsize_y
        self.scope, self.seed, self.mag_fixed, self.magnetization = scope, seed, mag_fixed, magnetization
        cid = getattr(cell, "cell_id", _lib.CELL_TENSGRU2D)
        self.engine = Engine(cid, systemsize_x, systemsize_y, num_units=units[0], complex_head=True, u1=bool(mag_fixed),
                             magnetization=int(magnetization), seed=seed)
        self._step = 0
        self._groups = {}

    def symmetry_group(self, name, group_character, spinparity_value=None):
        """Device remap tables of one of the reference's groups (MDComplexRNN.py:320-577), built once and cached."""
        key = (name, group_character, spinparity_value)
        if key not in self._groups:
            perm, flip, chi = heisenberg_tables(self.Nx, self.Ny, name, group_character, spinparity_value)
            self._groups[key] = self.engine.symmetry_group(perm, flip, chi)
        return self._groups[key]

    def sample(self, numsamples, inputdim, uniforms=None, philox=None):
        """MDComplexRNN.py:80-173 -> int64 [numsamples, Nx, Ny]."""
        if isinstance(numsamples, Placeholder):
            raise TypeError("numsamples must be an int")
        if uniforms is None and philox is None:
            philox = (self.seed, self._step * numsamples)
            self._step += 1
        sd, off = philox if philox is not None else (0, 0)
        s, lp, ph = self.engine.sample(numsamples, uniforms=uniforms, seed=sd, row_offset=off)
        self.samples = s.long()
        return self.samples

    def log_amplitude_nosymmetry(self, samples, inputdim):
        """MDComplexRNN.py:177-281 -> (log_probs, log_phases), float64 [B] each."""
        self.log_probs, self.log_phases = self.engine.logpsi(torch.as_tensor(samples))
        return self.log_probs, self.log_phases

    def log_amplitude_nosym(self, samples, inputdim):
        """MDComplexRNN.py:284-286 -> complex128 [B]."""
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self.log_amplitude_nosym(s, inputdim), samples)
        lp, ph = self.log_amplitude_nosymmetry(samples, inputdim)
        return torch.complex(0.5 * lp, ph)

    def log_amplitudes_fromsymmetrygroup(self, list_samples, inputdim, group_character_signs):
        """MDComplexRNN.py:290-316: amplitude sqrt(mean_g P_g), phase arg(sum_g e^{i phi_g}/chi_g) with the real part
        regularised.  All G copies go through the device sweep in one batch."""
        G = len(list_samples)
        allc = torch.cat([torch.as_tensor(s).reshape(-1, self.Nx, self.Ny) for s in list_samples], dim=0)
        lp, ph = self.engine.logpsi(allc)
        logp, phase = self.engine.symmetrize(lp.view(G, -1), ph.view(G, -1), list(group_character_signs))   # fused device epilogue
        return torch.complex(0.5 * logp, phase)

    def _sym(self, name, samples, inputdim, group_character, spinparity_value=None):
        grp = self.symmetry_group(name, group_character, spinparity_value)
        if isinstance(samples, Placeholder):
            return Handle(self, lambda s: self._sym(name, s, inputdim, group_character, spinparity_value), samples, symmetry=grp)
        # the G copies are formed on the device through the group's index-remap tables, swept in one batch, and combined
        # by the fused epilogue (amplitude sqrt(mean_g P_g), regularised character-weighted phase)
        logp, phase, _, _, _ = self.engine.logpsi_sym(grp, torch.as_tensor(samples))
        return torch.complex(0.5 * logp, phase)


def _add_symmetry_methods():
    def plain(name):
        return lambda self, samples, inputdim, group_character: self._sym(name, samples, inputdim, group_character)

    def parity(name):
        return lambda self, samples, inputdim, group_character, spinparity_value: self._sym(name, samples, inputdim, group_character, spinparity_value)
    for name in _GROUPS:
        setattr(RNNwavefunction, f"log_amplitude_{name}", plain(name))
        pname = "log_amplitude_c2v_spinparity" if name == "c2vsym" else f"log_amplitude_{name}_spinparity"
        setattr(RNNwavefunction, pname, parity(name))
    RNNwavefunction.log_amplitude_c2vsym_spinparity = RNNwavefunction.log_amplitude_c2v_spinparity


_add_symmetry_methods()


def Triangular2DHeis_MatrixElements(J2, Nx, Ny, sigmap, sigmas, matrixelements):
    """2DHeisenberg/Helper_Functions.py:19-93 for ONE configuration (compatibility helper, host arrays);
    returns the number of entries.  The training path enumerates on the device (Get_Elocs below)."""
    s = to_numpy(sigmap).astype(np.int64)
    z = 2.0 * s - 1
    matrixelements[0] = 0.25 * (z[:-1] * z[1:]).sum() + 0.25 * (z[:, :-1] * z[:, 1:]).sum() + 0.25 * J2 * (z[:-1, :-1] * z[1:, 1:]).sum()
    sigmas[0] = s
    num = 1
    for ny in range(Ny):
        for nx in range(Nx):
            for dx, dy, w in ((1, 0, -0.5), (0, 1, -0.5), (1, 1, J2 / 2)):
                mx, my = nx + dx, ny + dy
                if mx < Nx and my < Ny and s[nx, ny] != s[mx, my]:
                    c = s.copy()
                    c[nx, ny], c[mx, my] = s[mx, my], s[nx, ny]
                    sigmas[num] = c
                    matrixelements[num] = w
                    num += 1
    return num


def Triangular2DHeis_LocalEnergies(J2, Nx, Ny, sigmasp, sigmas, H, sigmaH, matrixelements):
    """2DHeisenberg/Helper_Functions.py:95-119 -> (slices, total)."""
    slices, length = [], 0
    for n in range(sigmasp.shape[0]):
        num = Triangular2DHeis_MatrixElements(J2, Nx, Ny, sigmasp[n], sigmaH, matrixelements)
        slices.append(slice(length, length + num))
        H[slices[n]] = matrixelements[:num]
        sigmas[slices[n]] = sigmaH[:num]
        length += num
    return slices, length


def Get_Elocs(J2, Nx, Ny, samples, sigmas=None, H=None, log_amplitudes=None, sigmaH=None, matrixelements=None, inputs=None,
              log_amps=None, sess=None, RNN_symmetry="nosym"):
    """2DHeisenberg/Helper_Functions.py:152-177 -> complex128 ndarray [numsamples].  ``log_amps`` is the wavefunction
    or a handle from ``wf.log_amplitude_nosym(placeholder, 2)``; the scratch arrays are not needed."""
    wf = find_wf(log_amps, sess, inputs)
    grp = _group_for(wf, log_amps, RNN_symmetry)
    smp = torch.as_tensor(to_numpy(samples))
    if grp is not None:
        e = wf.engine.local_energies_sym(grp, _lib.HAM_HEIS2D, [float(J2)], smp)[0]
    else:
        e = wf.engine.local_energies(_lib.HAM_HEIS2D, [float(J2)], smp)
    return e.cpu().numpy()


def _group_for(wf, log_amps, RNN_symmetry):
    """The symmetry group the local energies must use: the one behind the handle `log_amps` (the reference passes the
    tensor of wf.log_amplitude_<RNN_symmetry>, Run_RNNWF_2DHeis.py:153-197), cross-checked against the flag."""
    grp = find_symmetry(log_amps)
    if RNN_symmetry != "nosym" and grp is None:
        raise ValueError(f"RNN_symmetry={RNN_symmetry!r}: pass the handle returned by wf.log_amplitude_{RNN_symmetry}"
                         "(placeholder, inputdim, group_character[, spinparity_value]) as log_amps")
    if RNN_symmetry == "nosym" and grp is not None:
        raise ValueError("log_amps is a symmetry-projected amplitude but RNN_symmetry='nosym'")
    return grp


def Get_Samples_and_Elocs(J2, Nx, Ny, samples=None, sigmas=None, H=None, log_amplitudes=None, sigmaH=None, matrixelements=None,
                          samples_tensor=None, inputs=None, log_amps=None, sess=None, RNN_symmetry="nosym"):
    """2DHeisenberg/Helper_Functions.py:123-150 -> (samples int64 ndarray, local_energies complex128 ndarray).
    Sampling, the saved hidden states and the connected-configuration sweep stay on the device in between."""
    wf = find_wf(log_amps, samples_tensor, sess)
    grp = _group_for(wf, log_amps, RNN_symmetry)
    ns = samples.shape[0] if samples is not None else samples_tensor.numsamples
    eng = wf.engine
    if grp is not None:
        s, _, _ = eng.sample(ns, seed=wf.seed, row_offset=wf._step * ns)
        wf._step += 1
        e = eng.local_energies_sym(grp, _lib.HAM_HEIS2D, [float(J2)], s)[0]
        return s.long().cpu().numpy(), e.cpu().numpy()
    saved = eng.new_saved(ns)
    s, lp, ph = eng.sample(ns, seed=wf.seed, row_offset=wf._step * ns, saved=saved)
    wf._step += 1
    e = eng.local_energies(_lib.HAM_HEIS2D, [float(J2)], s, lp, ph, saved=saved)
    return s.long().cpu().numpy(), e.cpu().numpy()
