"""Symmetry groups of the reference's symmetric wavefunction methods as index-remap tables.

Every group element is a permutation of the lattice sites, optionally followed by a global spin flip; the tables are
built once per (wavefunction, group) on the host by pushing an index grid through the reference's own sequence of array
operations, and live on the device afterwards (``Engine.symmetry_group``).  The copies are then formed where they are
consumed -- ``rnnwf_symmetry_apply`` / ``rnnwf_local_energies_sym`` -- instead of materialising rot90 / flip / roll
tensors on the host side.

Reference: 2DHeisenberg/MDComplexRNN.py:320-577 (rotsym, rotreducedsym, c2sym, c2vsym, c2dsym, c4vsym, each with a
spin-parity variant), RydbergKagome_2DGRU/MDRNNWavefunction.py:242-291 (c6sym via rotate60, translationsym via tf.roll),
1DTFIM/RNNwavefunction_paritysym.py:122-145 (chain reversal)."""
import numpy as np
import torch

# group elements in the order the reference appends them to list_samples; tf.image.rot90(k) == rot90(k, dims=(1, 2))
_ROT = lambda k: (lambda s: torch.rot90(s, k, dims=(1, 2)))
HEIS_OPS = {
    "id": lambda s: s, "r1": _ROT(-1), "r2": _ROT(-2), "r3": _ROT(-3),
    "fx": lambda s: torch.flip(s, dims=[1]), "fy": lambda s: torch.flip(s, dims=[2]),
    "t": lambda s: s.transpose(1, 2), "tr2": lambda s: torch.rot90(s, -2, dims=(1, 2)).transpose(1, 2),
}
_C2V = {"A1": [1, 1, 1, 1], "A2": [1, 1, -1, -1], "B1": [1, -1, 1, -1], "B2": [1, -1, -1, 1]}
HEIS_GROUPS = {
    "rotsym": (["id", "r1", "r2", "r3"], {"A": [1, 1, 1, 1], "B": [1, -1, 1, -1]}),
    "rotreducedsym": (["id", "r1"], {"A": [1, 1], "B": [1, -1]}),
    "c2sym": (["id", "r2"], {"A": [1, 1], "B": [1, -1]}),
    "c2vsym": (["id", "r2", "fx", "fy"], _C2V),
    "c2dsym": (["id", "r2", "t", "tr2"], _C2V),
    "c4vsym": (["id", "r1", "r2", "r3", "fx", "fy", "t", "tr2"],
               {"A1": [1] * 8, "A2": [1, 1, 1, 1, -1, -1, -1, -1], "B1": [1, -1, 1, -1, 1, 1, -1, -1], "B2": [1, -1, 1, -1, -1, -1, 1, 1]}),
}


def heisenberg_tables(Nx, Ny, group, group_character, spinparity_value=None):
    """-> (perm [G, Nx*Ny], flip [G], characters [G]).  With a spin parity the spin-flipped copies follow in the same
    order with characters spinparity_value * chi (the intended product; upstream's list arithmetic drops them for
    spinparity_value = -1 in all variants but c4vsym, see oracle.wavefunctions.heis_symmetry_copies)."""
    if Nx != Ny and group in ("rotsym", "rotreducedsym", "c2dsym", "c4vsym"):
        raise ValueError(f"{group} needs a square lattice")
    ops, table = HEIS_GROUPS[group]
    idx = torch.arange(Nx * Ny).view(1, Nx, Ny)
    perm = [HEIS_OPS[o](idx).reshape(-1).numpy() for o in ops]
    chi = list(table[group_character])
    flip = [0] * len(ops)
    if spinparity_value is not None:
        perm, flip, chi = perm + perm, flip + [1] * len(ops), chi + [spinparity_value * c for c in chi]
    return np.stack(perm).astype(np.int32), np.asarray(flip, dtype=np.int8), chi


def rotate60_index(Lx, Ly):
    """Source index of every atom under one rotate60 (RydbergKagome_2DGRU/MDRNNWavefunction.py:281-291), flat
    [Lx*Ly*3]: out[i, j, k] = in[src]."""
    idx = np.arange(Lx * Ly * 3).reshape(Lx, Ly, 3)
    out = np.empty_like(idx)
    for i in range(Lx):
        for j in range(Ly):
            out[i, j, 0] = idx[(i + j + 1) % Lx, (Ly - i) % Ly, 1]
            out[i, j, 1] = idx[(i + j) % Lx, (Ly - i) % Ly, 2]
            out[i, j, 2] = idx[(i + j + 1) % Lx, (Ly - (i + 1)) % Ly, 0]
    return out.reshape(-1)


def rydberg_tables(Lx, Ly, symmetry):
    """-> (perm [G, Lx*Ly*3], None, None): the six rotations (c6sym) or all Lx*Ly translations (translationsym)."""
    n = Lx * Ly * 3
    if symmetry == "c6sym":
        r = rotate60_index(Lx, Ly)
        perm = [np.arange(n)]
        for _ in range(5):
            perm.append(perm[-1][r])            # s_{k+1}[i] = s_k[r[i]] = s[perm_k[r[i]]]
    elif symmetry == "translationsym":
        idx = np.arange(n).reshape(Lx, Ly, 3)
        perm = [np.roll(idx, shift=(rx, ry), axis=(0, 1)).reshape(-1) for rx in range(Lx) for ry in range(Ly)]
    else:
        raise ValueError(symmetry)
    return np.stack(perm).astype(np.int32), None, None


def parity_tables(N):
    """Chain and its mirror image (1DTFIM/RNNwavefunction_paritysym.py:125)."""
    return np.stack([np.arange(N), np.arange(N)[::-1]]).astype(np.int32), None, None
