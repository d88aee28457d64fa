"""Generate the golden fixtures in this directory from the UNMODIFIED reference.

Runs only inside the build container (needs /root/reference; TensorFlow is absent there, so the
module-level ``import tensorflow`` of the reference's training scripts is satisfied with a MagicMock --
none of the host-side NumPy functions exercised here touches it).  The reference's local-energy
functions are driven through a duck-typed ``sess`` whose ``run(handle, feed_dict)`` evaluates a
closed-form, seeded stand-in for log psi (``fake_logpsi`` below), so the fixtures pin the
enumeration / matrix elements / reduction code, not a network.

    python tests/golden/make_golden.py      # rewrites tests/golden/*.npz
"""
import importlib.util
import os
import sys
from unittest.mock import MagicMock

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def _load(subdir, fname, modname):
    sys.modules.setdefault("tensorflow", MagicMock())
    d = os.path.join(REF, subdir)
    sys.path.insert(0, d)
    try:
        spec = importlib.util.spec_from_file_location(modname, os.path.join(d, fname))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(d)
        for k in ("RNNwavefunction", "ComplexRNNwavefunction"):
            sys.modules.pop(k, None)
    return mod


def fake_params(nsites, seed):
    rng = np.random.RandomState(seed)
    return dict(a=rng.uniform(-0.4, 0.4, nsites), B=rng.uniform(-0.1, 0.1, (nsites, nsites)),
                c=rng.uniform(-1.0, 1.0, nsites))


def fake_logpsi(fp, configs):
    """Deterministic stand-in for the network: (logp, phase) of each configuration."""
    x = configs.reshape(configs.shape[0], -1).astype(np.float64)
    logp = x @ fp["a"] + np.einsum("bi,ij,bj->b", x, fp["B"], x) - 3.0
    phase = np.sin(x @ fp["c"])
    return logp, phase


class FakeSess:
    def __init__(self, fp, complex_):
        self.fp, self.complex_ = fp, complex_

    def run(self, handle, feed_dict=None):
        (cfg,) = feed_dict.values()
        lp, ph = fake_logpsi(self.fp, np.asarray(cfg))
        return 0.5 * lp + 1j * ph if self.complex_ else lp


def u1_samples(rng, ns, shape):
    n = int(np.prod(shape))
    out = np.zeros((ns, n), dtype=np.int32)
    for b in range(ns):
        out[b, rng.permutation(n)[: n // 2]] = 1
    return out.reshape((ns,) + tuple(shape))


def main():
    rng = np.random.RandomState(1234)

    # E1 -- 1D TFIM
    m = _load("1DTFIM", "TrainingRNN_1DTFIM.py", "ref_tfim1d")
    ns, N = 7, 9
    samples = rng.randint(0, 2, (ns, N)).astype(np.int32)
    Jz = rng.uniform(0.5, 1.5, N)
    fp = fake_params(N, 1)
    queue = np.zeros((N + 1, ns, N), dtype=np.int32)
    lp = np.zeros((N + 1) * ns)
    E = m.Ising_local_energies(Jz, 0.7, samples, queue, None, "ph", lp, FakeSess(fp, False))
    np.savez(os.path.join(HERE, "tfim1d.npz"), samples=samples, Jz=Jz, Bx=0.7, E=E, **{"fp_" + k: v for k, v in fp.items()})

    # E2 -- 2D TFIM
    m = _load("2DTFIM_2DRNN", "Training2DRNN_2DTFIM.py", "ref_tfim2d")
    ns, Nx, Ny = 5, 4, 4
    samples = rng.randint(0, 2, (ns, Nx, Ny)).astype(np.int32)
    Jz = rng.uniform(0.5, 1.5, (Nx, Ny))
    fp = fake_params(Nx * Ny, 2)
    queue = np.zeros((Nx * Ny + 1, ns, Nx, Ny), dtype=np.int32)
    lp = np.zeros((Nx * Ny + 1) * ns)
    E = m.Ising2D_local_energies(Jz, 3.0, Nx, Ny, samples, queue, None, "ph", lp, FakeSess(fp, False))
    np.savez(os.path.join(HERE, "tfim2d.npz"), samples=samples, Jz=Jz, Bx=3.0, E=E, **{"fp_" + k: v for k, v in fp.items()})

    # E3 -- 1D J1-J2 (open boundaries, no Marshall sign: the defaults the scripts run with)
    m = _load("J1J2", "TrainingRNN_J1J2.py", "ref_j1j2")
    ns, N = 6, 10
    samples = u1_samples(rng, ns, (N,))
    J1, J2, Bz = np.ones(N), 0.2 * np.ones(N), np.zeros(N)
    fp = fake_params(N, 3)
    sigmas = np.zeros((2 * N * ns, N), dtype=np.int32)
    H = np.zeros(2 * N * ns, dtype=np.float32)
    sigmaH = np.zeros((2 * N, N), dtype=np.int32)
    melems = np.zeros(2 * N, dtype=np.float32)
    slices, length = m.J1J2Slices(J1, J2, Bz, samples, sigmas, H, sigmaH, melems, False)
    la = FakeSess(fp, True).run(None, {"ph": sigmas[:length]})
    E = np.array([H[s].dot(np.exp(la[s] - la[s][0])) for s in slices])  # J1J2/TrainingRNN_J1J2.py:277-279
    np.savez(os.path.join(HERE, "j1j2.npz"), samples=samples, J1=J1, J2=J2, Bz=Bz, E=E,
             sigmas=sigmas[:length], H=H[:length], counts=np.array([s.stop - s.start for s in slices]),
             **{"fp_" + k: v for k, v in fp.items()})

    # E4 -- 2D Heisenberg, square (J2=0) and triangular (J2=1)
    m = _load("2DHeisenberg", "Helper_Functions.py", "ref_heis")
    for J2v, tag in ((0.0, "square"), (1.0, "triangular")):
        ns, Nx, Ny = 5, 4, 4
        N = Nx * Ny
        samples = u1_samples(rng, ns, (Nx, Ny))
        fp = fake_params(N, 4)
        sigmas = np.zeros(((4 * N + 1) * ns, Nx, Ny), dtype=np.int32)
        H = np.zeros((4 * N + 1) * ns)
        la = np.zeros((4 * N + 1) * ns, dtype=np.complex128)
        sigmaH = np.zeros((4 * N + 1, Nx, Ny), dtype=np.int32)
        melems = np.zeros(4 * N + 1)
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            E = m.Get_Elocs(J2v, Nx, Ny, samples, sigmas, H, la, sigmaH, melems, "ph", None, FakeSess(fp, True), "nosym")
            slices, length = m.Triangular2DHeis_LocalEnergies(J2v, Nx, Ny, samples, sigmas, H, sigmaH, melems)
        np.savez(os.path.join(HERE, f"heis2d_{tag}.npz"), samples=samples, J2=J2v, E=E, sigmas=sigmas[:length],
                 H=H[:length], counts=np.array([s.stop - s.start for s in slices]),
                 **{"fp_" + k: v for k, v in fp.items()})

    # E5 -- Rydberg / Kagome
    m = _load("RydbergKagome_2DGRU", "Helping_Functions.py", "ref_ryd")
    ns, Lx, Ly = 4, 3, 3
    pos = m.generate_kagome(Lx, Ly)
    nb, dist = m.neighbouring_atoms(pos, 2.001, Lx, Ly)
    samples = rng.randint(0, 2, (ns, Lx, Ly, 3)).astype(np.int64)
    fp = fake_params(Lx * Ly * 3, 5)
    N = 3 * Lx * Ly
    queue = np.zeros((N + 1, ns, Lx, Ly, 3), dtype=np.int64)
    lp = np.zeros((N + 1) * ns)
    m.numsampsteps = 1  # undefined global upstream
    V = 1.7 ** 6
    E = m.Rydberg_local_energies(V, 3.3, 1.0, nb, dist, samples, queue, None, "ph", lp, FakeSess(fp, False))
    np.savez(os.path.join(HERE, "rydberg.npz"), samples=samples, V=V, delta=3.3, Omega=1.0, E=E, positions=pos,
             nb0=np.array(nb[0]), nb1=np.array(nb[1]), nb2=np.array(nb[2]),
             d0=np.array(dist[0]), d1=np.array(dist[1]), d2=np.array(dist[2]),
             **{"fp_" + k: v for k, v in fp.items()})
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
