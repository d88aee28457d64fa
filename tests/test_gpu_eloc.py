"""GPU parity of the fused local-energy path (on-device enumeration of connected configurations,
batched sweep, in-kernel reduction) against the oracle's restatement of the reference's NumPy
estimators (oracle/local_energies.py, itself pinned to the reference by tests/golden/).

Tolerance: 1e-5 relative (FP32 mode, BASELINE.json north_star)."""
import numpy as np
import pytest
import torch

from oracle import local_energies as ole
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_logpsi_fn, assert_rel

RTOL = 1e-5


def _randomize(eng, seed):
    g = torch.Generator().manual_seed(seed)
    p = eng.params.cpu()
    eng.set_weights(p + 0.15 * (torch.rand(p.shape, generator=g) - 0.5))


def _samples(eng, ns, seed, saved=None):
    g = torch.Generator().manual_seed(seed)
    u = torch.rand(ns, eng.nsteps, generator=g, dtype=torch.float32)
    return eng.sample(ns, uniforms=u, saved=saved)


def _check(e_gpu, e_ref):
    assert_rel(e_gpu.cpu().numpy(), e_ref, RTOL, "E_loc fp32")


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
@pytest.mark.parametrize("layers", [1, 2])
def test_tfim1d(cuda, reuse, layers):
    N, ns = 20, 300
    eng = Engine(_lib.CELL_GRU1D, N, 1, num_units=50, num_layers=layers, seed=1)
    _randomize(eng, 2)
    saved = eng.new_saved(ns) if reuse else None
    s, lp, ph = _samples(eng, ns, 3, saved)
    Jz = np.linspace(0.7, 1.3, N - 1)
    e, cnt = eng.local_energies(_lib.HAM_TFIM1D, np.concatenate([[1.0], Jz]), s, lp, ph, saved=saved, return_count=True)
    ref = ole.ising1d_local_energies(np.append(Jz, 0.0), 1.0, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    _check(e.real, ref)
    assert int(cnt) == N * ns
    assert float(e.imag.abs().max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
def test_tfim2d(cuda, reuse):
    Nx, Ny, ns = 5, 4, 200
    eng = Engine(_lib.CELL_VANILLA2D, Nx, Ny, num_units=50, seed=4)
    _randomize(eng, 5)
    saved = eng.new_saved(ns) if reuse else None
    s, lp, ph = _samples(eng, ns, 6, saved)
    Jz = np.random.RandomState(0).uniform(0.5, 1.5, (Nx, Ny))
    e = eng.local_energies(_lib.HAM_TFIM2D, np.concatenate([[3.0], Jz.reshape(-1)]), s, lp, ph, saved=saved)
    ref = ole.ising2d_local_energies(Jz, 3.0, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    _check(e.real, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
@pytest.mark.parametrize("periodic,marshall", [(False, False), (True, True)])
def test_j1j2(cuda, reuse, periodic, marshall):
    N, ns = 20, 250
    eng = Engine(_lib.CELL_GRU1D, N, 1, num_units=50, complex_head=True, u1=True, seed=7)
    _randomize(eng, 8)
    saved = eng.new_saved(ns) if reuse else None
    s, lp, ph = _samples(eng, ns, 9, saved)
    J1, J2, Bz = np.ones(N), 0.2 * np.ones(N), np.linspace(-0.1, 0.1, N)
    ham = np.concatenate([[float(periodic), float(marshall)], J1, J2, Bz])
    e, cnt = eng.local_energies(_lib.HAM_J1J2, ham, s, lp, ph, saved=saved, return_count=True)
    ref, counts = ole.j1j2_local_energies(J1, J2, Bz, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)),
                                          periodic=periodic, Marshall_sign=marshall)
    _check(e, ref)
    assert int(cnt) == int(counts.sum() - ns)  # the oracle's count includes the diagonal entry of each sample


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
@pytest.mark.parametrize("J2", [0.0, 1.0])
@pytest.mark.parametrize("cell", [_lib.CELL_TENSGRU2D, _lib.CELL_TENS2D])
def test_heisenberg2d(cuda, reuse, J2, cell):
    Nx, Ny, ns = 4, 4, 150
    eng = Engine(cell, Nx, Ny, num_units=24, complex_head=True, u1=True, seed=10)
    _randomize(eng, 11)
    saved = eng.new_saved(ns) if reuse else None
    s, lp, ph = _samples(eng, ns, 12, saved)
    e, cnt = eng.local_energies(_lib.HAM_HEIS2D, [J2], s, lp, ph, saved=saved, return_count=True)
    ref, counts = ole.heis2d_local_energies(J2, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    _check(e, ref)
    if J2 != 0.0:  # the reference also lists the zero-weight diagonal-bond exchanges when J2 == 0
        assert int(cnt) == int(counts.sum() - ns)


@pytest.mark.gpu
@pytest.mark.parametrize("reuse", [False, True])
def test_rydberg(cuda, reuse):
    Lx, Ly, ns = 3, 3, 120
    eng = Engine(_lib.CELL_GRU4, Lx, Ly, num_units=40, seed=13)
    _randomize(eng, 14)
    saved = eng.new_saved(ns) if reuse else None
    s, lp, ph = _samples(eng, ns, 15, saved)
    V, delta, Omega, Rc = 1.7 ** 6, 3.3, 1.0, 2.001
    e = eng.local_energies(_lib.HAM_RYDBERG, [V, delta, Omega, Rc], s, lp, ph, saved=saved)
    pos = ole.generate_kagome(Lx, Ly)
    nb, dist = ole.neighbouring_atoms(pos, Rc, Lx, Ly)
    ref = ole.rydberg_local_energies(V, delta, Omega, nb, dist, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(eng, oracle_params(eng)))
    _check(e.real, ref)


@pytest.mark.gpu
def test_golden_fixture_through_gpu_enumeration(cuda):
    """The reference's own connected-configuration list for its 4x4 fixture (tests/golden/heis2d_*.npz)
    has the same multiset of non-zero (configuration, matrix element) pairs as the device enumeration:
    checked through E_loc with a trained-weights-free network (any weights) against the oracle fed the
    reference's list."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "heis2d_triangular.npz"))
    samples = z["samples"]
    eng = Engine(_lib.CELL_TENSGRU2D, 4, 4, num_units=16, complex_head=True, u1=True, seed=21)
    _randomize(eng, 22)
    f = oracle_logpsi_fn(eng, oracle_params(eng))
    lp, ph = f(z["sigmas"])
    la = 0.5 * lp + 1j * ph
    off, ref = 0, []
    for k in z["counts"]:
        ref.append(z["H"][off:off + k].dot(np.exp(la[off:off + k] - la[off])))
        off += k
    e = eng.local_energies(_lib.HAM_HEIS2D, [float(z["J2"])], torch.from_numpy(samples))
    _check(e, np.asarray(ref))
