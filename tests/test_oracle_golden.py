"""CPU: the oracle's local-energy restatements against the fixtures produced by the UNMODIFIED reference
functions (tests/golden/make_golden.py), plus oracle self-consistency (normalisation, U(1), known answers)."""
import os

import numpy as np
import torch

from oracle import local_energies as ole
from oracle import wavefunctions as owf
from oracle import costs

G = os.path.join(os.path.dirname(__file__), "golden")


def _fake(z):
    fp = {k[3:]: z[k] for k in z.files if k.startswith("fp_")}

    def f(configs):
        x = configs.reshape(configs.shape[0], -1).astype(np.float64)
        return x @ fp["a"] + np.einsum("bi,ij,bj->b", x, fp["B"], x) - 3.0, np.sin(x @ fp["c"])
    return f


def test_tfim1d_golden():
    z = np.load(os.path.join(G, "tfim1d.npz"))
    E = ole.ising1d_local_energies(z["Jz"], float(z["Bx"]), z["samples"], _fake(z))
    np.testing.assert_allclose(E, z["E"], rtol=1e-13, atol=1e-13)


def test_tfim2d_golden():
    z = np.load(os.path.join(G, "tfim2d.npz"))
    E = ole.ising2d_local_energies(z["Jz"], float(z["Bx"]), z["samples"], _fake(z))
    np.testing.assert_allclose(E, z["E"], rtol=1e-13, atol=1e-13)


def test_j1j2_golden():
    z = np.load(os.path.join(G, "j1j2.npz"))
    E, counts = ole.j1j2_local_energies(z["J1"], z["J2"], z["Bz"], z["samples"], _fake(z))
    assert (counts == z["counts"]).all()
    np.testing.assert_allclose(E, z["E"], rtol=0, atol=1e-6)   # the reference keeps H in float32 (J1J2/TrainingRNN_J1J2.py:236)
    c, h = ole.j1j2_connected(z["J1"], z["J2"], z["Bz"], z["samples"][0])
    assert (c == z["sigmas"][:len(h)]).all()
    np.testing.assert_allclose(h, z["H"][:len(h)], atol=1e-7)


def test_heis2d_golden():
    for tag in ("square", "triangular"):
        z = np.load(os.path.join(G, f"heis2d_{tag}.npz"))
        E, counts = ole.heis2d_local_energies(float(z["J2"]), z["samples"], _fake(z))
        assert (counts == z["counts"]).all()
        np.testing.assert_allclose(E, z["E"], rtol=1e-13, atol=1e-13)
        off = 0
        for s, k in zip(z["samples"], z["counts"]):
            c, h = ole.heis2d_connected(float(z["J2"]), s)
            assert (c == z["sigmas"][off:off + k]).all()
            np.testing.assert_allclose(h, z["H"][off:off + k], atol=0)
            off += k


def test_rydberg_golden():
    z = np.load(os.path.join(G, "rydberg.npz"))
    Lx, Ly = z["samples"].shape[1:3]
    pos = ole.generate_kagome(Lx, Ly)
    np.testing.assert_allclose(pos, z["positions"], atol=0)
    nb, d = ole.neighbouring_atoms(pos, 2.001, Lx, Ly)
    for k in range(3):
        assert (np.array(nb[k]) == z[f"nb{k}"]).all()
        np.testing.assert_allclose(np.array(d[k]), z[f"d{k}"], atol=1e-15)
    assert [len(x) for x in nb] == [14, 14, 14]
    E = ole.rydberg_local_energies(float(z["V"]), float(z["delta"]), float(z["Omega"]), nb, d, z["samples"], _fake(z))
    np.testing.assert_allclose(E, z["E"], rtol=1e-12, atol=1e-12)


def _params_1d(nh, complex_, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: (torch.rand(*s, generator=g, dtype=torch.float64) - 0.5)
    P = {"gru0.Wg": r(2 + nh, 2 * nh), "gru0.bg": r(2 * nh), "gru0.Wci": r(2, nh), "gru0.Wch": r(nh, nh),
         "gru0.bci": r(nh), "gru0.bch": r(nh), "dense.W": r(nh, 2), "dense.b": r(2)}
    if complex_:
        P["dense_phase.W"], P["dense_phase.b"] = r(nh, 2), r(2)
    return P


def test_oracle_normalisation_and_u1():
    N, nh = 8, 6
    allcfg = torch.as_tensor(((np.arange(2 ** N)[:, None] >> np.arange(N)[None, :]) & 1))
    P = _params_1d(nh, False)
    lp = owf.rnn1d(P, N, [nh], samples=allcfg)["logp"]
    assert abs(float(torch.exp(lp).sum()) - 1) < 1e-12
    Pc = _params_1d(nh, True, seed=1)
    r = owf.rnn1d(Pc, N, [nh], samples=allcfg, complex_=True, u1=True)
    p = torch.exp(r["logp"])
    in_sector = allcfg.sum(1) == N // 2
    assert abs(float(p[in_sector].sum()) - 1) < 1e-12          # all weight in the Sz=0 sector
    assert float(p[~in_sector].max()) == 0.0
    u = torch.rand(200, N, generator=torch.Generator().manual_seed(2), dtype=torch.float32)
    s = owf.rnn1d(Pc, N, [nh], uniforms=u, numsamples=200, complex_=True, u1=True)["samples"]
    assert (s.sum(1) == N // 2).all()


def test_oracle_sampling_follows_distribution():
    N, nh, ns = 4, 5, 40000
    P = _params_1d(nh, False, seed=3)
    u = torch.rand(ns, N, generator=torch.Generator().manual_seed(4), dtype=torch.float32)
    s = owf.rnn1d(P, N, [nh], uniforms=u, numsamples=ns)["samples"]
    allcfg = torch.as_tensor(((np.arange(2 ** N)[:, None] >> np.arange(N)[None, :]) & 1))
    p = torch.exp(owf.rnn1d(P, N, [nh], samples=allcfg)["logp"]).numpy()
    idx = (s.numpy() * (1 << np.arange(N))[None, :]).sum(1)
    freq = np.bincount(idx, minlength=2 ** N) / ns
    assert np.abs(freq - p).max() < 4 * np.sqrt(p.max() / ns) + 1e-3


def test_tfim_known_answer_exact_ground_state():
    """E_loc of the exact ground state is the ED energy for every configuration (Tutorial_1DTFIM.ipynb cell 8:
    E0 = -12.38148999965476 for N=10, h=1, OBC)."""
    N = 10
    dim = 2 ** N
    cfg = ((np.arange(dim)[:, None] >> np.arange(N)[None, :]) & 1)
    H = np.zeros((dim, dim))
    s = 2.0 * cfg - 1
    H[np.arange(dim), np.arange(dim)] = -(s[:, :-1] * s[:, 1:]).sum(1)
    for i in range(N):
        H[np.arange(dim), np.arange(dim) ^ (1 << i)] += -1.0
    w, v = np.linalg.eigh(H)
    assert abs(w[0] - (-12.38148999965476)) < 1e-9
    psi = np.abs(v[:, 0])
    lut = {tuple(c): 2 * np.log(a) for c, a in zip(cfg, psi)}
    f = lambda c: (np.array([lut[tuple(x)] for x in c]), np.zeros(len(c)))
    sel = cfg[np.random.RandomState(0).choice(dim, 50, replace=False)]
    E = ole.ising1d_local_energies(np.ones(N), 1.0, sel, f)
    np.testing.assert_allclose(E, w[0], rtol=1e-9)


def test_j1j2_known_answer_exact_ground_state():
    """Tutorial_1DJ1J2.ipynb cell 8: E0 = -3.9855798336170905 for N=10, J2=0.2, OBC."""
    N = 10
    dim = 2 ** N
    cfg = ((np.arange(dim)[:, None] >> np.arange(N)[None, :]) & 1)
    H = np.zeros((dim, dim))
    J1, J2, Bz = np.ones(N), 0.2 * np.ones(N), np.zeros(N)
    index = {tuple(c): i for i, c in enumerate(cfg)}
    for i, c in enumerate(cfg):
        cc, hh = ole.j1j2_connected(J1, J2, Bz, c)
        for c2, h in zip(cc, hh):
            H[i, index[tuple(c2)]] += h
    w, v = np.linalg.eigh(H)
    assert abs(w[0] - (-3.9855798336170905)) < 1e-9


def test_seeds_match_autograd_of_costs():
    g = torch.Generator().manual_seed(5)
    ns = 17
    lp = torch.randn(ns, generator=g, dtype=torch.float64, requires_grad=True)
    ph = torch.randn(ns, generator=g, dtype=torch.float64, requires_grad=True)
    E = torch.complex(torch.randn(ns, generator=g, dtype=torch.float64), torch.randn(ns, generator=g, dtype=torch.float64))
    costs.cost_complex(lp, ph, E, T=0.7).backward()
    sl, sp = costs.seeds_complex(lp.detach(), E, T=0.7)
    np.testing.assert_allclose(lp.grad.numpy(), sl.numpy(), atol=1e-14)
    np.testing.assert_allclose(ph.grad.numpy(), sp.numpy(), atol=1e-14)
    lp2 = torch.randn(ns, generator=g, dtype=torch.float64, requires_grad=True)
    costs.cost_real(lp2, E.real).backward()
    np.testing.assert_allclose(lp2.grad.numpy(), costs.seeds_real(E.real).numpy(), atol=1e-14)
