"""CPU: ``oracle/wavefunctions.py`` against fixtures produced by EXECUTING the unmodified reference wavefunction classes
under an eager NumPy stand-in for TensorFlow (tests/golden/tf_shim.py, tests/golden/make_wf_golden.py).

This pins the oracle's network half to the reference's own code -- sweep order, neighbour dictionaries, U(1) masks, head
combination, symmetry-group lists, rotate60 -- at 1e-12.  Still restated from knowledge of TF 1.13.1 (inside the shim,
SURVEY App. B) and therefore NOT pinned: CudnnCompatibleGRUCell, layers.Dense, multinomial / random.categorical.
"""
import os

import numpy as np
import pytest
import torch

from oracle import wavefunctions as owf

G = os.path.join(os.path.dirname(__file__), "golden")
TOL = dict(rtol=1e-12, atol=1e-12)
_GRU = {"gates/kernel": "Wg", "gates/bias": "bg", "candidate/input_projection/kernel": "Wci",
        "candidate/input_projection/bias": "bci", "candidate/hidden_projection/kernel": "Wch",
        "candidate/hidden_projection/bias": "bch"}


def _load(name):
    z = np.load(os.path.join(G, name + ".npz"))
    P = {}
    for k in z.files:
        if not k.startswith("var__"):
            continue
        tfname = k[5:].replace("__", "/")
        assert tfname.startswith("RNNwavefunction/")
        rest = tfname[len("RNNwavefunction/"):]
        v = torch.as_tensor(z[k], dtype=torch.float64)
        if rest.startswith("multi_rnn_cell/"):
            _, cell, _, leaf = rest.split("/", 3)
            P[f"gru{int(cell.split('_')[1])}.{_GRU[leaf]}"] = v
        elif rest.startswith("wf_dense"):
            head, leaf = rest.split("/")
            P[("dense_phase" if head == "wf_dense_phase" else "dense") + (".W" if leaf == "kernel" else ".b")] = v
        else:
            assert rest.endswith("_rnn_0"), rest
            P[rest[:-len("_rnn_0")]] = v
    return z, P


def _t(a):
    return torch.as_tensor(np.asarray(a)).long()


@pytest.mark.parametrize("name", ["wf_gru1d", "wf_gru1d_3layer", "wf_gru1d_raster2d"])
def test_1d_prnn(name):
    """1DTFIM/RNNwavefunction.py:35-118 (and 2DTFIM_1DRNN/RNNwavefunction.py, the same GRU over the raster index)."""
    z, P = _load(name)
    N = int(z["N"]) if "N" in z.files else int(z["Nx"]) * int(z["Ny"])
    units = [int(u) for u in z["units"]]
    r = owf.rnn1d(P, N, units, uniforms=torch.as_tensor(z["uniforms"]), numsamples=z["uniforms"].shape[0])
    assert r["margin"].min() > 1e-9
    assert (r["samples"].numpy() == z["samples"]).all()
    lp = owf.rnn1d(P, N, units, samples=_t(z["configs"]))["logp"]
    np.testing.assert_allclose(lp.numpy(), z["logp"], **TOL)


def test_1d_prnn_paritysym():
    """1DTFIM/RNNwavefunction_paritysym.py:80-145."""
    z, P = _load("wf_gru1d_paritysym")
    lp = owf.rnn1d_paritysym_logp(P, int(z["N"]), [int(u) for u in z["units"]], _t(z["configs"]))
    np.testing.assert_allclose(lp.numpy(), z["logp"], **TOL)


def test_1d_crnn_u1():
    """J1J2/ComplexRNNwavefunction.py:45-169: sampling in the Sz=0 sector and complex log-amplitudes."""
    z, P = _load("wf_cgru1d_u1")
    N, units = int(z["N"]), [int(u) for u in z["units"]]
    r = owf.rnn1d(P, N, units, uniforms=torch.as_tensor(z["uniforms"]), numsamples=z["uniforms"].shape[0], complex_=True, u1=True)
    assert (r["samples"].numpy() == z["samples"]).all()
    assert (z["samples"].sum(1) == N // 2).all()
    r = owf.rnn1d(P, N, units, samples=_t(z["configs"]), complex_=True, u1=True)
    la = torch.complex(0.5 * r["logp"], r["phase"]).numpy()
    # the reference sums complex logs of amplitude * e^{i phase}: its imaginary part is the phase sum with every term
    # wrapped to (-pi, pi] (|phase_n| < pi, so no term wraps)
    np.testing.assert_allclose(la, z["log_amplitude"], **TOL)


def test_2d_prnn_vanilla():
    """2DTFIM_2DRNN/MDRNNcell.py:51-66, RNNwavefunction.py:35-200."""
    z, P = _load("wf_vanilla2d")
    Nx, Ny, nh = int(z["Nx"]), int(z["Ny"]), int(z["nh"])
    r = owf.rnn2d(P, Nx, Ny, nh, "vanilla", uniforms=torch.as_tensor(z["uniforms"]), numsamples=z["uniforms"].shape[0])
    assert (r["samples"].numpy() == z["samples"]).all()
    lp = owf.rnn2d(P, Nx, Ny, nh, "vanilla", samples=_t(z["configs"]))["logp"]
    np.testing.assert_allclose(lp.numpy(), z["logp"], **TOL)


@pytest.mark.parametrize("name,cell", [("wf_tens2d_u1", "tens"), ("wf_tensgru2d_u1", "tensgru"), ("wf_tensgru2d_free", "tensgru"),
                                       ("wf_tensgru2d_sz1", "tensgru")])
def test_2d_crnn(name, cell):
    """2DHeisenberg/MDTensorized{,GRU}cell.py, MDComplexRNN.py:69-286 with mag_fixed on / off and Sz = 1."""
    z, P = _load(name)
    Nx, Ny, nh = int(z["Nx"]), int(z["Ny"]), int(z["nh"])
    kw = dict(complex_=True, mag_fixed=bool(z["mag_fixed"]), magnetization=int(z["magnetization"]))
    r = owf.rnn2d(P, Nx, Ny, nh, cell, uniforms=torch.as_tensor(z["uniforms"]), numsamples=z["uniforms"].shape[0], **kw)
    assert (r["samples"].numpy() == z["samples"]).all()
    if kw["mag_fixed"]:
        assert (z["samples"].reshape(len(z["samples"]), -1).sum(1) == (Nx * Ny + kw["magnetization"]) // 2).all()
    r = owf.rnn2d(P, Nx, Ny, nh, cell, samples=_t(z["configs"]), **kw)
    np.testing.assert_allclose(torch.complex(0.5 * r["logp"], r["phase"]).numpy(), z["log_amplitude"], **TOL)


def test_2d_crnn_symmetry_groups():
    """Every log_amplitude_*sym* variant of 2DHeisenberg/MDComplexRNN.py:290-577 the generator recorded."""
    z, P = _load("wf_tensgru2d_u1")
    Nx, Ny, nh = int(z["Nx"]), int(z["Ny"]), int(z["nh"])
    cfg = _t(z["configs"])
    keys = [k for k in z.files if k.startswith("sym__")]
    assert len(keys) >= 20
    for k in keys:
        parts = k[5:].split("_")
        group = parts[0] if parts[0] != "c2v" else "c2vsym"
        if "spinparity" in parts:
            ch, sp = parts[-2], {"p1": 1, "m1": -1}[parts[-1]]
        else:
            ch, sp = parts[-1], None
        # as shipped: chi stays a Python list in all spin-parity variants but c4vsym, so spinparity_value = -1 drops the
        # flipped copies from the phase sum (oracle docstring); the fixture is the reference as shipped
        lst, signs = owf.heis_symmetry_copies(cfg, group, ch, sp, legacy_list_bug=True)
        r = owf.rnn2d(P, Nx, Ny, nh, "tensgru", samples=torch.cat(lst), complex_=True, mag_fixed=True)
        lp_g, ph_g = r["logp"].view(len(lst), -1), r["phase"].view(len(lst), -1)
        la = owf.log_amplitudes_from_symmetry_group(lp_g, ph_g, signs).numpy()
        # where the character-weighted sum of e^{i phi_g} cancels exactly (a configuration that maps onto itself under
        # an element of character -1) its real part is rounding noise and sign(Re) -- hence the phase -- is arbitrary
        ch_ = torch.as_tensor(signs, dtype=torch.float64).reshape(-1, 1)
        ok = ((torch.cos(ph_g[:len(signs)]) / ch_).sum(0).abs() > 1e-9).numpy()
        assert ok.sum() >= len(ok) // 2, k
        np.testing.assert_allclose(la.real, z[k].real, err_msg=k, **TOL)
        np.testing.assert_allclose(la.imag[ok], z[k].imag[ok], err_msg=k, **TOL)
        # intended semantics (what the product implements): identical wherever the defect does not bite
        lst2, signs2 = owf.heis_symmetry_copies(cfg, group, ch, sp)
        if len(signs2) == len(signs):
            assert signs2 == signs
        else:
            assert sp == -1 and group != "c4vsym" and len(signs2) == 2 * len(signs)
            la2 = owf.log_amplitudes_from_symmetry_group(lp_g, ph_g, signs2)
            np.testing.assert_allclose(la2.real.numpy(), z[k].real, **TOL)


@pytest.mark.parametrize("name", ["wf_rydberg_2x2", "wf_rydberg_3x3", "wf_rydberg_4x4"])
def test_rydberg(name):
    """RydbergKagome_2DGRU/MDGRUcell.py:48-59, MDRNNWavefunction.py:12-34,73-291 (periodic wrap on 2, 3 and 4 cells)."""
    z, P = _load(name)
    Lx, Ly, nh = int(z["Lx"]), int(z["Ly"]), int(z["nh"])
    r = owf.rydberg2d(P, Lx, Ly, nh, uniforms=torch.as_tensor(z["uniforms"]), numsamples=z["uniforms"].shape[0])
    assert (r["samples"].numpy() == z["samples"]).all()
    cfg = _t(z["configs"])
    np.testing.assert_allclose(owf.rydberg2d(P, Lx, Ly, nh, samples=cfg)["logp"].numpy(), z["logp"], **TOL)
    assert (owf.rotate60(cfg).numpy() == z["rotate60"]).all()
    assert ((cfg[..., 0] + 2 * cfg[..., 1] + 4 * cfg[..., 2]).numpy() == z["binary_to_intarray"]).all()
    c = torch.arange(8)
    assert (torch.stack([c % 2, (c // 2) % 2, (c // 4) % 2], dim=1).numpy() == z["intarray_to_binary"]).all()
    for sym in ("c6sym", "translationsym"):
        lst = owf.rydberg_symmetry_copies(cfg, sym)
        lp = owf.rydberg2d(P, Lx, Ly, nh, samples=torch.cat(lst))["logp"].view(len(lst), -1)
        np.testing.assert_allclose(owf.log_probability_symmetry(lp).numpy(), z["sym__" + sym], err_msg=sym, **TOL)
