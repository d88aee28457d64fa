"""Generate the wavefunction fixtures ``wf_*.npz`` by EXECUTING the unmodified reference wavefunction classes.

Runs only inside the build container (needs /root/reference).  TensorFlow is absent there; the classes are imported
under ``tf_shim`` (an eager NumPy stand-in, see its docstring for exactly which three TF built-ins remain
restatements).  For every model family the reference's own ``.sample`` (driven by recorded uniforms),
``.log_probability`` / ``.log_amplitude*`` and the symmetry-projected variants are run on seeded weights; weights
(under their TF variable names), inputs and outputs are committed, and ``tests/test_oracle_golden.py`` holds
``oracle/wavefunctions.py`` to them at 1e-12.

    python tests/golden/make_wf_golden.py      # rewrites tests/golden/wf_*.npz
"""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import tf_shim  # noqa: E402

REF = "/root/reference"
_LOCAL_MODULES = ("RNNwavefunction", "ComplexRNNwavefunction", "MDRNNcell", "MDTensorizedcell", "MDTensorizedGRUcell",
                  "Tensordot2", "MDComplexRNN", "MDGRUcell", "MDRNNWavefunction", "RNNwavefunction_paritysym")


def load(subdir, fname):
    """Import one reference file by path; sibling imports (``from Tensordot2 import tensordot``) resolve inside its own
    directory, and the generic module names are dropped again so the next package gets its own copies."""
    tf_shim.install()
    d = os.path.join(REF, subdir)
    sys.path.insert(0, d)
    try:
        spec = importlib.util.spec_from_file_location("ref_" + subdir + "_" + fname[:-3], os.path.join(d, fname))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(d)
        for k in _LOCAL_MODULES:
            sys.modules.pop(k, None)
    return mod


def variables():
    return {"var__" + k.replace("/", "__"): v.copy() for k, v in tf_shim.VARIABLES.items()}


def perturb_biases(rng):
    """Zero-initialised biases (and the all-ones gate bias) would hide a swapped or dropped bias: jitter them, keeping
    the values float32-representable."""
    for k, v in tf_shim.VARIABLES.items():
        if k.endswith("bias"):
            v += rng.uniform(-0.3, 0.3, v.shape).astype(np.float32)


def save(name, **arrays):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
    print(f"{name}.npz: " + ", ".join(f"{k}{tuple(np.shape(v))}" for k, v in arrays.items() if not k.startswith("var__")))


def u1_configs(rng, ns, shape):
    n = int(np.prod(shape))
    out = np.zeros((ns, n), dtype=np.int32)
    for b in range(ns):
        out[b, rng.permutation(n)[: n // 2]] = 1
    return out.reshape((ns,) + tuple(shape))


def main():
    rng = np.random.RandomState(4321)
    tf = tf_shim.install()

    # ---- W1-W3: 1D pRNN, 1 and 3 layers (1DTFIM/RNNwavefunction.py) ---------------------------------------------------
    for tag, units in (("gru1d", [7]), ("gru1d_3layer", [6, 6, 6])):
        m = load("1DTFIM", "RNNwavefunction.py")
        tf_shim.reset(11)
        N, ns = 9, 12
        wf = m.RNNwavefunction(N, cell=tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell, units=units, seed=111)
        wf.log_probability(np.zeros((1, N), dtype=np.int32), 2)     # creates the variables
        perturb_biases(rng)
        u = rng.uniform(size=(ns, N))
        tf_shim.set_uniforms(u)
        samples = np.asarray(wf.sample(ns, 2))
        tf_shim.set_uniforms(None)
        extra = rng.randint(0, 2, (5, N))
        cfg = np.concatenate([samples, extra]).astype(np.int32)
        logp = np.asarray(wf.log_probability(cfg, 2))
        save("wf_" + tag, N=N, units=np.asarray(units), uniforms=u, samples=samples, configs=cfg, logp=logp, **variables())

    # ---- W4: parity-symmetrised 1D pRNN (1DTFIM/RNNwavefunction_paritysym.py) ------------------------------------------
    m = load("1DTFIM", "RNNwavefunction_paritysym.py")
    tf_shim.reset(12)
    N, ns = 8, 10
    wf = m.RNNwavefunction(N, cell=tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell, units=[6], seed=111)
    cfg = rng.randint(0, 2, (ns, N)).astype(np.int32)
    wf.log_probability(cfg[:1], 2)
    perturb_biases(rng)
    save("wf_gru1d_paritysym", N=N, units=np.asarray([6]), configs=cfg, logp=np.asarray(wf.log_probability(cfg, 2)), **variables())

    # ---- W5-W7: 1D cRNN with the U(1) mask (J1J2/ComplexRNNwavefunction.py) -------------------------------------------
    m = load("J1J2", "ComplexRNNwavefunction.py")
    tf_shim.reset(13)
    N, ns = 10, 14
    wf = m.RNNwavefunction(N, cell=tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell, units=[8], seed=111)
    wf.log_amplitude(u1_configs(rng, 1, (N,)), 2)
    perturb_biases(rng)
    u = rng.uniform(size=(ns, N))
    tf_shim.set_uniforms(u)
    samples = np.asarray(wf.sample(ns, 2))
    tf_shim.set_uniforms(None)
    cfg = np.concatenate([samples, u1_configs(rng, 6, (N,))]).astype(np.int32)
    la = np.asarray(wf.log_amplitude(cfg, 2))
    save("wf_cgru1d_u1", N=N, units=np.asarray([8]), uniforms=u, samples=samples, configs=cfg, log_amplitude=la, **variables())

    # ---- 2DTFIM_1DRNN: the same GRU raster-scanned over Nx x Ny (float64 upstream) -------------------------------------
    m = load("2DTFIM_1DRNN", "RNNwavefunction.py")
    tf_shim.reset(14)
    Nx, Ny, ns = 3, 3, 8
    wf = m.RNNwavefunction(Nx, Ny, cell=tf.contrib.cudnn_rnn.CudnnCompatibleGRUCell, units=[5], seed=111)
    wf.log_probability(np.zeros((1, Nx * Ny), dtype=np.int32), 2)
    perturb_biases(rng)
    u = rng.uniform(size=(ns, Nx * Ny))
    tf_shim.set_uniforms(u)
    samples = np.asarray(wf.sample(ns, 2))
    tf_shim.set_uniforms(None)
    cfg = np.concatenate([samples, rng.randint(0, 2, (4, Nx * Ny))]).astype(np.int32)
    save("wf_gru1d_raster2d", Nx=Nx, Ny=Ny, units=np.asarray([5]), uniforms=u, samples=samples, configs=cfg,
         logp=np.asarray(wf.log_probability(cfg, 2)), **variables())

    # ---- W8-W9: 2D pRNN, vanilla MD cell (2DTFIM_2DRNN) ----------------------------------------------------------------
    cellmod = load("2DTFIM_2DRNN", "MDRNNcell.py")
    m = load("2DTFIM_2DRNN", "RNNwavefunction.py")
    tf_shim.reset(15)
    L, nh, ns = 4, 6, 10
    wf = m.RNNwavefunction(L, L, cell=cellmod.MDRNNcell, units=[nh], seed=111)
    u = rng.uniform(size=(ns, L * L))
    tf_shim.set_uniforms(u)
    samples = np.asarray(wf.sample(ns, 2))
    tf_shim.set_uniforms(None)
    cfg = np.concatenate([samples, rng.randint(0, 2, (4, L, L))]).astype(np.int32)
    # wf_dense is created by .sample; jitter its bias and evaluate (samples are re-derived in the test from the final weights)
    perturb_biases(rng)
    tf_shim.set_uniforms(u)
    samples = np.asarray(wf.sample(ns, 2))
    tf_shim.set_uniforms(None)
    cfg[:ns] = samples
    save("wf_vanilla2d", Nx=L, Ny=L, nh=nh, uniforms=u, samples=samples, configs=cfg,
         logp=np.asarray(wf.log_probability(cfg, 2)), **variables())

    # ---- W10-W15: 2D complex RNN, tensorized and tensorized-GRU cells, with and without U(1) (2DHeisenberg) ------------
    heis = load("2DHeisenberg", "MDComplexRNN.py")
    cells = {"tens": load("2DHeisenberg", "MDTensorizedcell.py").MDRNNcell,
             "tensgru": load("2DHeisenberg", "MDTensorizedGRUcell.py").MDRNNcell}
    for tag, cellname, mag_fixed, mag in (("tens2d_u1", "tens", True, 0), ("tensgru2d_u1", "tensgru", True, 0),
                                          ("tensgru2d_free", "tensgru", False, 0), ("tensgru2d_sz1", "tensgru", True, 2)):
        tf_shim.reset(16)
        L, nh, ns = 4, 5, 10
        wf = heis.RNNwavefunction(L, L, cell=cells[cellname], units=[nh], seed=111, mag_fixed=mag_fixed, magnetization=mag)
        wf.log_amplitude_nosym(np.zeros((1, L, L), dtype=np.int32), 2)
        perturb_biases(rng)
        u = rng.uniform(size=(ns, L * L))
        tf_shim.set_uniforms(u)
        samples = np.asarray(wf.sample(ns, 2))
        tf_shim.set_uniforms(None)
        cfg = samples.astype(np.int32)
        if not mag_fixed:
            cfg = np.concatenate([cfg, rng.randint(0, 2, (4, L, L))]).astype(np.int32)
        la = np.asarray(wf.log_amplitude_nosym(cfg, 2))
        out = dict(Nx=L, Ny=L, nh=nh, mag_fixed=int(mag_fixed), magnetization=mag, uniforms=u, samples=samples, configs=cfg,
                   log_amplitude=la)
        if tag == "tensgru2d_u1":
            # every symmetry-projected variant of the class (MDComplexRNN.py:320-577), both characters where there are two
            sym = {
                "rotsym_A": lambda: wf.log_amplitude_rotsym(cfg, 2, "A"),
                "rotsym_B": lambda: wf.log_amplitude_rotsym(cfg, 2, "B"),
                "rotsym_spinparity_A_p1": lambda: wf.log_amplitude_rotsym_spinparity(cfg, 2, "A", 1),
                "rotsym_spinparity_B_m1": lambda: wf.log_amplitude_rotsym_spinparity(cfg, 2, "B", -1),
                "rotreducedsym_A": lambda: wf.log_amplitude_rotreducedsym(cfg, 2, "A"),
                "rotreducedsym_B": lambda: wf.log_amplitude_rotreducedsym(cfg, 2, "B"),
                "rotreducedsym_spinparity_A_m1": lambda: wf.log_amplitude_rotreducedsym_spinparity(cfg, 2, "A", -1),
                "c2vsym_A1": lambda: wf.log_amplitude_c2vsym(cfg, 2, "A1"),
                "c2vsym_B2": lambda: wf.log_amplitude_c2vsym(cfg, 2, "B2"),
                "c2v_spinparity_A2_m1": lambda: wf.log_amplitude_c2v_spinparity(cfg, 2, "A2", -1),
                "c2sym_A": lambda: wf.log_amplitude_c2sym(cfg, 2, "A"),
                "c2sym_B": lambda: wf.log_amplitude_c2sym(cfg, 2, "B"),
                "c2sym_spinparity_A_p1": lambda: wf.log_amplitude_c2sym_spinparity(cfg, 2, "A", 1),
                "c2dsym_A1": lambda: wf.log_amplitude_c2dsym(cfg, 2, "A1"),
                "c2dsym_B1": lambda: wf.log_amplitude_c2dsym(cfg, 2, "B1"),
                "c2dsym_spinparity_A1_m1": lambda: wf.log_amplitude_c2dsym_spinparity(cfg, 2, "A1", -1),
                "c4vsym_A1": lambda: wf.log_amplitude_c4vsym(cfg, 2, "A1"),
                "c4vsym_B2": lambda: wf.log_amplitude_c4vsym(cfg, 2, "B2"),
                "c4vsym_spinparity_A1_p1": lambda: wf.log_amplitude_c4vsym_spinparity(cfg, 2, "A1", 1),
                "c4vsym_spinparity_B1_m1": lambda: wf.log_amplitude_c4vsym_spinparity(cfg, 2, "B1", -1),
            }
            for k, f in sym.items():
                try:
                    out["sym__" + k] = np.asarray(f())
                except Exception as e:      # a variant the reference cannot run as shipped is recorded, not hidden
                    print(f"  {k}: reference raised {type(e).__name__}: {e}")
        save("wf_" + tag, **out, **variables())

    # ---- W16-W19: Rydberg 2D GRU over Kagome unit cells (RydbergKagome_2DGRU) -------------------------------------------
    ryd = load("RydbergKagome_2DGRU", "MDRNNWavefunction.py")
    rcell = load("RydbergKagome_2DGRU", "MDGRUcell.py").MDRNNcell
    for tag, Lx, Ly in (("rydberg_3x3", 3, 3), ("rydberg_2x2", 2, 2), ("rydberg_4x4", 4, 4)):
        tf_shim.reset(17)
        nh, ns = 6, 9
        wf = ryd.MDRNNWavefunction(Lx, Ly, cell=rcell, units=[nh], inputdim=2, seed=111)
        wf.log_probability(np.zeros((1, Lx, Ly, 3), dtype=np.int64), 2)
        perturb_biases(rng)
        u = rng.uniform(size=(ns, Lx * Ly))
        tf_shim.set_uniforms(u)
        samples = np.asarray(wf.sample(ns, 2))
        tf_shim.set_uniforms(None)
        cfg = np.concatenate([samples, rng.randint(0, 2, (4, Lx, Ly, 3))]).astype(np.int64)
        out = dict(Lx=Lx, Ly=Ly, nh=nh, uniforms=u, samples=samples, configs=cfg, logp=np.asarray(wf.log_probability(cfg, 2)),
                   rotate60=np.asarray(ryd.rotate60(tf_shim._t(cfg))),
                   binary_to_intarray=np.asarray(ryd.binary_to_intarray(tf_shim._t(cfg))),
                   intarray_to_binary=np.asarray(ryd.intarray_to_binary(tf_shim._t(np.arange(8, dtype=np.int64)))))
        for symname in ("c6sym", "translationsym"):
            out["sym__" + symname] = np.asarray(wf.log_probability_symmetry(tf_shim._t(cfg), 2, symname))
        save("wf_" + tag, **out, **variables())


if __name__ == "__main__":
    main()
