"""Weight interchange with the reference's TensorFlow checkpoints.

The engine keeps its parameters as one flat float32 vector whose tensors already have the TF layouts (gate kernel
`[d_in + nh, 2 nh]`, MD cells `W[nh, 2 nh, 4]`, ...; `rnnwf_param_info`).  This module only maps the engine's tensor
names to the variable names `tf.train.Saver` writes ("Expected variable names"), so a dictionary
`{tf_variable_name: ndarray}` -- e.g. produced on a machine that has TensorFlow with
`{v.name[:-2]: sess.run(v) for v in tf.global_variables()}` -- can be loaded, and our weights exported the same way.
Reference anchors: 1DTFIM/TrainingRNN_1DTFIM.py:146,220-228; 2DHeisenberg/Run_RNNWF_2DHeis.py:119-125,250-268,342-346.
"""
import re

import numpy as np
import torch

_GRU = {"Wg": "gates/kernel", "bg": "gates/bias", "Wci": "candidate/input_projection/kernel",
        "bci": "candidate/input_projection/bias", "Wch": "candidate/hidden_projection/kernel",
        "bch": "candidate/hidden_projection/bias"}


def tf_variable_name(name, complex_head, scope="RNNwavefunction", gru1d=True):
    """Engine tensor name -> TF variable name (without the ':0' suffix).  The amplitude head is 'wf_dense_ampl' only in
    the 1D complex RNN (J1J2/ComplexRNNwavefunction.py:42); every other model -- the complex 2D Heisenberg one included
    (2DHeisenberg/MDComplexRNN.py:63) -- calls it 'wf_dense'."""
    m = re.fullmatch(r"gru(\d+)\.(\w+)", name)
    if m:
        return f"{scope}/multi_rnn_cell/cell_{m.group(1)}/cudnn_compatible_gru_cell/{_GRU[m.group(2)]}"
    if name.startswith("dense"):
        head, leaf = name.split(".")
        leaf = "kernel" if leaf == "W" else "bias"
        if head == "dense_phase":
            return f"{scope}/wf_dense_phase/{leaf}"
        return f"{scope}/{'wf_dense_ampl' if (complex_head and gru1d) else 'wf_dense'}/{leaf}"
    return f"{scope}/{name}_rnn_0"          # MD cells: Wh, Uh, Wv, Uv, b / W, b, Wg, bg, Wmerge


def state_dict(engine, scope="RNNwavefunction"):
    """{tf_variable_name: float32 ndarray in TF layout} of the engine's current parameters."""
    views = engine.views()
    return {tf_variable_name(n, engine.complex_head, scope, engine.cell == 0): v.detach().cpu().numpy().copy() for n, v in views.items()}


def load_state_dict(engine, sd, scope="RNNwavefunction", strict=True):
    """Load `{tf_variable_name: array}` (names with or without ':0').  Shapes must match the TF layouts; the 2D
    cells' parameters do not depend on the lattice size, so a checkpoint of a smaller lattice warm-starts a larger
    one as in Run_RNNWF_2DHeis.py:119-125."""
    sd = {k[:-2] if k.endswith(":0") else k: v for k, v in sd.items()}
    flat = engine.params.detach().cpu().clone()
    views = {n: flat[o:o + int(np.prod(s))].view(*s) for n, o, s in engine.tensors}
    missing = []
    for n, v in views.items():
        key = tf_variable_name(n, engine.complex_head, scope, engine.cell == 0)
        if key not in sd:
            missing.append(key)
            continue
        a = torch.as_tensor(np.asarray(sd[key]), dtype=torch.float32)
        if tuple(a.shape) != tuple(v.shape):
            raise ValueError(f"{key}: checkpoint shape {tuple(a.shape)} != expected {tuple(v.shape)}")
        v.copy_(a)
    if missing and strict:
        raise KeyError(f"variables missing from the checkpoint: {missing}")
    engine.set_weights(flat)
    return missing


def save_npz(engine, path, scope="RNNwavefunction", **extra):
    """Portable checkpoint: TF-named tensors plus optional histories (meanEnergy=..., varEnergy=...)."""
    np.savez(path, **{k.replace("/", "__"): v for k, v in state_dict(engine, scope).items()},
             **{"extra__" + k: np.asarray(v) for k, v in extra.items()})


def load_npz(engine, path, scope="RNNwavefunction", strict=True):
    z = np.load(path)
    sd = {k.replace("__", "/"): z[k] for k in z.files if not k.startswith("extra__")}
    load_state_dict(engine, sd, scope, strict)
    return {k[len("extra__"):]: z[k] for k in z.files if k.startswith("extra__")}
