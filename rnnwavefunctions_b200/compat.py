"""Shims for the TF graph-mode tail of the reference's signatures.

The reference's local-energy functions end in ``(log_probs_tensor, samples_placeholder, log_probs, sess)``:
symbolic handles evaluated later with ``sess.run(handle, feed_dict={placeholder: ndarray})``
(1DTFIM/TrainingRNN_1DTFIM.py:13,65).  Here a wavefunction method called on a :class:`Placeholder`
returns a :class:`Handle` bound to the wavefunction; :class:`Session.run` evaluates it eagerly on the GPU,
and the local-energy functions recognise a Handle and take the fused device path instead of the
reference's chunked host loop.
"""
import numpy as np
import torch


class Placeholder:
    def __init__(self, dtype=None, shape=None, name=None):
        self.dtype, self.shape, self.name = dtype, shape, name


def placeholder(dtype=None, shape=None, name=None):
    return Placeholder(dtype, shape, name)


class Handle:
    """A deferred wavefunction call: fn(samples) -> tensor."""

    def __init__(self, wf, fn, placeholder=None, kind="log", symmetry=None):
        self.wf, self.fn, self.placeholder, self.kind = wf, fn, placeholder, kind
        self.symmetry = symmetry      # engine.SymmetryGroup when the call is a symmetry-projected amplitude, else None

    def __call__(self, samples):
        return self.fn(samples)


class Session:
    """``sess.run(handle, feed_dict)`` -> NumPy array, like the reference's ``tf.Session``."""

    def run(self, fetch, feed_dict=None):
        if isinstance(fetch, (list, tuple)):
            return [self.run(f, feed_dict) for f in fetch]
        if isinstance(fetch, Handle):
            if fetch.kind == "sample":
                return fetch.fn(None).cpu().numpy()
            feed = None
            if feed_dict:
                feed = feed_dict.get(fetch.placeholder) if fetch.placeholder in feed_dict else next(iter(feed_dict.values()))
            return fetch.fn(torch.as_tensor(np.asarray(feed))).cpu().numpy()
        if torch.is_tensor(fetch):
            return fetch.cpu().numpy()
        return fetch


def find_wf(*candidates):
    for c in candidates:
        if isinstance(c, Handle):
            return c.wf
        if hasattr(c, "engine"):
            return c
    raise TypeError("pass the wavefunction (or a handle returned by one of its methods) where the reference "
                    "takes log_probs_tensor / log_amps / sess")


def find_symmetry(*candidates):
    """The SymmetryGroup behind a handle of a symmetric wavefunction method (None for the plain amplitudes).  The
    reference hands its local-energy functions the tensor of whichever method was chosen
    (Run_RNNWF_2DHeis.py:153-197, Run_RNNWF_Rydberg.py:155); the handle carries the same choice here."""
    for c in candidates:
        if isinstance(c, Handle):
            return c.symmetry
    return None


def to_numpy(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)
