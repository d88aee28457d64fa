"""Learning-rate / temperature schedules and gradient transforms of the reference's training scripts.

Pure functions of the iteration counter, so they can be applied on every rank identically (the optimiser runs
redundantly on all ranks after the gradient all-reduce).  Each function cites the reference lines it restates.
"""
import torch


def lr_tfim2d_1drnn(it, lr):
    """2DTFIM_1DRNN/Training1DRNN_2DTFIM.py:229 -- lr_decayed = 1 / (1/lr + it/10)."""
    return 1.0 / ((1.0 / lr) + (it / 10.0))


def lr_tfim2d_2drnn(it, lr):
    """2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:228 -- lr * (1 + it/5000)^-1."""
    return lr * (1.0 + it / 5000.0) ** (-1)


def lr_heisenberg(it, lr, lrthreshold, lrdecaytime, lrthreshold_conv, lrdecaytime_conv, Nwarmup, Nannealing, Ntrain):
    """2DHeisenberg/Run_RNNWF_2DHeis.py:294-297 -- max(thr, lr/(1+it/tau)) until the end of annealing, then the
    convergence phase thr_conv / (1 + (it - end)/tau_conv).  `it` is the 0-based step index of the reference loop."""
    end = Nwarmup + Nannealing * Ntrain
    if it + 1 <= end:
        return max(lrthreshold, lr / (1.0 + it / lrdecaytime))
    return lrthreshold_conv / (1.0 + (it - end) / lrdecaytime_conv)


def temperature_heisenberg(it, T0, Nwarmup, Nannealing, Ntrain):
    """2DHeisenberg/Run_RNNWF_2DHeis.py:308-318 -- T0 during warm-up, piecewise-constant linear annealing
    T0 (1 - floor((it+1-Nwarmup)/Ntrain)/Nannealing), then 0."""
    if T0 == 0:
        return 0.0
    if it + 1 <= Nwarmup:
        return T0
    if it + 1 <= Nwarmup + Nannealing * Ntrain:
        return T0 * (1.0 - ((it + 1 - Nwarmup) // Ntrain) / Nannealing)
    return 0.0


def lr_rydberg(it, lr, lrthreshold, numwarmup, Nannealing, Neq):
    """RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:247-250 -- lr until the end of annealing, lrthreshold afterwards
    (`it` is the 1-based step index of the reference loop)."""
    return lrthreshold if it >= numwarmup + Nannealing * Neq else lr


class RydbergTemperature:
    """RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:240,252-258 -- stateful: T only changes on steps with it % Neq == 0
    inside the annealing window and drops to 0 strictly after it."""

    def __init__(self, T0, numwarmup, Nannealing, Neq):
        self.T0, self.numwarmup, self.Nannealing, self.Neq = T0, numwarmup, Nannealing, Neq
        self.T = T0

    def __call__(self, it):
        end = self.numwarmup + self.Nannealing * self.Neq
        if self.numwarmup <= it < end:
            if it % self.Neq == 0:
                self.T = self.T0 * (1.0 - (it - self.numwarmup) / (self.Nannealing * self.Neq))
        elif it > end:
            self.T = 0.0
        return self.T


def clip_by_value_(grad, clipvalue):
    """tf.clip_by_value(g, -c, c) (Run_RNNWF_2DHeis.py:204)."""
    return grad.clamp_(-clipvalue, clipvalue)


def sign_sgd_step_(params, grad, lr, seed, t):
    """The `normalized_gradients` optimiser of Run_RNNWF_2DHeis.py:206-213 as one fused device kernel
    (rnnwf_sign_sgd_step): params -= lr * sign(grad) * u, u the library's counter-based uniforms at (seed, element, t) --
    every rank applies the same update from the same (seed, t).  Device tensors, float32; call Engine.set_weights after."""
    import ctypes as C
    from ._lib import lib, check
    g = grad if (grad.dtype == torch.float32 and grad.is_contiguous()) else grad.float().contiguous()
    check(lib.rnnwf_sign_sgd_step(C.c_void_p(params.data_ptr()), C.c_void_p(g.data_ptr()), params.numel(), float(lr), int(seed), int(t),
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)), "rnnwf_sign_sgd_step")
    return params


def normalized_gradient(grad, generator=None):
    """sign(g) * U(0,1), used with plain SGD (Run_RNNWF_2DHeis.py:207-213).  The uniforms come from `generator`
    (seed it identically on every rank: the update must be the same everywhere)."""
    u = torch.rand(grad.shape, dtype=torch.float64, device=grad.device, generator=generator)
    return torch.sign(grad) * u.to(grad.dtype)
