"""Shared helpers of the parity tests: drive the oracle with the engine's own parameters."""
import numpy as np
import torch

from oracle import wavefunctions as owf
from rnnwavefunctions_b200 import _lib

CELL_NAME = {_lib.CELL_VANILLA2D: "vanilla", _lib.CELL_TENS2D: "tens", _lib.CELL_TENSGRU2D: "tensgru"}


def oracle_params(eng, requires_grad=False):
    P = {k: v.detach().double().cpu().clone() for k, v in eng.views().items()}
    if requires_grad:
        for v in P.values():
            v.requires_grad_(True)
    return P


def oracle_run(eng, P, samples=None, uniforms=None, numsamples=None):
    """Oracle twin of eng.sample / eng.logpsi -> dict(samples, logp, phase[, margin])."""
    if samples is not None:
        samples = torch.as_tensor(np.asarray(samples)).long()
    if eng.cell == _lib.CELL_GRU1D:
        return owf.rnn1d(P, eng.nx * eng.ny, [eng.nh] * eng.num_layers, samples=samples, uniforms=uniforms,
                         numsamples=numsamples, complex_=eng.complex_head, u1=eng.u1)
    if eng.cell == _lib.CELL_GRU4:
        return owf.rydberg2d(P, eng.nx, eng.ny, eng.nh, samples=samples, uniforms=uniforms, numsamples=numsamples)
    return owf.rnn2d(P, eng.nx, eng.ny, eng.nh, CELL_NAME[eng.cell], samples=samples, uniforms=uniforms,
                     numsamples=numsamples, complex_=eng.complex_head, mag_fixed=eng.u1,
                     magnetization=eng.cfg.magnetization)


def oracle_logpsi_fn(eng, P):
    def f(cfg):
        with torch.no_grad():
            r = oracle_run(eng, P, samples=cfg)
        return r["logp"].numpy(), r["phase"].numpy()
    return f


def philox_uniforms(seed, rows, nsteps, row_offset=0):
    """NumPy twin of csrc/common.cuh philox_uniform (Philox4x32-10, counter (row_lo,row_hi,t,0), word 0)."""
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    r = (np.arange(rows, dtype=np.uint64) + np.uint64(row_offset))[:, None] * np.ones((1, nsteps), dtype=np.uint64)
    c0 = (r & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    c1 = (r >> np.uint64(32)).astype(np.uint64)
    c2 = (np.arange(nsteps, dtype=np.uint64)[None, :] * np.ones((rows, 1), dtype=np.uint64))
    c3 = np.zeros_like(c0)
    k0, k1 = np.uint64(seed & 0xFFFFFFFF), np.uint64((seed >> 32) & 0xFFFFFFFF)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask, lo1, (hi0 ^ c3 ^ k1) & mask, lo0
        k0, k1 = (k0 + np.uint64(W0)) & mask, (k1 + np.uint64(W1)) & mask
    return ((c0 >> np.uint64(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)


def rel_err(x, ref):
    """The one bound the parity tests use for log-amplitudes and local energies (north_star: 1e-5 relative in the FP32
    regime, 1e-3 in the reduced-precision regime): per-sample relative error with a floor at the batch mean of |ref|,
        err_b = |x_b - ref_b| / max(|ref_b|, <|ref|>).
    Returns (max, rms) and prints them so the GPU log records the measured numbers."""
    x, ref = np.asarray(x), np.asarray(ref)
    floor = np.abs(ref).mean()
    err = np.abs(x - ref) / np.maximum(np.abs(ref), floor)
    return float(err.max()), float(np.sqrt((err ** 2).mean()))


def assert_rel(x, ref, tol, what=""):
    mx, rms = rel_err(x, ref)
    print(f"[parity] {what}: max rel err {mx:.3e}, rms {rms:.3e} (bound {tol:g}, n={np.size(ref)})")
    assert mx <= tol, f"{what}: max relative error {mx:.3e} > {tol:g} (rms {rms:.3e})"
