"""Accuracy of the tensor-core regime against the FP32 FFMA regime on the same weights and samples.

    python tools/tc_accuracy.py [--nx 10 --ny 10 --nh 100 --ns 2000 --scale 1.0]

Reports the differences in log-probability / phase of the samples and in the local energies (the quantities the
reduced-precision tolerance of 1e-3 applies to).  --scale multiplies the initial weights (trained networks have
larger weights than the Glorot initialisation).
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib  # noqa: E402
from rnnwavefunctions_b200.engine import Engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=10)
    ap.add_argument("--ny", type=int, default=10)
    ap.add_argument("--nh", type=int, default=100)
    ap.add_argument("--ns", type=int, default=2000)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--j2", type=float, default=1.0)
    args = ap.parse_args()
    kw = dict(cell=_lib.CELL_TENSGRU2D, nx=args.nx, ny=args.ny, num_units=args.nh, complex_head=True, u1=True, device="cuda:0")
    ref = Engine(precision="fp32", **kw)
    tc = Engine(precision="tc", **kw)
    ref.init_params(seed=5)
    ref.set_weights(ref.params * args.scale)
    tc.set_weights(ref.params)
    cfg, lp, ph = ref.sample(args.ns, seed=11)
    lp2, ph2 = tc.logpsi(cfg)
    e1 = ref.local_energies(_lib.HAM_HEIS2D, [args.j2], cfg, lp, ph)
    e2 = tc.local_energies(_lib.HAM_HEIS2D, [args.j2], cfg, lp2, ph2)
    scale = e1.abs().mean().item()
    print(f"max |dlogP| = {(lp - lp2).abs().max().item():.3e}   max |dphase| = {(ph - ph2).abs().max().item():.3e}")
    print(f"E_loc: mean |E| = {scale:.4f}   max |dE|/mean|E| = {(e1 - e2).abs().max().item() / scale:.3e}   "
          f"rms |dE|/mean|E| = {(e1 - e2).abs().pow(2).mean().sqrt().item() / scale:.3e}")
    print(f"<E>/N: fp32 {e1.real.mean().item() / (args.nx * args.ny):.6f}   tc {e2.real.mean().item() / (args.nx * args.ny):.6f}")


if __name__ == "__main__":
    main()
