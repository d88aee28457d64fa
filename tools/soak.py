"""Soak run: many VMC steps of the headline configuration (and of C1 / C5-4x4 / C3) back to back; every step's energy must be
finite and the run must not hang (run it under `timeout`).   python tools/soak.py [steps] [configuration 0..3]"""
import os, sys, time, math, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib, vmc
from rnnwavefunctions_b200.engine import Engine
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
only = slice(int(sys.argv[2]), int(sys.argv[2]) + 1) if len(sys.argv) > 2 else slice(None)      # optional: one configuration (0..3)
for name, eng, ham, hp, ns, n in (
        ("C4 tc", Engine(_lib.CELL_TENSGRU2D, 10, 10, num_units=100, complex_head=True, u1=True, precision="tc"), _lib.HAM_HEIS2D, [1.0], 10000, steps),
        ("C1", Engine(_lib.CELL_GRU1D, 20, 1, num_units=50), _lib.HAM_TFIM1D, [1.0] + [1.0] * 19, 500, 20 * steps),
        ("C5 4x4 tc", Engine(_lib.CELL_GRU4, 4, 4, num_units=256, precision="tc"), _lib.HAM_RYDBERG, [1.7 ** 6, 3.3, 1.0, 2.001], 12500, steps),
        ("C3 tc3", Engine(_lib.CELL_VANILLA2D, 12, 12, num_units=50, precision="tc3"), _lib.HAM_TFIM2D, [3.0] + [1.0] * 144, 10000, steps))[only]:
    V = vmc.VMC(eng, ham, hp, ns, lr=1e-3, seed=7)
    t0 = time.time()
    es = []
    for it in range(n):
        m, var = V.step()
        if it % max(1, n // 20) == 0 or it == n - 1:
            es.append(complex(m))
            assert math.isfinite(es[-1].real) and math.isfinite(float(var)), (name, it, es[-1], var)
    torch.cuda.synchronize()
    print(f"{name}: {n} steps in {time.time() - t0:.1f} s, <E> {es[0].real:.4f} -> {es[-1].real:.4f}", flush=True)
