"""Randomised cross-check of the regimes on odd shapes: for random (cell, lattice, width, batch) the tensor-core regimes
(tc, tc3) against the FFMA regime on the same weights -- sampling with given uniforms, teacher-forced log psi, E_loc with
prefix reuse (and, for the vanilla cell, without it), gradient.  Prints one line per case; exits non-zero at the end if
any case violated its bound.   python tools/fuzz_regimes.py [cases] [seed] [cell: tensgru | tens | vanilla | gru4]
(with `vanilla` the widths are drawn from 5..64, the range of the two-tile sweep of sweep_van2.cuh)"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine

ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 24
rng = np.random.RandomState(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
CELLS = [("tensgru", _lib.CELL_TENSGRU2D, True, _lib.HAM_HEIS2D), ("tens", _lib.CELL_TENS2D, True, _lib.HAM_HEIS2D),
         ("vanilla", _lib.CELL_VANILLA2D, False, _lib.HAM_TFIM2D), ("gru4", _lib.CELL_GRU4, False, _lib.HAM_RYDBERG)]
only = sys.argv[3] if len(sys.argv) > 3 else None
if only:
    CELLS = [c for c in CELLS if c[0] == only]
bad = 0
for case in range(ncases):
    name, cell, cplx, ham = CELLS[rng.randint(len(CELLS))]
    nx, ny = int(rng.randint(1, 7)), int(rng.randint(1, 7))
    if name == "gru4":
        nx, ny = min(nx, 4), min(ny, 4)
    nh = int(rng.randint(5, 105 if name != "gru4" else 200))
    if only == "vanilla":
        nh = int(rng.randint(5, 65))
    ns = int(rng.choice([1, 7, 64, 129, 300, 1000, 2500]))
    u1 = cplx and (nx * ny) % 2 == 0
    kw = dict(num_units=nh, complex_head=cplx, u1=u1, seed=int(rng.randint(1 << 30)))
    ref = Engine(cell, nx, ny, precision="fp32", **kw)
    p = ref.params.cpu() + 0.1 * (torch.rand(ref.num_params, generator=torch.Generator().manual_seed(case)) - 0.5)
    ref.set_weights(p)
    hp = {_lib.HAM_HEIS2D: [1.0], _lib.HAM_TFIM2D: [2.0] + [1.0] * (nx * ny), _lib.HAM_RYDBERG: [1.7 ** 6, 3.3, 1.0, 2.001]}[ham]
    u = torch.rand(ns, ref.nsteps, generator=torch.Generator().manual_seed(case + 1000), dtype=torch.float32)
    sv0 = ref.new_saved(ns)
    s0, lp0, ph0 = ref.sample(ns, uniforms=u, saved=sv0)
    e0 = ref.local_energies(ham, hp, s0, lp0, ph0, saved=sv0)
    seeds = torch.randn(ns, generator=torch.Generator().manual_seed(case + 2000), dtype=torch.float64)
    g0 = ref.backward(s0, sv0, seeds, seeds if cplx else None).double()
    line = f"{case:3d} {name:8s} {nx}x{ny} nh={nh:3d} ns={ns:5d}"
    for prec, tol, gtol in (("tc3", 2e-5, 2e-2), ("tc", 3e-3, 2e-2)):
        try:
            eng = Engine(cell, nx, ny, precision=prec, **kw)
        except ValueError as exc:
            line += f" | {prec}: n/a"
            continue
        eng.set_weights(p)
        sv = eng.new_saved(ns)
        lp, ph = eng.logpsi(s0, saved=sv)
        e = eng.local_energies(ham, hp, s0, lp, ph, saved=sv)
        if name == "vanilla":      # the same without prefix reuse must agree with the reused one to rounding
            e_nr = eng.local_energies(ham, hp, s0, lp, ph, saved=None)
            if float((e_nr - e).abs().max()) > 10 * tol * (float(e0.abs().mean()) + 1e-30):
                bad += 1
                line += " | no-reuse E_loc differs  <-- VIOLATION"
        g = eng.backward(s0, sv, seeds, seeds if cplx else None).double()
        scale = float(lp0.abs().mean()) + 1e-30
        elp = float((lp - lp0).abs().max()) / scale
        ee = float((e - e0).abs().max()) / (float(e0.abs().mean()) + 1e-30)
        eg = float((g - g0).abs().max()) / (float(g0.abs().max()) + 1e-30)
        ok = elp < tol and ee < 10 * tol and eg < gtol and bool(torch.isfinite(g).all())
        bad += not ok
        line += f" | {prec}: logp {elp:.1e} eloc {ee:.1e} grad {eg:.1e}{'' if ok else '  <-- VIOLATION'}"
        del eng
    print(line, flush=True)
    del ref
    torch.cuda.empty_cache()
sys.exit(1 if bad else 0)
