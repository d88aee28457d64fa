"""Timing experiment: the backward of C4 (tc regime) with parts of the weight-gradient kernel switched off
(RNNWF_TC_DEBUG bits 16 no MMAs / 17 no operand loads / 18 no masked copies in dw_tc_kernel, 19..21 in bwd_tc_kernel; results of such runs are invalid)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
ns = int(sys.argv[1]) if len(sys.argv) > 1 and int(sys.argv[1]) > 1000 else 10000
quick = len(sys.argv) > 1 and int(sys.argv[1]) <= 1000
eng = Engine(cell=_lib.CELL_TENSGRU2D, nx=10, ny=10, num_units=100, complex_head=True, u1=True, precision="tc")
saved = eng.new_saved(ns)
s, lp, ph = eng.sample(ns, seed=1, saved=saved)
seeds = torch.randn(ns, dtype=torch.float64, device="cuda")
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
if quick:      # one backward only (profiling runs)
    eng.backward(s, saved, seeds, seeds); torch.cuda.synchronize(); sys.exit(0)
for flags, what in ((0, "everything"), (1 << 16, "no MMAs"), (1 << 17, "no operand loads"), (1 << 18, "no masked copies"),
                    (7 << 16, "none of the three"), (1 << 19, "sweep: no MMAs"), (1 << 20, "sweep: no dpT stores"),
                    (1 << 21, "sweep: no dh update"), (7 << 19, "sweep: none"), (63 << 16, "all off")):
    os.environ["RNNWF_TC_DEBUG"] = str(flags)
    ms = t(lambda: eng.backward(s, saved, seeds, seeds))
    print(f"{what:22s} {ms:.3f} ms")
