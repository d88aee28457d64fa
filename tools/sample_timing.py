"""Timing experiment: the sampling sweep of C4 with individual saved-state stores switched off (RNNWF_TC_DEBUG bits 2/4/8;
results of such runs are invalid)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
eng = Engine(cell=_lib.CELL_TENSGRU2D, nx=10, ny=10, num_units=100, complex_head=True, u1=True, precision="tc")
saved = eng.new_saved(ns)
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for flags, what in ((0, "all stores"), (2, "no himg"), (4, "no hist"), (8, "no gates"), (14, "no saved stores"), (None, "saved=None")):
    os.environ["RNNWF_TC_DEBUG"] = str(flags or 0)
    ms = t(lambda: eng.sample(ns, saved=None if flags is None else saved))
    print(f"{what:18s} {ms:.3f} ms")
