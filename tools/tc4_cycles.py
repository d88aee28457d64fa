"""Cycle accounting of the tensor-core sweep of the Rydberg GRU (debug counters inside sweep_tc4_kernel).

    RNNWF_TC_PROF=1 python -c "import __graft_entry__ as g; g.build()"      # builds librnnwf_prof.so with the counters
    python tools/tc4_cycles.py [--nx 4 --ny 4 --nh 256 --ns 12500]

Prints CUDA-event timings of sample / local energies and the mean per-CTA cycle counters of one local-energy launch
(slots are described next to `prof4[]` in csrc/sweep_tc4.cuh)."""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
_PROF = os.path.join(ROOT, "rnnwavefunctions_b200", "librnnwf_prof.so")
if os.path.exists(_PROF):
    os.environ.setdefault("RNNWF_LIB_PATH", _PROF)
from rnnwavefunctions_b200 import _lib  # noqa: E402
from rnnwavefunctions_b200.engine import Engine  # noqa: E402

NAMES = ["mma:acc_free", "mma:stage_full", "mma:x_chunk", "mma:issue", "ld:stage_empty", "mma:total", "ld:ring_chunk", "ld:total",
         "epi:acc_full", "epi:pass", "epi:tail", "tile:head_wait", "epi:total", "tile:symbols", "tile:gather", "tile:publish"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=4)
    ap.add_argument("--ny", type=int, default=4)
    ap.add_argument("--nh", type=int, default=256)
    ap.add_argument("--ns", type=int, default=12500)
    ap.add_argument("--trace", type=int, default=0, help="print this many events of CTA 0's timeline (M: MMA 1 pass start / 2 pass "
                    "issued; L: loader 10+4c wait for ring chunk c / 11+4c got it; E: epilogue 20+p accumulators ready / 30+p chunk released / 40 pass done)")
    ap.add_argument("--trace-from", type=int, default=300000)
    ap.add_argument("--debug-flags", type=int, nargs="*", default=[], help="also time E_loc with RNNWF_TC_DEBUG set to each of these")
    a = ap.parse_args()
    eng = Engine(_lib.CELL_GRU4, a.nx, a.ny, num_units=a.nh, precision="tc")
    saved = eng.new_saved(a.ns)
    s, lp, ph = eng.sample(a.ns, saved=saved)
    ham = [1.7 ** 6, 3.3, 1.0, 2.001]
    eloc = lambda: eng.local_energies(_lib.HAM_RYDBERG, ham, s, lp, ph, saved=saved)  # noqa: E731
    eloc()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        eloc()
    e1.record()
    torch.cuda.synchronize()
    print(f"eloc {e0.elapsed_time(e1) / 3:.3f} ms")
    for flags in a.debug_flags:          # timing experiments (results invalid): bit 0 halves the weight stream, bit 1 the hidden-state stream
        os.environ["RNNWF_TC_DEBUG"] = str(flags)
        eloc()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            eloc()
        e1.record()
        torch.cuda.synchronize()
        print(f"eloc with RNNWF_TC_DEBUG={flags}: {e0.elapsed_time(e1) / 3:.3f} ms (results invalid)")
    os.environ["RNNWF_TC_DEBUG"] = "0"
    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    TRACE_LEN, ROLES = 4096, 3
    prof = torch.zeros(nsm * 16 + ROLES * TRACE_LEN, dtype=torch.int64, device="cuda:0")
    lib = _lib.lib
    lib.rnnwf_debug_set_tc_prof.argtypes = [ctypes.c_void_p]
    lib.rnnwf_debug_set_tc_prof.restype = None
    lib.rnnwf_debug_set_tc_prof(ctypes.c_void_p(prof.data_ptr()))
    eloc()
    torch.cuda.synchronize()
    lib.rnnwf_debug_set_tc_prof(ctypes.c_void_p(0))
    m = prof[:nsm * 16].view(nsm, 16).double().mean(0).tolist()
    if a.trace:
        tr = prof[nsm * 16:].view(ROLES, TRACE_LEN).cpu().numpy()
        ev = []
        for r in range(ROLES):
            for w in tr[r]:
                if w:
                    ev.append((int(w) & ((1 << 40) - 1), "MLE"[r], int(w) >> 40))
        ev.sort()
        for t, r, tag in [e for e in ev if e[0] >= a.trace_from][:a.trace]:
            print(f"    {t:9d} {r} {tag}")
    for n, v in zip(NAMES, m):
        if n != "-":
            print(f"  {n:16s} {v / 1e6:8.3f} Mcycles")


if __name__ == "__main__":
    main()
