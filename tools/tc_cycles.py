"""Cycle accounting of the tensor-core sweep (debug counters inside sweep_tc_kernel), C4 workload by default.

    python tools/tc_cycles.py [--nx 10 --ny 10 --nh 100 --ns 10000]

Prints per-phase timings (CUDA events) of sample / local energies / backward and the mean per-CTA cycle counters
of one local-energy launch.  Counter slots are described next to `prof[]` in csrc/sweep_tc.cuh.
"""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# counters exist only in the -DRNNWF_TC_PROF build:  RNNWF_TC_PROF=1 python -c "import __graft_entry__ as g; g.build()"
_PROF = os.path.join(ROOT, "rnnwavefunctions_b200", "librnnwf_prof.so")
if os.path.exists(_PROF) and not os.environ.get("RNNWF_NO_PROF"):
    os.environ.setdefault("RNNWF_LIB_PATH", _PROF)
from rnnwavefunctions_b200 import _lib  # noqa: E402
from rnnwavefunctions_b200.engine import Engine  # noqa: E402

CTRL = ["bar_A", "bar_B", "bar_free", "mma_issue", "ld:stage_wait", "total", "ld:issue", "-"]
EPI = ["bar_mma", "pass", "next_half", "tail", "total", "hd:wait", "hd:heads", "-"]


def timed(fn, n=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=10)
    ap.add_argument("--ny", type=int, default=10)
    ap.add_argument("--nh", type=int, default=100)
    ap.add_argument("--ns", type=int, default=10000)
    ap.add_argument("--lattice", default="triangular")
    ap.add_argument("--debug-flags", type=int, default=0, help="also time E_loc with RNNWF_TC_DEBUG set to this")
    ap.add_argument("--trace", type=int, default=0, help="print this many events of CTA 0's timeline")
    ap.add_argument("--trace-from", type=int, default=200000, help="first cycle of the printed window")
    ap.add_argument("--what", default="eloc", choices=["eloc", "sample"], help="which launch carries the counters / the trace")
    args = ap.parse_args()
    eng = Engine(cell=_lib.CELL_TENSGRU2D, nx=args.nx, ny=args.ny, num_units=args.nh, complex_head=True, u1=True,
                 device="cuda:0", precision="tc")
    eng.init_params(seed=3)
    saved = eng.new_saved(args.ns)
    t_s, (cfg, lp, ph) = timed(lambda: eng.sample(args.ns, saved=saved))
    hp = [1.0 if args.lattice == "triangular" else 0.0]
    eloc = lambda: eng.local_energies(_lib.HAM_HEIS2D, hp, cfg, lp, ph, saved=saved)  # noqa: E731
    t_e, e = timed(eloc)
    if args.debug_flags:
        os.environ["RNNWF_TC_DEBUG"] = str(args.debug_flags)
        t_x, _ = timed(eloc)
        os.environ["RNNWF_TC_DEBUG"] = "0"
        print(f"eloc with RNNWF_TC_DEBUG={args.debug_flags}: {t_x:.2f} ms (timing experiment, results invalid)")
    sl = torch.randn(args.ns, device="cuda:0", dtype=torch.float64) / args.ns
    sp = torch.randn(args.ns, device="cuda:0", dtype=torch.float64) / args.ns
    t_b, _ = timed(lambda: eng.backward(cfg, saved, sl, sp))
    print(f"sample {t_s:.2f} ms  eloc {t_e:.2f} ms  backward {t_b:.2f} ms   E/N = {e.real.mean().item() / (args.nx * args.ny):.5f}")

    lib = _lib.lib
    nsm = torch.cuda.get_device_properties(0).multi_processor_count
    TRACE_LEN, ROLES = 1024, 8
    prof = torch.zeros(nsm * 16 + ROLES * TRACE_LEN, dtype=torch.int64, device="cuda:0")
    lib.rnnwf_debug_set_tc_prof.argtypes = [ctypes.c_void_p]
    lib.rnnwf_debug_set_tc_prof.restype = None
    lib.rnnwf_debug_set_tc_prof(ctypes.c_void_p(prof.data_ptr()))
    if args.what == "sample":
        eng.sample(args.ns, saved=saved)
    else:
        eloc()
    torch.cuda.synchronize()
    lib.rnnwf_debug_set_tc_prof(ctypes.c_void_p(0))
    trace = prof[nsm * 16:].view(ROLES, TRACE_LEN).cpu().numpy()
    m = prof[:nsm * 16].view(nsm, 16).double().mean(0).tolist()
    print("library:", _lib.LIB_PATH)
    print("mma/loader warp (Mcycles): " + "  ".join(f"{n}={v / 1e6:.1f}" for n, v in zip(CTRL, m[:8]) if n != "-"))
    print("epilogue/head (Mcycles): " + "  ".join(f"{n}={v / 1e6:.1f}" for n, v in zip(EPI, m[8:]) if n != "-"))
    if args.trace:
        names = ["mma", "loader", "epi1", "epi16", "epi8", "head"]
        ev = []
        for r, nm in enumerate(names):
            for w in trace[r]:
                if w:
                    ev.append((int(w) & ((1 << 40) - 1), nm, int(w) >> 40))
        ev.sort()
        ev = [e for e in ev if e[0] >= args.trace_from][:args.trace]
        tags = {1: "A_ready", 2: "B_ready", 3: "acc_free", 4: "issued", 5: "stage_free", 6: "copy_issued", 11: "half_built",
                12: "acc_full", 13: "acc_drained", 14: "passes_done", 15: "A_arrived", 16: "part_arrived", 20: "part_ready",
                21: "heads_done"}
        t0 = ev[0][0] if ev else 0
        for t, nm, tag in ev:
            print(f"{t - t0:8d}  {nm:7s} {tags.get(tag, tag)}")


if __name__ == "__main__":
    main()

