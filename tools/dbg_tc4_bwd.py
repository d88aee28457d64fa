"""Debug helper: gradient of the tensor-core Rydberg backward vs the FP32 regime on the same saved samples."""
import sys
import torch
sys.path.insert(0, ".")
import __graft_entry__ as g
g.build()
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine

def run(nx, ny, nh, ns, poison=True):
    if poison:
        junk = torch.full((3 << 27,), float("nan"), dtype=torch.float32, device="cuda"); del junk
    eng = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=91, precision="tc")
    f32 = Engine(_lib.CELL_GRU4, nx, ny, num_units=nh, seed=91)
    f32.set_weights(eng.params)
    u = torch.rand(ns, eng.nsteps, generator=torch.Generator().manual_seed(9), dtype=torch.float32)
    sv = eng.new_saved(ns); s, lp, ph = eng.sample(ns, uniforms=u, saved=sv)
    sv2 = f32.new_saved(ns); f32.logpsi(s, saved=sv2)
    E = torch.randn(ns, generator=torch.Generator().manual_seed(10), dtype=torch.float64)
    seeds = (E - E.mean()) / ns
    ga = eng.backward(s, sv, seeds).cpu(); gb = f32.backward(s, sv2, seeds).cpu()
    out = []
    for (n, a), (_, b) in zip(eng.views(ga).items(), f32.views(gb).items()):
        nan = int(torch.isnan(a).sum())
        err = float((a - b).abs().nan_to_num(0).max() / (b.abs().max() + 1e-30))
        out.append(f"{n}: nan={nan}/{a.numel()} relerr={err:.2e}")
    print((nx, ny, nh, ns), "; ".join(out), flush=True)

for cfg in [(3, 4, 130, 300), (3, 4, 130, 3000), (3, 4, 130, 19500), (3, 4, 128, 19500), (4, 3, 256, 19500), (3, 4, 192, 2000)]:
    run(*cfg)
