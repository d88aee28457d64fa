import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, torch
from oracle import local_energies as ole
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
from helpers import oracle_params, oracle_logpsi_fn, rel_err
for jit in (0.15, 0.0):
    engs={}
    for prec in ("fp32","tc3","tc"):
        eng = Engine(_lib.CELL_VANILLA2D, 12, 12, num_units=50, seed=71, precision=prec)
        g = torch.Generator().manual_seed(72); p = eng.params.cpu()
        eng.set_weights(p + jit * (torch.rand(p.shape, generator=g) - 0.5))
        engs[prec]=eng
    ns=150
    u = torch.rand(ns, 144, generator=torch.Generator().manual_seed(4), dtype=torch.float32)
    s, lp, ph = engs["fp32"].sample(ns, uniforms=u)
    Jz=np.ones(144)
    ref = ole.ising2d_local_energies(Jz.reshape(12,12), 3.0, s.cpu().numpy().astype(np.int64), oracle_logpsi_fn(engs["fp32"], oracle_params(engs["fp32"])))
    for prec,eng in engs.items():
        l2,_=eng.logpsi(s)
        e = eng.local_energies(_lib.HAM_TFIM2D, np.concatenate([[3.0], Jz]), s, l2, None).cpu().numpy().real
        print(jit, prec, "E_loc max/rms", rel_err(e, ref), "mean|E|", np.abs(ref).mean(), "max|E|", np.abs(ref).max())
