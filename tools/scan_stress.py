"""Stress of the single-launch chained scan of the enumeration (many blocks): local energies and the number of connected
configurations must equal the three-launch scan's (RNNWF_SCAN3=1) bit for bit."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import _lib
from rnnwavefunctions_b200.engine import Engine
import numpy as np
ok = True
for name, eng, ham, hp, ns in (
        ("J1J2 N=20", Engine(_lib.CELL_GRU1D, 20, 1, num_units=16, complex_head=True, u1=True, seed=3), _lib.HAM_J1J2, [0.0, 0.0] + [1.0] * 20 + [0.2] * 20 + [0.0] * 20, 150000),
        ("Heis 10x10", Engine(_lib.CELL_TENSGRU2D, 10, 10, num_units=16, complex_head=True, u1=True, seed=4, precision="tc"), _lib.HAM_HEIS2D, [1.0], 30000)):
    s, lp, ph = eng.sample(ns, seed=5)
    res = []
    for rep in range(3):
        for env in (None, "1"):
            if env: os.environ["RNNWF_SCAN3"] = env
            else: os.environ.pop("RNNWF_SCAN3", None)
            e, cnt = eng.local_energies(ham, hp, s, lp, ph, return_count=True)
            res.append((e.clone(), int(cnt)))
    same = all(r[1] == res[0][1] for r in res) and all(float((r[0] - res[0][0]).abs().max()) < 1e-9 * float(res[0][0].abs().max()) for r in res)
    print(name, "ns", ns, "children", res[0][1], "consistent over 3 x {chained, three-launch}:", same)
    ok &= same
sys.exit(0 if ok else 1)
