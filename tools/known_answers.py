"""End-to-end check against the known answers the reference's tree holds: full VMC trainings through
the reference-facing API with the reference's tutorial / script hyper-parameters, compared with exact diagonalisation, DMRG
and the energies the reference's own notebooks print.  Writes gpurun_out/known_answers.json.

  python tools/known_answers.py            # ~1 minute on one B200
"""
import contextlib, io, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rnnwavefunctions_b200 import TFIM1D, J1J2

def run(name, fn, exact, ref_result, tail):
    t0 = time.time()
    with contextlib.redirect_stdout(io.StringIO()):
        meanE, varE = fn()
    e = np.real(np.asarray(meanE, dtype=np.complex128))
    out = dict(case=name, steps=len(e) - 1, seconds=round(time.time() - t0, 1), exact=exact, reference_notebook=ref_result,
               final_mean=float(e[-tail:].mean()), final_sem=float(e[-tail:].std() / np.sqrt(tail)),
               final_var=float(np.mean(varE[-tail:])), rel_err_vs_exact=float(abs(e[-tail:].mean() - exact) / abs(exact)))
    print(json.dumps(out))
    return out

res = []
# Tutorials/1DTFIM/Tutorial_1DTFIM.ipynb: N = 10, 10 units, 1 layer, 200 samples, 1000 steps, lr 5e-3, seed 111; ED -12.38148999965476
res.append(run("1DTFIM N=10 tutorial", lambda: TFIM1D.run_1DTFIM(numsteps=1000, systemsize=10, num_units=10, Bx=1, num_layers=1,
                                                                  numsamples=200, learningrate=5e-3, seed=111),
               -12.38148999965476, "-12.38084 +- 0.00074", 100))
# 1DTFIM/TrainingRNN_1DTFIM.py defaults (C1): N = 20, 50 units, 500 samples, lr 5e-3; DMRG -25.1077971081
res.append(run("1DTFIM N=20 (C1)", lambda: TFIM1D.run_1DTFIM(numsteps=4000, systemsize=20, num_units=50, Bx=1, num_layers=1,
                                                              numsamples=500, learningrate=5e-3, seed=111),
               -25.1077971081, None, 200))
# Tutorials/J1J2/Tutorial_1DJ1J2.ipynb: N = 10, J2 = 0.2, no Marshall sign, 10 units, 200 samples, 3000 steps, lr 5e-4, seed 111; ED -3.9855798336170905
res.append(run("J1J2 N=10 J2=0.2 tutorial", lambda: J1J2.run_J1J2(numsteps=3000, systemsize=10, J1_=1.0, J2_=0.2, Marshall_sign=False,
                                                                   num_units=10, num_layers=1, numsamples=200, learningrate=5e-4, seed=111),
               -3.9855798336170905, "-3.96471 +- 0.0020", 200))
# the same chain with the Marshall sign prior, and without it but trained longer (the sign structure has to be learned)
res.append(run("J1J2 N=10 J2=0.2, Marshall sign", lambda: J1J2.run_J1J2(numsteps=3000, systemsize=10, J1_=1.0, J2_=0.2, Marshall_sign=True,
                                                                         num_units=10, num_layers=1, numsamples=200, learningrate=5e-4, seed=111),
               -3.9855798336170905, None, 200))
res.append(run("J1J2 N=10 J2=0.2, no Marshall sign, 10^4 steps", lambda: J1J2.run_J1J2(numsteps=10000, systemsize=10, J1_=1.0, J2_=0.2, Marshall_sign=False,
                                                                                        num_units=10, num_layers=1, numsamples=200, learningrate=5e-4, seed=111),
               -3.9855798336170905, None, 200))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/known_answers.json", "w"), indent=1)
