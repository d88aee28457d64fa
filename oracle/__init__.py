"""TEST INFRASTRUCTURE ONLY -- CPU float64 restatement of mhibatallah/RNNWavefunctions' VMC hot path.

Nothing under ``oracle/`` is part of the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` leg may import it, and only as the
checker (or as the timed CPU baseline), never as the thing shipped.

Parity status.  The *network* part of the reference lives in TensorFlow 1.13.1 (not vendored, not installable
here) and the reference ships no tests / golden vectors / weights.  Pins:
  * the reference's wavefunction classes are EXECUTED unmodified under an eager NumPy stand-in for TensorFlow
    (``tests/golden/tf_shim.py``); ``tests/golden/make_wf_golden.py`` commits weights, uniforms, samples,
    log-probabilities / log-amplitudes and every symmetry-projected variant as ``tests/golden/wf_*.npz`` and
    ``tests/test_oracle_wf_golden.py`` holds ``oracle/wavefunctions.py`` to them at 1e-12.  That pins the sweep order,
    neighbour bookkeeping, masks, head combination, group lists and rotate60 to the reference's own code.  NOT pinned
    (restated inside the shim from knowledge of the TF sources): ``CudnnCompatibleGRUCell``,
    ``layers.Dense``, ``multinomial`` / ``random.categorical``;
  * the host-side NumPy local-energy code of the reference runs here unmodified; its outputs on seeded inputs are
    committed under ``tests/golden/`` (``tests/golden/make_golden.py``) and the restatements in
    ``oracle/local_energies.py`` are checked against them;
  * parameter counts 422 / 444 (Tutorial notebooks) pin the GRU / Dense parameterisation;
  * exact-diagonalisation energies pin the Hamiltonians end to end (brute-force <psi|H|psi>).
"""
