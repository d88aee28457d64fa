"""rnnwavefunctions_b200 -- B200-native VMC hot path of RNN wavefunctions (sampling, batched
log-amplitudes of connected configurations, local energies, backward of log psi) behind the
API of mhibatallah/RNNWavefunctions.  Host code is Python/PyTorch calling the C-ABI library
``librnnwf.so`` (hand-written sm_100a CUDA); there is no TF, Triton or CPU fallback.

Sub-modules mirror the reference's packages:
    TFIM1D         <- 1DTFIM/            (RNNwavefunction, Ising_local_energies, run_1DTFIM)
    J1J2           <- J1J2/              (RNNwavefunction, J1J2MatrixElements, J1J2Slices, run_J1J2)
    TFIM2D         <- 2DTFIM_2DRNN/      (RNNwavefunction, MDRNNcell, Ising2D_local_energies, run_2DTFIM)
    Heisenberg2D   <- 2DHeisenberg/      (RNNwavefunction, MDRNNcell(s), Get_Samples_and_Elocs, Get_Elocs)
    RydbergKagome  <- RydbergKagome_2DGRU/ (MDRNNWavefunction, MDRNNcell, Rydberg_local_energies, ...)
"""
__version__ = "0.1.0"
