/*
 * rnnwf.h -- C-ABI of the B200-native RNN-wavefunction VMC hot path (librnnwf.so).
 *
 * The reference (mhibatallah/RNNWavefunctions) is pure Python on TensorFlow 1.13 graph mode and has no
 * FFI/plugin layer of its own: its boundary for this path is the Python class / function API
 *     RNNwavefunction(...).sample / .log_probability / .log_amplitude      (L1)
 *     Ising_local_energies / J1J2Slices / Get_Elocs / Rydberg_local_energies   (L2)
 *     optimizer.compute_gradients(cost)                                     (backward, via TF autodiff)
 * Each entry point below names the reference interface it replaces (file:line under the reference root).
 * The Python package `rnnwavefunctions_b200` mirrors those reference names one-to-one and calls these
 * symbols through ctypes; INTEGRATION.md shows the stub a maintainer of the reference would add.
 *
 * Conventions
 *   - every pointer marked `dev` is a device pointer on the CUDA device current at the call; the caller
 *     (PyTorch in our package) owns every buffer; the library owns only the opaque handle (step tables,
 *     packed weights);
 *   - all entry points are asynchronous on `stream` (a cudaStream_t passed as void*), allocate nothing
 *     (create/destroy excepted) and return 0 on success or a negative RNNWF_E* code; the message of the
 *     last error on the calling thread is returned by rnnwf_last_error();
 *   - configurations are int8 occupation numbers in the reference's array layout:
 *       1D [B, N];  2D [B, Nx, Ny];  Rydberg [B, Lx, Ly, 3];
 *   - log-probabilities / phases are float64 (the reference returns float64 / complex128);
 *     log psi = 0.5*logp + i*phase for the complex models.
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with RNNWF_ECUDA.
 */
#ifndef RNNWF_H_
#define RNNWF_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rnnwf_model* rnnwf_handle;

/* cell families */
enum {
  RNNWF_CELL_GRU1D     = 0, /* stacked CudnnCompatibleGRUCell: 1DTFIM/RNNwavefunction.py:32, J1J2/ComplexRNNwavefunction.py:40 */
  RNNWF_CELL_VANILLA2D = 1, /* 2DTFIM_2DRNN/MDRNNcell.py:51-66 */
  RNNWF_CELL_TENS2D    = 2, /* 2DHeisenberg/MDTensorizedcell.py:40-49 */
  RNNWF_CELL_TENSGRU2D = 3, /* 2DHeisenberg/MDTensorizedGRUcell.py:49-61 */
  RNNWF_CELL_GRU4      = 4  /* RydbergKagome_2DGRU/MDGRUcell.py:48-59 (periodic, 4 neighbours, 8-way softmax) */
};

/* error codes */
enum { RNNWF_OK = 0, RNNWF_EINVAL = -1, RNNWF_ECUDA = -2, RNNWF_ENOMEM = -3, RNNWF_ESTATE = -4 };

typedef struct {
  int32_t cell;          /* RNNWF_CELL_* */
  int32_t complex_head;  /* 0: softmax head only (positive wavefunction); 1: + pi*softsign phase head */
  int32_t nx, ny;        /* lattice; 1D chains: nx = N, ny = 1; Rydberg: unit cells Lx, Ly */
  int32_t num_units;     /* hidden units per layer */
  int32_t num_layers;    /* GRU1D only (MultiRNNCell depth); 1 otherwise */
  int32_t u1;            /* 1: U(1) projection (J1J2/ComplexRNNwavefunction.py:85-93; 2DHeisenberg/MDComplexRNN.py:69-76) */
  int32_t magnetization; /* 2*Sz target of the projection */
  int32_t reserved[8];   /* reserved[0] = precision: 0 FP32 FFMA sweep (1e-5 parity), 1 tcgen05 tensor-core sweep
                          * (FP16 operands, FP32 accumulate in TMEM; 1e-3 parity bound), 2 tcgen05 sweep in split
                          * precision (FP16 hi+lo operand pairs, three MMAs per K step; 1e-5 parity).  1 and 2 exist for
                          * the 2D cells (vanilla RNN, tensorized, tensorized GRU, four-neighbour GRU) */
} rnnwf_config;

const char* rnnwf_last_error(void);
int rnnwf_version(void);

/* Replaces the graph construction of RNNwavefunction.__init__ (1DTFIM/RNNwavefunction.py:8-33,
 * J1J2/ComplexRNNwavefunction.py:16-43, 2DTFIM_2DRNN/RNNwavefunction.py, 2DHeisenberg/MDComplexRNN.py:28-65,
 * RydbergKagome_2DGRU/MDRNNWavefunction.py:38-68): builds the lattice sweep tables on the current device. */
int rnnwf_create(const rnnwf_config* cfg, rnnwf_handle* out);
int rnnwf_destroy(rnnwf_handle h);

/* Parameter vector: one flat float32 array holding the TF variables in creation order and TF layout
 * (SURVEY App. B).  rnnwf_param_info enumerates them (name without scope prefix, offset, shape). */
int64_t rnnwf_count_params(const rnnwf_config* cfg); /* host-only: no device needed (notebook checks: 422 / 444 parameters) */
int64_t rnnwf_num_params(rnnwf_handle h);
int rnnwf_num_tensors(rnnwf_handle h);
int rnnwf_param_info(rnnwf_handle h, int index, char* name, int name_cap, int64_t* offset, int32_t* ndim, int64_t shape[4]);

/* Replaces sess.run(init) / saver.restore: (re)packs `params` (dev, float32[num_params]) into the
 * kernels' weight layout.  Must be called after every parameter update. */
int rnnwf_set_weights(rnnwf_handle h, const float* params_dev, void* stream);

/* Size of the scratch workspace (bytes) the sweep kernels need for batches of any size. */
int64_t rnnwf_workspace_bytes(rnnwf_handle h);
/* Size of the per-sample "saved state" (hidden-state history + gate activations + running sums)
 * that sample()/logpsi() can record for `rows` configurations and backward()/eloc() consume. */
int64_t rnnwf_saved_bytes(rnnwf_handle h, int64_t rows);

/* wf.sample(numsamples, inputdim)  -- 1DTFIM/RNNwavefunction.py:35-74, J1J2/ComplexRNNwavefunction.py:45-103,
 * 2DTFIM_2DRNN/RNNwavefunction.py:35-118, 2DHeisenberg/MDComplexRNN.py:80-173,
 * RydbergKagome_2DGRU/MDRNNWavefunction.py:73-153.
 * Inverse-CDF draws (TF CPU multinomial semantics) from `uniforms` (dev float32 [rows, nsteps], u[b,t] with
 * t the generation-order step) or, when uniforms == NULL, from Philox4x32-10 keyed by `seed` with counter
 * (row_offset + b, t).  Also returns log P and the phase of every sample (the forward pass the reference
 * repeats in log_probability).  `saved` may be NULL. */
int rnnwf_sample(rnnwf_handle h, int64_t rows, const float* uniforms_dev, uint64_t seed, uint64_t row_offset,
                 int8_t* samples_dev, double* logp_dev, double* phase_dev, void* saved_dev,
                 void* workspace_dev, int64_t workspace_bytes, void* stream);

/* wf.log_probability / wf.log_amplitude(samples, inputdim) -- 1DTFIM/RNNwavefunction.py:76-118,
 * J1J2/ComplexRNNwavefunction.py:105-169, 2DTFIM_2DRNN/RNNwavefunction.py:120-200,
 * 2DHeisenberg/MDComplexRNN.py:177-286, RydbergKagome_2DGRU/MDRNNWavefunction.py:157-238. */
int rnnwf_logpsi(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, double* logp_dev, double* phase_dev,
                 void* saved_dev, void* workspace_dev, int64_t workspace_bytes, void* stream);

/* Local energies.  All of them: enumerate the connected configurations of every sample on the device,
 * evaluate their log-amplitudes in one batched sweep and reduce E_loc in-kernel.
 *   eloc_dev   : float64 [rows, 2] (re, im)  -- im == 0 for the positive wavefunctions
 *   logp/phase : log psi of the samples themselves (from rnnwf_sample / rnnwf_logpsi)
 *   nconn_dev  : optional int64 [1], number of configurations with non-zero matrix element evaluated
 *   `saved`    : optional saved state of the samples -> connected configurations restart from the first
 *                changed site (prefix reuse); NULL -> full sweeps.
 * Hamiltonian parameter block `ham` (host pointer, float64):
 *   RNNWF_HAM_TFIM1D  {Bx, Jz[0..N-2]}                 Ising_local_energies   1DTFIM/TrainingRNN_1DTFIM.py:13-75
 *   RNNWF_HAM_TFIM2D  {Bx, Jz[nx*Ny+ny] ...}           Ising2D_local_energies 2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:13-83
 *   RNNWF_HAM_J1J2    {periodic, marshall, J1[N], J2[N], Bz[N]}  J1J2MatrixElements/J1J2Slices J1J2/TrainingRNN_J1J2.py:12-127,277-279
 *   RNNWF_HAM_HEIS2D  {J2}                             Triangular2DHeis_* / Get_Elocs 2DHeisenberg/Helper_Functions.py:19-177
 *   RNNWF_HAM_RYDBERG {V, delta, Omega, Rcutoff}       generate_kagome / neighbouring_atoms / Rydberg_local_energies
 *                                                      RydbergKagome_2DGRU/Helping_Functions.py:33-153 */
enum { RNNWF_HAM_TFIM1D = 0, RNNWF_HAM_TFIM2D = 1, RNNWF_HAM_J1J2 = 2, RNNWF_HAM_HEIS2D = 3, RNNWF_HAM_RYDBERG = 4 };
int64_t rnnwf_eloc_workspace_bytes(rnnwf_handle h, int hamiltonian, int64_t rows);
int rnnwf_local_energies(rnnwf_handle h, int hamiltonian, const double* ham, int ham_len,
                         int64_t rows, const int8_t* samples_dev, const double* logp_dev, const double* phase_dev,
                         const void* saved_dev, double* eloc_dev, int64_t* nconn_dev,
                         void* workspace_dev, int64_t workspace_bytes, void* stream);

/* Backward of log psi: grad[k] = sum_b seed_logp[b] * dlogP_b/dtheta_k + seed_phase[b] * dphase_b/dtheta_k,
 * i.e. exactly what TF autodiff of the surrogate costs produces (1DTFIM/TrainingRNN_1DTFIM.py:157-161,
 * J1J2/TrainingRNN_J1J2.py:197-200, 2DHeisenberg/Run_RNNWF_2DHeis.py:199-217,
 * RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-176) once the caller has formed the seeds from E_loc
 * (SURVEY 3.4).  `saved` must come from rnnwf_sample / rnnwf_logpsi on the same `samples` and weights.
 * grad_dev: float32 [num_params] in the parameter layout (overwritten).  seed_phase may be NULL. */
int64_t rnnwf_backward_workspace_bytes(rnnwf_handle h, int64_t rows);
int rnnwf_backward(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, const void* saved_dev,
                   const double* seed_logp_dev, const double* seed_phase_dev, float* grad_dev,
                   void* workspace_dev, int64_t workspace_bytes, void* stream);

/* No counterpart in the reference.  rnnwf_logpsi for hosts that cannot give the library any device
 * scratch: a process whose runtime has already claimed the whole GPU (TensorFlow 1.x does by default), or a
 * bare ctypes script that holds nothing but its input and output arrays.  The sweep's history ring then lives
 * in a fixed RNNWF_STATIC_SCRATCH_BYTES array inside the library image (each GPU that loads the library keeps
 * it until exit).  Same inputs, outputs and results as rnnwf_logpsi without `saved`; no workspace argument;
 * the call allocates nothing.  Limit: one thread block's ring (ring_len x 32 rows x padded units x 4 bytes)
 * must fit in the array, otherwise the call fails with RNNWF_ENOMEM and says so; fewer thread blocks run
 * concurrently than with a caller workspace, so it is slower on large batches.  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress) -- make it outside any stream capture.  All
 * calls on a device share the array: calls that overlap on the GPU (different streams) race. */
#define RNNWF_STATIC_SCRATCH_BYTES (16u << 20)
int rnnwf_logpsi_static(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, double* logp_dev, double* phase_dev,
                        void* stream);

/* Optimiser step right after the gradient all-reduce: tf.train.AdamOptimizer.apply_gradients as the training scripts
 * use it (1DTFIM/TrainingRNN_1DTFIM.py:163-165, 2DHeisenberg/Run_RNNWF_2DHeis.py:201-217,
 * RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:161-177), one fused pass over the flat parameter vector:
 *   g <- clip_by_value(g, -clip, clip) if clip > 0;  m <- b1 m + (1-b1) g;  v <- b2 v + (1-b2) g^2;
 *   theta <- theta - lr sqrt(1-b2^t)/(1-b1^t) * m / (sqrt(v) + eps),   t = 1, 2, ... the step number.
 * All arrays float32 [n] on the device; m and v are the caller's optimiser state (zero before the first step).
 * Call rnnwf_set_weights afterwards to repack. */
int rnnwf_adam_step(float* params_dev, const float* grad_dev, float* m_dev, float* v_dev, int64_t n, double lr,
                    double beta1, double beta2, double eps, int64_t t, double clip, void* stream);

/* The other optimiser of 2DHeisenberg/Run_RNNWF_2DHeis.py:206-213 (`normalized_gradients`): plain gradient descent on
 *   sign(g_i) * u_i,   u_i uniform in [0, 1)   ->   theta_i <- theta_i - lr * sign(g_i) * u_i,
 * one fused pass.  The reference draws u with tf.random_uniform (a stream nobody can reproduce); here u_i is the library's
 * counter-based generator at (seed, element i, step t), so every rank of a data-parallel job applies the same update from
 * the same (seed, t) without communicating.  Call rnnwf_set_weights afterwards to repack. */
int rnnwf_sign_sgd_step(float* params_dev, const float* grad_dev, int64_t n, double lr, uint64_t seed, int64_t t, void* stream);

/* The step between the local energies and the backward: energy moments and the per-sample seeds of the reference's
 * surrogate costs, on the device (replaces the np.mean / np.var of the training loops and the cost lines
 * 1DTFIM/TrainingRNN_1DTFIM.py:157,207-208; J1J2/TrainingRNN_J1J2.py:197,281-282; 2DHeisenberg/Run_RNNWF_2DHeis.py:199,301-302;
 * RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-159).
 *   rnnwf_energy_moments: moments_dev[8] = {rows, sum Re E, sum Im E, sum (Re E)^2, sum (Im E)^2, sum logP, 0, 0} of this shard
 *     (eloc_dev float64 [rows][2]; logp_dev may be NULL).  Under data parallelism the caller all-reduces the 8 doubles.
 *   rnnwf_cost_seeds: from the (global) moments,  seed_logp[b] = (Re(E_b - <E>) + T (logP_b - <logP>)) / n  (the T term only for
 *     the complex cost with annealing or the free-energy cost), seed_phase[b] = 2 Im(E_b - <E>) / n (complex models);
 *     stats_dev[4] (optional) = {Re <E>, Im <E>, var E = var Re + var Im, <logP>}. */
int rnnwf_energy_moments(int64_t rows, const double* eloc_dev, const double* logp_dev, double* moments_dev, void* stream);
int rnnwf_cost_seeds(int64_t rows, const double* eloc_dev, const double* logp_dev, const double* moments_dev, int complex_head, double T,
                     int free_energy, double* seed_logp_dev, double* seed_phase_dev, double* stats_dev, void* stream);

/* Symmetry-projected amplitudes (2DHeisenberg/MDComplexRNN.py:290-316 log_amplitudes_fromsymmetrygroup and the rotsym /
 * c2v / c4v ... variants :320-577 built on it; RydbergKagome_2DGRU/MDRNNWavefunction.py:242-279 log_probability_symmetry):
 * the caller evaluates the G transformed copies of its `rows` configurations in one rnnwf_logpsi batch (copy-major:
 * logp[g*rows + b]); the epilogue returns  out_logp[b] = log(mean_g P_g)  and  out_phase[b] = arg(reg(Re) + i Im) of
 * sum_g e^{i phi_g} / chi_g,  reg(x) = sign(x) sqrt(x^2 + 1e-8).  chi: G host doubles (the group characters; NULL = all 1).
 * phase_dev / out_phase_dev may be NULL (positive wavefunctions).  rnnwf_symmetry_seeds is its backward: given
 * d cost / d out_logp and d cost / d out_phase per row it fills the per-copy seeds rnnwf_backward takes for the G*rows batch
 * (the sign inside reg is under stop-gradient, as in the reference). */
int rnnwf_symmetry_epilogue(int G, int64_t rows, const double* logp_dev, const double* phase_dev, const double* chi,
                            double* out_logp_dev, double* out_phase_dev, void* stream);
int rnnwf_symmetry_seeds(int G, int64_t rows, const double* logp_dev, const double* phase_dev, const double* chi,
                         const double* s_logp_dev, const double* s_phase_dev, double* seed_logp_dev, double* seed_phase_dev,
                         void* stream);

/* The symmetry group as device-resident index-remap tables, built once per (wavefunction, group): element g maps a
 * configuration s to  s_g[i] = s[perm[g][i]] ^ flip[g]  -- tf.image.rot90 / [:, ::-1] / transpose / abs(1 - s) of
 * 2DHeisenberg/MDComplexRNN.py:320-577, rotate60 and tf.roll of RydbergKagome_2DGRU/MDRNNWavefunction.py:246-291, the chain
 * reversal of 1DTFIM/RNNwavefunction_paritysym.py:125.  perm: host int32 [G][nsites] in the configuration's array layout
 * (each row a permutation), flip: host int8 [G] or NULL, chi: host double [G] group characters or NULL (all 1; also used
 * for the positive wavefunctions).  G <= 64. */
typedef struct rnnwf_symmetry* rnnwf_symmetry_handle;
int rnnwf_symmetry_create(rnnwf_handle h, int G, const int32_t* perm, const int8_t* flip, const double* chi, rnnwf_symmetry_handle* out);
int rnnwf_symmetry_destroy(rnnwf_symmetry_handle s);
int rnnwf_symmetry_order(rnnwf_symmetry_handle s);
/* list_samples of the reference's symmetric methods, formed on the device: out_dev int8 [G][rows][nsites] (copy-major, the
 * layout rnnwf_symmetry_epilogue / rnnwf_symmetry_seeds take after one rnnwf_logpsi over G*rows configurations). */
int rnnwf_symmetry_apply(rnnwf_symmetry_handle s, int64_t rows, const int8_t* in_dev, int8_t* out_dev, void* stream);
/* Local energies of a symmetry-projected wavefunction: Get_Elocs(..., RNN_symmetry != "nosym")
 * (2DHeisenberg/Helper_Functions.py:131-144,158-177 with log_amps = wf.log_amplitude_<group>), Rydberg_local_energies with
 * log_probability_symmetry (RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:155,262), Ising_local_energies with the parity-symmetric
 * log_probability (1DTFIM/RNNwavefunction_paritysym.py:122-145).  Enumerates the connected configurations on the device,
 * forms the G copies of the samples and of every connected configuration through the remap tables, evaluates all of them in
 * one batched sweep (full sweeps: a transformed copy shares no prefix with its parent), combines each group of G and
 * reduces E_loc.  Also returns the symmetrised log psi of the samples (out_logp, out_phase; phase NULL for positive
 * wavefunctions).  Same `ham` blocks and eloc / nconn outputs as rnnwf_local_energies. */
int64_t rnnwf_eloc_sym_workspace_bytes(rnnwf_handle h, rnnwf_symmetry_handle s, int hamiltonian, int64_t rows);
int rnnwf_local_energies_sym(rnnwf_handle h, rnnwf_symmetry_handle s, int hamiltonian, const double* ham, int ham_len, int64_t rows,
                             const int8_t* samples_dev, double* out_logp_dev, double* out_phase_dev, double* eloc_dev,
                             int64_t* nconn_dev, void* workspace_dev, int64_t workspace_bytes, void* stream);

/* Kernel-regime report for the profiler notes: fills `buf` with a one-line description of the tile
 * shape / pipe (FFMA or tcgen05) the handle's sweeps run in. */
int rnnwf_describe(rnnwf_handle h, char* buf, int cap);

/* Measurement helpers (no reference counterpart): number of kernels launched by this library so far in
 * this process, and an FP32-FFMA throughput probe (blocks x 512 threads x iters x 16 FMA chains) used by
 * bench.py as the roofline denominator of the FFMA regime. */
long long rnnwf_launch_count(void);
int rnnwf_ffma_probe(float* out_dev, int blocks, int iters, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RNNWF_H_ */
