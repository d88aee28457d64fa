"""One VMC step of the reference's training loops, on the device:

    sample -> local energies (connected configurations, one batched sweep) -> energy moments
           -> backward of log psi seeded with (E_loc - <E>) -> optimiser update -> repack weights

mirroring 1DTFIM/TrainingRNN_1DTFIM.py:200-222, J1J2/TrainingRNN_J1J2.py:240-305,
2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:200-231, 2DHeisenberg/Run_RNNWF_2DHeis.py:288-349 and
RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:243-293.

Data parallelism (new; the reference is single-replica): every rank draws its own shard of the samples
(Philox counters offset by rank * numsamples_per_rank, so the union over ranks is the single-rank stream),
and the only collectives are an all-reduce of the float64 energy moments (needed before the backward
because the seed is (E_b - <E>_global)/ns_global) and an all-reduce of the float32 gradient.
"""
import math

import torch
import torch.distributed as dist

from . import _lib


def _dist_on():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def allreduce_(t):
    if _dist_on():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def energy_moments(eloc, extra=None):
    """Global moments from the local shard: returns (ns_global, mean E (complex), var, mean(extra))."""
    e = torch.view_as_real(eloc) if eloc.is_complex() else torch.stack([eloc, torch.zeros_like(eloc)], dim=-1)
    m = torch.zeros(6, dtype=torch.float64, device=e.device)
    m[0] = e.shape[0]
    m[1] = e[:, 0].sum()
    m[2] = e[:, 1].sum()
    m[3] = (e[:, 0] ** 2).sum()
    m[4] = (e[:, 1] ** 2).sum()
    if extra is not None:
        m[5] = extra.sum()
    allreduce_(m)
    n = m[0]
    mean_re, mean_im = m[1] / n, m[2] / n
    var = m[3] / n - mean_re ** 2 + m[4] / n - mean_im ** 2   # np.var of a complex array = var(re) + var(im)
    return n, torch.complex(mean_re, mean_im), var, m[5] / n


def seeds(eloc, mean_e, n_global, complex_head, logp=None, mean_logp=None, T=0.0, free_energy=False):
    """Per-sample backward seeds of the reference's surrogate costs.

    real:     d cost/d logP_b = (E_b - <E>)/ns                       (1DTFIM/TrainingRNN_1DTFIM.py:157)
    free en.: E -> E + T*logP                                        (RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-159)
    complex:  d/d logP_b = Re(E_b - <E>)/ns + T (logP_b - <logP>)/ns,  d/d phase_b = 2 Im(E_b - <E>)/ns
                                                                     (J1J2/TrainingRNN_J1J2.py:197, 2DHeisenberg/Run_RNNWF_2DHeis.py:199)
    """
    d = eloc - mean_e
    if complex_head:
        s_logp = d.real / n_global
        if T != 0.0:
            s_logp = s_logp + T * (logp - mean_logp) / n_global
        return s_logp.contiguous(), (2.0 * d.imag / n_global).contiguous()
    s = d.real if d.is_complex() else d
    if free_energy and T != 0.0:
        s = s + T * (logp - mean_logp)
    return (s / n_global).contiguous(), None


class TFAdam:
    """tf.train.AdamOptimizer semantics: lr_t = lr*sqrt(1-b2^t)/(1-b1^t); theta -= lr_t*m/(sqrt(v)+eps)."""

    def __init__(self, numel, device, beta1=0.9, beta2=0.999, eps=1e-8):
        self.m = torch.zeros(numel, dtype=torch.float32, device=device)
        self.v = torch.zeros(numel, dtype=torch.float32, device=device)
        self.b1, self.b2, self.eps, self.t = beta1, beta2, eps, 0

    def step(self, params, grad, lr, clip=0.0):
        self.t += 1
        if params.is_cuda:
            # device parameters: one fused kernel of the C-ABI library (rnnwf_adam_step)
            import ctypes as C
            from ._lib import lib, check
            g = grad if (grad.dtype == torch.float32 and grad.is_contiguous()) else grad.float().contiguous()
            check(lib.rnnwf_adam_step(C.c_void_p(params.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(self.m.data_ptr()),
                                      C.c_void_p(self.v.data_ptr()), params.numel(), float(lr), self.b1, self.b2, self.eps,
                                      self.t, float(clip or 0.0), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                  "rnnwf_adam_step")
            return
        # host tensors (the gloo tests of the data-parallel logic): the same arithmetic as torch ops
        if clip:
            grad = grad.clamp(-clip, clip)
        self.m.mul_(self.b1).add_(grad, alpha=1 - self.b1)
        self.v.mul_(self.b2).addcmul_(grad, grad, value=1 - self.b2)
        lr_t = lr * math.sqrt(1 - self.b2 ** self.t) / (1 - self.b1 ** self.t)
        params.addcdiv_(self.m, self.v.sqrt().add_(self.eps), value=-lr_t)


class VMC:
    """Sample -> E_loc -> gradient -> update for one engine and one Hamiltonian."""

    def __init__(self, engine, ham, ham_params, numsamples, lr=1e-3, seed=111, T=0.0, free_energy=False,
                 prefix_reuse=True, clip=None, symmetry=None):
        self.eng, self.ham, self.ham_params = engine, ham, ham_params
        self.ns = int(numsamples)
        self.rank = dist.get_rank() if _dist_on() else 0
        self.world = dist.get_world_size() if _dist_on() else 1
        self.seed, self.T, self.free_energy, self.lr, self.clip = seed, T, free_energy, lr, clip
        # symmetry: an engine.SymmetryGroup -> the wavefunction is the symmetry-projected one (local energies with
        # rnnwf_local_energies_sym, gradient through the G copies of every sample and rnnwf_symmetry_seeds)
        self.sym = symmetry
        self.saved = engine.new_saved(self.ns * (symmetry.G if symmetry is not None else 1))
        self.prefix_reuse = prefix_reuse
        self.opt = TFAdam(engine.num_params, engine.device)
        self.it = 0
        self.last = {}

    def sample_and_energies(self, uniforms=None):
        eng = self.eng
        offset = (self.it * self.world + self.rank) * self.ns   # disjoint Philox rows per step and rank
        if self.sym is not None:
            s, _, _ = eng.sample(self.ns, uniforms=uniforms, seed=self.seed, row_offset=offset)
            e, lp, ph, cnt = eng.local_energies_sym(self.sym, self.ham, self.ham_params, s, return_count=True)
            return s, lp, ph, e, cnt
        s, lp, ph = eng.sample(self.ns, uniforms=uniforms, seed=self.seed, row_offset=offset, saved=self.saved)
        e, cnt = eng.local_energies(self.ham, self.ham_params, s, lp, ph,
                                    saved=self.saved if self.prefix_reuse else None, return_count=True)
        return s, lp, ph, e, cnt

    def _moments_and_seeds(self, lp, e):
        """Energy moments (all-reduced) and the backward seeds: two library kernels around the one collective
        (rnnwf_energy_moments, rnnwf_cost_seeds) instead of a chain of eager tensor ops."""
        import ctypes as C
        from ._lib import lib, check
        eng = self.eng
        ns = e.shape[0]
        dev = eng.device
        er = torch.view_as_real(e.to(torch.complex128)).contiguous() if e.is_complex() else torch.stack([e.double(), torch.zeros_like(e, dtype=torch.float64)], dim=-1).contiguous()
        lpc = None if lp is None else lp.to(device=dev, dtype=torch.float64).contiguous()
        mom = torch.empty(8, dtype=torch.float64, device=dev)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        pt = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        check(lib.rnnwf_energy_moments(ns, pt(er), pt(lpc), pt(mom), st), "rnnwf_energy_moments")
        allreduce_(mom)
        sl = torch.empty(ns, dtype=torch.float64, device=dev)
        sp = torch.empty(ns, dtype=torch.float64, device=dev) if eng.complex_head else None
        stats = torch.empty(4, dtype=torch.float64, device=dev)
        check(lib.rnnwf_cost_seeds(ns, pt(er), pt(lpc), pt(mom), int(eng.complex_head), float(self.T), int(bool(self.free_energy)),
                                   pt(sl), pt(sp), pt(stats), st), "rnnwf_cost_seeds")
        return sl, sp, torch.view_as_complex(stats[0:2]), stats[2]

    def gradient(self, s, lp, ph, e):
        eng = self.eng
        if e.is_cuda:
            sl, sp, mean_e, var = self._moments_and_seeds(lp, e)
        else:
            extra = lp if (self.T != 0.0) else None
            n, mean_e, var, mean_lp = energy_moments(e, extra)
            sl, sp = seeds(e, mean_e, n, eng.complex_head, logp=lp, mean_logp=mean_lp, T=self.T, free_energy=self.free_energy)
        if self.sym is not None:
            # d cost / d (symmetrised log P, phase) -> seeds of the G copies -> one backward over the G * ns batch
            _, _, lp_g, ph_g, copies = eng.logpsi_sym(self.sym, s, saved=self.saved)
            cplx = eng.complex_head
            sl_g, sp_g = eng.symmetry_seeds(lp_g, ph_g if cplx else None, self.sym.characters, sl, sp)
            grad = eng.backward(copies, self.saved, sl_g.reshape(-1), None if sp_g is None else sp_g.reshape(-1))
        else:
            grad = eng.backward(s, self.saved, sl, sp)
        allreduce_(grad)
        return grad, mean_e[0] if mean_e.dim() else mean_e, var

    def step(self, uniforms=None, lr=None, update=True):
        s, lp, ph, e, cnt = self.sample_and_energies(uniforms)
        grad, mean_e, var = self.gradient(s, lp, ph, e)
        if update:
            if self.clip is not None:
                grad.clamp_(-self.clip, self.clip)
            self.opt.step(self.eng.params, grad, self.lr if lr is None else lr)
            self.eng.set_weights()
        self.it += 1
        self.last = dict(samples=s, logp=lp, phase=ph, eloc=e, nconn=cnt, grad=grad)
        return mean_e, var


def estimate_energy(engine, ham, ham_params, numsamples_estimation, numbatches, nsites, seed=111, row_offset=0,
                    prefix_reuse=True, return_samples=False):
    """Final-estimate pass: `numsamples_estimation` fresh samples in batches of `numbatches`, their
    local energies, and  meanE/N, sqrt(var/n)/N  exactly as 2DHeisenberg/Run_RNNWF_2DHeis.py:356-387 and
    RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:299-322 compute them (population variance, n = all samples of all ranks).
    Pure reuse of the sampling and E_loc kernels; under torch.distributed every rank draws disjoint Philox rows and the
    moments are all-reduced.  Returns dict(mean, err, var, n[, eloc, samples]) with per-site mean / error."""
    nb = int(numsamples_estimation) // int(numbatches)
    rank = dist.get_rank() if _dist_on() else 0
    world = dist.get_world_size() if _dist_on() else 1
    saved = engine.new_saved(numbatches) if prefix_reuse else None
    elocs, samples = [], []
    for i in range(nb):
        off = row_offset + (i * world + rank) * int(numbatches)
        s, lp, ph = engine.sample(numbatches, seed=seed, row_offset=off, saved=saved)
        elocs.append(engine.local_energies(ham, ham_params, s, lp, ph, saved=saved))
        if return_samples:
            samples.append(s)
    e = torch.cat(elocs)
    n, mean_e, var, _ = energy_moments(e)
    out = dict(mean=float(mean_e.real) / nsites,
               err=math.sqrt(float(var) / float(n)) / nsites, var=float(var), n=int(n), eloc=e)
    if return_samples:
        out["samples"] = torch.cat(samples)
    return out


def site_densities(samples):
    """Rydberg ground-state densities of the final-estimate pass (Run_RNNWF_Rydberg.py:331-336): per-site mean and
    standard error, overall mean density and its error (std of the per-sample mean density / sqrt(n))."""
    x = samples.to(torch.float64)
    n = x.shape[0]
    dens = x.mean(0)
    dens_err = x.std(0, unbiased=False) / math.sqrt(n)
    per_sample = x.reshape(n, -1).mean(1)
    return dens, dens_err, float(x.mean()), float(per_sample.std(unbiased=False)) / math.sqrt(n)
