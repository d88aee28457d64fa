"""CPU, world_size 2 over gloo: the data-parallel host logic of vmc.py (global moments, seeds, gradient
all-reduce) gives on two shards what one process gives on the whole batch."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rnnwavefunctions_b200 import vmc


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, E_all, lp_all, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = E_all.shape[0] // world
    E, lp = E_all[rank * n:(rank + 1) * n], lp_all[rank * n:(rank + 1) * n]
    ntot, mean_e, var, mean_lp = vmc.energy_moments(E, lp)
    sl, sp = vmc.seeds(E, mean_e, ntot, True, logp=lp, mean_logp=mean_lp, T=0.4)
    g = torch.stack([sl.sum(), sp.sum(), (sl * lp).sum()])   # stand-in for the per-rank gradient
    vmc.allreduce_(g)
    out[rank] = dict(n=float(ntot), mean=complex(mean_e), var=float(var), sl=sl.numpy(), sp=sp.numpy(), g=g.numpy())
    dist.destroy_process_group()


def test_two_rank_moments_and_seeds_match_single_process():
    gen = torch.Generator().manual_seed(0)
    ns = 64
    E = torch.complex(torch.randn(ns, generator=gen, dtype=torch.float64), torch.randn(ns, generator=gen, dtype=torch.float64))
    lp = torch.randn(ns, generator=gen, dtype=torch.float64)
    ntot, mean_e, var, mean_lp = vmc.energy_moments(E, lp)
    sl, sp = vmc.seeds(E, mean_e, ntot, True, logp=lp, mean_logp=mean_lp, T=0.4)
    assert abs(float(var) - float(np.var(E.numpy()))) < 1e-12
    from oracle import costs
    sl_ref, sp_ref = costs.seeds_complex(lp, E, T=0.4)
    np.testing.assert_allclose(sl.numpy(), sl_ref.numpy(), atol=1e-14)
    np.testing.assert_allclose(sp.numpy(), sp_ref.numpy(), atol=1e-14)

    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, E, lp, out), nprocs=2, join=True)
    assert out[0]["n"] == ns and abs(out[0]["mean"] - complex(mean_e)) < 1e-13 and abs(out[0]["var"] - float(var)) < 1e-12
    np.testing.assert_allclose(np.concatenate([out[0]["sl"], out[1]["sl"]]), sl.numpy(), atol=1e-14)
    np.testing.assert_allclose(np.concatenate([out[0]["sp"], out[1]["sp"]]), sp.numpy(), atol=1e-14)
    g_ref = np.array([float(sl.sum()), float(sp.sum()), float((sl * lp).sum())])
    np.testing.assert_allclose(out[0]["g"], g_ref, atol=1e-13)
    np.testing.assert_allclose(out[1]["g"], g_ref, atol=1e-13)


def test_tf_adam_semantics():
    """tf.train.AdamOptimizer: epsilon outside the bias correction."""
    p = torch.tensor([1.0, -2.0])
    g = torch.tensor([0.5, 0.25])
    opt = vmc.TFAdam(2, "cpu")
    opt.step(p, g, lr=0.1)
    m, v = 0.1 * g, 0.001 * g * g
    lr_t = 0.1 * np.sqrt(1 - 0.999) / (1 - 0.9)
    want = torch.tensor([1.0, -2.0]) - lr_t * m / (v.sqrt() + 1e-8)
    np.testing.assert_allclose(p.numpy(), want.numpy(), rtol=1e-6)
