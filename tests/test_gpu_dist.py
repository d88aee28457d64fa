"""GPU, multi-process: the data-parallel VMC step over NCCL gives, on 2 ranks with half the samples each, the
same energies and (to FP32 summation order) the same gradient as one rank with all the samples -- the Philox
row offsets make the union of the shards the single-rank sample stream.  Needs >= 2 GPUs (skipped otherwise)."""
import os
import socket

import numpy as np
import pytest
import torch


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from rnnwavefunctions_b200 import _lib, vmc
    from rnnwavefunctions_b200.engine import Engine
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    eng = Engine(_lib.CELL_TENSGRU2D, 4, 4, num_units=16, complex_head=True, u1=True, seed=5)
    v = vmc.VMC(eng, _lib.HAM_HEIS2D, [1.0], 256 // world, seed=9)
    mean_e, var = v.step(update=False)
    out[rank] = dict(mean=complex(mean_e), var=float(var), grad=v.last["grad"].cpu().numpy(),
                     samples=v.last["samples"].cpu().numpy())
    dist.destroy_process_group()


@pytest.mark.gpu
def test_two_gpu_step_matches_single_gpu(cuda):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from rnnwavefunctions_b200 import _lib, vmc
    from rnnwavefunctions_b200.engine import Engine
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    eng = Engine(_lib.CELL_TENSGRU2D, 4, 4, num_units=16, complex_head=True, u1=True, seed=5)
    v = vmc.VMC(eng, _lib.HAM_HEIS2D, [1.0], 256, seed=9)
    mean_e, var = v.step(update=False)
    s1 = v.last["samples"].cpu().numpy()
    assert (np.concatenate([out[0]["samples"], out[1]["samples"]]) == s1).all()      # bit-identical sample stream
    assert abs(out[0]["mean"] - complex(mean_e)) < 1e-9 and abs(out[0]["var"] - float(var)) < 1e-9
    g1 = v.last["grad"].cpu().numpy()
    for r in (0, 1):
        np.testing.assert_allclose(out[r]["grad"], g1, rtol=0, atol=1e-6 * np.abs(g1).max() + 1e-9)
