"""CPU, world_size 2 over gloo: the N>1 path of the trainer -- cameras sharded by rank
(synth.shard_cameras), every rank accumulating with the GLOBAL S = 2 x total cameras, one all-reduce
(sum) of the packed running means [var | loc | sh | scale | opacity | rot] -- must reproduce the
single-rank running means (accumulateGradients is linear, reference src/Trainer.cu:47-77).
The per-view work runs on the CPU oracle here; the same exchange runs over NCCL in bench.py."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import golden_cases as GC
from oracle_train import oracle_running_means

gsb = GC.gsb


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = GC.load("trainer_t2")
    c, cloud, _teacher, project, cams = GC.trainer_inputs("t2")
    n1 = project.sphere1.count
    mine = gsb.synth.shard_cameras(cams, n1, rank, world)
    idx = [cams.index(x) for x in mine]
    avg = oracle_running_means(c, cloud, mine, g["truthW"][idx], g["truthB"][idx], samples=2 * len(cams))
    packed = torch.from_numpy(np.concatenate(avg))
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)          # the one exchange step of an iteration
    if rank == 0:
        q.put(packed.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_reproduce_single_rank_running_means():
    g = GC.load("trainer_t2")
    assert g is not None
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    packed = q.get(timeout=300)
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    ref = np.concatenate([g["avg_" + k] for k in ("var", "loc", "sh", "scale", "opacity", "rot")]).astype(np.float64)
    rel = np.linalg.norm(packed.astype(np.float64) - ref) / np.linalg.norm(ref)
    assert rel <= 1e-3, rel
    # and the single-rank oracle result agrees to float rounding (only the summation order differs)
    c, cloud, _t, project, cams = GC.trainer_inputs("t2")
    single = np.concatenate(oracle_running_means(c, cloud, cams, g["truthW"], g["truthB"]))
    np.testing.assert_allclose(packed, single, rtol=1e-4, atol=1e-6)
