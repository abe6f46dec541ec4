"""world_size-2 gloo test (CPU) of the N>1 path: shard by clip, compute locally, gather labels and
scores.  The local compute is the CPU oracle here (a checker standing in for the GPU kernel); what is
under test is the sharding arithmetic and the gather order."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_TOTAL, N_SAMPLES = 13, 4096  # odd on purpose: shards differ by one clip


def _clips():
    rng = np.random.default_rng(5)
    return rng.uniform(-0.4, 0.4, (N_TOTAL, N_SAMPLES)).astype(np.float32)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import dsp_b200
    from sharded_torch import gather_clip_results
    from oracle_lib import CepstrumOracle
    lo, hi = dsp_b200.shard_range(N_TOTAL, rank, world)
    labels, _, score, _ = CepstrumOracle(16000).infer_batch(_clips()[lo:hi])
    al, as_ = gather_clip_results(torch.from_numpy(labels), torch.from_numpy(score), N_TOTAL, dist)
    np.save(os.path.join(out_dir, "l%d.npy" % rank), al.numpy())
    np.save(os.path.join(out_dir, "s%d.npy" % rank), as_.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_and_gather(tmp_path):
    from oracle_lib import CepstrumOracle
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    want_l, _, want_s, _ = CepstrumOracle(16000).infer_batch(_clips())
    for r in range(2):
        assert np.array_equal(np.load(tmp_path / ("l%d.npy" % r)), want_l)
        assert np.array_equal(np.load(tmp_path / ("s%d.npy" % r)), want_s)
