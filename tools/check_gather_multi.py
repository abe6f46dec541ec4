#!/usr/bin/env python3
"""Fused NVLink gather against the NCCL all-gather it replaces, one process per GPU:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/check_gather_multi.py [clips_per_gpu]
Every rank runs its shard of the synthetic set twice -- plain kernel + two NCCL all-gathers, and the gather kernel
that stores into every rank's arrays through peer mappings -- checks that its complete arrays are identical, and
rank 0 prints one JSON line with the step times (CUDA events, max over ranks)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import dsp_b200
from sharded_torch import PeerGather

C = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
NS, STEPS = 16000, 5
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
sj = dsp_b200.ScrubJay(local, 16000)
stream = torch.cuda.current_stream()
sj.set_stream(stream.cuda_stream)
clips = torch.empty((C, NS), device=dev)
sj.synth_clips_dev(clips, rank * C, C, NS)
lab = torch.empty(C, dtype=torch.int32, device=dev); sc = torch.empty(C, device=dev)
all_l = torch.empty(world * C, dtype=torch.int32, device=dev); all_s = torch.empty(world * C, device=dev)


def step_nccl():
    sj.infer_batch_dev(clips, C, NS, lab, None, sc, None)
    dist.all_gather_into_tensor(all_l, lab)
    dist.all_gather_into_tensor(all_s, sc)


g = PeerGather(sj, C, dist)
g.labels.fill_(-1); g.score.fill_(float("nan"))
torch.cuda.synchronize(); dist.barrier()


def step_p2p():
    sj.infer_batch_gather_dev(clips, C, NS)


def timed(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(STEPS):
        fn()
    e1.record(stream)
    torch.cuda.synchronize(); dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / STEPS], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


ms_nccl = timed(step_nccl)
ms_p2p = timed(step_p2p)
ok = bool(torch.equal(g.labels, all_l)) and bool(torch.equal(g.score, all_s))
t = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"n_gpus": world, "clips_per_gpu": C, "identical_on_all_ranks": bool(t.item()),
                      "ms_per_step_nccl_allgather": ms_nccl, "ms_per_step_fused_peer_stores": ms_p2p,
                      "clips_per_s_nccl": world * C / ms_nccl * 1e3, "clips_per_s_fused": world * C / ms_p2p * 1e3,
                      "positives": int((all_l == 1).sum().item())}))
g.close()
sj.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if t.item() else 1)
