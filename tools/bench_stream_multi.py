#!/usr/bin/env python3
"""BASELINE config 5 at its named size: long continuous 16 kHz streams (1 h per channel, 4096 channels) through the
sliding-frame cepstrum + SVM, channels sharded over the GPUs of one box (512 channels = 118 GB per GPU at 8 GPUs), no
data-path collective, the per-window labels and scores gathered by the kernel itself (peer stores over NVLink,
tools/sharded_torch.py: PeerGather; SJ_GATHER=nccl runs two NCCL all-gathers after the kernel instead).
usage (one rank per GPU): python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
          tools/bench_stream_multi.py [channels=4096] [seconds=3600] [win=16] [step=16] [reps=3]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import dsp_b200

nch_total = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
secs = int(sys.argv[2]) if len(sys.argv) > 2 else 3600
W = int(sys.argv[3]) if len(sys.argv) > 3 else 16
S = int(sys.argv[4]) if len(sys.argv) > 4 else 16
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
lo, hi = dsp_b200.shard_range(nch_total, rank, world)  # channels owned by this rank
nch = hi - lo
sj = dsp_b200.ScrubJay(local, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
ns = secs * 16000
d = torch.empty((nch * secs, 16000), dtype=torch.float32, device=dev)  # channel c = rows [c*secs, (c+1)*secs)
sj.synth_clips_dev(d, lo * secs, nch * secs, 16000)
nw = sj.stream_windows(ns, W, S)
lab = torch.empty((nch, nw), dtype=torch.int32, device=dev)
sc = torch.empty((nch, nw), dtype=torch.float32, device=dev)
torch.cuda.synchronize()
peer = None
if world > 1 and os.environ.get("SJ_GATHER", "p2p") == "p2p":
    from sharded_torch import PeerGather
    assert nch_total % world == 0, "fused gather: equal channel counts per rank"
    peer = PeerGather(sj, nch * nw, dist)
best = None
for i in range(reps + 1):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if peer is not None:
        sj.stream_infer_gather_dev(d, nch, ns, ns, win_frames=W, step_frames=S)
    else:
        sj.stream_infer_dev(d, nch, ns, ns, lab, None, sc, None, W, S)
    if world > 1 and peer is None:  # the path's only exchange: per-window labels and scores, 8 B per window
        gl = [torch.empty_like(lab) for _ in range(world)]
        gs = [torch.empty_like(sc) for _ in range(world)]
        dist.all_gather(gl, lab)
        dist.all_gather(gs, sc)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if i > 0:
        best = float(t.item()) if best is None else min(best, float(t.item()))
if peer is not None:  # complete arrays on every rank: own slots equal a plain run, total count equals the sum over ranks
    sj.stream_infer_dev(d, nch, ns, ns, lab, None, sc, None, W, S)
    torch.cuda.synchronize()
    dist.barrier()
    assert torch.equal(peer.labels[rank * nch * nw:(rank + 1) * nch * nw], lab.reshape(-1))
    assert torch.equal(peer.score[rank * nch * nw:(rank + 1) * nch * nw], sc.reshape(-1))
pos = lab.sum().to(torch.float64)
if world > 1:
    dist.all_reduce(pos)
if peer is not None:
    assert int(peer.labels.sum().item()) == int(pos.item())
if rank == 0:
    hops = (ns + 1023) // 1024
    print(json.dumps({"workload": "config 5: %d channels x %d s @16 kHz, window %d frames step %d, %d GPU(s)" % (nch_total, secs, W, S, world),
                      "n_gpus": world, "channels_per_gpu": nch, "bytes_per_gpu": nch * ns * 4, "ms": best,
                      "windows_per_s": nch_total * nw / best * 1e3, "frames_per_s": nch_total * hops / best * 1e3,
                      "audio_seconds_per_s": nch_total * secs / best * 1e3, "audio_hours_per_s": nch_total * secs / 3600.0 / best * 1e3,
                      "positive_windows": int(pos.item()), "gather": "fused peer stores (NVLink)" if peer is not None else ("nccl all-gather" if world > 1 else "none"),
                      "timing": "CUDA events incl. the label/score gather, max over ranks"}))
if peer is not None:
    peer.close()
if world > 1:
    dist.destroy_process_group()
