"""Clip-sharded multi-GPU driver logic: clips are independent units, so rank r
of W processes the contiguous range shard_range(n, r, W) with no data-path collective; the only
exchange is the final gather of (label int32, score fp32) = 8 B per clip.

`torch.distributed` is plumbing here (NCCL on GPUs, gloo in the CPU tests)."""
from . import shard_range


def gather_clip_results(labels, score, n_total, dist=None):
    """labels/score: this rank's 1-D tensors for its shard (sizes may differ by one between ranks).
    Returns (all_labels, all_score) of length n_total in global clip order, on every rank."""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        assert labels.numel() == n_total
        return labels, score
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    cap = max(hi - lo for lo, hi in sizes)
    lo, hi = sizes[rank]
    assert labels.numel() == hi - lo and score.numel() == hi - lo

    def pad(t):
        out = torch.zeros(cap, dtype=t.dtype, device=t.device)
        out[:t.numel()] = t
        return out

    gl = torch.empty(world * cap, dtype=labels.dtype, device=labels.device)
    gs = torch.empty(world * cap, dtype=score.dtype, device=score.device)
    dist.all_gather_into_tensor(gl, pad(labels))
    dist.all_gather_into_tensor(gs, pad(score))
    all_l = torch.cat([gl[r * cap:r * cap + (b - a)] for r, (a, b) in enumerate(sizes)])
    all_s = torch.cat([gs[r * cap:r * cap + (b - a)] for r, (a, b) in enumerate(sizes)])
    return all_l, all_s
