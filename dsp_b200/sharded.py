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


class _DevView:
    """Wraps a raw device address for torch.as_tensor (CUDA array interface v3)."""

    def __init__(self, addr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (addr, False), "version": 3,
                                         "strides": None}


def gather_layout(n_per_rank, rank, world):
    """Weak-scaling layout of the fused gather: every rank owns `n_per_rank` consecutive result slots;
    -> (total slots, first slot of `rank`)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return n_per_rank * world, n_per_rank * rank


class PeerGather:
    """Result arrays of the fused multi-GPU gather (include/scrubjay_cuda.h, scrubjay_gather_bind): every rank
    allocates labels[int32] + score[fp32] for ALL clips of the job, exchanges the CUDA IPC handles through
    torch.distributed (plumbing), maps the other ranks' arrays and binds the list; afterwards
    ScrubJay.infer_batch_gather_dev stores each result into every rank's arrays from inside the kernel.

    exchange(list_of_bytes) -> list over ranks; the default uses dist.all_gather_object."""

    def __init__(self, sj, n_per_rank, dist, exchange=None):
        self.sj, self.dist = sj, dist
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        if self.world > 8:
            raise ValueError("at most 8 peers (one NVSwitch domain)")
        self.total, self.base = gather_layout(n_per_rank, self.rank, self.world)
        self.l_addr, l_h = sj.peer_alloc(self.total * 4)
        self.s_addr, s_h = sj.peer_alloc(self.total * 4)
        if exchange is None:
            def exchange(item):
                out = [None] * self.world
                dist.all_gather_object(out, item)
                return out
        handles = exchange((l_h, s_h))
        self.opened = []
        pl, ps = [], []
        for r, (hl, hs) in enumerate(handles):
            if r == self.rank:
                pl.append(self.l_addr)
                ps.append(self.s_addr)
            else:
                a, b = sj.peer_open(hl), sj.peer_open(hs)
                self.opened += [a, b]
                pl.append(a)
                ps.append(b)
        self.peer_labels, self.peer_score = pl, ps
        sj.gather_bind(pl, ps, self.base)
        self._views = None

    def _view(self):
        if self._views is None:
            import torch
            dev = torch.device("cuda", self.sj.device)
            self._views = (torch.as_tensor(_DevView(self.l_addr, self.total, "<i4"), device=dev),
                           torch.as_tensor(_DevView(self.s_addr, self.total, "<f4"), device=dev))
        return self._views

    @property
    def labels(self):
        """this rank's complete label array [world * n_per_rank] int32 as a torch view (no copy)"""
        return self._view()[0]

    @property
    def score(self):
        return self._view()[1]

    def close(self):
        """Collective: every rank must have stopped writing before anybody frees (the caller synchronises its
        device first)."""
        self.dist.barrier()
        self.sj.gather_bind([], None, 0)
        for a in self.opened:
            self.sj.peer_close(a)
        self.opened = []
        self.dist.barrier()
        self._views = None
        self.sj.peer_free(self.l_addr)
        self.sj.peer_free(self.s_addr)
