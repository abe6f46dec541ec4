"""Clip-sharded multi-GPU driver logic: clips are independent units, so rank r
of W processes the contiguous range shard_range(n, r, W) with no data-path collective; the only
exchange is the final gather of (label int32, score fp32) = 8 B per clip.

`torch.distributed` is plumbing here (NCCL on GPUs, gloo in the CPU tests).  This module is bench / test
infrastructure and lives outside the product package on purpose: the product (libscrubjay_cuda + the ctypes mirror
dsp_b200) has no torch in it; a C host exchanges the 64-byte handles itself (examples/multi_gpu_gather.c)."""
from dsp_b200 import shard_range


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
        if exchange is None:
            def exchange(item):
                out = [None] * self.world
                dist.all_gather_object(out, item)
                return out
        # Every step below is collective: a rank that fails locally still takes part in both exchanges (with its
        # error text), so that all ranks raise together instead of one of them waiting for the others for ever.
        self.l_addr = self.s_addr = None
        self.opened = []
        mine = err = None
        try:
            self.l_addr, l_h = sj.peer_alloc(self.total * 4)
            self.s_addr, s_h = sj.peer_alloc(self.total * 4)
            mine = (l_h, s_h)
        except Exception as e:
            err = "rank %d: %s" % (self.rank, e)
        handles = exchange(mine if err is None else ("error", err))
        bad = [h[1] for h in handles if h[0] == "error"]
        pl, ps = [], []
        if not bad:
            try:
                for r, (hl, hs) in enumerate(handles):
                    if r == self.rank:
                        pl.append(self.l_addr)
                        ps.append(self.s_addr)
                    else:
                        a = sj.peer_open(hl)
                        self.opened.append(a)
                        b = sj.peer_open(hs)
                        self.opened.append(b)
                        pl.append(a)
                        ps.append(b)
                sj.gather_bind(pl, ps, self.base)
            except Exception as e:
                err = "rank %d: %s" % (self.rank, e)
        bad += [m for m in exchange(err) if m]
        if bad:
            self._release()
            raise RuntimeError("peer gather unavailable: " + "; ".join(bad))
        self.peer_labels, self.peer_score = pl, ps
        self._views = None

    def _release(self):
        for a in self.opened:
            try:
                self.sj.peer_close(a)
            except Exception:
                pass
        self.opened = []
        for a in (self.l_addr, self.s_addr):
            if a:
                try:
                    self.sj.peer_free(a)
                except Exception:
                    pass
        self.l_addr = self.s_addr = None

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
        self.dist.barrier()  # nobody maps these arrays any more
        self._views = None
        self._release()
