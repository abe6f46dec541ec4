"""Host logic of the fused multi-GPU gather (tools/sharded_torch.py: PeerGather) on CPU, world_size 2 over gloo.
The CUDA side (peer allocations, IPC mapping, the kernel's stores) is replaced by a recording stand-in that
"stores" through /dev/shm files, so what is under test is the handle exchange, the rank order of the target
list, the slot layout and the close protocol.  The kernel side is tests/test_gather_gpu.py."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_PER_RANK = 7


class FakeCtx:
    """peer_alloc -> a numpy memmap in a shared directory, handle = its file name; peer_open maps the file."""

    def __init__(self, rank, shm):
        self.rank, self.shm, self.device = rank, shm, 0
        self.maps, self.n, self.bound, self.closed, self.freed = {}, 0, None, [], []

    def peer_alloc(self, nbytes, want_handle=True):
        name = "r%d_%d.bin" % (self.rank, self.n)  # the 64-byte handle carries only the file name
        self.n += 1
        m = np.memmap(os.path.join(self.shm, name), dtype=np.uint8, mode="w+", shape=(nbytes,))
        addr = 0x1000 * (len(self.maps) + 1) + self.rank * 0x100000
        self.maps[addr] = m
        return addr, name.encode().ljust(64, b"\0")[:64]

    def peer_open(self, handle):
        name = bytes(handle).rstrip(b"\0").decode()
        m = np.memmap(os.path.join(self.shm, name), dtype=np.uint8, mode="r+")
        addr = 0x9000000 + 0x1000 * len(self.maps)
        self.maps[addr] = m
        return addr

    def peer_close(self, addr):
        self.closed.append(addr)

    def peer_free(self, addr):
        self.freed.append(addr)

    def gather_bind(self, pl, ps, base):
        self.bound = (list(pl), list(ps) if ps is not None else None, base)

    def store(self, i, label, score):  # what lane p of the kernel does for result i
        pl, ps, base = self.bound
        for a, b in zip(pl, ps):
            self.maps[a].view(np.int32)[base + i] = label
            self.maps[b].view(np.float32)[base + i] = score


def _worker(rank, world, port, shm):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sharded_torch import PeerGather
    ctx = FakeCtx(rank, shm)
    g = PeerGather(ctx, N_PER_RANK, dist)
    assert (g.total, g.base) == (N_PER_RANK * world, N_PER_RANK * rank)
    pl, ps, base = ctx.bound
    assert len(pl) == world and pl[rank] == g.l_addr and ps[rank] == g.s_addr and base == g.base
    for i in range(N_PER_RANK):
        ctx.store(i, 100 * rank + i, float(rank) + i / 16.0)
    for m in ctx.maps.values():
        m.flush()
    dist.barrier()
    lab = np.array(ctx.maps[g.l_addr].view(np.int32))
    sco = np.array(ctx.maps[g.s_addr].view(np.float32))
    want_l = np.concatenate([100 * r + np.arange(N_PER_RANK) for r in range(world)])
    want_s = np.concatenate([r + np.arange(N_PER_RANK) / 16.0 for r in range(world)]).astype(np.float32)
    assert np.array_equal(lab, want_l) and np.array_equal(sco, want_s)
    own = sorted([g.l_addr, g.s_addr])
    g.close()
    assert ctx.bound == ([], None, 0)
    assert len(ctx.closed) == 2 * (world - 1) and sorted(ctx.freed) == own
    dist.barrier()
    dist.destroy_process_group()


def _worker_fail(rank, world, port, shm):
    """rank 1 cannot allocate: both ranks must raise (nobody may be left waiting in the exchange)."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sharded_torch import PeerGather

    class Broken(FakeCtx):
        def peer_alloc(self, nbytes, want_handle=True):
            if self.rank == 1:
                raise RuntimeError("out of memory (simulated)")
            return super().peer_alloc(nbytes, want_handle)

    class NoOpen(FakeCtx):
        def peer_open(self, handle):
            if self.rank == 0:
                raise RuntimeError("peer access unsupported (simulated)")
            return super().peer_open(handle)

    for cls, what in ((Broken, "out of memory"), (NoOpen, "peer access unsupported")):
        ctx = cls(rank, shm)
        with pytest.raises(RuntimeError, match=what):
            PeerGather(ctx, N_PER_RANK, dist)
        assert len(ctx.freed) == (0 if (cls is Broken and rank == 1) else 2)  # what was allocated is released
        dist.barrier()
    dist.destroy_process_group()


def test_two_rank_failure_is_collective(tmp_path):
    port = 33500 + os.getpid() % 2000
    mp.spawn(_worker_fail, args=(2, port, str(tmp_path)), nprocs=2, join=True)


def test_two_rank_peer_gather_host_logic(tmp_path):
    port = 31500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)


def test_gather_layout():
    from sharded_torch import gather_layout
    assert gather_layout(1000, 0, 1) == (1000, 0)
    assert gather_layout(1000, 3, 8) == (8000, 3000)
    with pytest.raises(ValueError):
        gather_layout(10, 2, 2)
