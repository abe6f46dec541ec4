"""Fused result gather (scrubjay_gather_bind / scrubjay_infer_batch_gather_dev) on one GPU: the kernel must
store every clip's label and score into EVERY bound target at out_base + i, bit-identical to the plain call,
and touch nothing else.  Two ranks' arrays are stood in for by two allocations on the same device (the
multi-process NVLink form is tools/check_gather_multi.py, run under torchrun)."""
import numpy as np
import pytest
import torch

import dsp_b200

pytestmark = pytest.mark.gpu
N, NS = 301, 16000


@pytest.fixture(scope="module")
def sj():
    s = dsp_b200.ScrubJay(0, 16000)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    yield s
    s.close()


def _plain(sj, clips, n):
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    conf = torch.empty(n, device="cuda")
    sc = torch.empty(n, device="cuda")
    sj.infer_batch_dev(clips, n, NS, lab, conf, sc, None)
    torch.cuda.synchronize()
    return lab, conf, sc


def test_every_target_gets_every_result(sj):
    clips = torch.empty((N, NS), device="cuda")
    sj.synth_clips_dev(clips, 0, N, NS)
    lab, conf, sc = _plain(sj, clips, N)
    base, total = 5, N + 11
    tl = [torch.full((total,), -7, dtype=torch.int32, device="cuda") for _ in range(3)]
    ts = [torch.full((total,), -7.0, device="cuda") for _ in range(3)]
    sj.gather_bind(tl, ts, base)
    conf2 = torch.empty(N, device="cuda")
    sj.infer_batch_gather_dev(clips, N, NS, conf2)
    torch.cuda.synchronize()
    for l, s in zip(tl, ts):
        assert torch.equal(l[base:base + N], lab) and torch.equal(s[base:base + N], sc)
        assert bool((l[:base] == -7).all()) and bool((l[base + N:] == -7).all())
        assert bool((s[:base] == -7).all()) and bool((s[base + N:] == -7).all())
    assert torch.equal(conf2, conf)
    # score targets are optional, one by one
    tl2 = [torch.zeros(total, dtype=torch.int32, device="cuda") for _ in range(2)]
    sj.gather_bind(tl2, [ts[0], None], 0)
    ts[0].fill_(-7.0)
    sj.infer_batch_gather_dev(clips, N, NS)
    torch.cuda.synchronize()
    assert torch.equal(tl2[0][:N], lab) and torch.equal(tl2[1][:N], lab) and torch.equal(ts[0][:N], sc)
    sj.gather_bind([], None, 0)


def test_bind_errors(sj):
    sj.gather_bind([], None, 0)
    clips = torch.zeros((2, NS), device="cuda")
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.infer_batch_gather_dev(clips, 2, NS)          # nothing bound
    t = torch.zeros(4, dtype=torch.int32, device="cuda")
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.gather_bind([t] * (dsp_b200.MAX_PEERS + 1), None, 0)
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.gather_bind([t], None, -1)


def test_peer_alloc_round_trip(sj):
    addr, h = sj.peer_alloc(4096)
    assert addr and len(h) == dsp_b200.PEER_HANDLE_BYTES and any(h)
    from sharded_torch import _DevView
    v = torch.as_tensor(_DevView(addr, 1024, "<i4"), device="cuda")
    v.fill_(3)
    assert int(v.sum().item()) == 3072
    del v
    sj.peer_free(addr)


def test_stream_mode_gather(sj):
    nch, secs, W, S = 6, 9, 16, 8
    ns = secs * NS
    d = torch.empty((nch * secs, NS), device="cuda")
    sj.synth_clips_dev(d, 100, nch * secs, NS)
    nw = sj.stream_windows(ns, W, S)
    lab = torch.empty((nch, nw), dtype=torch.int32, device="cuda")
    sc = torch.empty((nch, nw), device="cuda")
    sj.stream_infer_dev(d, nch, ns, ns, lab, None, sc, None, W, S)
    base = 3 * nw  # as if three channels of other ranks came first
    tl = [torch.full((nch * nw + base + 4,), -7, dtype=torch.int32, device="cuda") for _ in range(2)]
    ts = [torch.full((nch * nw + base + 4,), -7.0, device="cuda") for _ in range(2)]
    sj.gather_bind(tl, ts, base)
    sj.stream_infer_gather_dev(d, nch, ns, ns, win_frames=W, step_frames=S)
    torch.cuda.synchronize()
    for l, s in zip(tl, ts):
        assert torch.equal(l[base:base + nch * nw], lab.reshape(-1)) and torch.equal(s[base:base + nch * nw], sc.reshape(-1))
        assert bool((l[:base] == -7).all()) and bool((l[base + nch * nw:] == -7).all())
    sj.gather_bind([], None, 0)
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.stream_infer_gather_dev(d, nch, ns, ns, win_frames=W, step_frames=S)
