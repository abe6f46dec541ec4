"""Frame-parallel small-batch path (scrubjay_set_latency_mode, default on): the frames of a clip are spread over
warps / SMs and a second kernel forms the statistics in frame order.  It must be BIT-identical to the one-kernel
path: features, label, confidence and score, for float and int16 clips, ragged lengths, the host-pointer call and the
fused-gather call."""
import numpy as np
import pytest
import torch

import dsp_b200

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sj():
    s = dsp_b200.ScrubJay(0, 16000)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    yield s
    s.set_latency_mode(True)
    s.close()


def _run(sj, clips, n, ns, on, pcm16=False):
    sj.set_latency_mode(on)
    lab = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    conf = torch.zeros(n, device="cuda")
    sc = torch.zeros(n, device="cuda")
    ft = torch.zeros((n, 40), device="cuda")
    l0 = sj.launch_count
    (sj.infer_batch_pcm16_dev if pcm16 else sj.infer_batch_dev)(clips, n, ns, lab, conf, sc, ft)
    torch.cuda.synchronize()
    return lab, conf, sc, ft, sj.launch_count - l0


@pytest.mark.parametrize("n,ns", [(1, 16000), (3, 16000), (40, 16000), (500, 16000), (888, 16000), (7, 1025), (5, 2048),
                                  (4, 5000), (2, 20001), (9, 48000)])
def test_bit_identical_to_one_kernel_path(sj, n, ns):
    base = torch.empty((n, 48000), device="cuda")
    sj.synth_clips_dev(base, 17, n, 48000)
    clips = base[:, :ns].contiguous()
    a = _run(sj, clips, n, ns, False)
    b = _run(sj, clips, n, ns, True)
    assert a[4] == 1 and b[4] == 2, "latency mode must take the two-kernel path for %d clips" % n
    for x, y in zip(a[:4], b[:4]):
        assert torch.equal(x, y)


def test_large_batches_and_single_frame_clips_keep_the_one_kernel_path(sj):
    for n, ns in ((889, 16000), (4, 1024), (4, 700)):
        clips = torch.empty((n, ns), device="cuda")
        sj.synth_clips_dev(clips, 0, n, ns)
        assert _run(sj, clips, n, ns, True)[4] == 1


def test_pcm16_and_silence(sj):
    n, ns = 6, 16000
    f = torch.empty((n, ns), device="cuda")
    sj.synth_clips_dev(f, 3, n, ns)
    f[2].zero_()
    f[4, 9000:].zero_()
    pcm = torch.clamp(torch.round(f * 32768.0), -32768, 32767).to(torch.int16)
    a = _run(sj, pcm, n, ns, False, pcm16=True)
    b = _run(sj, pcm, n, ns, True, pcm16=True)
    for x, y in zip(a[:4], b[:4]):
        assert torch.equal(x, y)
    c = _run(sj, f, n, ns, True)
    assert torch.equal(c[3][2], _run(sj, f, n, ns, False)[3][2])  # the log floor of an all-zero clip


def test_host_call_and_gather_call(sj):
    n, ns = 5, 16000
    d = torch.empty((n, ns), device="cuda")
    sj.synth_clips_dev(d, 50, n, ns)
    host = d.cpu().numpy()
    sj.set_latency_mode(False)
    l0, c0, s0 = sj.infer_batch(host)
    sj.set_latency_mode(True)
    l1, c1, s1 = sj.infer_batch(host)
    assert np.array_equal(l0, l1) and np.array_equal(c0, c1) and np.array_equal(s0, s1)
    f1 = sj.mfcc_stats(host[0])
    sj.set_latency_mode(False)
    assert np.array_equal(f1, sj.mfcc_stats(host[0]))
    sj.set_latency_mode(True)
    tl = [torch.full((n + 2,), -7, dtype=torch.int32, device="cuda") for _ in range(2)]
    ts = [torch.full((n + 2,), -7.0, device="cuda") for _ in range(2)]
    sj.gather_bind(tl, ts, 1)
    sj.infer_batch_gather_dev(d, n, ns)
    torch.cuda.synchronize()
    for l, s in zip(tl, ts):
        assert np.array_equal(l[1:n + 1].cpu().numpy(), l0) and np.array_equal(s[1:n + 1].cpu().numpy(), s0)
        assert int(l[0]) == -7 and int(l[n + 1]) == -7
    sj.gather_bind([], None, 0)


def test_zero_copy_path_of_one_to_three_pinned_clips_equals_the_copy_path():
    """One to three clips in PINNED host memory take the zero-copy path of scrubjay_infer_batch (the kernels read the clips
    through their device alias and write into mapped host memory); a pageable copy of the same clips takes the copy path.
    Same bits, for float and int16 clips, with and without features, and for a ragged length; four clips take the copy path
    either way."""
    import torch
    sj = dsp_b200.ScrubJay(0, 16000)
    for n, ns in ((1, 16000), (1, 12345), (2, 16000), (3, 7777), (4, 16000)):
        d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
        sj.set_stream(torch.cuda.current_stream().cuda_stream)
        sj.synth_clips_dev(d, 4711, n, ns)
        torch.cuda.synchronize()
        pinned = d.cpu().pin_memory()
        pageable = pinned.numpy().copy()
        a = sj.infer_batch(pinned.data_ptr(), want_feats=True, n_clips=n, n_samples=ns)
        b = sj.infer_batch(pageable, want_feats=True)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        a2 = sj.infer_batch(pinned.data_ptr(), n_clips=n, n_samples=ns)
        assert np.array_equal(a2[0], b[0]) and np.array_equal(a2[2], b[2])
        pcm = (pinned.numpy() * 32767.0).astype(np.int16)
        pp = torch.from_numpy(pcm).pin_memory()
        c1 = sj.infer_batch_pcm16(pp.data_ptr(), n_clips=n, n_samples=ns)
        c2 = sj.infer_batch_pcm16(pcm)
        for x, y in zip(c1, c2):
            assert np.array_equal(x, y)
    sj.close()


def test_mfcc_stats_of_a_pinned_clip_equals_the_copy_path():
    import torch
    sj = dsp_b200.ScrubJay(0, 16000)
    for ns in (16000, 5000, 40001):
        d = torch.empty((1, ns), dtype=torch.float32, device="cuda")
        sj.set_stream(torch.cuda.current_stream().cuda_stream)
        sj.synth_clips_dev(d, 99, 1, ns)
        torch.cuda.synchronize()
        pinned = d.cpu().pin_memory()
        a = sj.mfcc_stats(pinned.data_ptr(), n_samples=ns)
        b = sj.mfcc_stats(pinned.numpy().copy().reshape(-1))
        assert np.array_equal(a, b)
    sj.close()
