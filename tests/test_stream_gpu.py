"""GPU parity tests of the stream mode (BASELINE config 5: long continuous channels, sliding-frame cepstrum +
SVM) through the C ABI, against the oracle's restatement (tests/oracle_lib.py: stream_infer) on the same
bytes.  Same tolerances as tests/test_cepstrum_gpu.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FEAT_RTOL, FEAT_ATOL = 1e-4, 1e-5


def _ctx():
    import torch
    import dsp_b200
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    return sj


def _channels(sj, n_channels, n_samples, first=0):
    """continuous channels cut from the synthetic generator: channel c = clips first+c*k .. concatenated"""
    import torch
    k = (n_samples + 15999) // 16000
    d = torch.empty((n_channels * k, 16000), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, first, n_channels * k, 16000)
    torch.cuda.synchronize()
    return d.view(n_channels, k * 16000), k * 16000


def _run(sj, d, n_channels, n_samples, stride, W, S, prev=None):
    import torch
    nw = sj.stream_windows(n_samples, W, S)
    labels = torch.full((n_channels, nw), -1, dtype=torch.int32, device="cuda")
    conf = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    score = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    feats = torch.empty((n_channels, nw, 40), dtype=torch.float32, device="cuda")
    sj.stream_infer_dev(d, n_channels, n_samples, stride, labels, conf, score, feats, W, S, prev)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), conf.cpu().numpy(), score.cpu().numpy(), feats.cpu().numpy()


def _check(gpu, host, n_samples, W, S, prev=None):
    from oracle_lib import CepstrumOracle
    ora = CepstrumOracle(16000)
    gl, gc, gs, gf = gpu
    for c in range(host.shape[0]):
        ol, oc, os_, of = ora.stream_infer(host[c, :n_samples], W, S, None if prev is None else prev[c])
        assert of.shape == gf[c].shape
        err = np.abs(gf[c] - of) / (FEAT_RTOL * np.abs(of) + FEAT_ATOL)
        assert err.max() <= 1.0, "channel %d: feature error %.3g x tolerance" % (c, err.max())
        assert np.allclose(gs[c], os_, rtol=1e-4, atol=1e-6)
        decided = np.abs(os_) >= 1e-6
        assert np.array_equal(gl[c][decided], ol[decided])
        assert np.abs(gc[c] - oc)[decided].max() <= 5e-3


@pytest.mark.parametrize("W,S,chunk", [(16, 16, 0), (16, 16, 1), (16, 1, 0), (16, 1, 7), (8, 4, 3), (1, 1, 0), (5, 2, 0)])
def test_stream_windows_match_oracle(W, S, chunk):
    sj = _ctx()
    n_channels, n_samples = 6, 16000 * 6 + 517  # ragged tail: the last hop is zero-padded
    d, stride = _channels(sj, n_channels, n_samples + 16000)
    sj.set_stream_chunk(chunk)
    gpu = _run(sj, d, n_channels, n_samples, stride, W, S)
    assert gpu[0].shape[1] == ((n_samples + 1023) // 1024 - W) // S + 1 and (gpu[0] >= 0).all()
    _check(gpu, d.cpu().numpy(), n_samples, W, S)


def test_first_tumbling_window_equals_clip_mode():
    """window 0 of a stream whose length is a whole number of hops = the clip path on the same 16 hops"""
    import torch
    sj = _ctx()
    n_channels, n_samples = 4, 16384 * 3
    d, stride = _channels(sj, n_channels, n_samples + 16000, first=100)
    gl, gc, gs, gf = _run(sj, d, n_channels, n_samples, stride, 16, 16)
    clips = d[:, :16384].contiguous()
    labels = torch.empty(n_channels, dtype=torch.int32, device="cuda")
    feats = torch.empty((n_channels, 40), dtype=torch.float32, device="cuda")
    sj.infer_batch_dev(clips, n_channels, 16384, labels, None, None, feats)
    torch.cuda.synchronize()
    assert np.array_equal(gf[:, 0], feats.cpu().numpy())


def test_carried_prev_hop_continues_the_stream():
    """a stream processed as two time chunks (second one given the hop before it, overlapping by win - step
    hops) yields the same windows as one pass"""
    import torch
    sj = _ctx()
    W, S = 16, 4
    n_channels, hops = 3, 80
    n_samples = hops * 1024
    d, stride = _channels(sj, n_channels, n_samples, first=300)
    whole = _run(sj, d, n_channels, n_samples, stride, W, S)
    cut = 40  # second chunk starts at hop `cut`; its first window must start at a multiple of S
    prev = d[:, (cut - 1) * 1024:cut * 1024].contiguous()
    part = _run(sj, d[:, cut * 1024:], n_channels, n_samples - cut * 1024, stride, W, S, prev)
    j0 = cut // S
    assert part[3].shape[1] == whole[3].shape[1] - j0
    assert np.array_equal(part[3], whole[3][:, j0:])
    assert np.array_equal(part[0], whole[0][:, j0:])
    _check(part, d[:, cut * 1024:].cpu().numpy(), n_samples - cut * 1024, W, S, prev.cpu().numpy())


def test_stream_rejects_bad_windows():
    import torch
    import dsp_b200
    sj = _ctx()
    d = torch.zeros((1, 32768), dtype=torch.float32, device="cuda")
    lab = torch.empty(64, dtype=torch.int32, device="cuda")
    for W, S in [(0, 1), (17, 1), (4, 5), (4, 0)]:
        with pytest.raises(dsp_b200.ScrubJayError):
            sj.stream_infer_dev(d, 1, 32768, 32768, lab, None, None, None, W, S)
    assert sj.stream_windows(1024 * 15, 16, 16) == 0
    sj.stream_infer_dev(d, 1, 1024 * 15, 32768, lab, None, None, None, 16, 16)  # no window fits: a no-op


@pytest.mark.parametrize("W,S,n_samples,carry", [(16, 16, 16384 * 5, False), (16, 1, 16000 * 4 + 517, False),
                                                 (8, 4, 1024 * 60 + 8, True), (16, 4, 1024 * 48 + 3, False)])
def test_pcm16_stream_equals_float_stream(W, S, n_samples, carry):
    """int16 PCM channels give bit-identical windows to the float entry point on s / 32768 (the conversion is exact
    in fp32), TMA path (rows 16-byte aligned) and the ragged plain-load path alike"""
    import torch
    sj = _ctx()
    n_channels = 5
    g = torch.Generator(device="cpu").manual_seed(1234 + W * 100 + S)
    stride = n_samples + (8 - n_samples % 8) % 8 + (0 if n_samples % 8 == 0 else 3)  # odd stride on the ragged cases
    pcm = torch.randint(-32768, 32768, (n_channels, stride), dtype=torch.int16, generator=g)
    t = torch.arange(stride, dtype=torch.float32)
    tone = (6000 * torch.sin(2 * np.pi * 3100 * t / 16000)).to(torch.int16)
    pcm = (pcm // 8 + tone).to(torch.int16).cuda()
    prev16 = torch.randint(-3000, 3000, (n_channels, 1024), dtype=torch.int16, generator=g).cuda() if carry else None
    f = pcm.to(torch.float32) / 32768.0
    prevf = None if prev16 is None else prev16.to(torch.float32) / 32768.0
    ref = _run(sj, f, n_channels, n_samples, stride, W, S, prevf)
    nw = sj.stream_windows(n_samples, W, S)
    labels = torch.full((n_channels, nw), -1, dtype=torch.int32, device="cuda")
    conf = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    score = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    feats = torch.empty((n_channels, nw, 40), dtype=torch.float32, device="cuda")
    sj.stream_infer_pcm16_dev(pcm, n_channels, n_samples, stride, labels, conf, score, feats, W, S, prev16)
    torch.cuda.synchronize()
    assert np.array_equal(feats.cpu().numpy(), ref[3])
    assert np.array_equal(score.cpu().numpy(), ref[2])
    assert np.array_equal(conf.cpu().numpy(), ref[1])
    assert np.array_equal(labels.cpu().numpy(), ref[0])
