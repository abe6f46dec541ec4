"""GPU tests of the int16 PCM ingest: the pcm16 entry points must be BIT-identical to the float entry points
fed with s / 32768 (the conversion is exact), and the decode kernel must follow the two reference channel
policies (classifier.c:293-297 channel 0; aubio's mean of channels)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx():
    import torch
    import dsp_b200
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    return sj


def _pcm(sj, n, ns, first=0):
    import torch
    k = (ns + 15999) // 16000
    d = torch.empty((n * k, 16000), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, first, n * k, 16000)
    x = d.view(n, k * 16000)[:, :ns]
    return torch.clamp(torch.round(x * 32768.0), -32768, 32767).to(torch.int16).contiguous()


def _both(sj, pcm):
    import torch
    n, ns = pcm.shape
    out = []
    for is16 in (True, False):
        labels = torch.empty(n, dtype=torch.int32, device="cuda")
        conf = torch.empty(n, dtype=torch.float32, device="cuda")
        score = torch.empty(n, dtype=torch.float32, device="cuda")
        feats = torch.empty((n, 40), dtype=torch.float32, device="cuda")
        if is16:
            sj.infer_batch_pcm16_dev(pcm, n, ns, labels, conf, score, feats)
        else:
            x = (pcm.to(torch.float32) / 32768.0).contiguous()
            sj.infer_batch_dev(x, n, ns, labels, conf, score, feats)
        torch.cuda.synchronize()
        out.append([t.cpu().numpy() for t in (labels, conf, score, feats)])
    return out


@pytest.mark.parametrize("ns", [16000, 16384, 16001, 1000, 2051, 40000])
def test_pcm16_bit_identical_to_float_path(ns):
    sj = _ctx()
    pcm = _pcm(sj, 96, ns)
    a, b = _both(sj, pcm)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_pcm16_matches_oracle_and_host_path():
    from oracle_lib import CepstrumOracle
    sj = _ctx()
    pcm = _pcm(sj, 128, 16000, first=700)
    a, _ = _both(sj, pcm)
    host = pcm.cpu().numpy()
    hl, hc, hs, hf = sj.infer_batch_pcm16(host, want_feats=True)
    assert np.array_equal(hl, a[0]) and np.array_equal(hf, a[3]) and np.array_equal(hs, a[2])
    ol, oc, os_, of = CepstrumOracle(16000).infer_batch(host.astype(np.float32) / 32768.0)
    err = np.abs(a[3] - of) / (1e-4 * np.abs(of) + 1e-5)
    assert err.max() <= 1.0
    assert np.array_equal(a[0], ol)


def test_unaligned_pcm16_pointer_takes_the_plain_load_path():
    import torch
    sj = _ctx()
    pcm = _pcm(sj, 8, 16000)
    flat = torch.empty(8 * 16000 + 8, dtype=torch.int16, device="cuda")
    flat[3:3 + 8 * 16000] = pcm.view(-1)
    a, _ = _both(sj, pcm)
    labels = torch.empty(8, dtype=torch.int32, device="cuda")
    feats = torch.empty((8, 40), dtype=torch.float32, device="cuda")
    sj.infer_batch_pcm16_dev(flat[3:].data_ptr(), 8, 16000, labels, None, None, feats)
    torch.cuda.synchronize()
    assert np.array_equal(feats.cpu().numpy(), a[3])


def test_decode_policies():
    import torch
    import dsp_b200
    sj = _ctx()
    rng = np.random.default_rng(5)
    for ch in (1, 2, 3):
        x = rng.integers(-32768, 32768, size=(5000, ch), dtype=np.int16)
        d = torch.from_numpy(x).cuda()
        out = torch.empty(5000, dtype=torch.float32, device="cuda")
        sj.decode_pcm16_dev(d, 5000, ch, dsp_b200.MIX_FIRST, out)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), x[:, 0].astype(np.float32) / np.float32(32768))
        sj.decode_pcm16_dev(d, 5000, ch, dsp_b200.MIX_MEAN, out)
        torch.cuda.synchronize()
        acc = np.zeros(5000, np.float32)
        for c in range(ch):
            acc = acc + x[:, c].astype(np.float32) / np.float32(32768)
        assert np.array_equal(out.cpu().numpy(), acc / np.float32(ch))


def _ingest_oracle():
    import ctypes
    from oracle_lib import ORACLE_DIR, build_oracle
    path = os.path.join(ORACLE_DIR, "libsj_oracle_f32.so")
    if not os.path.exists(path):
        build_oracle()
    L = ctypes.CDLL(path)
    L.sj_oracle_adc12.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p]
    L.sj_oracle_upsample_linear.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    return L


def test_adc12_decode_matches_firmware_arithmetic():
    """sync/sync.cpp:236-241: every one of the 4096 ADC codes (and out-of-range counts, which wrap like the int16 cast)"""
    import torch
    sj = _ctx()
    L = _ingest_oracle()
    adc = np.concatenate([np.arange(4096), [4096, 5000, 65535]]).astype(np.uint16)
    want = np.empty(adc.size, np.float32)
    L.sj_oracle_adc12(adc.ctypes.data, adc.size, want.ctypes.data)
    d = torch.from_numpy(adc.view(np.int16)).cuda()
    out = torch.empty(adc.size, dtype=torch.float32, device="cuda")
    sj.decode_adc12_dev(d, adc.size, out)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
    assert want[0] == -1.0 and want[2048] == 0.0 and want[4095] == np.float32(2047 * 16 / 32768.0)


@pytest.mark.parametrize("old,new", [(9000, 16000), (16000, 16000), (1, 5), (7, 2), (12345, 48000)])
def test_upsample_linear_bit_identical(old, new):
    """microphone/src/microphone.cpp:55-70"""
    import torch
    sj = _ctx()
    L = _ingest_oracle()
    rng = np.random.default_rng(old + new)
    x = rng.uniform(-1, 1, (6, old)).astype(np.float32)
    want = np.empty((6, new), np.float32)
    for i in range(6):
        L.sj_oracle_upsample_linear(x[i].ctypes.data, old, want[i].ctypes.data, new)
    d = torch.from_numpy(x).cuda()
    out = torch.empty((6, new), dtype=torch.float32, device="cuda")
    sj.upsample_linear_dev(d, 6, old, new, out)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
