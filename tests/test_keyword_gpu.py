"""GPU parity tests of the keyword ("stop") detector through the C ABI against the oracle (pinned to the
unmodified reference) and the committed reference golden fixture.
Tolerances: MFCCs are DCT sums of dB values (|values| up to ~600): |gpu - ref| <= 1e-4*|ref| + 5e-3.  The
absolute term is the reference's own accuracy: its FFT builds twiddles by a float recurrence (mfcc.c:58-85), and
against a float64 evaluation of the same formulas it is off by up to 3.2e-3 on these signals (measured with
numpy; the kernel uses exact twiddles and sits closer to the float64 values than the reference does).
Probability 1e-4 absolute; frame counts identical."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "keyword_golden.npz")
CASES = {"a": (6, 80000, 1), "b": (4, 16000, 2), "c": (2, 400, 3), "d": (2, 399, 4), "e": (2, 81000, 5), "f": (3, 23456, 6)}


def _ctx():
    import torch
    import dsp_b200
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.keyword_load()
    return sj


def _gpu(sj, x):
    import torch
    n, ns = x.shape
    d = torch.from_numpy(x).cuda()
    prob = torch.empty(n, dtype=torch.float32, device="cuda")
    mfcc = torch.full((n, 500, 13), 7.0, dtype=torch.float32, device="cuda")
    nfr = torch.empty(n, dtype=torch.int32, device="cuda")
    sj.keyword_classify_batch_dev(d, n, ns, prob, mfcc, nfr)
    torch.cuda.synchronize()
    return prob.cpu().numpy(), mfcc.cpu().numpy(), nfr.cpu().numpy()


def _close(a, b):
    err = np.abs(a - b) / (1e-4 * np.abs(b) + 5e-3)
    return err.max() if err.size else 0.0


@pytest.mark.parametrize("tag", sorted(CASES))
def test_matches_reference_golden(tag):
    from keyword_lib import test_signals as make_signals
    g = np.load(GOLD)
    n, ns, seed = CASES[tag]
    sj = _ctx()
    prob, mfcc, nfr = _gpu(sj, make_signals(n, ns, seed))
    assert np.array_equal(nfr, g["nfr_" + tag])
    assert _close(mfcc, g["mfcc_" + tag]) <= 1.0, _close(mfcc, g["mfcc_" + tag])
    assert np.abs(prob - g["prob_" + tag]).max() <= 1e-4


def test_batch_matches_oracle_and_host_path():
    from keyword_lib import KeywordOracle, test_signals as make_signals
    sj = _ctx()
    x = make_signals(300, 32000, 21)
    prob, mfcc, nfr = _gpu(sj, x)
    ora = KeywordOracle()
    worst = 0.0
    for i in range(0, 300, 7):
        p, mf, nf = ora.classify(x[i])
        assert nf == nfr[i]
        worst = max(worst, _close(mfcc[i], mf))
        assert abs(p - prob[i]) <= 1e-4
    assert worst <= 1.0, worst
    assert np.abs(ora.classify_batch(x) - prob).max() <= 1e-4
    assert np.array_equal(sj.keyword_classify_batch(x), prob)


def test_drop_in_symbols_and_errors(monkeypatch):
    import ctypes
    import dsp_b200
    from keyword_lib import MODEL, KeywordOracle, test_signals as make_signals
    x = make_signals(1, 48000, 5)[0]
    monkeypatch.setenv("SCRUBJAY_KEYWORD_MODEL", MODEL)
    L = dsp_b200.load_library()
    p = L.classify_signal(x.ctypes.data, x.size)
    assert abs(p - KeywordOracle().classify(x)[0]) <= 1e-4
    assert L.classify_signal_binary(x.ctypes.data, x.size, ctypes.c_float(0.5)) == int(p > 0.5)
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.keyword_classify_batch(x[None, :])  # no model loaded
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.keyword_load(b"not a model")
