"""CPU tests: pin the oracle (oracle/cepstrum_oracle.c) against the committed golden vectors.
 * SVM half: sklearn outputs from the reference's own scrubjay_svm.joblib (tools/make_golden.py).
 * MFCC half: an independent float64 numpy/scipy restatement (the reference holds no golden vectors
   for aubio's arithmetic: "parity unpinned")."""
import json
import os

import numpy as np
import pytest

from oracle_lib import CepstrumOracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def oracle():
    return CepstrumOracle(16000, "f64")


def _seeded_clip(seed, n):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 16000.0
    x = 0.25 * rng.uniform(-1, 1, n)
    f0, f1 = rng.uniform(500, 7000, 2)
    x += 0.3 * np.sin(2 * np.pi * (f0 * t + 0.5 * (f1 - f0) * t * t))
    return x.astype(np.float32)


def test_svm_matches_sklearn_golden(oracle):
    g = json.load(open(os.path.join(GOLD, "svm_golden.json")))
    X = np.array(g["X"], np.float32)
    dec, pred, proba = np.array(g["decision_function"]), np.array(g["predict"]), np.array(g["predict_proba"])
    assert 0.05 < pred.mean() < 0.95, "golden set must straddle the decision boundary"
    for i in range(X.shape[0]):
        label, conf, score = oracle.predict(X[i])
        assert abs(score - dec[i]) <= 1e-6 + 1e-6 * abs(dec[i])
        if abs(dec[i]) > 1e-6:
            assert label == pred[i]
            assert abs(conf - proba[i, label]) < 1e-6


def test_mfcc_matches_numpy_golden(oracle):
    g = np.load(os.path.join(GOLD, "mfcc_golden.npz"))
    for k, (seed, n) in enumerate(zip(g["seeds"], g["lens"])):
        feat, frames = oracle.mfcc_stats(_seeded_clip(int(seed), int(n)), want_frames=True)
        assert frames.shape[0] == (n + 1023) // 1024
        assert np.allclose(feat, g["feats"][k], rtol=1e-5, atol=2e-6), np.abs(feat - g["feats"][k]).max()
        if k == 0:
            assert np.allclose(frames, g["frames0"], rtol=1e-5, atol=2e-6)
        if k == 2:
            assert np.allclose(frames, g["frames2"], rtol=1e-5, atol=2e-6)


def test_mfcc_real_clips_golden(oracle):
    """two 1 s / 16 kHz clips of the reference's donut-classifier/try folder (channel 0)"""
    g = np.load(os.path.join(GOLD, "mfcc_golden.npz"))
    for pcm, want in zip(g["real_pcm"], g["real_feats"]):
        feat = oracle.mfcc_stats(pcm.astype(np.float32) / 32768.0)
        assert np.allclose(feat, want, rtol=1e-5, atol=5e-6), np.abs(feat - want).max()


def test_f32_port_close_to_f64(oracle):
    o32 = CepstrumOracle(16000, "f32")
    x = _seeded_clip(11, 16000)
    a, b = oracle.mfcc_stats(x), o32.mfcc_stats(x)
    assert np.allclose(a, b, rtol=1e-4, atol=1e-5)


def test_frame_count_and_empty(oracle):
    """do/while of cepstrum/scrubjay_infer.c:39-53: ceil(n/1024) frames; empty -> -1 (:55-59)"""
    assert oracle.mfcc_stats(np.zeros(0, np.float32)) is None
    for n, want in [(1, 1), (1023, 1), (1024, 1), (1025, 2), (16000, 16), (16384, 16)]:
        _, fr = oracle.mfcc_stats(np.ones(n, np.float32) * 0.1, want_frames=True)
        assert fr.shape[0] == want


def test_silence_hits_log_floor(oracle):
    feat, fr = oracle.mfcc_stats(np.zeros(4096, np.float32), want_frames=True)
    assert abs(fr[0, 0] - np.log10(np.float32(2e-42)) * np.sqrt(40)) < 1e-3
    assert np.abs(fr[:, 1:]).max() < 1e-4 and np.all(feat[20:] == 0)


def test_slaney_bank_shape(oracle):
    fb = oracle.fbank()
    assert np.count_nonzero(fb) == 1654
    nz = np.nonzero(fb.sum(0))[0]
    assert nz[0] == 18 and nz[-1] == 877
    assert (fb > 0).sum(0).max() == 2


def test_slaney_bank_against_an_independent_implementation(oracle):
    """Second opinion on the restated aubio filterbank (the reference's own library is absent): torchaudio's
    Slaney-scale, Slaney-normalised bank -- a different code lineage (librosa / Auditory Toolbox) -- between the same edge
    frequencies.  aubio keeps the toolbox's constants (13 linear filters from 133.33 Hz, then a ratio of 1.0711703, which
    puts the knee at 999.8 Hz), librosa's scale has its knee at exactly 1000 Hz, so centres differ by up to ~0.2 Hz and
    weights by a few 1e-5; what must agree exactly is the structure: which bins each filter touches."""
    import torchaudio as ta  # mandatory: a missing second opinion must fail, not skip
    fb = oracle.fbank()
    lo = 133.3333
    top = (lo + 12 * 66.66666666) * 1.0711703 ** 29  # upper edge of filter 39
    ref = ta.functional.melscale_fbanks(n_freqs=1025, f_min=lo, f_max=float(top), n_mels=40, sample_rate=16000,
                                        norm="slaney", mel_scale="slaney").numpy().T
    assert np.array_equal(ref > 0, fb > 0)
    assert np.abs(ref - fb).max() < 0.005 * fb.max()
    assert abs(ref.sum() - fb.sum()) < 1e-4 * fb.sum()
