"""GPU parity tests of the GENERIC (table-driven, `FAST = false`) instantiation of the cepstrum kernel: every sample
rate other than 16 kHz -- in particular the 96 kHz / 44.1 kHz material of the reference's own cepstrum/testing and
cepstrum/data folders, whose native rate scrubjay_infer.c:28-30 hands to new_aubio_mfcc -- and every non-default
switch of scrubjay_cepstrum_params.  Float, int16 and stream instantiations, clip lengths incl. ragged ones, against
the CPU oracle built for the same rate / the same parameter block, on the same bytes.  Tolerances as in
tests/test_cepstrum_gpu.py (north_star: 1e-4 relative)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FEAT_RTOL, FEAT_ATOL = 1e-4, 1e-5
RATES = [8000, 22050, 44100, 48000, 96000]


def _ctx(sr, params=None, blob=None):
    import torch
    import dsp_b200
    sj = dsp_b200.ScrubJay(device=0, samplerate=sr, params=params, model_blob=blob)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    return sj


def _clips(sr, n, ns, seed):
    """noise floor + a chirp placed inside the Slaney bank's range (133 Hz .. 6.85 kHz) of this rate, ~6 % silent clips
    and silent tails (exercise the log floor), like the 16 kHz generator"""
    rng = np.random.default_rng(seed)
    t = np.arange(ns) / float(sr)
    x = 0.25 * rng.uniform(-1, 1, (n, ns))
    top = min(6800.0, 0.45 * sr)
    for c in range(n):
        f0, f1 = rng.uniform(150.0, top, 2)
        x[c] += rng.uniform(0.05, 0.5) * np.sin(2 * np.pi * (f0 * t + 0.5 * (f1 - f0) * t * t / max(t[-1], 1e-9)))
        if c % 16 == 5:
            x[c] = 0.0
        if c % 16 == 9:
            x[c, ns // 3:] = 0.0
    return (0.7 * x).astype(np.float32)


def _run_dev(sj, clips_np, pcm16=False):
    import torch
    n, ns = clips_np.shape
    d = torch.from_numpy(clips_np).cuda()
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    conf = torch.empty(n, dtype=torch.float32, device="cuda")
    score = torch.empty(n, dtype=torch.float32, device="cuda")
    feats = torch.empty((n, 40), dtype=torch.float32, device="cuda")
    (sj.infer_batch_pcm16_dev if pcm16 else sj.infer_batch_dev)(d, n, ns, labels, conf, score, feats)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), conf.cpu().numpy(), score.cpu().numpy(), feats.cpu().numpy()


def _check(gpu, ora, what="", atol_scale=1.0):
    gl, gc, gs, gf = gpu
    ol, oc, os_, of = ora
    err = np.abs(gf - of) / (FEAT_RTOL * np.abs(of) + FEAT_ATOL * atol_scale)
    assert err.max() <= 1.0, "%s: feature error %.3g x tolerance at %s" % (what, err.max(), np.unravel_index(err.argmax(), err.shape))
    assert np.allclose(gs, os_, rtol=1e-4, atol=1e-6), (what, np.abs(gs - os_).max())
    decided = np.abs(os_) >= 1e-6
    assert np.array_equal(gl[decided], ol[decided]), what
    assert np.abs(gc - oc)[decided].max() <= 5e-3, what


@pytest.mark.parametrize("sr", RATES)
def test_generic_rate_float_clips_match_oracle(sr):
    """one second of audio at the rate (cepstrum/scrubjay_infer.c:28-30: the file's own rate reaches aubio) and ragged
    lengths: TMA and non-TMA (length not a multiple of 4) forms of the FAST = false instantiation"""
    from oracle_lib import CepstrumOracle
    sj, ora = _ctx(sr), CepstrumOracle(sr)
    for ns in (sr, 4100, 1025, 20481, 999):
        clips = _clips(sr, 48, ns, seed=sr + ns)
        # large batch path (one warp per clip) and the frame-parallel small-batch path
        for lat in (False, True):
            sj.set_latency_mode(lat)
            _check(_run_dev(sj, clips), ora.infer_batch(clips), "sr %d ns %d latency_mode %s" % (sr, ns, lat))


@pytest.mark.parametrize("sr", RATES)
def test_generic_rate_pcm16_clips(sr):
    """int16 instantiation at the rate: bit-identical to the float path on s / 32768, and within tolerance of the oracle"""
    from oracle_lib import CepstrumOracle
    sj, ora = _ctx(sr), CepstrumOracle(sr)
    sj.set_latency_mode(False)
    for ns in (sr, 5003):
        x = _clips(sr, 40, ns, seed=7 * sr + ns)
        pcm = np.clip(np.round(x * 32768.0), -32768, 32767).astype(np.int16)
        xf = (pcm.astype(np.float32) / np.float32(32768.0)).astype(np.float32)
        g16, gf = _run_dev(sj, pcm, pcm16=True), _run_dev(sj, xf)
        for a, b in zip(g16, gf):
            assert np.array_equal(a, b), "pcm16 path differs from the float path at sr %d" % sr
        _check(g16, ora.infer_batch(xf), "pcm16 sr %d ns %d" % (sr, ns))


@pytest.mark.parametrize("sr", [22050, 96000])
@pytest.mark.parametrize("W,S", [(16, 16), (8, 3)])
def test_generic_rate_stream_mode(sr, W, S):
    """STREAM instantiation at the rate against the oracle's stream restatement"""
    import torch
    from oracle_lib import CepstrumOracle
    sj, ora = _ctx(sr), CepstrumOracle(sr)
    n_channels, n_samples = 5, 1024 * 40 + 333
    host = _clips(sr, n_channels, n_samples, seed=sr + W)
    d = torch.from_numpy(host).cuda()
    nw = sj.stream_windows(n_samples, W, S)
    labels = torch.full((n_channels, nw), -1, dtype=torch.int32, device="cuda")
    conf = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    score = torch.empty((n_channels, nw), dtype=torch.float32, device="cuda")
    feats = torch.empty((n_channels, nw, 40), dtype=torch.float32, device="cuda")
    sj.stream_infer_dev(d, n_channels, n_samples, n_samples, labels, conf, score, feats, W, S)
    torch.cuda.synchronize()
    for c in range(n_channels):
        o = ora.stream_infer(host[c], W, S)
        g = (labels[c].cpu().numpy(), conf[c].cpu().numpy(), score[c].cpu().numpy(), feats[c].cpu().numpy())
        _check(g, o, "stream sr %d channel %d" % (sr, c))


def _switches():
    import dsp_b200 as d
    return [
        ("symmetric_window", dict(window=d.WINDOW_HANN_SYMMETRIC)),
        ("power_spectrum", dict(spectrum=d.SPECTRUM_POWER)),
        ("log_floor_1e-10", dict(log_floor=1e-10)),
        ("dct_unnormalised", dict(dct_norm=d.DCT_UNNORMALISED)),
        ("no_half_swap", dict(half_swap=0)),
        ("all_feature_switches", dict(window=d.WINDOW_HANN_SYMMETRIC, spectrum=d.SPECTRUM_POWER, log_floor=1e-20,
                                      dct_norm=d.DCT_UNNORMALISED, half_swap=0)),
    ]


@pytest.mark.parametrize("idx", range(6))
@pytest.mark.parametrize("sr", [16000, 48000])
def test_each_parameter_switch_matches_the_oracle_with_the_same_switch(idx, sr):
    """scrubjay_cepstrum_params: one POD consumed by the table builder and by the oracle; every non-default value is
    checked against the oracle built with the same block, and must CHANGE the features where it should (a switch that
    silently does nothing would pass a parity test against itself)"""
    import dsp_b200
    from oracle_lib import CepstrumOracle
    name, kw = _switches()[idx]
    prm = dsp_b200.CepstrumParams(**kw)
    sj, ora = _ctx(sr, params=prm), CepstrumOracle(sr, params=prm)
    sj.set_latency_mode(False)
    clips = _clips(sr, 64, sr, seed=100 + idx)
    g, o = _run_dev(sj, clips), ora.infer_batch(clips)
    # the unnormalised DCT is 2 / sqrt(2 / 40) = 8.9 times the orthonormal one: the absolute part of the tolerance
    # (which covers coefficients near zero) scales with it
    _check(g, o, "%s @ %d" % (name, sr), atol_scale=9.0 if "dct_norm" in kw else 1.0)
    base = _run_dev(_ctx(sr), clips)
    loud = np.abs(clips).max(axis=1) > 0
    changed = np.abs(g[3][loud] - base[3][loud]).max()
    if name == "no_half_swap":  # phase only: |X| and therefore every feature is unchanged (the kernel ignores the flag)
        assert changed == 0.0
    elif name == "log_floor_1e-10":  # only silent stretches reach the floor
        assert np.abs(g[3][~loud] - base[3][~loud]).max() > 1.0
    else:
        # (the symmetric Hann differs from the periodic one by ~1/2047 of a sample: ~1e-3 on the features)
        assert changed > 3e-4, "%s changed nothing" % name


@pytest.mark.parametrize("rule", [0, 1, 2])
def test_label_rules_of_survey_b3(rule):
    """(ii) d > 0 (run.py:46, default), (i) d >= 0 (vote), (iii) argmax of the probability pair: checked on the golden
    SVM vectors (support vectors +- eps, boundary-straddling points) against the oracle with the same rule, and against
    sklearn's predict / predict_proba of the reference's joblib where the rule is sklearn's"""
    import json
    import os
    import torch
    import dsp_b200
    from oracle_lib import CepstrumOracle
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "svm_golden.json")))
    X = np.array(g["X"], np.float32)
    prm = dsp_b200.CepstrumParams(label_rule=rule)
    sj, ora = _ctx(16000, params=prm), CepstrumOracle(16000, params=prm)
    n_gold = X.shape[0]
    # the rules differ only for |d| < ~0.015: add points that straddle d = 0 on segments between golden vectors of
    # opposite sign (bisected with the oracle), spread over +-0.02 around the crossing
    dec_g = np.array(g["decision_function"])
    neg, pos = np.nonzero(dec_g < -0.05)[0], np.nonzero(dec_g > 0.05)[0]
    extra = []
    for a, b in zip(neg[:40], pos[:40]):
        lo, hi = 0.0, 1.0
        for _ in range(30):
            mid = 0.5 * (lo + hi)
            if ora.predict(((1 - mid) * X[a] + mid * X[b]).astype(np.float32))[2] > 0:
                hi = mid
            else:
                lo = mid
        slope = (dec_g[b] - dec_g[a])
        for dd in np.linspace(-0.02, 0.02, 41):
            t = hi + dd / slope
            extra.append(((1 - t) * X[a] + t * X[b]).astype(np.float32))
    X = np.concatenate([X, np.array(extra, np.float32)])
    n = X.shape[0]
    d = torch.from_numpy(X).cuda()
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    conf = torch.empty(n, dtype=torch.float32, device="cuda")
    score = torch.empty(n, dtype=torch.float32, device="cuda")
    sj.predict_batch_dev(d, n, labels, conf, score)
    torch.cuda.synchronize()
    gl, gs = labels.cpu().numpy(), score.cpu().numpy()
    res = [ora.predict(x) for x in X]
    ol = np.array([r[0] for r in res])
    os_ = np.array([r[2] for r in res])
    dec, proba = np.array(g["decision_function"]), np.array(g["predict_proba"])
    assert np.allclose(gs, os_, rtol=1e-4, atol=2e-6)
    near = np.abs(os_) < 0.02
    assert near.sum() > 500, "the added points must populate the band where the rules differ"
    decided = np.abs(os_) > 1e-5  # a float score this close to 0 may round to the other side
    assert np.array_equal(gl[decided], ol[decided]), "GPU labels differ from the oracle under rule %d" % rule
    gold = slice(0, n_gold)
    if rule == 2:
        # the iterative probability sits at exactly (0.5, 0.5) for -0.015 < d < 0.001: there the
        # argmax is a tie -> class 0, where rule (ii) says 1 for 0 < d
        sure = np.abs(proba[:, 1] - proba[:, 0]) > 1e-6
        assert np.array_equal(gl[gold][sure], (proba[sure, 1] > proba[sure, 0]).astype(int))
        assert (gl[decided] != (os_[decided] > 0)).sum() > 0, "rule (iii) must differ from rule (ii) on this set"
    else:
        sure = np.abs(dec) > 1e-5
        assert np.array_equal(gl[gold][sure], np.array(g["predict"])[sure])  # sklearn model.predict (rule ii = rule i off d = 0)
