"""GPU parity tests of the cepstrum path: libscrubjay_cuda (through its C ABI) vs the CPU oracle on
the same bytes.  Tolerances (north_star: "within 1e-4 relative"):
  features  |gpu - oracle| <= 1e-4*|oracle| + 1e-5
  score     |gpu - oracle| <= 1e-4*|oracle| + 1e-6
  conf      5e-3 abs (libsvm's iterative probability has a 0.005/k stopping rule)
  labels    identical, except clips whose |score| < 1e-6 (reported, none expected)
"""
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


def _run_dev(sj, clips_np):
    import torch
    n, ns = clips_np.shape
    d = torch.from_numpy(clips_np).cuda()
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    conf = torch.empty(n, dtype=torch.float32, device="cuda")
    score = torch.empty(n, dtype=torch.float32, device="cuda")
    feats = torch.empty((n, 40), dtype=torch.float32, device="cuda")
    sj.infer_batch_dev(d, n, ns, labels, conf, score, feats)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), conf.cpu().numpy(), score.cpu().numpy(), feats.cpu().numpy()


def _check(gpu, ora):
    gl, gc, gs, gf = gpu
    ol, oc, os_, of = ora
    err = np.abs(gf - of) / (FEAT_RTOL * np.abs(of) + FEAT_ATOL)
    assert err.max() <= 1.0, "feature error %.3g x tolerance at %s" % (err.max(), np.unravel_index(err.argmax(), err.shape))
    assert np.allclose(gs, os_, rtol=1e-4, atol=1e-6), np.abs(gs - os_).max()
    decided = np.abs(os_) >= 1e-6
    assert np.array_equal(gl[decided], ol[decided])
    assert np.abs(gc - oc)[decided].max() <= 5e-3


def test_synthetic_clips_match_oracle():
    import torch
    from oracle_lib import CepstrumOracle
    sj = _ctx()
    n, ns = 512, 16000
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 0, n, ns)
    torch.cuda.synchronize()
    clips = d.cpu().numpy()
    assert np.abs(clips).max() < 1.0 and (np.abs(clips).max(axis=1) == 0).any(), "generator must emit silent clips too"
    _check(_run_dev(sj, clips), CepstrumOracle(16000).infer_batch(clips))


@pytest.mark.parametrize("ns", [1, 5, 640, 1023, 1024, 1025, 2048, 4100, 15999, 16000, 16001, 20480])
def test_ragged_lengths(ns):
    """frame count follows the do/while of cepstrum/scrubjay_infer.c:39-53, incl. the TMA-less path
    (lengths that are not a multiple of 4)"""
    from oracle_lib import CepstrumOracle
    rng = np.random.default_rng(ns)
    clips = rng.uniform(-0.5, 0.5, (37, ns)).astype(np.float32)
    _check(_run_dev(_ctx(), clips), CepstrumOracle(16000).infer_batch(clips))


def test_silence_and_floor():
    """an all-zero frame gives exactly log10(2e-42) in all 40 bands (aubio SAFE_LOG10)"""
    from oracle_lib import CepstrumOracle
    clips = np.zeros((4, 16000), np.float32)
    clips[1, :3000] = np.random.default_rng(1).uniform(-0.3, 0.3, 3000)
    clips[2, 9000:] = np.random.default_rng(2).uniform(-0.3, 0.3, 7000)
    clips[3] = 1e-15 * np.random.default_rng(3).uniform(-1, 1, 16000)  # tiny but well above the floor
    g = _run_dev(_ctx(), clips)
    o = CepstrumOracle(16000).infer_batch(clips)
    assert abs(g[3][0, 0] - np.log10(np.float32(2e-42)) * np.sqrt(40)) < 1e-2
    assert np.all(g[3][0, 20:] == 0)
    _check(g, o)


def test_host_api_matches_device_api():
    import dsp_b200
    rng = np.random.default_rng(7)
    clips = rng.uniform(-0.5, 0.5, (300, 16000)).astype(np.float32)
    sj = _ctx()
    gl, gc, gs, gf = _run_dev(sj, clips)
    hl, hc, hs, hf = sj.infer_batch(clips, want_feats=True)
    assert np.array_equal(gl, hl) and np.array_equal(gs, hs) and np.array_equal(gf, hf) and np.array_equal(gc, hc)
    one = sj.mfcc_stats(clips[5])
    assert np.array_equal(one, gf[5])
    label, conf = sj.predict(gf[5])
    assert label == gl[5] and conf == gc[5]
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.mfcc_stats(np.zeros(0, np.float32))  # "Empty file" -> -1, scrubjay_infer.c:55-59


def test_predict_batch_matches_sklearn_golden():
    import json
    import os
    import torch
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "svm_golden.json")))
    X = np.array(g["X"], np.float32)
    sj = _ctx()
    d = torch.from_numpy(X).cuda()
    n = X.shape[0]
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    conf = torch.empty(n, dtype=torch.float32, device="cuda")
    score = torch.empty(n, dtype=torch.float32, device="cuda")
    sj.predict_batch_dev(d, n, labels, conf, score)
    torch.cuda.synchronize()
    dec = np.array(g["decision_function"])
    assert np.allclose(score.cpu().numpy(), dec, rtol=1e-4, atol=2e-6)
    decided = np.abs(dec) > 1e-5
    assert np.array_equal(labels.cpu().numpy()[decided], np.array(g["predict"])[decided])
    proba = np.array(g["predict_proba"])
    pc = proba[np.arange(n), labels.cpu().numpy()]
    assert np.abs(conf.cpu().numpy() - pc)[decided].max() < 5e-3


def test_round_trip_properties_large():
    """size-independent properties on a large batch: determinism, and permutation equivariance
    (clips are independent units: shuffling clips shuffles results bit-for-bit)"""
    import torch
    sj = _ctx()
    n, ns = 20000, 16000
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 12345, n, ns)
    outs = []
    perm = torch.randperm(n, device="cuda")
    for src in (d, d, d[perm].contiguous()):
        labels = torch.empty(n, dtype=torch.int32, device="cuda")
        score = torch.empty(n, dtype=torch.float32, device="cuda")
        feats = torch.empty((n, 40), dtype=torch.float32, device="cuda")
        sj.infer_batch_dev(src, n, ns, labels, None, score, feats)
        torch.cuda.synchronize()
        outs.append((labels, score, feats))
    assert torch.equal(outs[0][2], outs[1][2]) and torch.equal(outs[0][1], outs[1][1])
    assert torch.equal(outs[0][2][perm], outs[2][2]) and torch.equal(outs[0][0][perm], outs[2][0])
    assert torch.isfinite(outs[0][2]).all()


def _balanced_blob(feats, n_sv=48, seed=3):
    """An SJSV model blob fitted to the synthetic clips so that the fused path sees BOTH labels with many
    clips close to the decision boundary (the shipped model puts every synthetic clip deep in class 0, which
    makes "identical labels" trivial there): Scaler = the features' own mean/std, support vectors = random
    clips' standardised features with alternating coefficient signs, rho = minus the median score."""
    import struct
    rng = np.random.default_rng(seed)
    mean = feats.mean(axis=0).astype(np.float32)
    std = feats.std(axis=0).astype(np.float32)
    std[std < 1e-6] = 1.0
    z = ((feats - mean) * (1.0 / std)).astype(np.float32)
    pick = rng.choice(len(feats), n_sv, replace=False)
    sv = z[pick]
    coef = np.where(np.arange(n_sv) % 2 == 0, 1.0, -1.0).astype(np.float32) * rng.uniform(0.5, 1.5, n_sv).astype(np.float32)
    gamma = np.float32(0.025)
    d2 = ((z[:, None, :] - sv[None, :, :]) ** 2).sum(-1)
    s = (np.exp(-gamma * d2) * coef).sum(-1)
    rho = np.float32(-np.median(s))
    hdr = struct.pack("<6I4f", 0x56534A53, 1, 40, n_sv, n_sv // 2, n_sv - n_sv // 2, float(gamma), float(rho), -2.2435925, 0.017861439)
    body = np.concatenate([mean, (1.0 / std).astype(np.float32), coef, sv.reshape(-1)]).astype("<f4").tobytes()
    return hdr + body


def test_labels_identical_on_a_model_that_splits_the_synthetic_clips():
    import torch
    import dsp_b200
    from oracle_lib import CepstrumOracle
    sj = _ctx()
    n, ns = 2048, 16000
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 4000, n, ns)
    torch.cuda.synchronize()
    clips = d.cpu().numpy()
    _, _, _, feats = _run_dev(sj, clips)
    loud = np.abs(clips).max(axis=1) > 0
    blob = _balanced_blob(feats[loud])
    sj2 = dsp_b200.ScrubJay(device=0, samplerate=16000, model_blob=blob)
    sj2.set_stream(torch.cuda.current_stream().cuda_stream)
    gl, gc, gs, gf = _run_dev(sj2, clips)
    ora = CepstrumOracle(16000)
    of = np.stack([ora.mfcc_stats(c) for c in clips])
    res = [ora.predict(f, blob) for f in of]
    ol = np.array([r[0] for r in res]); oc = np.array([r[1] for r in res], np.float32); os_ = np.array([r[2] for r in res], np.float32)
    frac = gl[loud].mean()
    assert 0.25 < frac < 0.75, "the fitted model must split the clips (got %.2f positive)" % frac
    assert (np.abs(os_) < 0.05).sum() > 50, "and put many of them near the boundary"
    # the score inherits the feature error (1e-4 relative on standardised features); a label may only differ
    # where |score| is inside that band
    assert np.abs(gs - os_).max() <= 2e-4
    decided = np.abs(os_) >= 2e-4
    assert np.array_equal(gl[decided], ol[decided])
    assert (~decided).sum() <= 0.02 * n
    assert np.abs(gc - oc)[decided].max() <= 5e-3


def test_failed_host_call_leaves_nothing_behind_for_the_next_one():
    """a host-pointer call that fails half-way (chunks still in flight) must not drain its pending results
    into the arrays of the NEXT call (stale data / out-of-bounds host writes when the new arrays are smaller).  The
    failure is injected before chunk 2 of a 3-chunk batch (scrubjay_debug_inject_failure)."""
    import dsp_b200
    sj = _ctx()
    ns = 1024                                  # 256 MiB chunks = 65536 clips of 1024 samples
    big = np.random.default_rng(5).uniform(-0.5, 0.5, (65536 * 2 + 100, ns)).astype(np.float32)
    sj.debug_inject_failure(2)
    with pytest.raises(dsp_b200.ScrubJayError, match="injected failure"):
        sj.infer_batch(big)
    small = np.random.default_rng(6).uniform(-0.5, 0.5, (10, ns)).astype(np.float32)
    guard = np.full(64, -7, np.int32)           # the labels array of the next call sits inside a guarded block
    labels = guard[20:30]
    conf = np.empty(10, np.float32); score = np.empty(10, np.float32)
    sj._check(sj._L.scrubjay_infer_batch(sj._h, small.ctypes.data, 10, ns, labels.ctypes.data, conf.ctypes.data,
                                         score.ctypes.data, None))
    gl, gc, gs, _ = _run_dev(sj, small)
    assert np.array_equal(labels, gl) and np.array_equal(score, gs) and np.array_equal(conf, gc)
    assert (guard[:20] == -7).all() and (guard[30:] == -7).all(), "the failed call's results leaked into the next call"
    # and the pipeline still works for a multi-chunk batch afterwards
    hl, hc, hs = sj.infer_batch(big[:65536 + 50])
    dl, dc, ds, _ = _run_dev(sj, big[:65536 + 50])
    assert np.array_equal(hl, dl) and np.array_equal(hs, ds)
