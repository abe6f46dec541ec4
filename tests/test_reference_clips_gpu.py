"""BASELINE config 1 on the GPU: the reference's own clips (cepstrum/testing/, iterated by scrubjay_infer.c:15,156-177; 96 kHz
stereo) through scrubjay_wav_parse -> scrubjay_decode_pcm16_dev(SJ_MIX_MEAN) -> scrubjay_infer_batch_dev at the file's native
rate (:28-30) -- the generic-rate instantiation -- against the oracle on the samples the GPU decoded and against the
committed golden (independent numpy float64 features; sklearn decision values of the reference's joblib)."""
import json
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_reference_clips_features_scores_labels():
    import reference_clips as rc
    from oracle_lib import CepstrumOracle
    assert rc.list_clips(), "oracle/_ref/cepstrum_testing/ is empty: `make -C oracle` copies the reference's WAVs there"
    gold = {r["file"]: r for r in json.load(open(rc.GOLDEN))["rows"]}
    rows = rc.run_gpu_rows()
    assert len([r for r in rows if r["file"][:2] in ("P_", "N_")]) == 5
    agree = 0
    for r in rows:
        g = gold[r["file"]]
        assert r["samplerate"] == g["samplerate"] and r["samples"] == g["samples"]
        # the device decode is aubio's down-mix restated (mean of the channels of s / 32768), bit for bit
        host, _, _ = rc.read_wav_mean(os.path.join(rc.clip_dir(), r["file"]))
        assert np.array_equal(r["mono"], host), r["file"]
        ora = CepstrumOracle(r["samplerate"])
        of = ora.mfcc_stats(r["mono"])
        err = np.abs(r["features"] - of) / (1e-4 * np.abs(of) + 1e-5)
        assert err.max() <= 1.0, (r["file"], err.max())
        ol, oc, os_ = ora.predict(of)
        assert abs(r["score"] - os_) <= 1e-4 * abs(os_) + 1e-6
        assert r["label"] == ol and abs(r["conf"] - oc) <= 5e-3
        # golden: independent float64 numpy features, sklearn decision value / label of the reference's joblib
        want = np.array(g["features_numpy_f64"], np.float32)
        assert (np.abs(r["features"] - want) <= 1e-4 * np.abs(want) + 2e-5).all(), (r["file"], np.abs(r["features"] - want).max())
        assert abs(r["score"] - g["sklearn_decision_function"]) < 2e-5
        assert r["label"] == g["sklearn_predict"]
        agree += r["file"][:2] in ("P_", "N_") and r["label"] == rc.truth_of(r["file"])
    # the reference's own outcome on its five files (librosa-trained SVM fed aubio-scale features): 2 of 5
    assert agree == 2
