"""BASELINE config 1 on the CPU side: the reference's own cepstrum test material (cepstrum/testing/*.wav -- the folder
scrubjay_infer.c:15 iterates -- and the 16-bit WAVs of cepstrum/data/{neg,pos}) through the oracle, against
  * an INDEPENDENT float64 numpy/scipy end-to-end restatement (np.fft.rfft, scipy dct; tools/make_golden.py:mfcc_stats_np)
    committed as tests/golden/reference_clips.json, and
  * sklearn's decision_function / predict on the reference's own scrubjay_svm.joblib for those features.
The feature arithmetic stays "restated aubio, parity unpinned" (no aubio here); what these files pin is the oracle against
a second implementation ON THE REFERENCE'S MATERIAL at its native 96 kHz / 44.1 kHz rates, and the SVM half against the
reference's model.  The WAVs are copied to oracle/_ref/cepstrum_testing/ by oracle/Makefile (git-ignored)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

from oracle_lib import CepstrumOracle  # noqa: E402
import reference_clips as rc           # noqa: E402


def _golden():
    return {r["file"]: r for r in json.load(open(rc.GOLDEN))["rows"]}


def test_golden_holds_the_five_testing_files_with_their_prefix_truth():
    g = _golden()
    testing = [f for f in g if f[:2] in ("P_", "N_")]
    assert sorted(testing) == ["N_1060-control.wav", "N_1809v2-not-sj.wav", "P_1363v2-sj-short.WAV", "P_1363v2-sj.WAV",
                               "P_2287-sj.wav"]
    assert all(g[f]["samplerate"] == 96000 and g[f]["channels"] == 2 for f in testing)
    # what the reference's own C path would print: the SVM was trained on librosa (dB-power) MFCCs and is fed aubio-scale
    # features here, so the decision value sits at the intercept for every file: all five "Not"
    assert [g[f]["sklearn_predict"] for f in sorted(testing)] == [0, 0, 0, 0, 0]
    assert sum(g[f]["sklearn_predict"] == g[f]["truth"] for f in testing) == 2


def test_oracle_svm_equals_sklearn_on_the_reference_clip_features():
    """no WAV needed: the oracle's Scaler + SVM on the committed features vs sklearn on the reference's joblib"""
    ora = CepstrumOracle(16000)
    for name, r in _golden().items():
        label, conf, d = ora.predict(np.array(r["features_oracle"], np.float32))
        assert abs(d - r["sklearn_decision_function"]) < 2e-6, name
        assert label == r["sklearn_predict"], name
        assert abs(conf - r["sklearn_predict_proba"][label]) < 1e-5, name


def test_oracle_matches_the_independent_numpy_restatement_on_the_reference_wavs():
    files = rc.list_clips()
    assert files, ("the reference's cepstrum WAVs are missing: run `make -C oracle` in a container that has /root/reference "
                   "(copies them to oracle/_ref/cepstrum_testing/)")
    g = _golden()
    seen = 0
    for path in files:
        name = os.path.basename(path)
        assert name in g, "golden is stale: rerun tools/reference_clips.py --make-golden"
        x, sr, ch = rc.read_wav_mean(path)
        assert (sr, ch, x.size) == (g[name]["samplerate"], g[name]["channels"], g[name]["samples"])
        feat, fr = CepstrumOracle(sr).mfcc_stats(x, want_frames=True)
        assert fr.shape[0] == g[name]["frames"] == (x.size + 1023) // 1024
        want = np.array(g[name]["features_numpy_f64"], np.float32)
        assert np.allclose(feat, want, rtol=1e-5, atol=1e-5), (name, np.abs(feat - want).max())
        assert np.array_equal(feat, np.array(g[name]["features_oracle"], np.float32)), name
        seen += 1
    assert seen >= 5
