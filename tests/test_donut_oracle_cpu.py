"""CPU tests pinning the donut port (oracle/donut_oracle.c) against (a) the reference's own IIR
known-answer artefact donut-classifier/_postbutter.txt, (b) golden outputs of the UNMODIFIED
reference builds (tests/golden/donut_golden.npz, tools/make_golden.py) and (c) -- when oracle/_ref
is present -- the reference builds themselves, live."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from donut_lib import DonutPort, DonutRef32, DonutRef64, have_ref  # noqa: E402
from synth_np import synth_clips  # noqa: E402

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "donut_golden.npz"))


def test_iir_matches_postbutter_kat():
    """_postbutter.txt = butter_bandpass_filter with the 48 kHz-design coefficients of
    classifier.c:323-341 on channel 0 of testing/1060-control.wav; printed with 7 significant digits"""
    x = G["kat_pcm"].astype(np.float64) / 32768.0
    y = DonutPort("f64").filter_coef(x, G["kat_b"], G["kat_a"])
    want = G["kat_out"]
    assert np.abs(y - want).max() <= 1e-8 + 1e-6 * np.abs(want).max()
    assert np.abs(want).max() > 1e-3


def test_port_f32_labels_match_reference_golden():
    clips = synth_clips(int(G["synth_first"]), int(G["synth_n"]), 16000)
    labels, detail = DonutPort("f32").classify_batch(clips, variant=1)
    assert np.array_equal(labels, G["ref32_labels"])
    assert 0 < labels.sum() < labels.size


def test_port_f64_midpoints_and_real_clips_match_reference_golden():
    clips = synth_clips(int(G["synth_first"]), int(G["synth_n"]), 16000)
    _, detail = DonutPort("f64").classify_batch(clips, variant=0)
    assert np.array_equal(detail[:, 0].astype(np.int32), G["ref64_nmid"])
    real = G["real_pcm"].astype(np.float32) / 32768.0
    l32, _ = DonutPort("f32").classify_batch(real, variant=1)
    assert np.array_equal(l32, G["real_ref32_labels"])
    p64 = DonutPort("f64")
    for c, iir_tail, spec in zip(real, G["real_ref64_iir0"], G["real_ref64_spec"]):
        assert np.array_equal(p64.filter(c, 0)[-64:], iir_tail)  # bit-exact: same operation order
        S = p64.spectrogram(p64.filter(c, 1))[40:44, 30:34]
        assert np.allclose(S, spec, rtol=1e-9, atol=0)


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (reference absent)")
def test_port_against_live_reference():
    clips = synth_clips(7, 24, 16000)
    p32, p64, r32, r64 = DonutPort("f32"), DonutPort("f64"), DonutRef32(), DonutRef64()
    for c in clips[:4]:
        for w, (lo, hi) in enumerate([(1000, 3000), (3000, 7500)]):
            assert np.array_equal(p32.filter(c, w), r32.filter(c, lo, hi))
            assert np.array_equal(p64.filter(c, w), r64.filter(c, lo, hi))
        f = p64.filter(c, 1)
        assert np.allclose(p64.spectrogram(f), r64.spectrogram(f), rtol=1e-9, atol=1e-30)
        f = p32.filter(c, 1)
        assert np.array_equal(p32.spectrogram(f), r32.spectrogram(f))  # same FFT arithmetic as PlainFFT: bit-exact
    labels, _ = p32.classify_batch(clips, variant=1)
    assert labels.tolist() == [r32.classify(c) for c in clips]


def test_short_clip_is_rejected():
    labels, detail = DonutPort("f64").classify_batch(np.zeros((2, 100), np.float32), variant=0)
    assert labels.tolist() == [0, 0] and detail[:, 0].tolist() == [0, 0]


def test_port_f64_verdict_pinned_to_the_unmodified_reference_functions():
    """The f64 verdict, one step closer than midpoint counts: the per-file body of main (classifier.c:84-198) driven
    through the UNMODIFIED butter_bandpass_filter / compute_spectrogram / find_midpoints / sum_intense of
    oracle/_ref/libdonut_ref_f64.so (DonutRef64.classify_main) against the port, on "call"-flavoured synthetic clips
    where the desktop variant says 1 on a sizeable fraction (the mixed flavour never does) and on mixed clips."""
    from donut_lib import DonutPort, DonutRef64, have_ref
    from synth_np import synth_clips
    assert have_ref(), "oracle/_ref missing: run `make -C oracle` where /root/reference exists"
    clips = np.concatenate([synth_clips(300, 192, 16000, flavour=1), synth_clips(300, 64, 16000, flavour=0)])
    ref, port = DonutRef64(), DonutPort("f64")
    pl, pd = port.classify_batch(clips, 0)
    res = [ref.classify_main(c) for c in clips]
    rl = np.array([r[0] for r in res], np.int32)
    assert np.array_equal(pl, rl), "port labels differ from the reference-driven main on %d clips" % (pl != rl).sum()
    assert 0.2 < rl[:192].mean() < 0.8, "the call flavour must give both f64 verdicts (got %.2f positive)" % rl[:192].mean()
    for (label, mids, sums), d in zip(res, pd):
        assert len(mids) == int(d[0])
        if sums:
            assert np.abs(np.array(sums[-1]) - d[1:]).max() <= 1e-4 * max(1.0, np.abs(d[1:]).max())
