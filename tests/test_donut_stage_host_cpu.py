"""CPU tests of the host-side pieces of the donut stage API (no GPU calls): coefficient lookup, frame-count
arithmetic, the transition matrix used by the scan mode, and the checkers built from the unmodified reference
(classifier.cpp, classifier.c, classifier.ino + oracle/shim/ino) that the GPU tests rely on."""
import ctypes
import os
import sys

import numpy as np
import pytest

import dsp_b200
from donut_lib import DonutPort, DonutRef32, DonutRef64, DonutRefIno, have_ref, have_ref_ino

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


def test_butter_bandpass_equals_reference_tables():
    for band in ((1000.0, 3000.0), (3000.0, 7500.0)):
        b, a = dsp_b200.butter_bandpass(*band)
        bd, ad = dsp_b200.butter_bandpass(*band, design=dsp_b200.BUTTER_16K, dtype=np.float64)
        assert np.array_equal(b, bd.astype(np.float32)) and np.array_equal(a, ad.astype(np.float32))
        if have_ref():
            r32, r64 = DonutRef32(), DonutRef64()
            rb, ra = np.zeros(9, np.float32), np.zeros(9, np.float32)
            assert r32.bb_fn(band[0], band[1], rb.ctypes.data, ra.ctypes.data)
            assert np.array_equal(b, rb) and np.array_equal(a, ra)
            rb, ra = np.zeros(9), np.zeros(9)
            assert r64.L.butter_bandpass(band[0], band[1], rb.ctypes.data, ra.ctypes.data)
            assert np.array_equal(bd, rb) and np.array_equal(ad, ra)
    assert dsp_b200.butter_bandpass(1000, 7500) is None                  # "invalid bandpass range" (classifier.cpp:186-190)
    assert dsp_b200.butter_bandpass(1000, 3000, design=7, dtype=np.float64) is None
    # the commented-out 48 kHz-design set is the one behind the reference's _postbutter.txt artefact
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "donut_golden.npz"))
    b, a = dsp_b200.butter_bandpass(1000, 3000, design=dsp_b200.BUTTER_48K_DESIGN, dtype=np.float64)
    assert np.array_equal(b, g["kat_b"]) and np.array_equal(a, g["kat_a"])


def test_frame_and_second_counts_follow_the_reference_arithmetic():
    L = dsp_b200.load_library()
    for ns in (0, 10, 32, 33, 200, 255, 256, 479, 480, 16000, 21536, 239390):
        for nperseg in (256, 64):
            hop = nperseg - nperseg // 8
            want = int((ns - nperseg) / hop) + 1  # C division truncates toward zero (classifier.cpp:234)
            assert dsp_b200.spectrogram_frames(ns, nperseg) == max(want, 0), (ns, nperseg)
    assert dsp_b200.spectrogram_frames(16000, 128) == 0
    assert L.scrubjay_donut_stream_seconds(16000 * 7 + 15999, 16000) == 7
    assert L.scrubjay_donut_stream_seconds(100, 16000) == 0 and L.scrubjay_donut_stream_seconds(100, 0) == 0
    if have_ref():
        x = np.zeros(200, np.float32)
        assert DonutRef32().spectrogram(x).shape == (129, dsp_b200.spectrogram_frames(200, 256))


def test_port_filter_with_any_coefficients_equals_reference():
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(3)
    x = rng.uniform(-1, 1, 3000)
    for design in (0, 1):
        b, a = dsp_b200.butter_bandpass(3000, 7500, design=design, dtype=np.float64)
        assert np.array_equal(DonutPort("f64").filter_coef(x, b, a), DonutRef64().filter_coef(x, b, a))
    b, a = dsp_b200.butter_bandpass(1000, 3000)
    xf = x.astype(np.float32)
    assert np.array_equal(DonutPort("f32").filter_coef(xf, b, a), DonutRef32().filter_coef(xf, b, a))


def test_ino_checker_runs_the_unmodified_sketch():
    """classifier.ino compiled against oracle/shim/ino: setup() on a silent second reports no midpoint and no bird;
    on synthetic clips it reports one count and three sums per midpoint, and its spectrogram has 33 x 285 cells"""
    if not have_ref_ino():
        pytest.skip("oracle/_ref/libdonut_ref_ino.so not built")
    import synth_np
    ref = DonutRefIno()
    lab, nm, sums = ref.classify(np.zeros(16000, np.float32))
    assert (lab, nm, sums) == (0, 0, [])
    clips = synth_np.synth_clips(0, 6, 16000)
    seen = 0
    for c in clips:
        lab, nm, sums = ref.classify(c)
        assert lab in (0, 1) and nm >= 0 and len(sums) <= max(nm, 0)
        if lab:
            above, middle, below = sums[-1]
            assert middle < 75 and above > 300 and below > 100  # classifier.ino:134
        seen += nm
    assert seen > 0
    assert ref.spectrogram(clips[0]).shape == (33, 285)


def test_ingest_oracle_restatements():
    from oracle_lib import ORACLE_DIR, build_oracle
    path = os.path.join(ORACLE_DIR, "libsj_oracle_f32.so")
    if not os.path.exists(path):
        build_oracle()
    L = ctypes.CDLL(path)
    L.sj_oracle_adc12.argtypes = [ctypes.c_void_p, ctypes.c_long, ctypes.c_void_p]
    L.sj_oracle_upsample_linear.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    adc = np.array([0, 2047, 2048, 2049, 4095], np.uint16)
    out = np.empty(5, np.float32)
    L.sj_oracle_adc12(adc.ctypes.data, 5, out.ctypes.data)
    assert np.array_equal(out, np.array([-1.0, -16 / 32768, 0.0, 16 / 32768, 2047 * 16 / 32768], np.float32))
    x = np.array([0.0, 1.0, 0.0], np.float32)
    y = np.empty(5, np.float32)
    L.sj_oracle_upsample_linear(x.ctypes.data, 3, y.ctypes.data, 5)
    assert np.allclose(y, [0.0, 0.5, 1.0, 0.5, 0.0])
