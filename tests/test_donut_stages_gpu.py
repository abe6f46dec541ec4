"""GPU parity tests of the donut STAGE functions (sync/lib/classifier.h:14-18) through the C ABI.

Checkers: the UNMODIFIED reference builds in oracle/_ref (classifier.cpp float, classifier.c double, classifier.ino
for the 64-sample window) and the reference's own known-answer artefact _postbutter.txt (first 6000 samples in
tests/golden/donut_golden.npz).  Bar: the filter is bit-identical in SJ_IIR_EXACT mode for any coefficient set, the
float spectrogram is bit-identical (PlainFFT replayed), the double spectrogram agrees to 1e-12 of the frame's
largest cell (FFTW's operation order is not reproducible), the scan mode agrees with the exact mode to 1e-11."""
import os

import numpy as np
import pytest

from donut_lib import DonutPort, DonutRef32, DonutRef64, DonutRefIno, have_ref, have_ref_ino

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "donut_golden.npz")


def _ctx():
    import dsp_b200
    return dsp_b200.ScrubJay(device=0, samplerate=16000)


def _signals(n, ns, seed, dtype):
    rng = np.random.default_rng(seed)
    t = np.arange(ns) / 16000.0
    x = 0.2 * rng.uniform(-1, 1, (n, ns))
    for i in range(n):
        f0, f1 = rng.uniform(500, 7000, 2)
        x[i] += 0.5 * np.sin(2 * np.pi * (f0 * t + 0.5 * (f1 - f0) * t * t / t[-1]))
    return x.astype(dtype)


@pytest.mark.parametrize("band", [(1000.0, 3000.0), (3000.0, 7500.0)])
def test_filter_f32_bit_identical_to_reference(band):
    import dsp_b200
    sj = _ctx()
    x = _signals(40, 5003, 1, np.float32)  # length not a multiple of 8: exercises the scalar tail
    b, a = dsp_b200.butter_bandpass(*band)
    y = sj.butter_filter_batch(x, b, a)
    ref = DonutRef32() if have_ref() else None
    port = DonutPort("f32")
    for i in range(x.shape[0]):
        want = ref.filter(x[i], *band) if ref else port.filter_coef(x[i], b, a)
        assert np.array_equal(y[i], want), "signal %d differs" % i


@pytest.mark.parametrize("design", [0, 1])
def test_filter_f64_bit_identical_to_reference(design):
    import dsp_b200
    sj = _ctx()
    x = _signals(24, 4096, 2, np.float64)
    for band in ((1000.0, 3000.0), (3000.0, 7500.0)):
        b, a = dsp_b200.butter_bandpass(*band, design=design, dtype=np.float64)
        y = sj.butter_filter_batch(x, b, a)
        chk = DonutRef64() if have_ref() else DonutPort("f64")
        for i in range(x.shape[0]):
            assert np.array_equal(y[i], chk.filter_coef(x[i], b, a))


def test_filter_postbutter_kat_exact_and_scan():
    """_postbutter.txt: butter_bandpass_filter with the 48 kHz-design 1000-3000 Hz set (classifier.c:323-341) on
    ch0 of testing/1060-control.wav / 32768, tolerance 1e-8 absolute (the file prints 8 decimals)."""
    import dsp_b200
    g = np.load(GOLD)
    x = (g["kat_pcm"].astype(np.float64) / 32768.0)[None, :]
    b, a = dsp_b200.butter_bandpass(1000, 3000, design=dsp_b200.BUTTER_48K_DESIGN, dtype=np.float64)
    assert np.array_equal(b, g["kat_b"]) and np.array_equal(a, g["kat_a"])
    sj = _ctx()
    y = sj.butter_filter_batch(x, b, a)[0]
    assert np.abs(y - g["kat_out"]).max() < 1e-8
    # The scan re-orders the arithmetic.  This coefficient set (designed for 48 kHz, |a| up to 46, b ~ 2e-4) keeps a
    # delay line ~1e4 times larger than its output and amplifies any rounding difference by ~1e5: the scan lands
    # ~1e-6 away from the sequential result -- which is why the reference's own artefact can only be met to 1e-8
    # by the sequential operation order, and why the clip path does not use a scan.
    ys = sj.butter_filter_batch(x, b, a, mode=dsp_b200.IIR_SCAN)[0]
    assert np.abs(ys - g["kat_out"]).max() < 1e-5


def test_filter_state_carry_equals_one_call():
    import torch
    import dsp_b200
    sj = _ctx()
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    for dtype, prec in ((np.float32, dsp_b200.F32), (np.float64, dsp_b200.F64)):
        x = _signals(37, 3000, 3, dtype)
        b, a = dsp_b200.butter_bandpass(3000, 7500, dtype=dtype)
        whole = sj.butter_filter_batch(x, b, a)
        tdt = torch.float32 if dtype == np.float32 else torch.float64
        d = torch.from_numpy(x).cuda()
        out = torch.empty_like(d)
        state = torch.zeros((37, 8), dtype=tdt, device="cuda")
        cuts = [0, 1001, 1008, 2500, 3000]
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            sj.butter_filter_batch_dev(d[:, lo:].data_ptr(), out[:, lo:].data_ptr(), 37, hi - lo, 3000, prec, b, a, state)
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), whole)


def test_filter_scan_long_streams():
    import dsp_b200
    sj = _ctx()
    x = _signals(3, 100000 + 77, 4, np.float64)
    for band in ((1000.0, 3000.0), (3000.0, 7500.0)):
        b, a = dsp_b200.butter_bandpass(*band, dtype=np.float64)
        exact = sj.butter_filter_batch(x, b, a)
        scan = sj.butter_filter_batch(x, b, a, mode=dsp_b200.IIR_SCAN)
        assert np.abs(scan - exact).max() < 1e-11 * max(1.0, np.abs(exact).max())
    xf = x.astype(np.float32)
    b, a = dsp_b200.butter_bandpass(3000, 7500)
    ef, sf = sj.butter_filter_batch(xf, b, a), sj.butter_filter_batch(xf, b, a, mode=dsp_b200.IIR_SCAN)
    assert np.abs(sf - ef).max() < 2e-4 * np.abs(ef).max()  # float: block boundaries perturb at the 1e-7 * filter gain level


@pytest.mark.parametrize("ns", [16000, 5000, 480, 256, 200])
def test_spectrogram_f32_bit_identical_to_reference(ns):
    import dsp_b200
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    sj = _ctx()
    x = _signals(5, ns, 5 + ns, np.float32)
    b, a = dsp_b200.butter_bandpass(3000, 7500)
    y = sj.butter_filter_batch(x, b, a)
    S = sj.spectrogram_batch(y, 256, 16000)
    ref = DonutRef32()
    for i in range(y.shape[0]):
        want = ref.spectrogram(y[i])
        assert S[i].shape == want.shape
        assert np.array_equal(S[i], want), "signal %d: %d cells differ" % (i, (S[i] != want).sum())


def test_spectrogram_f64_matches_reference():
    import dsp_b200
    sj = _ctx()
    x = _signals(4, 16000, 6, np.float64)
    b, a = dsp_b200.butter_bandpass(3000, 7500, dtype=np.float64)
    y = sj.butter_filter_batch(x, b, a)
    S = sj.spectrogram_batch(y, 256, 16000)
    chk = DonutRef64() if have_ref() else DonutPort("f64")
    for i in range(4):
        want = chk.spectrogram(y[i])
        assert S[i].shape == want.shape
        assert np.abs(S[i] - want).max() <= 1e-12 * want.max()


@pytest.mark.parametrize("ns", [16000, 1000, 64])
def test_spectrogram_nperseg64_bit_identical_to_ino(ns):
    """classifier.ino:245-395 (64-sample windows), the unmodified sketch compiled against oracle/shim/ino"""
    import dsp_b200
    if not have_ref_ino():
        pytest.skip("oracle/_ref/libdonut_ref_ino.so not built")
    sj = _ctx()
    x = _signals(3, ns, 7 + ns, np.float32)
    S = sj.spectrogram_batch(x, 64, 16000)
    ref = DonutRefIno()
    assert S.shape[1:] == (33, dsp_b200.spectrogram_frames(ns, 64))
    for i in range(3):
        assert np.array_equal(S[i], ref.spectrogram(x[i]))


def test_reference_symbols_match_reference():
    """butter_bandpass / butter_bandpass_filter / compute_spectrogram / find_midpoints / sum_intense under their
    own names (sync/lib/classifier.h:14-18) against the unmodified classifier.cpp"""
    import torch
    import dsp_b200
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    ref = DonutRef32()
    sj = _ctx()
    d = torch.empty((64, 16000), dtype=torch.float32, device="cuda")
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.synth_clips_dev(d, 300, 64, 16000)
    torch.cuda.synchronize()
    clips = d.cpu().numpy()
    assert dsp_b200.butter_bandpass(1000, 2000) is None
    b, a = dsp_b200.butter_bandpass(3000, 7500)
    n_with_mid = 0
    for i in range(0, 64, 4):
        x = clips[i]
        y = dsp_b200.butter_bandpass_filter(x, b, a)
        assert np.array_equal(y, ref.filter(x, 3000.0, 7500.0))
        fr, tm, S = dsp_b200.compute_spectrogram(y, 16000)
        want = ref.spectrogram(y)
        assert np.array_equal(S, want)
        assert np.array_equal(fr, np.arange(129, dtype=np.float32) * np.float32(16000) / np.float32(256))
        mids = dsp_b200.find_midpoints(x)
        wmids = ref.midpoints(x)
        assert mids is not None and mids.shape == wmids.shape
        assert np.allclose(mids, wmids, rtol=0, atol=1e-6)
        n_with_mid += len(wmids) > 0
        # sum_intense on a dB plane with NaN holes, box around a midpoint (or mid-clip)
        with np.errstate(divide="ignore", invalid="ignore"):
            plane = np.where(S > 0, 10 * np.log10(S.astype(np.float64) / 1e-12), np.nan).astype(np.float32)
        plane[::3, ::5] = np.nan
        mp = float(wmids[0]) if len(wmids) else 0.5
        for lo, hi, hr in ((5000, 7000, 0.18), (2500, 5000, 0.05), (500, 2500, 0.18), (9000, 9500, 0.01)):
            got = dsp_b200.sum_intense(lo, hi, hr, fr, tm, plane, mp)
            assert got == ref.sum_intense(lo, hi, hr, fr, tm, plane, mp)
    assert n_with_mid > 0


def test_midpoints_batch_matches_reference_f64():
    import torch
    import dsp_b200
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    sj = _ctx()
    n = 48
    d = torch.empty((n, 16000), dtype=torch.float32, device="cuda")
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.synth_clips_dev(d, 1000, n, 16000)
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    detail = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    mids = torch.empty((n, dsp_b200.DONUT_MAX_MID), dtype=torch.float32, device="cuda")
    sj.donut_midpoints_batch_dev(d, n, 16000, dsp_b200.DONUT_DESKTOP_F64, labels, detail, mids)
    torch.cuda.synchronize()
    clips, mids, detail = d.cpu().numpy(), mids.cpu().numpy(), detail.cpu().numpy()
    ref = DonutRef64()
    seen = 0
    for i in range(n):
        want = ref.midpoints(clips[i])
        assert int(detail[i, 0]) == len(want)
        k = min(len(want), dsp_b200.DONUT_MAX_MID)
        assert np.allclose(mids[i, :k], want[:k], rtol=0, atol=1e-6)
        assert np.isnan(mids[i, k:]).all()
        seen += k
    assert seen > 0


def test_stage_argument_errors():
    import dsp_b200
    sj = _ctx()
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.spectrogram_batch(np.zeros((1, 16000), np.float32), nperseg=128)
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.spectrogram_batch(np.zeros((1, 10), np.float32), nperseg=256)  # (10 - 256) / 224 + 1 = 0 frames
    assert sj.butter_filter_batch(np.zeros((0, 100), np.float32), np.zeros(9), np.ones(9)).shape == (0, 100)


def _synth_dev(sj, first, n, ns=16000):
    import torch
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.synth_clips_dev(d, first, n, ns)
    torch.cuda.synchronize()
    return d


def _long(sj, d, variant, max_mid=64):
    import torch
    n, ns = d.shape
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    detail = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    mids = torch.empty((n, max_mid), dtype=torch.float32, device="cuda")
    sj.donut_classify_long_dev(d, n, ns, variant, labels, detail, mids, max_mid)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), detail.cpu().numpy(), mids.cpu().numpy()


@pytest.mark.parametrize("variant,tol", [(0, 1e-6), (1, 2e-3), (2, 1e-6), (3, 2e-3)])
def test_staged_classifier_equals_fused_kernel(variant, tol):
    """the staged pipeline (any length) against the fused kernel on one-second clips: same midpoints, labels and band
    sums (the float variants share the bit-exact IIR and FFT; sums differ only by the order of a double accumulation)"""
    import torch
    sj = _ctx()
    n = 192
    d = _synth_dev(sj, 2000, n)
    fl = torch.empty(n, dtype=torch.int32, device="cuda")
    fd = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    fm = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    sj.donut_midpoints_batch_dev(d, n, 16000, variant, fl, fd, fm)
    torch.cuda.synchronize()
    fl, fd, fm = fl.cpu().numpy(), fd.cpu().numpy(), fm.cpu().numpy()
    ll, ld, lm = _long(sj, d, variant, 8)
    assert np.array_equal(ld[:, 0], fd[:, 0])
    assert np.array_equal(np.isnan(lm), np.isnan(fm)) and np.allclose(np.nan_to_num(lm), np.nan_to_num(fm), rtol=0, atol=1e-6)
    assert np.abs(ld[:, 1:] - fd[:, 1:]).max() <= tol * max(1.0, np.abs(fd[:, 1:]).max())
    assert np.array_equal(ll, fl)
    assert (fd[:, 0] > 0).any()


def test_ino_variant_matches_unmodified_sketch():
    """SJ_DONUT_INO_F32 against classifier.ino's own setup() (oracle/_ref/libdonut_ref_ino.so): verdict, number of
    midpoints and the three band sums of the last midpoint the sketch evaluated"""
    import dsp_b200
    if not have_ref_ino():
        pytest.skip("oracle/_ref/libdonut_ref_ino.so not built")
    sj = _ctx()
    n = 96
    d = _synth_dev(sj, 5000, n)
    labels, detail, mids = _long(sj, d, dsp_b200.DONUT_INO_F32, 16)
    clips = d.cpu().numpy()
    ref = DonutRefIno()
    pos = 0
    for i in range(n):
        lab, nm, sums = ref.classify(clips[i])
        assert int(detail[i, 0]) == nm, i
        if sums:
            assert np.allclose(detail[i, 1:], sums[-1], rtol=1e-4, atol=2e-2), (i, detail[i], sums[-1])
        assert labels[i] == lab, (i, detail[i], sums)
        pos += lab
    assert 0 < pos < n
    # the batch entry point routes this variant to the staged pipeline by itself
    l2, d2 = sj.donut_classify_batch(clips, dsp_b200.DONUT_INO_F32)
    assert np.array_equal(l2, labels) and np.array_equal(d2, detail)


@pytest.mark.parametrize("secs", [2, 5])
def test_long_signals_match_unmodified_reference(secs):
    """whole recordings (classifier.c:37-210 works on complete files): classify(float*, n) of the unmodified
    classifier.cpp and find_midpoints of classifier.c on signals longer than the fused kernel's 1.3 s"""
    import dsp_b200
    if not have_ref():
        pytest.skip("oracle/_ref not built")
    sj = _ctx()
    n = 12
    d = _synth_dev(sj, 7000, n * secs).view(n, secs * 16000)
    x = d.cpu().numpy()
    l32, d32, m32 = _long(sj, d, dsp_b200.DONUT_SYNC_F32, 64)
    l64, d64, m64 = _long(sj, d, dsp_b200.DONUT_DESKTOP_F64, 64)
    r32, r64 = DonutRef32(), DonutRef64()
    for i in range(n):
        assert l32[i] == r32.classify(x[i]), i
        w32 = r32.midpoints(x[i])
        assert int(d32[i, 0]) == len(w32) and np.allclose(m32[i, :len(w32)], w32, rtol=0, atol=1e-6)
        w64 = r64.midpoints(x[i])
        assert int(d64[i, 0]) == len(w64) and np.allclose(m64[i, :len(w64)], w64, rtol=0, atol=1e-6)
    # the host entry point and the reference's own symbol accept the long signal as well
    lh, dh = sj.donut_classify_batch(x, dsp_b200.DONUT_SYNC_F32)
    assert np.array_equal(lh, l32) and np.array_equal(dh, d32)
    assert dsp_b200.classify(x[0]) == l32[0]


def test_long_recording_with_more_than_64_clusters_matches_the_unmodified_reference():
    """the staged path used to keep 64 midpoints at most, and the drop-in find_midpoints() refused inputs above
    21 536 samples although classify() takes them.  A 60 s recording with 100 bursts in the 1000-3000 Hz band (0.25 s on,
    0.35 s off): every midpoint of the unmodified find_midpoints (sync/lib/classifier.cpp:433-598) must come back, through
    the reference's own symbol names, and classify() must agree with the unmodified classify()."""
    import dsp_b200
    from donut_lib import DonutRef32, have_ref
    assert have_ref(), "oracle/_ref not shipped"
    fs, secs = 16000, 60
    rng = np.random.default_rng(64)
    t = np.arange(fs * secs) / fs
    x = 0.002 * rng.uniform(-1, 1, t.size)
    for k in range(100):
        t0 = 0.1 + 0.6 * k
        m = (t >= t0) & (t < t0 + 0.25)
        f = 1400.0 + 12.0 * k
        x[m] += 0.4 * np.sin(2 * np.pi * f * (t[m] - t0)) + 0.1 * np.sin(2 * np.pi * (f + 4200.0) * (t[m] - t0))
    x = x.astype(np.float32)
    ref = DonutRef32()
    want = ref.midpoints(x)
    assert len(want) > 64, "the test signal must have more clusters than the old cap (%d)" % len(want)
    got = dsp_b200.find_midpoints(x)
    assert got is not None and len(got) == len(want), (None if got is None else len(got), len(want))
    assert np.allclose(got, want, rtol=0, atol=1e-6)
    assert dsp_b200.classify(x) == ref.classify(x)
    # the batched host entry point takes the same route (two copies of the recording)
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    labels, detail = sj.donut_classify_batch(np.stack([x, x]), dsp_b200.DONUT_SYNC_F32)
    assert labels.tolist() == [ref.classify(x)] * 2 and detail[:, 0].tolist() == [float(len(want))] * 2
