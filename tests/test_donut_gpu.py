"""GPU parity tests of the donut path through the C ABI.

Checkers: the UNMODIFIED reference builds in oracle/_ref (when shipped) and the port
oracle/donut_oracle.c that is pinned against them (tests/test_donut_oracle_cpu.py).

Parity bar.  The IIR stage is bit-identical to the reference (same operation order, no FMA).  The FFT
is a different algorithm from PlainFFT/FFTW, so PSD cells differ by rounding (~1e-7 relative in float,
~1e-15 in double); a cell sitting within that distance of a hard threshold (band 0.65/0.80, 70 dB) can
flip, which moves a band sum by one cell value (<= 0.85).  Hence:
  * midpoint counts identical,
  * band sums within 2.0 absolute (float variant) / 1e-6 relative (double variant),
  * labels identical on every clip whose sums are not within that slack of a verdict threshold
    (such clips are counted; the count must stay small)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
pytestmark = pytest.mark.gpu


def _ctx():
    import dsp_b200
    return dsp_b200.ScrubJay(device=0, samplerate=16000)


def _synth(sj, first, n, ns=16000):
    import torch
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.synth_clips_dev(d, first, n, ns)
    torch.cuda.synchronize()
    return d


def _gpu(sj, d, variant):
    import torch
    n, ns = d.shape
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    detail = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    sj.donut_classify_batch_dev(d, n, ns, variant, labels, detail)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), detail.cpu().numpy()


def _compare(gl, gd, ol, od, variant, sum_tol):
    thr = {0: (300, 75, 100), 1: (200, 100, 80), 2: (200, 50, 200), 3: (200, 100, 150)}[variant]
    assert np.array_equal(gd[:, 0], od[:, 0]), "midpoint counts differ on %d clips" % (gd[:, 0] != od[:, 0]).sum()
    has = od[:, 0] > 0
    diff = np.abs(gd[:, 1:] - od[:, 1:])
    assert not has.any() or diff[has].max() <= sum_tol, "band sums differ by %g" % diff[has].max()
    near = has & (np.abs(od[:, 1:] - np.array(thr)) <= sum_tol).any(axis=1)
    bad = (gl != ol) & ~near
    assert not bad.any(), "labels differ on %d clips away from any threshold" % bad.sum()
    assert near.mean() < 0.05
    return int((gl != ol).sum()), int(near.sum())


def test_f64_variant_matches_port_and_reference():
    import dsp_b200
    from donut_lib import DonutPort
    sj = _ctx()
    d = _synth(sj, 0, 1024)
    gl, gd = _gpu(sj, d, dsp_b200.DONUT_DESKTOP_F64)
    ol, od = DonutPort("f64").classify_batch(d.cpu().numpy(), variant=0)
    mism, near = _compare(gl, gd, ol, od, 0, sum_tol=1e-3)
    assert mism == 0
    assert (od[:, 0] > 0).mean() > 0.5, "workload must exercise the midpoint/band-sum branch"


def test_f32_variant_matches_port():
    import dsp_b200
    from donut_lib import DonutPort
    sj = _ctx()
    d = _synth(sj, 5000, 1024)
    gl, gd = _gpu(sj, d, dsp_b200.DONUT_SYNC_F32)
    ol, od = DonutPort("f32").classify_batch(d.cpu().numpy(), variant=1)
    # same FFT arithmetic as the reference (PlainFFT butterflies): PSD cells are bit-identical, what is
    # left is the band sum's accumulation order and last-bit differences of log10
    mism, near = _compare(gl, gd, ol, od, 1, sum_tol=0.01)
    assert mism == 0
    assert 0 < ol.sum() < ol.size, "need both verdicts in the sample"
    print("f32 variant: %d label mismatches, %d clips within slack of a threshold, of %d" % (mism, near, ol.size))


def test_f32_variant_matches_unmodified_reference_classify():
    """labels of the reference's own classify() (oracle/_ref/libdonut_ref_f32.so) on the same bytes"""
    import dsp_b200
    from donut_lib import DonutPort, DonutRef32, have_ref
    if not have_ref():
        pytest.skip("oracle/_ref not shipped")
    sj = _ctx()
    d = _synth(sj, 9000, 256)
    clips = d.cpu().numpy()
    gl, gd = _gpu(sj, d, dsp_b200.DONUT_SYNC_F32)
    r = DonutRef32()
    ref = np.array([r.classify(c) for c in clips], np.int32)
    _, od = DonutPort("f32").classify_batch(clips, variant=1)
    assert np.array_equal(gl, ref), "labels differ from the reference's classify() on %d clips" % (gl != ref).sum()
    assert 0 < ref.sum() < ref.size


def test_golden_real_clips_and_reference_symbol():
    """the four 1 s / 16 kHz recordings of donut-classifier/try: labels of the reference build, via the
    reference's own symbol name classify(float*, int)"""
    import dsp_b200
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "donut_golden.npz"))
    real = g["real_pcm"].astype(np.float32) / 32768.0
    got = [dsp_b200.classify(c) for c in real]
    assert got == g["real_ref32_labels"].tolist()
    sj = _ctx()
    labels, detail = sj.donut_classify_batch(real, dsp_b200.DONUT_SYNC_F32)
    assert labels.tolist() == got


@pytest.mark.parametrize("ns", [256, 479, 480, 3001, 7200, 16000, 20480])
def test_ragged_lengths(ns):
    import dsp_b200
    from donut_lib import DonutPort
    rng = np.random.default_rng(ns)
    t = np.arange(ns) / 16000.0
    clips = (0.2 * rng.uniform(-1, 1, (19, ns)) + 0.3 * np.sin(2 * np.pi * 2000 * t)[None, :]).astype(np.float32)
    sj = _ctx()
    for variant, prec, tol in ((0, "f64", 1e-3), (1, "f32", 0.01)):
        gl, gd = sj.donut_classify_batch(clips, variant)
        ol, od = DonutPort(prec).classify_batch(clips, variant)
        _compare(gl, gd, ol, od, variant, tol)


@pytest.mark.parametrize("ns", [479, 3001, 7200])
def test_ragged_lengths_large_batch_equals_small_batch(ns):
    """Large batches take the warp-specialised kernels (unaligned instantiation for odd lengths, aligned otherwise), small
    float batches the round-1 kernel: the first clips of a 9 500-clip batch must equal the same clips classified alone
    (float: another kernel, same bits; double: the same kernel in another grouping)."""
    import torch
    sj = _ctx()
    n = 9500
    d = _synth_calls(sj, 321, n, ns)
    for variant in (0, 1):
        l, det = _gpu(sj, d, variant)
        ls, ds = _gpu(sj, d[:300].contiguous(), variant)
        assert np.array_equal(ls, l[:300]) and np.array_equal(ds, det[:300], equal_nan=True)


def test_too_short_and_silent():
    import dsp_b200
    sj = _ctx()
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.donut_classify_batch(np.zeros((2, 100), np.float32), 1)
    labels, detail = sj.donut_classify_batch(np.zeros((3, 16000), np.float32), 1)
    assert labels.tolist() == [0, 0, 0] and detail[:, 0].tolist() == [0, 0, 0]


def test_permutation_equivariance_large():
    import torch
    import dsp_b200
    sj = _ctx()
    n = 8192 + 5
    d = _synth(sj, 777, n)
    perm = torch.randperm(n, device="cuda")
    a = _gpu(sj, d, dsp_b200.DONUT_SYNC_F32)
    b = _gpu(sj, d[perm].contiguous(), dsp_b200.DONUT_SYNC_F32)
    p = perm.cpu().numpy()
    assert np.array_equal(a[0][p], b[0]) and np.array_equal(a[1][p], b[1])


@pytest.mark.parametrize("variant,prec,tol", [(2, "f64", 1e-3), (3, "f32", 0.01)])
def test_other_reference_variants_match_port(variant, prec, tol):
    """revert.c / microphone/src/classifier.c (double, gap tolerance 0.15 s) and microphone/src/classifier.cpp
    (float): same kernels, the variant's own constants"""
    from donut_lib import DonutPort
    sj = _ctx()
    d = _synth(sj, 9000, 768)
    gl, gd = _gpu(sj, d, variant)
    ol, od = DonutPort(prec).classify_batch(d.cpu().numpy(), variant=variant)
    # the float variant replays PlainFFT bit for bit, so variant 3 holds the same bar as variant 1 (0.01 on the sums,
    # no label mismatch); only the accumulation order of the box sums differs from the reference's
    mism, near = _compare(gl, gd, ol, od, variant, sum_tol=tol)
    assert mism == 0
    assert (od[:, 0] > 0).mean() > 0.5


def test_stream_decision_track_equals_clip_path():
    """sync/sync.cpp:190-212: every full second of a stream is classified from zero state; track byte = label + 1"""
    import torch
    import dsp_b200
    sj = _ctx()
    ch, secs, tail, stride = 5, 7, 4321, 7 * 16000 + 4321 + 3  # odd stride: the unaligned-load instantiation
    clips = _synth(sj, 700, ch * secs)
    streams = torch.zeros((ch, stride), dtype=torch.float32, device="cuda")
    streams[:, :secs * 16000] = clips.view(ch, secs * 16000)
    streams[:, secs * 16000:] = 0.25
    for variant in (dsp_b200.DONUT_SYNC_F32, dsp_b200.DONUT_DESKTOP_F64):
        labels, detail = _gpu(sj, clips, variant)
        track = torch.zeros((ch, secs), dtype=torch.int8, device="cuda")
        det2 = torch.empty((ch, secs, 4), dtype=torch.float32, device="cuda")
        sj.donut_classify_streams_dev(streams, ch, secs * 16000 + tail, stride, variant, track, det2)
        torch.cuda.synchronize()
        assert np.array_equal(track.cpu().numpy().reshape(-1), labels.astype(np.int8) + 1)
        assert np.array_equal(det2.cpu().numpy().reshape(-1, 4), detail)
    # aligned stride takes the vector-load instantiation: same bits
    s2 = torch.zeros((ch, secs * 16000 + 16), dtype=torch.float32, device="cuda")
    s2[:, :secs * 16000] = clips.view(ch, secs * 16000)
    track2 = torch.zeros((ch, secs), dtype=torch.int8, device="cuda")
    sj.donut_classify_streams_dev(s2, ch, secs * 16000, secs * 16000 + 16, dsp_b200.DONUT_SYNC_F32, track2)
    torch.cuda.synchronize()
    labels, _ = _gpu(sj, clips, dsp_b200.DONUT_SYNC_F32)
    assert np.array_equal(track2.cpu().numpy().reshape(-1), labels.astype(np.int8) + 1)


def test_stream_decision_tracks_of_a_large_stream_array_equal_the_clip_path():
    """the same for 1 300 channels x 7 s = 9 100 clips: the warp-specialised kernels' stream addressing (odd and aligned row
    strides), against the clip path on contiguous clips"""
    import torch
    import dsp_b200
    sj = _ctx()
    ch, secs = 1300, 7
    clips = _synth_calls(sj, 5000, ch * secs)
    for stride in (secs * 16000 + 3, secs * 16000 + 16):
        streams = torch.zeros((ch, stride), dtype=torch.float32, device="cuda")
        streams[:, :secs * 16000] = clips.view(ch, secs * 16000)
        for variant in (dsp_b200.DONUT_SYNC_F32, dsp_b200.DONUT_DESKTOP_F64):
            labels, detail = _gpu(sj, clips, variant)
            track = torch.zeros((ch, secs), dtype=torch.int8, device="cuda")
            det2 = torch.empty((ch, secs, 4), dtype=torch.float32, device="cuda")
            sj.donut_classify_streams_dev(streams, ch, secs * 16000, stride, variant, track, det2)
            torch.cuda.synchronize()
            assert np.array_equal(track.cpu().numpy().reshape(-1), labels.astype(np.int8) + 1)
            assert np.array_equal(det2.cpu().numpy().reshape(-1, 4), detail, equal_nan=True)
        del streams


@pytest.mark.parametrize("variant", [0, 1])
def test_small_batch_grouping_equals_large_batch_grouping(variant):
    """Small batches are spread over the CTAs (down to one clip per CTA, absent clips' passes skipped), large ones
    use 16 clips per CTA: the same clips must give the same bits either way, for every group size in between."""
    sj = _ctx()
    big = _synth(sj, 4242, 14000)          # 16 clips per CTA
    ref_l, ref_d = _gpu(sj, big, variant)
    for n in (1, 2, 37, 700, 1500, 5000):  # 1, 1, 1, 1-2, 3, 8-ish clips per CTA
        l, d = _gpu(sj, big[:n].contiguous(), variant)
        assert np.array_equal(l, ref_l[:n]) and np.array_equal(d, ref_d[:n], equal_nan=True), "n = %d" % n


@pytest.mark.parametrize("variant", [0, 1])
def test_many_rounds_per_group_slot_equal_single_round_chunks(variant):
    """A CTA's group slot takes a new group while the epilogue warps still work on the previous one (two copies of the clip
    state and of the PSD scratch per slot): 30 000 call-flavoured clips are 4-5 groups per slot; the same
    clips in chunks that fit one round (and so never reuse a slot within a launch) must give the same bits."""
    sj = _ctx()
    n, chunk = 30000, 6000
    d = _synth_calls(sj, 90000, n)
    l, det = _gpu(sj, d, variant)
    for a in range(0, n, chunk):
        lc, dc = _gpu(sj, d[a:a + chunk].contiguous(), variant)
        assert np.array_equal(lc, l[a:a + chunk]) and np.array_equal(dc, det[a:a + chunk], equal_nan=True), "chunk at %d" % a
    assert (det[:, 0] > 0).mean() > 0.3  # many clips have midpoints: the epilogue had real work


def test_round1_kernels_behind_the_env_switch_agree_with_the_warp_specialised_ones(tmp_path):
    """SCRUBJAY_DONUT_LEGACY (read once per process) = 1 forces the round-1 kernels, = 0 the warp-specialised ones (by default
    small float batches take the former, everything else the latter): same bits in float (the same arithmetic in another
    schedule -- which is what allows the library to switch by batch size); in double the FFT is organised differently (4 lanes x FFT32 against 8 lanes x FFT16 + FFT8), so
    midpoint counts and labels are equal and the band sums agree to rounding."""
    import subprocess
    script = (
        "import sys, numpy as np, torch\n"
        "sys.path.insert(0, %r)\n"
        "import dsp_b200\n"
        "sj = dsp_b200.ScrubJay(device=0, samplerate=16000)\n"
        "sj.set_stream(torch.cuda.current_stream().cuda_stream)\n"
        "n = 3000\n"
        "d = torch.empty((n, 16000), dtype=torch.float32, device='cuda')\n"
        "sj.synth_clips_dev(d, 5150, n, 16000, flavour=dsp_b200.SYNTH_CALLS)\n"
        "out = {}\n"
        "for v in (0, 1):\n"
        "    l = torch.empty(n, dtype=torch.int32, device='cuda'); det = torch.empty((n, 4), device='cuda')\n"
        "    sj.donut_classify_batch_dev(d, n, 16000, v, l, det); torch.cuda.synchronize()\n"
        "    out['l%%d' %% v] = l.cpu().numpy(); out['d%%d' %% v] = det.cpu().numpy()\n"
        "np.savez(sys.argv[1], **out)\n" % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    res = {}
    for tag, env in (("ws", {"SCRUBJAY_DONUT_LEGACY": "0"}), ("r1", {"SCRUBJAY_DONUT_LEGACY": "1"})):  # 0 = never, 1 = always round 1
        f = str(tmp_path / (tag + ".npz"))
        subprocess.run([sys.executable, "-c", script, f], check=True, env=dict(os.environ, **env), timeout=300)
        res[tag] = np.load(f)
    a, b = res["ws"], res["r1"]
    assert np.array_equal(a["l1"], b["l1"]) and np.array_equal(a["d1"], b["d1"], equal_nan=True)
    assert np.array_equal(a["l0"], b["l0"]) and np.array_equal(a["d0"][:, 0], b["d0"][:, 0])
    assert np.allclose(a["d0"][:, 1:], b["d0"][:, 1:], rtol=1e-5, atol=1e-4)
    assert (a["d1"][:, 0] > 0).mean() > 0.3


def _synth_calls(sj, first, n, ns=16000):
    import torch
    import dsp_b200
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    sj.synth_clips_dev(d, first, n, ns, flavour=dsp_b200.SYNTH_CALLS)
    torch.cuda.synchronize()
    return d


def test_f64_verdict_pinned_to_unmodified_reference_functions_on_1024_clips():
    """the double variant's labels were pinned only through the port.  Here the GPU's midpoints, band
    sums and labels are held against the per-file body of main (classifier.c:84-198) driven through the UNMODIFIED
    butter_bandpass_filter / compute_spectrogram / find_midpoints / sum_intense of oracle/_ref/libdonut_ref_f64.so, on
    768 "call"-flavoured clips -- on which the desktop variant says 1 about half the time -- and 256 mixed clips."""
    import torch
    import dsp_b200
    from donut_lib import DonutRef64, have_ref
    assert have_ref(), "oracle/_ref not shipped"
    sj = _ctx()
    d = torch.cat([_synth_calls(sj, 2000, 768), _synth(sj, 2000, 256)])
    n = d.shape[0]
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    detail = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    mids = torch.empty((n, dsp_b200.DONUT_MAX_MID), dtype=torch.float32, device="cuda")
    sj.donut_midpoints_batch_dev(d, n, 16000, dsp_b200.DONUT_DESKTOP_F64, labels, detail, mids)
    torch.cuda.synchronize()
    gl, gd, gm = labels.cpu().numpy(), detail.cpu().numpy(), mids.cpu().numpy()
    ref = DonutRef64()
    thr = np.array([300.0, 75.0, 100.0])
    near = 0
    pos = 0
    for i, clip in enumerate(d.cpu().numpy()):
        label, rm, sums = ref.classify_main(clip)
        pos += label
        assert int(gd[i, 0]) == len(rm), "clip %d: %d midpoints, reference %d" % (i, int(gd[i, 0]), len(rm))
        k = min(len(rm), dsp_b200.DONUT_MAX_MID)
        assert np.allclose(gm[i, :k], rm[:k], rtol=0, atol=1e-6) and np.isnan(gm[i, k:]).all()
        if sums:
            last = np.array(sums[-1])
            assert np.abs(gd[i, 1:] - last).max() <= 1e-3, "clip %d: band sums %s vs reference %s" % (i, gd[i, 1:], last)
            margin = np.abs(np.array(sums) - thr).min()  # distance of any evaluated sum to its verdict threshold
            if margin <= 1e-3:
                near += 1
                continue
        assert gl[i] == label, "clip %d: label %d, reference %d" % (i, gl[i], label)
    assert near <= 2, "%d clips within 1e-3 of a verdict threshold" % near
    assert 0.25 < pos / 768.0 < 0.75, "the f64 verdict must fire on a sizeable fraction (got %d of 768 calls)" % pos


def test_f32_variant_on_call_clips_matches_unmodified_classify_with_margins():
    """labels of the reference's own classify() on the call flavour (both verdicts, many band sums near the thresholds);
    ADVICE r1: report the margin to the thresholds so "identical labels" is not luck -- the fused kernel sums the box
    in double in warp order, the reference in float in cell order (classifier.cpp:418-430): they differ by ~1e-4"""
    import dsp_b200
    from donut_lib import DonutPort, DonutRef32, have_ref
    assert have_ref(), "oracle/_ref not shipped"
    sj = _ctx()
    d = _synth_calls(sj, 6000, 512)
    clips = d.cpu().numpy()
    gl, gd = _gpu(sj, d, dsp_b200.DONUT_SYNC_F32)
    r = DonutRef32()
    ref = np.array([r.classify(c) for c in clips], np.int32)
    ol, od = DonutPort("f32").classify_batch(clips, variant=1)
    assert np.array_equal(ol, ref), "port differs from the unmodified classify()"
    has = od[:, 0] > 0
    margin = np.abs(od[has, 1:] - np.array([200.0, 100.0, 80.0])).min(axis=1)
    tight = has.copy()
    tight[has] = margin <= 1e-2
    assert np.array_equal(gl[~tight], ref[~tight]), "labels differ on %d clips" % (gl[~tight] != ref[~tight]).sum()
    assert tight.sum() <= 5
    assert np.abs(gd[has, 1:] - od[has, 1:]).max() <= 0.01
    assert 0.2 < ref.mean() < 0.8
    print("call clips: %d positive of %d, smallest margin to a verdict threshold %.4f, %d within 0.01"
          % (ref.sum(), ref.size, margin.min(), tight.sum()))
