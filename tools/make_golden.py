#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ (run in the build container; needs
/root/reference and sklearn).  Committed so that the GPU box, which has no /root/reference, can
check against them.

  svm_golden.json   feature vectors X (float32) and sklearn's decision_function / predict /
                    predict_proba from the reference's own cepstrum/scrubjay_svm.joblib
                    (the executable statement of the label rule, cepstrum/run.py:46).
  mfcc_golden.npz   per-frame MFCCs and mean/std features for seeded clips from an INDEPENDENT
                    float64 numpy/scipy restatement of aubio pvoc+mfcc (np.fft.rfft, scipy dct),
                    plus two real 16 kHz clips from the reference's donut-classifier/try folder
                    (channel 0, int16) with their features.
"""
import json
import os
import sys
import warnings
import wave

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def slaney_dense(sr, nfft=2048):
    """float32 restatement of aubio filterbank_mel.c (Slaney, 40 bands)"""
    f32 = np.float32
    freqs = np.zeros(42, f32)
    for i in range(13):
        freqs[i] = f32(133.3333) + f32(i) * f32(66.66666666)
    for i in range(29):
        freqs[13 + i] = freqs[12] * f32(np.power(f32(1.0711703), f32(i + 1), dtype=f32))
    nb = nfft // 2 + 1
    hz = (f32(sr) / f32(nfft)) * np.arange(nb, dtype=f32)
    fb = np.zeros((40, nb), f32)
    for m in range(40):
        lo, ce, up = freqs[m], freqs[m + 1], freqs[m + 2]
        h = f32(2.0) / (up - lo)
        b = 0
        while b < nb - 1:
            hit = hz[b] <= lo and hz[b + 1] > lo
            b += 1
            if hit:
                break
        rise = h / (ce - lo)
        while b < nb - 1:
            fb[m, b] = (hz[b] - lo) * rise
            crossed = hz[b + 1] >= ce
            b += 1
            if crossed:
                break
        down = h / (up - ce)
        while b < nb - 1:
            fb[m, b] += (up - hz[b]) * down
            if fb[m, b] < 0:
                fb[m, b] = 0
            if hz[b + 1] >= up:
                break
            b += 1
    return fb


def mfcc_stats_np(x, sr, symmetric_window=False, power=False, log_floor=2e-42, dct_ortho=True):
    """independent float64 restatement (np.fft.rfft, scipy dct) of aubio pvoc + mfcc and of the statistics of
    cepstrum/scrubjay_infer.c:36-67; the keyword arguments are the switches of scrubjay_cepstrum_params"""
    from scipy.fft import dct
    x = np.asarray(x, np.float64)
    n = x.size
    frames = (n + 1023) // 1024
    den = 2047 if symmetric_window else 2048
    w = (0.5 * (1 - np.cos(2 * np.pi * np.arange(2048) / den))).astype(np.float32).astype(np.float64)
    fb = slaney_dense(sr).astype(np.float64)
    xp = np.concatenate([np.zeros(1024), x, np.zeros(frames * 1024 - n)])
    out = []
    for t in range(frames):
        mag = np.abs(np.fft.rfft(xp[t * 1024:t * 1024 + 2048] * w)).astype(np.float32).astype(np.float64)
        if power:
            mag = (mag.astype(np.float32) * mag.astype(np.float32)).astype(np.float64)
        e = (fb @ mag).astype(np.float32)
        e = np.maximum(e, np.float32(log_floor))
        lg = np.log10(e.astype(np.float64)).astype(np.float32)
        out.append(dct(lg.astype(np.float64), type=2, norm="ortho" if dct_ortho else None)[:20].astype(np.float32))
    fr = np.array(out, np.float32)
    v = fr.astype(np.float64)
    mean = v.sum(0) / frames
    var = (v * v).sum(0) / frames - mean * mean
    feat = np.concatenate([mean.astype(np.float32), np.sqrt(np.maximum(var, 0).astype(np.float32))])
    return feat.astype(np.float32), fr


def seeded_clip(seed, n):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 16000.0
    x = 0.25 * rng.uniform(-1, 1, n)
    f0, f1 = rng.uniform(500, 7000, 2)
    x += 0.3 * np.sin(2 * np.pi * (f0 * t + 0.5 * (f1 - f0) * t * t))
    return x.astype(np.float32)


def read_wav_ch0(path):
    w = wave.open(path, "rb")
    a = np.frombuffer(w.readframes(w.getnframes()), np.int16).reshape(-1, w.getnchannels())
    return a[:, 0].copy(), w.getframerate()


def main():
    warnings.filterwarnings("ignore")
    import joblib
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    m = joblib.load(os.path.join(REF, "cepstrum", "scrubjay_svm.joblib"))
    sc, svc = m.steps[0][1], m.steps[1][1]
    rng = np.random.default_rng(20261104)
    X = [sc.mean_ + sc.scale_ * rng.normal(0, 1, (400, 40))]                       # training-scale
    X.append(sc.inverse_transform(svc.support_vectors_) + rng.normal(0, 1e-3, (55, 40)))  # SVs +- eps
    aub = np.concatenate([[-2.3], rng.normal(0, 0.05, 19), rng.uniform(0.05, 0.3, 20)])
    X.append(aub + rng.normal(0, 0.2, (400, 40)))                                  # aubio-scale features
    X = np.concatenate(X).astype(np.float32)
    Xd = X.astype(np.float64)
    g = {"X": X.tolist(), "decision_function": m.decision_function(Xd).tolist(),
         "predict": m.predict(Xd).astype(int).tolist(), "predict_proba": m.predict_proba(Xd).tolist()}
    json.dump(g, open(os.path.join(out, "svm_golden.json"), "w"))
    seeds = [1, 2, 3, 4]
    lens = [16000, 16000, 5000, 20480]
    feats, frames = [], []
    for s, n in zip(seeds, lens):
        f, fr = mfcc_stats_np(seeded_clip(s, n), 16000)
        feats.append(f)
        frames.append(fr)
    real = {}
    for name in ("good", "mic-1809bad"):
        p = os.path.join(REF, "donut-classifier", "try", name + ".wav")
        pcm, sr = read_wav_ch0(p)
        assert sr == 16000 and pcm.size >= 16000, (p, sr, pcm.size)
        pcm = pcm[:16000]
        f, _ = mfcc_stats_np(pcm.astype(np.float32) / 32768.0, sr)
        real[os.path.basename(p)] = (pcm, f)
    np.savez_compressed(os.path.join(out, "mfcc_golden.npz"), seeds=np.array(seeds), lens=np.array(lens),
                        feats=np.array(feats), frames0=frames[0], frames2=frames[2],
                        real_names=np.array(list(real.keys())),
                        real_pcm=np.array([v[0] for v in real.values()]),
                        real_feats=np.array([v[1] for v in real.values()]))
    print("wrote golden fixtures:", os.listdir(out))




def make_donut_golden():
    """donut fixtures: (a) the reference's own IIR known-answer vector donut-classifier/_postbutter.txt
    (first 6000 samples; input = channel 0 of donut-classifier/testing/1060-control.wav / 32768, filter
    = the commented-out 48 kHz-design coefficient set of classifier.c:323-341); (b) labels, midpoint
    counts and band sums of the UNMODIFIED reference builds (oracle/_ref) on seeded synthetic clips and
    on the reference's 1 s / 16 kHz WAVs."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from donut_lib import DonutPort, DonutRef32, DonutRef64
    from synth_np import synth_clips
    out = os.path.join(ROOT, "tests", "golden")
    pcm, sr = read_wav_ch0(os.path.join(REF, "donut-classifier", "testing", "1060-control.wav"))
    want = np.loadtxt(os.path.join(REF, "donut-classifier", "_postbutter.txt"), max_rows=6000)
    b48 = [0.00021314, 0., -0.00085255, 0., 0.00127883, 0., -0.00085255, 0., 0.00021314]
    a48 = [1., -7.12847885, 22.41882266, -40.62891245, 46.40780141, -34.21333503, 15.89913237, -4.25840048, 0.50337536]
    clips = synth_clips(1000, 96, 16000)
    r32, r64, p64 = DonutRef32(), DonutRef64(), DonutPort("f64")
    ref32_labels = np.array([r32.classify(c) for c in clips], np.int32)
    ref64_nmid = np.array([len(r64.midpoints(c)) for c in clips], np.int32)
    real = []
    for name in ("good", "mic-1809bad", "old", "without_filter"):
        p, _ = read_wav_ch0(os.path.join(REF, "donut-classifier", "try", name + ".wav"))
        real.append(p[:16000])
    real = np.array(real)
    realf = real.astype(np.float32) / 32768.0
    np.savez_compressed(os.path.join(out, "donut_golden.npz"), kat_pcm=pcm[:6000], kat_sr=sr, kat_out=want,
                        kat_b=np.array(b48), kat_a=np.array(a48), synth_first=1000, synth_n=96,
                        ref32_labels=ref32_labels, ref64_nmid=ref64_nmid, real_pcm=real,
                        real_ref32_labels=np.array([r32.classify(c) for c in realf], np.int32),
                        real_ref64_iir0=np.array([r64.filter(c, 1000, 3000)[-64:] for c in realf]),
                        real_ref64_spec=np.array([r64.spectrogram(r64.filter(c, 3000, 7500))[40:44, 30:34] for c in realf]))
    print("donut golden: ref32 positives %d/96, real labels %s" % (ref32_labels.sum(), [r32.classify(c) for c in realf]))


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    main()
    make_donut_golden()
