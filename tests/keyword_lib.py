"""ctypes access to the keyword-detector checkers.  Test infrastructure only.
  KeywordOracle  oracle/keyword_oracle.c (restatement, reads the SJKW blob)
  KeywordRef     oracle/_ref/libkeyword_ref.so = the UNMODIFIED reference sources (2fa/audio/word/c), when built"""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODEL = os.path.join(ROOT, "dsp_b200", "data", "keyword_model.bin")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libkeyword_ref.so")
MAXFR, NMFCC = 500, 13


def model_blob():
    return open(MODEL, "rb").read()


def test_signals(n_clips, n_samples, seed=11):
    """seeded speech-like test audio: noise floor + a few gated tones/chirps, some near-silent clips"""
    rng = np.random.default_rng(seed)
    t = np.arange(n_samples, dtype=np.float64) / 16000.0
    out = np.zeros((n_clips, n_samples), np.float32)
    for c in range(n_clips):
        x = rng.uniform(-1, 1, n_samples) * rng.choice([0.002, 0.02, 0.1])
        for _ in range(rng.integers(0, 4)):
            f0, f1 = rng.uniform(100, 4000, 2)
            a = rng.uniform(0.05, 0.4)
            on, off = sorted(rng.uniform(0, max(t[-1], 1e-3), 2)) if n_samples > 1 else (0, 0)
            gate = (t >= on) & (t <= off)
            x += a * np.sin(2 * np.pi * (f0 * t + 0.5 * (f1 - f0) * t * t / max(t[-1], 1e-3))) * gate
        if c % 7 == 3:
            x *= 0.0
        out[c] = np.clip(x, -0.999, 0.999).astype(np.float32)
    return out


class KeywordOracle:
    def __init__(self):
        from oracle_lib import ORACLE_DIR, build_oracle
        path = os.path.join(ORACLE_DIR, "libsj_oracle_f32.so")
        if not os.path.exists(path):
            build_oracle()
        L = ctypes.CDLL(path)
        vp = ctypes.c_void_p
        L.kw_oracle_classify.restype = ctypes.c_float
        L.kw_oracle_classify.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_int, vp, vp]
        L.kw_oracle_classify_batch.restype = ctypes.c_long
        L.kw_oracle_classify_batch.argtypes = [vp, ctypes.c_size_t, vp, ctypes.c_long, ctypes.c_int, vp]
        self.L = L
        self.blob = model_blob()

    def classify(self, x):
        x = np.ascontiguousarray(x, np.float32)
        mf = np.zeros((MAXFR, NMFCC), np.float32)
        nf = ctypes.c_int(0)
        p = self.L.kw_oracle_classify(self.blob, len(self.blob), x.ctypes.data, x.size, mf.ctypes.data, ctypes.byref(nf))
        return float(p), mf, nf.value

    def classify_batch(self, xs):
        xs = np.ascontiguousarray(xs, np.float32)
        prob = np.zeros(xs.shape[0], np.float32)
        self.L.kw_oracle_classify_batch(self.blob, len(self.blob), xs.ctypes.data, xs.shape[0], xs.shape[1], prob.ctypes.data)
        return prob


class KeywordRef:
    """the reference's own compute_mfcc / classify_signal"""

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        L = ctypes.CDLL(REF_SO)
        vp = ctypes.c_void_p
        L.compute_mfcc.restype = ctypes.c_int
        L.compute_mfcc.argtypes = [vp, ctypes.c_int, vp, ctypes.c_int]
        L.classify_signal.restype = ctypes.c_float
        L.classify_signal.argtypes = [vp, ctypes.c_int]
        self.L = L

    def classify(self, x):
        x = np.ascontiguousarray(x, np.float32)
        mf = np.zeros((MAXFR, NMFCC), np.float32)
        nf = self.L.compute_mfcc(x.ctypes.data, x.size, mf.ctypes.data, MAXFR)
        p = self.L.classify_signal(x.ctypes.data, x.size)
        return float(p), mf, nf
