"""ctypes access to the CPU oracle (oracle/*.c).  Test infrastructure only: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by dsp_b200."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
MODEL_BLOB = os.path.join(ROOT, "dsp_b200", "data", "scrubjay_svm.bin")


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


class CepstrumOracle:
    """oracle/cepstrum_oracle.c: restated aubio pvoc+mfcc (+ stats) and ORT Scaler+SVMClassifier."""

    def __init__(self, samplerate=16000, precision="f64", params=None):
        """params: a dsp_b200.CepstrumParams (the POD both the library and this oracle consume) or None = defaults"""
        path = os.path.join(ORACLE_DIR, "libsj_oracle_%s.so" % precision)
        if not os.path.exists(path):
            build_oracle()
        L = ctypes.CDLL(path)
        vp, c_long = ctypes.c_void_p, ctypes.c_long
        L.sj_oracle_new_ex.restype = vp
        L.sj_oracle_new_ex.argtypes = [ctypes.c_uint, vp]
        L.sj_oracle_predict_ex.restype = ctypes.c_int
        L.sj_oracle_predict_ex.argtypes = [vp, vp, ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float)]
        L.sj_oracle_free.argtypes = [vp]
        L.sj_oracle_mfcc_stats.restype = c_long
        L.sj_oracle_mfcc_stats.argtypes = [vp, vp, c_long, vp, vp]
        L.sj_oracle_predict.restype = ctypes.c_int
        L.sj_oracle_predict.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float)]
        L.sj_oracle_infer_batch.restype = c_long
        L.sj_oracle_infer_batch.argtypes = [vp, vp, vp, c_long, c_long, vp, vp, vp, vp]
        L.sj_oracle_fbank.restype = ctypes.POINTER(ctypes.c_float)
        L.sj_oracle_fbank.argtypes = [vp]
        self.L = L
        self.params = params
        self.h = L.sj_oracle_new_ex(samplerate, ctypes.addressof(params) if params is not None else None)
        self.label_rule = params.label_rule if params is not None else 0
        self.blob = open(MODEL_BLOB, "rb").read()

    def fbank(self):
        return np.ctypeslib.as_array(self.L.sj_oracle_fbank(self.h), shape=(40, 1025)).copy()

    def mfcc_stats(self, x, want_frames=False):
        x = np.ascontiguousarray(x, np.float32)
        out = np.zeros(40, np.float32)
        nfr = (x.size + 1023) // 1024
        fr = np.zeros((max(nfr, 1), 20), np.float32)
        n = self.L.sj_oracle_mfcc_stats(self.h, x.ctypes.data, x.size, out.ctypes.data, fr.ctypes.data)
        if n < 0:
            return None
        return (out, fr[:n]) if want_frames else out

    def predict(self, feat, blob=None):
        f = np.ascontiguousarray(feat, np.float32)
        c, s = ctypes.c_float(), ctypes.c_float()
        label = self.L.sj_oracle_predict_ex(blob or self.blob, f.ctypes.data, self.label_rule, ctypes.byref(c), ctypes.byref(s))
        return label, c.value, s.value

    def infer_batch(self, clips):
        clips = np.ascontiguousarray(clips, np.float32)
        n, ns = clips.shape
        labels = np.zeros(n, np.int32)
        conf = np.zeros(n, np.float32)
        score = np.zeros(n, np.float32)
        feats = np.zeros((n, 40), np.float32)
        r = self.L.sj_oracle_infer_batch(self.h, self.blob, clips.ctypes.data, n, ns, labels.ctypes.data,
                                         conf.ctypes.data, score.ctypes.data, feats.ctypes.data)
        assert r == n
        return labels, conf, score, feats

    def stream_infer(self, x, win_frames=16, step_frames=16, prev_hop=None):
        """Stream mode restated on the oracle: one channel `x`, the pvoc running on without a restart
        (per-frame MFCCs from sj_oracle_mfcc_stats over the whole channel), then for every window of
        `win_frames` frames advancing by `step_frames` the statistics of cepstrum/scrubjay_infer.c:61-67
        (double sums in frame order, float sqrt of the float-converted variance) and sj_oracle_predict.
        prev_hop (1024 samples) is the hop before x[0]: it is prepended and its own frame dropped."""
        x = np.ascontiguousarray(x, np.float32)
        if prev_hop is not None:
            x = np.concatenate([np.asarray(prev_hop, np.float32), x])
        _, fr = self.mfcc_stats(x, want_frames=True)
        if prev_hop is not None:
            fr = fr[1:]
        nw = 0 if len(fr) < win_frames else (len(fr) - win_frames) // step_frames + 1
        labels = np.zeros(nw, np.int32)
        conf = np.zeros(nw, np.float32)
        score = np.zeros(nw, np.float32)
        feats = np.zeros((nw, 40), np.float32)
        for j in range(nw):
            w = fr[j * step_frames:j * step_frames + win_frames].astype(np.float64)
            s, q = np.zeros(20), np.zeros(20)
            for v in w:  # running sums in frame order, like the reference's loop
                s += v
                q += v * v
            mean = s / win_frames
            var = np.maximum(q / win_frames - mean * mean, 0.0)
            feats[j, :20] = mean.astype(np.float32)
            feats[j, 20:] = np.sqrt(var.astype(np.float32))
            labels[j], conf[j], score[j] = self.predict(feats[j])
        return labels, conf, score, feats
