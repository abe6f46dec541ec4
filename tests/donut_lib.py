"""ctypes access to the donut checkers: the port (oracle/donut_oracle.c, inside libsj_oracle_*.so)
and the UNMODIFIED reference compiled into oracle/_ref/ (when built).  Test infrastructure only."""
import ctypes
import os

import numpy as np

from oracle_lib import ORACLE_DIR, build_oracle

REF32 = os.path.join(ORACLE_DIR, "_ref", "libdonut_ref_f32.so")
REF64 = os.path.join(ORACLE_DIR, "_ref", "libdonut_ref_f64.so")
REFINO = os.path.join(ORACLE_DIR, "_ref", "libdonut_ref_ino.so")
vp, c_int, c_long = ctypes.c_void_p, ctypes.c_int, ctypes.c_long


def _silence_stdout(fn, *a):
    """the reference prints from inside its hot functions (classifier.cpp:85,104-106): mute fd 1"""
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)
    os.dup2(null, 1)
    try:
        return fn(*a)
    finally:
        ctypes.CDLL(None).fflush(None)
        os.dup2(saved, 1)
        os.close(null)
        os.close(saved)


class DonutPort:
    def __init__(self, precision="f64"):
        path = os.path.join(ORACLE_DIR, "libsj_oracle_%s.so" % precision)
        if not os.path.exists(path):
            build_oracle()
        self.L = ctypes.CDLL(path)
        self.dt = np.float32 if precision == "f32" else np.float64
        self.L.sj_oracle_donut_batch.restype = c_long
        self.L.sj_oracle_donut_batch.argtypes = [vp, c_long, c_int, c_int, vp, vp]
        self.L.sj_oracle_donut_filter.argtypes = [vp, c_int, c_int, vp]
        self.L.sj_oracle_donut_filter_coef.argtypes = [vp, c_int, vp, vp, vp]
        self.L.sj_oracle_donut_spectrogram.argtypes = [vp, c_int, c_int, vp, c_int]

    def filter(self, x, which):
        x = np.ascontiguousarray(x, self.dt)
        out = np.empty_like(x)
        self.L.sj_oracle_donut_filter(x.ctypes.data, x.size, which, out.ctypes.data)
        return out

    def filter_coef(self, x, b, a):
        x = np.ascontiguousarray(x, self.dt)
        b = np.ascontiguousarray(b, self.dt)
        a = np.ascontiguousarray(a, self.dt)
        out = np.empty_like(x)
        self.L.sj_oracle_donut_filter_coef(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
        return out

    def spectrogram(self, sig):
        sig = np.ascontiguousarray(sig, self.dt)
        tb = (sig.size - 256) // 224 + 1
        S = np.empty((129, tb), self.dt)
        r = self.L.sj_oracle_donut_spectrogram(sig.ctypes.data, sig.size, 16000, S.ctypes.data, tb)
        assert r == tb
        return S

    def classify_batch(self, clips, variant):
        clips = np.ascontiguousarray(clips, np.float32)
        n, ns = clips.shape
        labels = np.zeros(n, np.int32)
        detail = np.zeros((n, 4), np.float32)
        self.L.sj_oracle_donut_batch(clips.ctypes.data, n, ns, variant, labels.ctypes.data, detail.ctypes.data)
        return labels, detail


class DonutRef32:
    """sync/lib/classifier.cpp compiled as-is (C++ mangled names)"""

    def __init__(self):
        self.L = ctypes.CDLL(REF32)
        self.classify_fn = getattr(self.L, "_Z8classifyPfi")
        self.classify_fn.argtypes = [vp, c_int]
        self.filter_fn = getattr(self.L, "_Z22butter_bandpass_filterPfiS_S_S_")
        self.filter_fn.argtypes = [vp, c_int, vp, vp, vp]
        self.bb_fn = getattr(self.L, "_Z15butter_bandpassffPfS_")
        self.bb_fn.argtypes = [ctypes.c_float, ctypes.c_float, vp, vp]
        self.bb_fn.restype = ctypes.c_bool
        self.spec_fn = getattr(self.L, "_Z19compute_spectrogramPfiiPS_S0_PS0_PiS2_")
        self.spec_fn.argtypes = [vp, c_int, c_int, vp, vp, vp, vp, vp]

        self.mid_fn = getattr(self.L, "_Z14find_midpointsPfiiPi")
        self.mid_fn.argtypes = [vp, c_int, c_int, vp]
        self.mid_fn.restype = ctypes.POINTER(ctypes.c_float)
        self.sum_fn = getattr(self.L, "_Z11sum_intensefffPfiS_iPS_f")
        self.sum_fn.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_float, vp, c_int, vp, c_int, vp, ctypes.c_float]
        self.sum_fn.restype = ctypes.c_float

    def classify(self, clip):
        x = np.ascontiguousarray(clip, np.float32).copy()
        return int(_silence_stdout(self.classify_fn, x.ctypes.data, x.size))

    def filter_coef(self, x, b, a):
        x = np.ascontiguousarray(x, np.float32).copy()
        b = np.ascontiguousarray(b, np.float32).copy()
        a = np.ascontiguousarray(a, np.float32).copy()
        out = np.empty_like(x)
        self.filter_fn(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
        return out

    def midpoints(self, clip):
        x = np.ascontiguousarray(clip, np.float32).copy()
        n = c_int()
        p = _silence_stdout(self.mid_fn, x.ctypes.data, x.size, 16000, ctypes.byref(n))
        return np.array([p[i] for i in range(n.value)], np.float32)

    def sum_intense(self, lower, upper, half_range, freqs, times, plane, midpoint):
        fr = np.ascontiguousarray(freqs, np.float32)
        tm = np.ascontiguousarray(times, np.float32)
        pl = np.ascontiguousarray(plane, np.float32)
        rows = (vp * pl.shape[0])(*[pl[i].ctypes.data for i in range(pl.shape[0])])
        return float(self.sum_fn(lower, upper, half_range, fr.ctypes.data, fr.size, tm.ctypes.data, tm.size,
                                 ctypes.cast(rows, vp), midpoint))

    def filter(self, x, lo, hi):
        x = np.ascontiguousarray(x, np.float32).copy()
        b = np.zeros(9, np.float32)
        a = np.zeros(9, np.float32)
        assert self.bb_fn(lo, hi, b.ctypes.data, a.ctypes.data)
        out = np.empty_like(x)
        self.filter_fn(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
        return out

    def spectrogram(self, sig):
        sig = np.ascontiguousarray(sig, np.float32).copy()
        fr, tm, S = vp(), vp(), vp()
        fb, tb = c_int(), c_int()
        self.spec_fn(sig.ctypes.data, sig.size, 16000, ctypes.byref(fr), ctypes.byref(tm), ctypes.byref(S),
                     ctypes.byref(fb), ctypes.byref(tb))
        rows = ctypes.cast(S, ctypes.POINTER(ctypes.POINTER(ctypes.c_float)))
        out = np.array([np.ctypeslib.as_array(rows[i], shape=(tb.value,)).copy() for i in range(fb.value)])
        return out


class DonutRef64:
    """donut-classifier/classifier.c compiled with -Dmain=ref_main and the oracle/shim headers"""

    def __init__(self):
        self.L = ctypes.CDLL(REF64)
        self.L.butter_bandpass.argtypes = [ctypes.c_double, ctypes.c_double, vp, vp]
        self.L.butter_bandpass.restype = ctypes.c_bool
        self.L.butter_bandpass_filter.argtypes = [vp, c_int, vp, vp, vp]
        self.L.compute_spectrogram.argtypes = [vp, c_int, c_int, vp, vp, vp, vp, vp]
        self.L.find_midpoints.argtypes = [vp, c_int, c_int, vp]
        self.L.find_midpoints.restype = ctypes.POINTER(ctypes.c_double)
        cd = ctypes.c_double
        self.L.sum_intense.argtypes = [cd, cd, cd, vp, c_int, vp, c_int, vp, cd]
        self.L.sum_intense.restype = cd

    def sum_intense(self, lower, upper, half_range, freqs, times, plane, midpoint):
        """the unmodified sum_intense (classifier.c:594-655) on a [freq][time] double plane (NaN = filtered out)"""
        fr = np.ascontiguousarray(freqs, np.float64)
        tm = np.ascontiguousarray(times, np.float64)
        pl = np.ascontiguousarray(plane, np.float64)
        rows = (vp * pl.shape[0])(*[pl[i].ctypes.data for i in range(pl.shape[0])])
        return float(self.L.sum_intense(lower, upper, half_range, fr.ctypes.data, fr.size, tm.ctypes.data, tm.size,
                                        ctypes.cast(rows, vp), midpoint))

    def classify_main(self, clip, lo_norm=0.70, hi_norm=0.85, thr=(300.0, 75.0, 100.0)):
        """The per-file body of main (classifier.c:84-198) driven through the UNMODIFIED functions of the reference
        build: butter_bandpass_filter, compute_spectrogram, find_midpoints and sum_intense are the reference's own
        code; only main's inline element-wise step (:104-156: dB, min / max, normalise, keep lo < v < hi) is restated
        here in float64 numpy.  Returns (label, midpoints, [(above, middle, below) per evaluated midpoint])."""
        x = np.ascontiguousarray(clip, np.float64)
        S = self.spectrogram(self.filter(x, 3000, 7500))
        import warnings
        with np.errstate(divide="ignore", invalid="ignore"), warnings.catch_warnings():
            warnings.simplefilter("ignore")  # an all-zero clip has no positive cell: min / max stay NaN, nothing passes
            dB = np.where(S > 0, 10.0 * np.log10(S / 1e-12), np.nan)
            mn, mx = np.nanmin(dB), np.nanmax(dB)
            nv = (dB - mn) / (mx - mn)
            plane = np.where((nv > lo_norm) & (nv < hi_norm), nv, np.nan)
        freqs = np.arange(S.shape[0]) * (16000.0 / 256.0)              # classifier.c:466-469
        times = (np.arange(S.shape[1]) * 224 + 128) / 16000.0          # :471-474
        mids = self.midpoints(x)
        sums, label = [], 0
        for m in mids:
            a = _silence_stdout(self.sum_intense, 5000, 7000, 0.18, freqs, times, plane, m)
            mi = _silence_stdout(self.sum_intense, 2500, 5000, 0.05, freqs, times, plane, m)
            b = _silence_stdout(self.sum_intense, 500, 2500, 0.18, freqs, times, plane, m)
            sums.append((a, mi, b))
            if mi < thr[1] and a > thr[0] and b > thr[2]:               # :184
                label = 1
                break
        return label, mids, sums

    def filter(self, x, lo, hi):
        x = np.ascontiguousarray(x, np.float64).copy()
        b = np.zeros(9)
        a = np.zeros(9)
        assert self.L.butter_bandpass(lo, hi, b.ctypes.data, a.ctypes.data)
        out = np.empty_like(x)
        self.L.butter_bandpass_filter(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
        return out

    def filter_coef(self, x, b, a):
        x = np.ascontiguousarray(x, np.float64).copy()
        b = np.ascontiguousarray(b, np.float64)
        a = np.ascontiguousarray(a, np.float64)
        out = np.empty_like(x)
        self.L.butter_bandpass_filter(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
        return out

    def spectrogram(self, sig):
        sig = np.ascontiguousarray(sig, np.float64).copy()
        fr, tm, S = vp(), vp(), vp()
        fb, tb = c_int(), c_int()
        self.L.compute_spectrogram(sig.ctypes.data, sig.size, 16000, ctypes.byref(fr), ctypes.byref(tm),
                                   ctypes.byref(S), ctypes.byref(fb), ctypes.byref(tb))
        rows = ctypes.cast(S, ctypes.POINTER(ctypes.POINTER(ctypes.c_double)))
        return np.array([np.ctypeslib.as_array(rows[i], shape=(tb.value,)).copy() for i in range(fb.value)])

    def midpoints(self, clip):
        x = np.ascontiguousarray(clip, np.float64).copy()
        n = c_int()
        p = _silence_stdout(self.L.find_midpoints, x.ctypes.data, x.size, 16000, ctypes.byref(n))
        return np.array([p[i] for i in range(n.value)])


class DonutRefIno:
    """donut-classifier/classifier.ino (64-sample windows) compiled as C++ against oracle/shim/ino"""

    def __init__(self):
        self.L = ctypes.CDLL(REFINO)
        self.spec_fn = getattr(self.L, "_Z19compute_spectrogramPfiiPS_S0_PS0_PiS2_")
        self.spec_fn.argtypes = [vp, c_int, c_int, vp, vp, vp, vp, vp]

    def spectrogram(self, sig, fs=16000):
        sig = np.ascontiguousarray(sig, np.float32).copy()
        fr, tm, S = vp(), vp(), vp()
        fb, tb = c_int(), c_int()
        self.spec_fn(sig.ctypes.data, sig.size, fs, ctypes.byref(fr), ctypes.byref(tm), ctypes.byref(S),
                     ctypes.byref(fb), ctypes.byref(tb))
        rows = ctypes.cast(S, ctypes.POINTER(ctypes.POINTER(ctypes.c_float)))
        return np.array([np.ctypeslib.as_array(rows[i], shape=(tb.value,)).copy() for i in range(fb.value)])


    def classify(self, clip):
        """the sketch's setup() on a one-second clip -> (label, number of midpoints, [(above, middle, below), ...])"""
        x = np.ascontiguousarray(clip, np.float32)
        assert x.size == 16000
        buf = (ctypes.c_float * 16000).in_dll(self.L, "data")
        ctypes.memmove(buf, x.ctypes.data, 16000 * 4)
        cnt = ctypes.c_int.in_dll(self.L, "sj_shim_serial_count")
        scrub = ctypes.c_int.in_dll(self.L, "sj_shim_serial_scrub")
        nums = (ctypes.c_double * 4096).in_dll(self.L, "sj_shim_serial_numbers")
        cnt.value = 0
        scrub.value = -1
        getattr(self.L, "_Z5setupv")()
        vals = [nums[i] for i in range(cnt.value)]
        n_mid = int(vals[0]) if vals else 0
        sums = [tuple(vals[1 + 3 * k:4 + 3 * k]) for k in range((len(vals) - 1) // 3)]
        return (1 if scrub.value == 1 else 0), n_mid, sums


def have_ref_ino():
    return os.path.exists(REFINO)


def have_ref():
    return os.path.exists(REF32) and os.path.exists(REF64)
