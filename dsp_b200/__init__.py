"""dsp_b200 -- host-side Python mirror of libscrubjay_cuda's C ABI (include/scrubjay_cuda.h).

The product is the C shared library `dsp_b200/libscrubjay_cuda.so` (hand-written sm_100a CUDA behind
plain-C entry points, see dsp_b200/csrc).  This module is only a thin ctypes binding so that tests
and bench.py can call the library the way a C host would; it holds no arithmetic and no fallback:
if the library is missing or no B200 is present the calls raise.

Reference interface mirrored (cornell-c2s2/dsp):
  mfcc_stats   cepstrum/scrubjay_infer.c:19-77   -> ScrubJay.mfcc_stats / mfcc_stats_batch_dev
  ort_init     cepstrum/scrubjay_infer.c:93-102  -> ScrubJay(...)
  ort_predict  cepstrum/scrubjay_infer.c:105-141 -> ScrubJay.predict / predict_batch_dev
  main loop    cepstrum/scrubjay_infer.c:156-177 -> ScrubJay.infer_batch / infer_batch_dev
"""
import ctypes
import os

import numpy as np

N_MFCC = 20
FEAT_DIM = 40
WIN_SIZE = 2048
HOP_SIZE = 1024

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SCRUBJAY_LIB") or os.path.join(_HERE, "libscrubjay_cuda.so")  # override: A/B builds of the same ABI
DEFAULT_MODEL = os.path.join(_HERE, "data", "scrubjay_svm.bin")
KEYWORD_MODEL = os.path.join(_HERE, "data", "keyword_model.bin")

_lib = None


F32, F64 = 0, 1                  # element type of the donut stage functions
IIR_EXACT, IIR_SCAN = 0, 1       # one recurrence per lane (bit-identical) / block-wise state-space scan over time
BUTTER_16K, BUTTER_48K_DESIGN = 0, 1
DONUT_MAX_MID = 8
MAX_PEERS = 8
NOALLOC_MAX_TOTAL_FRAMES = 16384
PEER_HANDLE_BYTES = 64


class ScrubJayError(RuntimeError):
    pass


def load_library():
    """dlopen the in-tree library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ScrubJayError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "or `make -C dsp_b200/csrc`" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, c_long, c_int, c_uint = ctypes.c_void_p, ctypes.c_long, ctypes.c_int, ctypes.c_uint
    L.scrubjay_create.argtypes = [ctypes.POINTER(vp), c_int, vp, ctypes.c_size_t, c_uint]
    L.scrubjay_create_ex.argtypes = [ctypes.POINTER(vp), c_int, vp, ctypes.c_size_t, c_uint, vp]
    L.scrubjay_cepstrum_params_default.argtypes = [vp]
    L.scrubjay_cepstrum_params_default.restype = None
    L.scrubjay_destroy.argtypes = [vp]
    L.scrubjay_destroy.restype = None
    L.scrubjay_last_error.argtypes = [vp]
    L.scrubjay_last_error.restype = ctypes.c_char_p
    L.scrubjay_set_stream.argtypes = [vp, vp]
    L.scrubjay_launch_count.argtypes = [vp]
    L.scrubjay_launch_count.restype = c_long
    L.scrubjay_last_kernel_ms.argtypes = [vp, ctypes.POINTER(ctypes.c_float)]
    L.scrubjay_mfcc_stats_batch_dev.argtypes = [vp, vp, c_long, c_int, vp]
    L.scrubjay_predict_batch_dev.argtypes = [vp, vp, c_long, vp, vp, vp]
    L.scrubjay_infer_batch_dev.argtypes = [vp, vp, c_long, c_int, vp, vp, vp, vp]
    L.scrubjay_infer_batch_pcm16_dev.argtypes = [vp, vp, c_long, c_int, vp, vp, vp, vp]
    L.scrubjay_infer_batch_noalloc_dev.argtypes = [vp, vp, c_long, c_int, vp, vp, vp, vp]
    L.scrubjay_peer_alloc.argtypes = [vp, ctypes.c_size_t, ctypes.POINTER(vp), vp]
    L.scrubjay_peer_free.argtypes = [vp, vp]
    L.scrubjay_peer_open.argtypes = [vp, vp, ctypes.POINTER(vp)]
    L.scrubjay_peer_close.argtypes = [vp, vp]
    L.scrubjay_gather_bind.argtypes = [vp, c_int, vp, vp, c_long]
    L.scrubjay_infer_batch_gather_dev.argtypes = [vp, vp, c_long, c_int, vp, vp]
    L.scrubjay_infer_batch_pcm16.argtypes = [vp, vp, c_long, c_int, vp, vp, vp, vp]
    L.scrubjay_wav_parse.argtypes = [vp, ctypes.c_size_t, vp]
    L.scrubjay_wav_parse_ex.argtypes = [vp, ctypes.c_size_t, vp, c_int]
    L.scrubjay_decode_pcm16_dev.argtypes = [vp, vp, c_long, c_int, c_int, vp]
    L.scrubjay_decode_adc12_dev.argtypes = [vp, vp, c_long, vp]
    L.scrubjay_upsample_linear_dev.argtypes = [vp, vp, c_long, c_int, c_int, vp]
    L.scrubjay_keyword_load.argtypes = [vp, vp, ctypes.c_size_t]
    L.scrubjay_keyword_classify_batch_dev.argtypes = [vp, vp, c_long, c_int, vp, vp, vp]
    L.scrubjay_keyword_classify_batch.argtypes = [vp, vp, c_long, c_int, vp]
    L.classify_signal.argtypes = [vp, c_int]
    L.classify_signal.restype = ctypes.c_float
    L.classify_signal_binary.argtypes = [vp, c_int, ctypes.c_float]
    L.scrubjay_stream_windows.argtypes = [c_long, c_int, c_int]
    L.scrubjay_stream_windows.restype = c_long
    L.scrubjay_stream_infer_dev.argtypes = [vp, vp, c_long, c_long, c_long, vp, c_int, c_int, vp, vp, vp, vp]
    L.scrubjay_stream_infer_pcm16_dev.argtypes = [vp, vp, c_long, c_long, c_long, vp, c_int, c_int, vp, vp, vp, vp]
    L.scrubjay_stream_infer_gather_dev.argtypes = [vp, vp, c_long, c_long, c_long, vp, c_int, c_int, vp, vp]
    L.scrubjay_stream_infer_pcm16_gather_dev.argtypes = [vp, vp, c_long, c_long, c_long, vp, c_int, c_int, vp, vp]
    L.scrubjay_set_stream_chunk.argtypes = [vp, c_int]
    L.scrubjay_set_latency_mode.argtypes = [vp, c_int]
    L.scrubjay_mfcc_stats.argtypes = [vp, vp, c_long, vp]
    L.scrubjay_predict.argtypes = [vp, vp, ctypes.POINTER(ctypes.c_float)]
    L.scrubjay_infer_batch.argtypes = [vp, vp, c_long, c_int, vp, vp, vp, vp]
    L.scrubjay_synth_clips_dev.argtypes = [vp, vp, c_long, c_long, c_int, c_uint]
    L.scrubjay_synth_clips_ex_dev.argtypes = [vp, vp, c_long, c_long, c_int, c_uint, c_int]
    L.scrubjay_donut_classify_batch_dev.argtypes = [vp, vp, c_long, c_int, c_int, vp, vp]
    L.scrubjay_donut_classify_batch.argtypes = [vp, vp, c_long, c_int, c_int, vp, vp]
    L.classify.argtypes = [vp, c_int]
    L.scrubjay_donut_midpoints_batch_dev.argtypes = [vp, vp, c_long, c_int, c_int, vp, vp, vp]
    L.scrubjay_donut_classify_long_dev.argtypes = [vp, vp, c_long, c_long, c_int, vp, vp, vp, c_int]
    L.scrubjay_donut_stream_seconds.argtypes = [c_long, c_int]
    L.scrubjay_donut_stream_seconds.restype = c_long
    L.scrubjay_donut_classify_streams_dev.argtypes = [vp, vp, c_long, c_long, c_long, c_int, c_int, vp, vp]
    L.butter_bandpass.argtypes = [ctypes.c_float, ctypes.c_float, vp, vp]
    L.butter_bandpass.restype = ctypes.c_bool
    L.scrubjay_butter_bandpass_f64.argtypes = [ctypes.c_double, ctypes.c_double, c_int, vp, vp]
    L.scrubjay_butter_filter_batch_dev.argtypes = [vp, vp, vp, c_long, c_long, c_long, c_int, vp, vp, vp, c_int]
    L.scrubjay_butter_filter_batch.argtypes = [vp, vp, vp, c_long, c_long, c_int, vp, vp, c_int]
    L.butter_bandpass_filter.argtypes = [vp, c_int, vp, vp, vp]
    L.butter_bandpass_filter.restype = None
    L.scrubjay_spectrogram_frames.argtypes = [c_long, c_int]
    L.scrubjay_spectrogram_frames.restype = c_long
    L.scrubjay_spectrogram_batch_dev.argtypes = [vp, vp, c_long, c_long, c_long, c_int, c_int, c_int, vp]
    L.scrubjay_spectrogram_batch.argtypes = [vp, vp, c_long, c_long, c_int, c_int, c_int, vp]
    fpp = ctypes.POINTER(ctypes.POINTER(ctypes.c_float))
    L.compute_spectrogram.argtypes = [vp, c_int, c_int, fpp, fpp, ctypes.POINTER(fpp), ctypes.POINTER(c_int), ctypes.POINTER(c_int)]
    L.compute_spectrogram.restype = None
    L.find_midpoints.argtypes = [vp, c_int, c_int, ctypes.POINTER(c_int)]
    L.find_midpoints.restype = ctypes.POINTER(ctypes.c_float)
    L.sum_intense.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_float, vp, c_int, vp, c_int, vp, ctypes.c_float]
    L.sum_intense.restype = ctypes.c_float
    L.scrubjay_debug_inject_failure.argtypes = [vp, c_long]
    L.scrubjay_host_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_size_t]
    L.scrubjay_host_free.argtypes = [vp]
    L.scrubjay_host_free.restype = None
    _lib = L
    return L


def _ptr(x):
    """Raw address of a device tensor (anything with data_ptr()), a numpy array, an int or None."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    raise TypeError("cannot take the address of %r" % type(x))


WINDOW_HANN_PERIODIC, WINDOW_HANN_SYMMETRIC = 0, 1
SPECTRUM_MAGNITUDE, SPECTRUM_POWER = 0, 1
DCT_ORTHO, DCT_UNNORMALISED = 0, 1
LABEL_PREDICT, LABEL_VOTE, LABEL_PROB_ARGMAX = 0, 1, 2


class CepstrumParams(ctypes.Structure):
    """scrubjay_cepstrum_params (include/scrubjay_cuda.h): the switchable aubio-isms / ORT-isms of the cepstrum path.
    CepstrumParams() holds the defaults (the restated aubio 0.4.x / onnxruntime behaviour); keyword arguments override."""
    _fields_ = [("window", ctypes.c_int32), ("spectrum", ctypes.c_int32), ("log_floor", ctypes.c_float),
                ("dct_norm", ctypes.c_int32), ("half_swap", ctypes.c_int32), ("label_rule", ctypes.c_int32),
                ("reserved", ctypes.c_int32 * 2)]

    def __init__(self, **kw):
        super().__init__()
        self.window, self.spectrum, self.log_floor = WINDOW_HANN_PERIODIC, SPECTRUM_MAGNITUDE, 2e-42
        self.dct_norm, self.half_swap, self.label_rule = DCT_ORTHO, 1, LABEL_PREDICT
        for k, v in kw.items():
            if k not in dict(self._fields_):
                raise TypeError("unknown cepstrum parameter %r" % k)
            setattr(self, k, v)


class ScrubJay:
    """One context per GPU (scrubjay_create / scrubjay_destroy)."""

    def __init__(self, device=0, samplerate=16000, model_blob=None, params=None):
        self._L = load_library()
        self._h = ctypes.c_void_p()
        if model_blob is None:
            blob, n = None, 0
        else:
            blob = bytes(model_blob)
            n = len(blob)
        rc = self._L.scrubjay_create_ex(ctypes.byref(self._h), device, blob, n, samplerate,
                                        ctypes.addressof(params) if params is not None else None)
        if rc != 0:
            msg = self._L.scrubjay_last_error(None).decode()
            self._h = ctypes.c_void_p()
            raise ScrubJayError("scrubjay_create failed (%d): %s" % (rc, msg))
        self.device = device
        self.samplerate = samplerate

    def close(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self._L.scrubjay_destroy(h)
            h.value = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter shutdown
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc < 0:
            raise ScrubJayError("libscrubjay_cuda error %d: %s" % (rc, self._L.scrubjay_last_error(self._h).decode()))
        return rc

    # ---- plumbing
    def set_stream(self, cuda_stream):
        """cuda_stream: integer cudaStream_t handle (e.g. torch.cuda.current_stream().cuda_stream)."""
        self._check(self._L.scrubjay_set_stream(self._h, cuda_stream))

    @property
    def launch_count(self):
        return self._L.scrubjay_launch_count(self._h)

    def last_kernel_ms(self):
        ms = ctypes.c_float()
        self._check(self._L.scrubjay_last_kernel_ms(self._h, ctypes.byref(ms)))
        return ms.value

    # ---- device-pointer entry points
    def mfcc_stats_batch_dev(self, d_clips, n_clips, n_samples, d_feats):
        self._check(self._L.scrubjay_mfcc_stats_batch_dev(self._h, _ptr(d_clips), n_clips, n_samples, _ptr(d_feats)))

    def predict_batch_dev(self, d_feats, n, d_labels, d_conf=None, d_score=None):
        self._check(self._L.scrubjay_predict_batch_dev(self._h, _ptr(d_feats), n, _ptr(d_labels), _ptr(d_conf),
                                                       _ptr(d_score)))

    def infer_batch_dev(self, d_clips, n_clips, n_samples, d_labels, d_conf=None, d_score=None, d_feats=None):
        self._check(self._L.scrubjay_infer_batch_dev(self._h, _ptr(d_clips), n_clips, n_samples, _ptr(d_labels),
                                                     _ptr(d_conf), _ptr(d_score), _ptr(d_feats)))

    def infer_batch_noalloc_dev(self, d_clips, n_clips, n_samples, d_labels, d_conf=None, d_score=None, d_feats=None):
        """scrubjay_infer_batch_noalloc_dev: small batches, scratch in the library's static device array, no allocation
        (n_clips * ceil(n_samples / 1024) <= NOALLOC_MAX_TOTAL_FRAMES)"""
        self._check(self._L.scrubjay_infer_batch_noalloc_dev(self._h, _ptr(d_clips), n_clips, n_samples, _ptr(d_labels),
                                                             _ptr(d_conf), _ptr(d_score), _ptr(d_feats)))

    # ---- fused multi-GPU gather (peer stores over NVLink instead of a collective after the kernel)
    def peer_alloc(self, nbytes, want_handle=True):
        """-> (device address, 64-byte IPC handle or None)"""
        p = ctypes.c_void_p()
        h = (ctypes.c_ubyte * PEER_HANDLE_BYTES)() if want_handle else None
        self._check(self._L.scrubjay_peer_alloc(self._h, nbytes, ctypes.byref(p), h))
        return p.value, (bytes(h) if want_handle else None)

    def peer_free(self, addr):
        self._check(self._L.scrubjay_peer_free(self._h, addr))

    def peer_open(self, handle):
        p = ctypes.c_void_p()
        buf = (ctypes.c_ubyte * PEER_HANDLE_BYTES).from_buffer_copy(bytes(handle))
        self._check(self._L.scrubjay_peer_open(self._h, buf, ctypes.byref(p)))
        return p.value

    def peer_close(self, addr):
        self._check(self._L.scrubjay_peer_close(self._h, addr))

    def gather_bind(self, peer_labels, peer_score=None, out_base=0):
        """peer_labels / peer_score: device addresses (or tensors), one per rank, in rank order."""
        n = len(peer_labels)
        la = (ctypes.c_void_p * max(n, 1))(*[_ptr(x) for x in peer_labels])
        sa = (ctypes.c_void_p * max(n, 1))(*[_ptr(x) for x in peer_score]) if peer_score is not None else None
        self._check(self._L.scrubjay_gather_bind(self._h, n, la, sa, out_base))

    def infer_batch_gather_dev(self, d_clips, n_clips, n_samples, d_conf=None, d_feats=None):
        self._check(self._L.scrubjay_infer_batch_gather_dev(self._h, _ptr(d_clips), n_clips, n_samples, _ptr(d_conf),
                                                            _ptr(d_feats)))

    def stream_windows(self, n_samples, win_frames=16, step_frames=16):
        return self._L.scrubjay_stream_windows(n_samples, win_frames, step_frames)

    def stream_infer_dev(self, streams, n_channels, n_samples, chan_stride, labels, conf=None, score=None, feats=None,
                         win_frames=16, step_frames=16, prev_hop=None):
        """scrubjay_stream_infer_dev: sliding-window cepstrum + SVM over continuous channels; outputs are
        [channel][window] device arrays with stream_windows(n_samples, win, step) windows per channel."""
        self._check(self._L.scrubjay_stream_infer_dev(self._h, _ptr(streams), n_channels, n_samples, chan_stride,
                                                      _ptr(prev_hop), win_frames, step_frames, _ptr(labels),
                                                      _ptr(conf), _ptr(score), _ptr(feats)))

    def stream_infer_pcm16_dev(self, streams, n_channels, n_samples, chan_stride, labels, conf=None, score=None,
                               feats=None, win_frames=16, step_frames=16, prev_hop=None):
        """scrubjay_stream_infer_pcm16_dev: stream_infer_dev on int16 PCM channels (s / 32768 inside the kernel)."""
        self._check(self._L.scrubjay_stream_infer_pcm16_dev(self._h, _ptr(streams), n_channels, n_samples, chan_stride,
                                                            _ptr(prev_hop), win_frames, step_frames, _ptr(labels),
                                                            _ptr(conf), _ptr(score), _ptr(feats)))

    def stream_infer_gather_dev(self, streams, n_channels, n_samples, chan_stride, conf=None, feats=None,
                                win_frames=16, step_frames=16, prev_hop=None, pcm16=False):
        """stream mode with the fused gather: labels / scores go to every target bound with gather_bind."""
        fn = self._L.scrubjay_stream_infer_pcm16_gather_dev if pcm16 else self._L.scrubjay_stream_infer_gather_dev
        self._check(fn(self._h, _ptr(streams), n_channels, n_samples, chan_stride, _ptr(prev_hop), win_frames,
                       step_frames, _ptr(conf), _ptr(feats)))

    def debug_inject_failure(self, chunk):
        """test hook: the next host-pointer batch call fails before enqueueing chunk number `chunk`"""
        self._check(self._L.scrubjay_debug_inject_failure(self._h, chunk))

    def set_latency_mode(self, on):
        """frame-parallel path for small batches (default on); False forces the one-kernel path"""
        self._check(self._L.scrubjay_set_latency_mode(self._h, 1 if on else 0))

    def set_stream_chunk(self, windows_per_item):
        self._check(self._L.scrubjay_set_stream_chunk(self._h, windows_per_item))

    def synth_clips_dev(self, d_out, first_clip, n_clips, n_samples, seed=0x5C2B, flavour=0):
        """flavour: SYNTH_MIXED (noise + chirps) or SYNTH_CALLS ("donut"-shaped calls: both donut verdicts occur)"""
        self._check(self._L.scrubjay_synth_clips_ex_dev(self._h, _ptr(d_out), first_clip, n_clips, n_samples, seed, flavour))

    def donut_classify_batch_dev(self, d_clips, n_clips, n_samples, variant, d_labels, d_detail=None):
        self._check(self._L.scrubjay_donut_classify_batch_dev(self._h, _ptr(d_clips), n_clips, n_samples, variant,
                                                              _ptr(d_labels), _ptr(d_detail)))

    def donut_classify_batch(self, clips, variant):
        """host clips [n][n_samples] -> (labels int32 [n], detail float32 [n][4])"""
        clips = np.ascontiguousarray(clips, dtype=np.float32)
        n, ns = clips.shape
        labels = np.empty(n, np.int32)
        detail = np.empty((n, 4), np.float32)
        self._check(self._L.scrubjay_donut_classify_batch(self._h, clips.ctypes.data, n, ns, variant,
                                                          labels.ctypes.data, detail.ctypes.data))
        return labels, detail

    def donut_midpoints_batch_dev(self, d_clips, n_clips, n_samples, variant, d_labels, d_detail, d_mids):
        self._check(self._L.scrubjay_donut_midpoints_batch_dev(self._h, _ptr(d_clips), n_clips, n_samples, variant,
                                                               _ptr(d_labels), _ptr(d_detail), _ptr(d_mids)))

    def donut_classify_long_dev(self, d_signals, n_signals, n_samples, variant, d_labels, d_detail=None, d_mids=None,
                                max_mid=64):
        """staged classifier: signals of any length, and DONUT_INO_F32 (64-sample windows)"""
        self._check(self._L.scrubjay_donut_classify_long_dev(self._h, _ptr(d_signals), n_signals, n_samples, variant,
                                                             _ptr(d_labels), _ptr(d_detail), _ptr(d_mids), max_mid))

    def donut_classify_streams_dev(self, d_streams, n_channels, n_samples, chan_stride, variant, d_track, d_detail=None,
                                   clip_samples=16000):
        """per-second decision track (int8: 1 = no bird, 2 = scrub jay; sync/sync.cpp:202-203) of continuous streams"""
        self._check(self._L.scrubjay_donut_classify_streams_dev(self._h, _ptr(d_streams), n_channels, n_samples, chan_stride,
                                                                clip_samples, variant, _ptr(d_track), _ptr(d_detail)))

    # ---- donut stage functions (sync/lib/classifier.h:14-18)
    def butter_filter_batch_dev(self, d_in, d_out, n_signals, n_samples, stride, precision, b, a, d_state=None,
                                mode=IIR_EXACT):
        dt = np.float64 if precision == F64 else np.float32
        b = np.ascontiguousarray(b, dtype=dt)
        a = np.ascontiguousarray(a, dtype=dt)
        self._check(self._L.scrubjay_butter_filter_batch_dev(self._h, _ptr(d_in), _ptr(d_out), n_signals, n_samples,
                                                             stride, precision, b.ctypes.data, a.ctypes.data,
                                                             _ptr(d_state), mode))

    def butter_filter_batch(self, signals, b, a, mode=IIR_EXACT):
        """host signals [n][n_samples] float32 or float64 -> filtered array of the same shape and type"""
        signals = np.ascontiguousarray(signals)
        if signals.dtype not in (np.float32, np.float64):
            signals = signals.astype(np.float32)
        precision = F64 if signals.dtype == np.float64 else F32
        b = np.ascontiguousarray(b, dtype=signals.dtype)
        a = np.ascontiguousarray(a, dtype=signals.dtype)
        n, ns = signals.shape
        out = np.empty_like(signals)
        self._check(self._L.scrubjay_butter_filter_batch(self._h, signals.ctypes.data, out.ctypes.data, n, ns, precision,
                                                         b.ctypes.data, a.ctypes.data, mode))
        return out

    def spectrogram_batch_dev(self, d_signals, n_signals, n_samples, stride, precision, nperseg, fs, d_sxx):
        self._check(self._L.scrubjay_spectrogram_batch_dev(self._h, _ptr(d_signals), n_signals, n_samples, stride,
                                                           precision, nperseg, fs, _ptr(d_sxx)))

    def spectrogram_batch(self, signals, nperseg=256, fs=16000):
        """host signals [n][n_samples] float32/float64 -> Sxx [n][nperseg/2+1][frames] (the reference's Sxx[f][t])"""
        signals = np.ascontiguousarray(signals)
        if signals.dtype not in (np.float32, np.float64):
            signals = signals.astype(np.float32)
        precision = F64 if signals.dtype == np.float64 else F32
        n, ns = signals.shape
        frames = spectrogram_frames(ns, nperseg)
        out = np.empty((n, nperseg // 2 + 1, frames), signals.dtype)
        self._check(self._L.scrubjay_spectrogram_batch(self._h, signals.ctypes.data, n, ns, precision, nperseg, fs,
                                                       out.ctypes.data))
        return out

    # ---- host-pointer entry points (numpy in, numpy out)
    def mfcc_stats(self, samples, n_samples=None):
        """samples: numpy float32 array, or a raw host address with n_samples"""
        out = np.empty(FEAT_DIM, np.float32)
        if isinstance(samples, int):
            self._check(self._L.scrubjay_mfcc_stats(self._h, samples, n_samples, out.ctypes.data))
            return out
        x = np.ascontiguousarray(samples, dtype=np.float32)
        self._check(self._L.scrubjay_mfcc_stats(self._h, x.ctypes.data, x.size, out.ctypes.data))
        return out

    def predict(self, feat):
        f = np.ascontiguousarray(feat, dtype=np.float32)
        assert f.size == FEAT_DIM
        conf = ctypes.c_float()
        label = self._check(self._L.scrubjay_predict(self._h, f.ctypes.data, ctypes.byref(conf)))
        return label, conf.value

    def infer_batch(self, clips, want_feats=False, n_clips=None, n_samples=None):
        """clips: numpy [n_clips][n_samples] float32, or a raw host address with n_clips/n_samples."""
        if isinstance(clips, np.ndarray):
            clips = np.ascontiguousarray(clips, dtype=np.float32)
            n_clips, n_samples = clips.shape
        labels = np.empty(n_clips, np.int32)
        conf = np.empty(n_clips, np.float32)
        score = np.empty(n_clips, np.float32)
        feats = np.empty((n_clips, FEAT_DIM), np.float32) if want_feats else None
        self._check(self._L.scrubjay_infer_batch(self._h, _ptr(clips), n_clips, n_samples, labels.ctypes.data,
                                                 conf.ctypes.data, score.ctypes.data, _ptr(feats)))
        return (labels, conf, score, feats) if want_feats else (labels, conf, score)

    def infer_batch_pcm16(self, clips, want_feats=False, n_clips=None, n_samples=None):
        """clips: numpy [n_clips][n_samples] int16 PCM, or a raw host address with n_clips/n_samples."""
        if isinstance(clips, np.ndarray):
            clips = np.ascontiguousarray(clips, dtype=np.int16)
            n_clips, n_samples = clips.shape
        labels = np.empty(n_clips, np.int32)
        conf = np.empty(n_clips, np.float32)
        score = np.empty(n_clips, np.float32)
        feats = np.empty((n_clips, FEAT_DIM), np.float32) if want_feats else None
        self._check(self._L.scrubjay_infer_batch_pcm16(self._h, _ptr(clips), n_clips, n_samples, labels.ctypes.data,
                                                       conf.ctypes.data, score.ctypes.data, _ptr(feats)))
        return (labels, conf, score, feats) if want_feats else (labels, conf, score)

    def infer_batch_pcm16_dev(self, clips, n_clips, n_samples, labels, conf=None, score=None, feats=None):
        self._check(self._L.scrubjay_infer_batch_pcm16_dev(self._h, _ptr(clips), n_clips, n_samples, _ptr(labels),
                                                           _ptr(conf), _ptr(score), _ptr(feats)))

    # ---- keyword ("stop") detector, 2fa/audio/word/c
    def keyword_load(self, blob=None):
        """scrubjay_keyword_load; default = the shipped export of the reference's model (data/keyword_model.bin)"""
        if blob is None:
            blob = open(KEYWORD_MODEL, "rb").read()
        blob = bytes(blob)
        self._check(self._L.scrubjay_keyword_load(self._h, blob, len(blob)))

    def keyword_classify_batch_dev(self, signals, n_clips, n_samples, prob, mfcc=None, n_frames=None):
        self._check(self._L.scrubjay_keyword_classify_batch_dev(self._h, _ptr(signals), n_clips, n_samples, _ptr(prob),
                                                                _ptr(mfcc), _ptr(n_frames)))

    def keyword_classify_batch(self, signals):
        """signals: numpy [n_clips][n_samples] float32 (host) -> probabilities"""
        signals = np.ascontiguousarray(signals, dtype=np.float32)
        prob = np.empty(signals.shape[0], np.float32)
        self._check(self._L.scrubjay_keyword_classify_batch(self._h, signals.ctypes.data, signals.shape[0],
                                                            signals.shape[1], prob.ctypes.data))
        return prob

    def decode_adc12_dev(self, adc, n, out):
        """12-bit ADC counts (uint16, device) -> float (sync/sync.cpp:236-241)"""
        self._check(self._L.scrubjay_decode_adc12_dev(self._h, _ptr(adc), n, _ptr(out)))

    def upsample_linear_dev(self, d_in, n_signals, old_size, new_size, d_out):
        """upsampleLinear (microphone/src/microphone.cpp:55-70), batched"""
        self._check(self._L.scrubjay_upsample_linear_dev(self._h, _ptr(d_in), n_signals, old_size, new_size, _ptr(d_out)))

    def decode_pcm16_dev(self, interleaved, frames, channels, policy, out):
        """interleaved int16 frames -> mono float (policy MIX_FIRST / MIX_MEAN)"""
        self._check(self._L.scrubjay_decode_pcm16_dev(self._h, _ptr(interleaved), frames, channels, policy, _ptr(out)))


MIX_FIRST, MIX_MEAN = 0, 1
SYNTH_MIXED, SYNTH_CALLS = 0, 1


class WavInfo(ctypes.Structure):
    _fields_ = [("audio_format", ctypes.c_int32), ("channels", ctypes.c_int32), ("sample_rate", ctypes.c_int32),
                ("bits_per_sample", ctypes.c_int32), ("data_offset", ctypes.c_int64), ("data_bytes", ctypes.c_int64),
                ("frames", ctypes.c_int64)]


def wav_parse(data, strict=True):
    """scrubjay_wav_parse(_ex) over a bytes object (host only, no GPU): WavInfo, or None if the reference's
    read_wav_file would reject the file (strict) / if no fmt + data chunk pair is found (strict=False: chunks in front
    of "fmt " are skipped like libsndfile / aubio_source do)"""
    buf = bytes(data)
    info = WavInfo()
    rc = load_library().scrubjay_wav_parse_ex(buf, len(buf), ctypes.byref(info), 1 if strict else 0)
    return info if rc == 0 else None


DONUT_DESKTOP_F64 = 0  # donut-classifier/classifier.c constants, double arithmetic
DONUT_SYNC_F32 = 1     # sync/lib/classifier.cpp constants, float arithmetic
DONUT_REVERT_F64 = 2   # donut-classifier/revert.c = microphone/src/classifier.c constants, double arithmetic
DONUT_MIC_F32 = 3      # microphone/src/classifier.cpp constants, float arithmetic
DONUT_INO_F32 = 4      # donut-classifier/classifier.ino constants, float arithmetic, 64-sample windows (staged path)


def classify(data):
    """int classify(float *data, int data_size) -- sync/lib/classifier.h:19, same symbol name"""
    x = np.ascontiguousarray(data, dtype=np.float32)
    return int(load_library().classify(x.ctypes.data, x.size))


def butter_bandpass(lowcut, highcut, design=BUTTER_16K, dtype=np.float32):
    """butter_bandpass (classifier.h:14 / classifier.c:319-409): (b, a) or None for an unknown band"""
    L = load_library()
    if dtype == np.float32 and design == BUTTER_16K:
        b, a = np.zeros(9, np.float32), np.zeros(9, np.float32)
        return (b, a) if L.butter_bandpass(lowcut, highcut, b.ctypes.data, a.ctypes.data) else None
    b, a = np.zeros(9, np.float64), np.zeros(9, np.float64)
    if L.scrubjay_butter_bandpass_f64(lowcut, highcut, design, b.ctypes.data, a.ctypes.data) != 0:
        return None
    return b.astype(dtype), a.astype(dtype)


def butter_bandpass_filter(data, b, a):
    """void butter_bandpass_filter(float*, int, float*, float*, float*) -- classifier.h:15, same symbol name"""
    x = np.ascontiguousarray(data, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    a = np.ascontiguousarray(a, dtype=np.float32)
    out = np.zeros_like(x)
    load_library().butter_bandpass_filter(x.ctypes.data, x.size, b.ctypes.data, a.ctypes.data, out.ctypes.data)
    return out


def spectrogram_frames(n_samples, nperseg=256):
    return int(load_library().scrubjay_spectrogram_frames(n_samples, nperseg))


def compute_spectrogram(signal, fs=16000):
    """compute_spectrogram (classifier.h:16) through the reference's own pointer-to-pointer interface:
    returns (frequencies, times, Sxx[f][t]) and frees the malloc'd arrays like the reference's caller does."""
    L = load_library()
    x = np.ascontiguousarray(signal, dtype=np.float32)
    fp = ctypes.POINTER(ctypes.c_float)
    fr, tm, S = fp(), fp(), ctypes.POINTER(fp)()
    nb, nt = ctypes.c_int(0), ctypes.c_int(0)
    L.compute_spectrogram(x.ctypes.data, x.size, fs, ctypes.byref(fr), ctypes.byref(tm), ctypes.byref(S),
                          ctypes.byref(nb), ctypes.byref(nt))
    if nb.value == 0:
        return None
    libc = ctypes.CDLL(None)
    libc.free.argtypes = [ctypes.c_void_p]
    freqs = np.ctypeslib.as_array(fr, (nb.value,)).copy()
    times = np.ctypeslib.as_array(tm, (nt.value,)).copy()
    sxx = np.stack([np.ctypeslib.as_array(S[i], (nt.value,)).copy() for i in range(nb.value)])
    for i in range(nb.value):
        libc.free(ctypes.cast(S[i], ctypes.c_void_p))
    for p in (fr, tm, S):
        libc.free(ctypes.cast(p, ctypes.c_void_p))
    return freqs, times, sxx


def find_midpoints(data, fs=16000):
    """float *find_midpoints(float*, int, int, int*) -- classifier.h:18; returns a float32 array or None"""
    L = load_library()
    x = np.ascontiguousarray(data, dtype=np.float32)
    n = ctypes.c_int(0)
    p = L.find_midpoints(x.ctypes.data, x.size, fs, ctypes.byref(n))
    if not p:
        return None
    out = np.ctypeslib.as_array(p, (max(n.value, 1),))[:n.value].copy()
    libc = ctypes.CDLL(None)
    libc.free.argtypes = [ctypes.c_void_p]
    libc.free(ctypes.cast(p, ctypes.c_void_p))
    return out


def sum_intense(lower, upper, half_range, frequencies, times, plane, midpoint):
    """float sum_intense(...) -- classifier.h:17; plane is [freq_bins][time_bins] float32 (NaN = filtered out)"""
    L = load_library()
    fr = np.ascontiguousarray(frequencies, dtype=np.float32)
    tm = np.ascontiguousarray(times, dtype=np.float32)
    pl = np.ascontiguousarray(plane, dtype=np.float32)
    rows = (ctypes.c_void_p * pl.shape[0])(*[pl[i].ctypes.data for i in range(pl.shape[0])])
    return float(L.sum_intense(lower, upper, half_range, fr.ctypes.data, fr.size, tm.ctypes.data, tm.size,
                               ctypes.cast(rows, ctypes.c_void_p), midpoint))


def shard_range(n_items, rank, world_size):
    """Contiguous clip range [lo, hi) owned by `rank`: clips are independent,
    so multi-GPU is pure sharding with no data-path collective."""
    lo = (n_items * rank) // world_size
    hi = (n_items * (rank + 1)) // world_size
    return lo, hi


def shard_range_weighted(n_items, rank, weights):
    """Contiguous clip range [lo, hi) of `rank` when the ranks take shares in proportion to `weights` (one per rank, e.g. each
    rank's measured host-to-device rate: the PCIe links of a box are not all alike).  The ranges
    partition [0, n_items) exactly; equal weights give shard_range()."""
    w = [float(x) for x in weights]
    if not w or min(w) < 0 or sum(w) <= 0:
        raise ValueError("weights must be non-negative with a positive sum")
    tot = sum(w)
    cut = lambda k: int(round(n_items * sum(w[:k]) / tot)) if k < len(w) else n_items
    return cut(rank), cut(rank + 1)
