"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/scrubjay_cuda.h declares, and fails loudly (no CPU fallback) when there is no GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

import dsp_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "scrubjay_cuda.h")).read()
    # every SJ_API declaration: the scrubjay_* entry points and the reference's own symbol names (classify, ...)
    names = set(re.findall(r"SJ_API\s+[\w\s]*?\**\s*\b(\w+)\s*\(", hdr))
    return sorted(names - {"__attribute__"})  # the SJ_API macro's own definition


def test_library_exports_every_declared_symbol():
    L = dsp_b200.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 40
    for ref_name in ("classify", "butter_bandpass", "butter_bandpass_filter", "compute_spectrogram", "find_midpoints",
                     "sum_intense", "classify_signal", "classify_signal_binary"):
        assert ref_name in syms, ref_name
    for s in syms:
        assert hasattr(L, s), "libscrubjay_cuda.so does not export %s" % s


def test_library_does_not_depend_on_torch_or_the_oracle():
    import subprocess
    out = subprocess.check_output(["ldd", dsp_b200.LIB_PATH]).decode()
    assert "torch" not in out and "oracle" not in out and "c10" not in out


def test_product_sources_never_touch_the_oracle():
    for d, _, files in os.walk(os.path.join(ROOT, "dsp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".inc")) and "svm_default" not in f:
                src = open(os.path.join(d, f)).read()
                assert "oracle_lib" not in src and "libsj_oracle" not in src and "oracle/" not in src, f


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(dsp_b200.ScrubJayError) as e:
        dsp_b200.ScrubJay(device=0)
    assert "no CPU fallback" in str(e.value) or "not sm_100" in str(e.value)


def test_sparse_mel_tables_equal_dense_oracle_bank():
    """the kernel's chunked sparse filterbank, expanded back to dense on the host, equals the
    oracle's dense aubio-style bank bit for bit (x0.5: the kernel carries |X| scaled by 2)"""
    from oracle_lib import CepstrumOracle
    L = dsp_b200.load_library()
    L.scrubjay_debug_mel_dense.argtypes = [ctypes.c_uint, ctypes.c_void_p]
    for sr in (16000, 8000, 22050, 44100, 48000):
        dense = np.zeros((40, 1025), np.float32)
        rc = L.scrubjay_debug_mel_dense(sr, dense.ctypes.data)
        assert rc == 0, (sr, rc)
        ref = CepstrumOracle(sr).fbank()
        assert np.array_equal(dense, 0.5 * ref), sr


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 1000000, 10000001):
        for w in (1, 2, 4, 8):
            r = [dsp_b200.shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_shard_range_weighted_partitions_exactly_and_follows_the_weights():
    for n in (0, 1, 7, 65536 * 8, 10_000_019):
        for w in ([1, 1, 1, 1], [23.2] * 4 + [35.3] * 4, [0, 1, 0, 3], [5]):
            r = [dsp_b200.shard_range_weighted(n, k, w) for k in range(len(w))]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(len(w) - 1))
            for k, (lo, hi) in enumerate(r):
                assert abs((hi - lo) - n * w[k] / sum(w)) <= 1
    assert [dsp_b200.shard_range_weighted(1000, k, [1, 1, 1]) for k in range(3)] == [(0, 333), (333, 667), (667, 1000)]
    with pytest.raises(ValueError):
        dsp_b200.shard_range_weighted(10, 0, [0, 0])
