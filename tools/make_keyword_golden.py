#!/usr/bin/env python3
"""Generate tests/golden/keyword_golden.npz with the UNMODIFIED reference keyword detector
(oracle/_ref/libkeyword_ref.so, built from /root/reference/2fa/audio/word/c by oracle/Makefile): seeded
signals of several lengths -> compute_mfcc output, frame count and classify_signal probability.  Run in the
build container (the reference tree is not on the GPU box); the fixture travels instead."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from keyword_lib import KeywordRef, test_signals  # noqa: E402

ref = KeywordRef()
out = {}
for tag, (n, ns, seed) in {"a": (6, 80000, 1), "b": (4, 16000, 2), "c": (2, 400, 3), "d": (2, 399, 4), "e": (2, 81000, 5),
                           "f": (3, 23456, 6)}.items():
    x = test_signals(n, ns, seed)
    res = [ref.classify(s) for s in x]
    out["prob_" + tag] = np.array([r[0] for r in res], np.float32)
    out["nfr_" + tag] = np.array([r[2] for r in res], np.int32)
    out["mfcc_" + tag] = np.stack([r[1] for r in res])
# the fixture holds no audio: tests regenerate the signals from (n, n_samples, seed)
meta = out
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "keyword_golden.npz"), **meta)
print({k: v.shape for k, v in meta.items()})
print({k: v for k, v in meta.items() if k.startswith("prob")})
