#!/usr/bin/env python3
"""One worker PROCESS of the donut CPU baseline (bench.py): the UNMODIFIED reference classify()
(sync/lib/classifier.cpp built into oracle/_ref/libdonut_ref_f32.so by oracle/Makefile) keeps static buffers
(classifier.cpp:230-232), so it cannot be shared between threads -- the baseline runs one process per core, as
BASELINE.md section 4 / SURVEY 8(d) specify.  argv: clips.npy first count.  Prints {"n", "seconds", "labels"}.
No torch, no CUDA, none of the product's code."""
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, first, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    clips = np.load(path, mmap_mode="r")
    x = np.ascontiguousarray(clips[first:first + count], dtype=np.float32)
    L = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libdonut_ref_f32.so"))
    fn = getattr(L, "_Z8classifyPfi")
    fn.argtypes = [ctypes.c_void_p, ctypes.c_int]
    real = os.dup(1)
    null = os.open(os.devnull, os.O_WRONLY)  # the reference prints from inside classify() (classifier.cpp:85,104-106)
    os.dup2(null, 1)
    labels = []
    t0 = time.perf_counter()
    for i in range(x.shape[0]):
        labels.append(int(fn(x[i].ctypes.data, x.shape[1])))
    dt = time.perf_counter() - t0
    ctypes.CDLL(None).fflush(None)
    os.dup2(real, 1)
    os.write(1, (json.dumps({"first": first, "n": len(labels), "seconds": dt, "labels": labels}) + "\n").encode())


if __name__ == "__main__":
    main()
