#!/usr/bin/env python3
"""Throughput of the stand-alone donut stage kernels on one GPU (samples/s, frames/s).
usage: bench_stages.py"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch, dsp_b200

sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
out = {}


def timed(fn, reps=3):
    ms = []
    for _ in range(reps + 1):
        fn()
        ms.append(sj.last_kernel_ms())
    return min(ms[1:])


# many streams, exact recurrence per lane (config 3: 16384 streams, here 10 s of each)
for name, tdt, prec, npdt in (("f32", torch.float32, dsp_b200.F32, np.float32), ("f64", torch.float64, dsp_b200.F64, np.float64)):
    n, ns = 16384, 160000
    x = (torch.rand((n, ns), dtype=torch.float32, device="cuda") - 0.5).to(tdt)
    y = torch.empty_like(x)
    b, a = dsp_b200.butter_bandpass(3000, 7500, dtype=npdt)
    ms = timed(lambda: sj.butter_filter_batch_dev(x, y, n, ns, ns, prec, b, a))
    es = 4 if prec == dsp_b200.F32 else 8
    out["iir_exact_" + name] = {"signals": n, "samples": ns, "ms": ms, "samples_per_s": n * ns / ms * 1e3,
                                "GB_per_s": 2 * n * ns * es / ms / 1e6}
    del x, y
# few long streams: sequential lane vs block-wise state-space scan
n, ns = 4, 16000 * 600
x = (torch.rand((n, ns), dtype=torch.float64, device="cuda") - 0.5)
y = torch.empty_like(x)
b, a = dsp_b200.butter_bandpass(3000, 7500, dtype=np.float64)
ms_e = timed(lambda: sj.butter_filter_batch_dev(x, y, n, ns, ns, dsp_b200.F64, b, a), 1)
ms_s = timed(lambda: sj.butter_filter_batch_dev(x, y, n, ns, ns, dsp_b200.F64, b, a, None, dsp_b200.IIR_SCAN))
out["iir_long_f64"] = {"signals": n, "samples": ns, "exact_ms": ms_e, "scan_ms": ms_s, "scan_samples_per_s": n * ns / ms_s * 1e3,
                       "speedup": ms_e / ms_s}
del x, y
# spectrogram of 1 s clips
for name, tdt, prec in (("f32", torch.float32, dsp_b200.F32), ("f64", torch.float64, dsp_b200.F64)):
    for nperseg in (256, 64):
        n, ns = 50000, 16000
        x = (torch.rand((n, ns), dtype=torch.float32, device="cuda") - 0.5).to(tdt)
        frames = dsp_b200.spectrogram_frames(ns, nperseg)
        S = torch.empty((n, nperseg // 2 + 1, frames), dtype=tdt, device="cuda")
        ms = timed(lambda: sj.spectrogram_batch_dev(x, n, ns, ns, prec, nperseg, 16000, S))
        out["spectrogram_%s_%d" % (name, nperseg)] = {"signals": n, "frames": frames, "ms": ms, "frames_per_s": n * frames / ms * 1e3}
        del x, S
# staged classifier: the .ino variant on 1 s clips, and whole recordings (10 min each)
for name, variant, n, ns in (("ino_f32_1s", dsp_b200.DONUT_INO_F32, 50000, 16000), ("sync_f32_1s_staged", dsp_b200.DONUT_SYNC_F32, 50000, 16000),
                             ("sync_f32_10min", dsp_b200.DONUT_SYNC_F32, 64, 9600000), ("desktop_f64_10min", dsp_b200.DONUT_DESKTOP_F64, 64, 9600000)):
    d = torch.empty((n * ns // 16000, 16000), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 0, n * ns // 16000, 16000)
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    det = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    ms = timed(lambda: sj.donut_classify_long_dev(d, n, ns, variant, lab, det, None, 64), 2)
    out["classify_" + name] = {"signals": n, "samples": ns, "ms": ms, "audio_seconds_per_s": n * (ns / 16000) / ms * 1e3,
                               "positives": int(lab.sum().item())}
    del d
print(json.dumps(out))
