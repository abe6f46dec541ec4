#!/usr/bin/env python3
"""Throughput of the donut path (BASELINE config 3) on one GPU: clips/s, frames/s, HBM-roofline fraction.
usage: bench_donut.py [n_clips] [variant: 1 = sync f32 (default), 0 = desktop f64] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, dsp_b200

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 1
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
d = torch.empty((n, 16000), dtype=torch.float32, device="cuda")
sj.synth_clips_dev(d, 0, n, 16000)
lab = torch.empty(n, dtype=torch.int32, device="cuda")
det = torch.empty((n, 4), device="cuda")
ms = []
for i in range(reps + 1):
    sj.donut_classify_batch_dev(d, n, 16000, variant, lab, det)
    ms.append(sj.last_kernel_ms())
ms = min(ms[1:])
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
gbs = n * 64004 / ms / 1e6
print(json.dumps({"workload": "donut classify (IIR x2 + spectrogram x2 + band sums), 1 s 16 kHz clips", "variant": variant,
                  "n_clips": n, "kernel_ms": ms, "clips_per_s": n / ms * 1e3, "frames_per_s": n / ms * 1e3 * 142,
                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                               "bytes_per_clip": 64004},
                  "positives": int(lab.sum().item()), "with_midpoints": int((det[:, 0] > 0).sum().item())}))
