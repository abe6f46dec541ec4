#!/usr/bin/env python3
"""Throughput of the stream mode (BASELINE config 5 shape, one GPU's share scaled down in time):
usage: bench_stream.py [n_channels=512] [seconds=600] [win=16] [step=16] [pcm16=0]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, dsp_b200

nch = int(sys.argv[1]) if len(sys.argv) > 1 else 512
secs = int(sys.argv[2]) if len(sys.argv) > 2 else 600
W = int(sys.argv[3]) if len(sys.argv) > 3 else 16
S = int(sys.argv[4]) if len(sys.argv) > 4 else 16
PCM = int(sys.argv[5]) if len(sys.argv) > 5 else 0
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
ns = secs * 16000
d = torch.empty((nch * secs, 16000), dtype=torch.float32, device="cuda")
sj.synth_clips_dev(d, 0, nch * secs, 16000)
if PCM:  # the same audio as int16 PCM: half the resident bytes per channel
    d16 = torch.empty((nch * secs, 16000), dtype=torch.int16, device="cuda")
    for r in range(0, nch * secs, 65536):
        d16[r:r + 65536].copy_((d[r:r + 65536] * 32768.0).clamp_(-32768, 32767))
    del d
    d = d16
nw = sj.stream_windows(ns, W, S)
lab = torch.empty((nch, nw), dtype=torch.int32, device="cuda")
sc = torch.empty((nch, nw), dtype=torch.float32, device="cuda")
ms = []
for i in range(4):
    (sj.stream_infer_pcm16_dev if PCM else sj.stream_infer_dev)(d, nch, ns, ns, lab, None, sc, None, W, S)
    ms.append(sj.last_kernel_ms())
ms = min(ms[1:])
hops = (ns + 1023) // 1024
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
gbs = (nch * ns * (2 if PCM else 4) + nch * nw * 8) / ms / 1e6
print(json.dumps({"workload": "stream mode%s: %d channels x %d s @16 kHz, window %d frames step %d" % (" (int16 PCM)" if PCM else "", nch, secs, W, S),
                  "kernel_ms": ms, "windows_per_s": nch * nw / ms * 1e3, "frames_per_s": nch * hops / ms * 1e3,
                  "audio_seconds_per_s": nch * secs / ms * 1e3,
                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak}}))
