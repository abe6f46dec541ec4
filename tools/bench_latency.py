#!/usr/bin/env python3
"""Small-batch latency of the cepstrum path (BASELINE config 1 is single-clip inference): wall time of the
host-pointer call scrubjay_infer_batch and kernel time of the device call for n = 1 .. 4096 clips of 1 s."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
NS = 16000
out = []
for n in (1, 2, 4, 16, 64, 256, 1024, 4096):
    d = torch.empty((n, NS), device="cuda")
    sj.synth_clips_dev(d, 0, n, NS)
    lab = torch.empty(n, dtype=torch.int32, device="cuda"); cf = torch.empty(n, device="cuda"); sc = torch.empty(n, device="cuda")
    ks = []
    for i in range(12):
        sj.infer_batch_dev(d, n, NS, lab, cf, sc, None)
        ks.append(sj.last_kernel_ms() * 1e3)
    host = torch.empty((n, NS), pin_memory=True); host.copy_(d); torch.cuda.synchronize()
    ws = []
    for i in range(12):
        t0 = time.perf_counter()
        sj.infer_batch(host.data_ptr(), n_clips=n, n_samples=NS)
        ws.append((time.perf_counter() - t0) * 1e6)
    out.append({"clips": n, "kernel_us": float(np.median(ks[2:])), "host_call_us": float(np.median(ws[2:]))})
    print(out[-1], flush=True)
print(json.dumps({"latency": out}))
