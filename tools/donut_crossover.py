#!/usr/bin/env python3
"""Kernel time of the donut classifier over batch sizes, for choosing between the round-1 kernels (lower latency for small
batches) and the warp-specialised ones (throughput).  usage: [SCRUBJAY_DONUT_LEGACY=1] python tools/donut_crossover.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
N = 40000
d = torch.empty((N, 16000), device="cuda"); sj.synth_clips_dev(d, 0, N, 16000)
lab = torch.empty(N, dtype=torch.int32, device="cuda")
for n in (1, 64, 256, 592, 1184, 2368, 4736, 7104, 9472, 14208, 20000, 40000):
    row = [n]
    for v in (dsp_b200.DONUT_SYNC_F32, dsp_b200.DONUT_DESKTOP_F64):
        ks = []
        for i in range(4):
            sj.donut_classify_batch_dev(d, n, 16000, v, lab); ks.append(sj.last_kernel_ms() * 1e3)
        row.append(round(float(np.median(ks[1:])), 1))
    print(*row, flush=True)
