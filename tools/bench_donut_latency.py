import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
d = torch.empty((64, 16000), device="cuda"); sj.synth_clips_dev(d, 0, 64, 16000)
h = d.cpu().numpy()
for n in (1, 16, 64):
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    for v, nm in ((dsp_b200.DONUT_SYNC_F32, "f32"), (dsp_b200.DONUT_DESKTOP_F64, "f64")):
        ks = []
        for i in range(6):
            sj.donut_classify_batch_dev(d, n, 16000, v, lab); ks.append(sj.last_kernel_ms() * 1e3)
        print(n, nm, "kernel_us", np.median(ks[1:]))
x = np.ascontiguousarray(h[0])
ws = []
for i in range(8):
    t0 = time.perf_counter(); dsp_b200.classify(x); ws.append((time.perf_counter() - t0) * 1e6)
print("classify() host call us", np.median(ws[2:]))
for n in (1, 2, 8, 32):
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    for v, nm in ((dsp_b200.DONUT_SYNC_F32, "f32"), (dsp_b200.DONUT_DESKTOP_F64, "f64")):
        ks = []
        for i in range(6):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            sj.donut_classify_long_dev(d, n, 16000, v, lab)
            torch.cuda.synchronize(); ks.append((time.perf_counter() - t0) * 1e6)
        print(n, nm, "staged path wall_us", np.median(ks[1:]))
