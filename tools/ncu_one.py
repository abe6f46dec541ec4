"""One short run of ONE kernel for ncu (and for the plain run that must precede it):
  python tools/ncu_one.py cepstrum|pcm16|donut_f32|donut_f64 [n_clips]
Two launches of the kernel on synthetic 1 s / 16 kHz clips (profile the second: ncu -s 1 -c 1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, dsp_b200
kind = sys.argv[1] if len(sys.argv) > 1 else "cepstrum"
n = int(sys.argv[2]) if len(sys.argv) > 2 else (200000 if kind in ("cepstrum", "pcm16") else 37888)
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
d = torch.empty((n, 16000), dtype=torch.float32, device="cuda")
sj.synth_clips_dev(d, 0, n, 16000)
lab = torch.empty(n, dtype=torch.int32, device="cuda")
sc, cf, det = torch.empty(n, device="cuda"), torch.empty(n, device="cuda"), torch.empty((n, 4), device="cuda")
if kind == "pcm16":
    p = torch.empty((n, 16000), dtype=torch.int16, device="cuda")
    for i in range(0, n, 50000):
        p[i:i + 50000] = torch.clamp(torch.round(d[i:i + 50000] * 32768.0), -32768, 32767).to(torch.int16)
    del d
for i in range(2):
    if kind == "cepstrum":
        sj.infer_batch_dev(d, n, 16000, lab, cf, sc, None)
    elif kind == "pcm16":
        sj.infer_batch_pcm16_dev(p, n, 16000, lab, cf, sc, None)
    else:
        sj.donut_classify_batch_dev(d, n, 16000, 1 if kind == "donut_f32" else 0, lab, det)
torch.cuda.synchronize()
ms = sj.last_kernel_ms()
print("ok", kind, n, "clips", ms, "ms", n / ms * 1e-3, "Mclips/s")
