import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
n = 100000
d = torch.empty((n, 16000), dtype=torch.float32, device='cuda')
sj.synth_clips_dev(d, 0, n, 16000)
lab = torch.empty(n, dtype=torch.int32, device='cuda'); det = torch.empty((n, 4), device='cuda')
for variant in (1, 0):
    for i in range(3):
        sj.donut_classify_batch_dev(d, n, 16000, variant, lab, det); ms = sj.last_kernel_ms()
        print('variant', variant, n, 'clips', ms, 'ms', n / ms * 1e3 / 1e6, 'Mclips/s', n * 64004 / ms / 1e6, 'GB/s', flush=True)
    print('positives', lab.sum().item(), 'with midpoints', (det[:, 0] > 0).sum().item())
