"""one launch set of the fused donut kernel for ncu: argv[1] = variant (1 float / 0 double), argv[2] = clips"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, dsp_b200
variant = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n = int(sys.argv[2]) if len(sys.argv) > 2 else 148 * 64 * 2
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
d = torch.empty((n, 16000), dtype=torch.float32, device='cuda')
sj.synth_clips_dev(d, 0, n, 16000)
lab = torch.empty(n, dtype=torch.int32, device='cuda'); det = torch.empty((n, 4), device='cuda')
for i in range(2):
    sj.donut_classify_batch_dev(d, n, 16000, variant, lab, det)
torch.cuda.synchronize()
print('ok', n, sj.last_kernel_ms(), 'ms', n / sj.last_kernel_ms() * 1e-3, 'Mclips/s')
