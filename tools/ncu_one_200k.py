import os, sys
sys.path.insert(0, os.getcwd())
import torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
n = 200000
d = torch.empty((n,16000), dtype=torch.float32, device='cuda')
sj.synth_clips_dev(d, 0, n, 16000)
lab = torch.empty(n, dtype=torch.int32, device='cuda'); sc = torch.empty(n, device='cuda'); cf = torch.empty(n, device='cuda')
for i in range(2):
    sj.infer_batch_dev(d, n, 16000, lab, cf, sc, None)
torch.cuda.synchronize()
print('ok', sj.last_kernel_ms())
