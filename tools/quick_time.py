import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, dsp_b200, time
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
for n in (1776*8, 200000):
    d = torch.empty((n,16000), dtype=torch.float32, device='cuda')
    sj.synth_clips_dev(d, 0, n, 16000)
    lab = torch.empty(n, dtype=torch.int32, device='cuda'); sc = torch.empty(n, device='cuda'); cf = torch.empty(n, device='cuda')
    for i in range(3):
        sj.infer_batch_dev(d, n, 16000, lab, cf, sc, None); ms = sj.last_kernel_ms()
        print(n, 'clips', ms, 'ms', n/ms*1e3/1e6, 'Mclips/s', n*64008/ms/1e6, 'GB/s', flush=True)
    print('labels mean', lab.float().mean().item(), 'score range', sc.min().item(), sc.max().item())
