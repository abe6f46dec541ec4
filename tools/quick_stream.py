import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
nch, hops = 512, 4688
ns = hops * 1024
d = torch.empty((nch * ns // 16000 + 1, 16000), dtype=torch.float32, device='cuda')
sj.synth_clips_dev(d, 0, d.shape[0], 16000)
flat = d.view(-1)[:nch * ns]
nw = sj.stream_windows(ns, 16, 16)
lab = torch.empty((nch, nw), dtype=torch.int32, device='cuda')
for k in (0, 1, 4, 16, 64, 293):
    sj.set_stream_chunk(k)
    for i in range(2):
        sj.stream_infer_dev(flat, nch, ns, ns, lab, None, None, None, 16, 16); ms = sj.last_kernel_ms()
    print('chunk', k, 'ms', ms, 'Mframes/s', nch * hops / ms / 1e3, flush=True)
# the same hops as clips of 16 hops
ncl = nch * hops // 16
lab2 = torch.empty(ncl, dtype=torch.int32, device='cuda')
for i in range(2):
    sj.infer_batch_dev(flat, ncl, 16384, lab2, None, None, None); ms = sj.last_kernel_ms()
print('clip mode 16384', 'ms', ms, 'Mframes/s', ncl * 16 / ms / 1e3)
