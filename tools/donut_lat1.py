import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np, torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
d = torch.empty((64, 16000), device="cuda"); sj.synth_clips_dev(d, 0, 64, 16000)
lab = torch.empty(64, dtype=torch.int32, device="cuda")
for n in (1, 64):
    row=[n]
    for v in (1, 0):
        ks=[]
        for i in range(5):
            sj.donut_classify_batch_dev(d, n, 16000, v, lab); ks.append(sj.last_kernel_ms()*1e3)
        row.append(round(float(np.median(ks[1:])),1))
    print(*row)
