import os, sys, time
sys.path.insert(0, os.getcwd())
import torch, dsp_b200, numpy as np
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
for n in (1, 4, 16):
    d = torch.empty((n, 16000), dtype=torch.float32, device='cuda')
    sj.synth_clips_dev(d, 0, n, 16000)
    h = d.cpu().pin_memory()
    lab = torch.empty(n, dtype=torch.int32, device='cuda'); sc = torch.empty(n, device='cuda'); cf = torch.empty(n, device='cuda')
    sj.infer_batch_dev(d, n, 16000, lab, cf, sc, None); torch.cuda.synchronize()
    ref = (lab.cpu().numpy().copy(), sc.cpu().numpy().copy())
    # results to mapped pinned memory, input from pinned host memory (UVA: the same address on the device)
    hl = torch.empty(n, dtype=torch.int32).pin_memory(); hs = torch.empty(n).pin_memory(); hc = torch.empty(n).pin_memory()
    class V:  # raw pointer views
        def __init__(s, t): s.t = t
        def data_ptr(s): return s.t.data_ptr()
    try:
        for rep in range(3):
            t0 = time.perf_counter()
            for i in range(200):
                sj.infer_batch_dev(h, n, 16000, hl, hc, hs, None)
                torch.cuda.current_stream().synchronize()
            t1 = time.perf_counter()
        print(n, 'zero-copy call %.1f us' % ((t1 - t0) / 200 * 1e6), 'equal', np.array_equal(hl.numpy(), ref[0]), np.array_equal(hs.numpy(), ref[1]), 'kernel ms', sj.last_kernel_ms())
    except Exception as e:
        print('zero-copy failed', e)
    hh = h.numpy()
    for rep in range(3):
        t0 = time.perf_counter()
        for i in range(200):
            r = sj.infer_batch(hh)
        t1 = time.perf_counter()
    print(n, 'host call      %.1f us' % ((t1 - t0) / 200 * 1e6), 'equal', np.array_equal(r[0], ref[0]), np.array_equal(r[2], ref[1]))
