"""Small invocation of every kernel (for compute-sanitizer runs and quick smoke checks)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, dsp_b200
sj = dsp_b200.ScrubJay(0, 16000)
sj.set_stream(torch.cuda.current_stream().cuda_stream)
n = 40
for ns in (16000, 16001, 1000, 2051):
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 0, n, ns)
    lab = torch.empty(n, dtype=torch.int32, device="cuda"); cf = torch.empty(n, device="cuda"); sc = torch.empty(n, device="cuda")
    ft = torch.empty((n, 40), device="cuda")
    sj.infer_batch_dev(d, n, ns, lab, cf, sc, ft)
    pcm = torch.clamp(torch.round(d * 32768.0), -32768, 32767).to(torch.int16).contiguous()
    sj.infer_batch_pcm16_dev(pcm, n, ns, lab, cf, sc, ft)
torch.cuda.synchronize()
print("cepstrum ok", lab.sum().item())
d = torch.empty((n, 16000), dtype=torch.float32, device="cuda"); sj.synth_clips_dev(d, 0, n, 16000)
flat = d.view(4, -1)
for W, S in ((16, 16), (16, 1), (5, 2)):
    nw = sj.stream_windows(flat.shape[1], W, S)
    lab = torch.empty((4, nw), dtype=torch.int32, device="cuda")
    sj.stream_infer_dev(flat, 4, flat.shape[1], flat.shape[1], lab, None, None, None, W, S)
torch.cuda.synchronize(); print("stream ok")
det = torch.empty((n, 4), device="cuda"); lab = torch.empty(n, dtype=torch.int32, device="cuda")
for v in range(4):
    sj.donut_classify_batch_dev(d, n, 16000, v, lab, det)
torch.cuda.synchronize(); print("donut ok", lab.sum().item())
sj.keyword_load()
sig = d.view(8, -1)
prob = torch.empty(8, device="cuda"); mf = torch.empty((8, 500, 13), device="cuda"); nf = torch.empty(8, dtype=torch.int32, device="cuda")
sj.keyword_classify_batch_dev(sig, 8, sig.shape[1], prob, mf, nf)
sj.keyword_classify_batch_dev(sig, 8, 399, prob, mf, nf)
torch.cuda.synchronize(); print("keyword ok", prob.tolist()[:3])
x = torch.randint(-30000, 30000, (3000, 2), dtype=torch.int16, device="cuda"); o = torch.empty(3000, device="cuda")
sj.decode_pcm16_dev(x, 3000, 2, 1, o); torch.cuda.synchronize(); print("decode ok")
# other sample rates: generic instantiation
for sr in (8000, 22050, 44100):
    s2 = dsp_b200.ScrubJay(0, sr); s2.set_stream(torch.cuda.current_stream().cuda_stream)
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    s2.infer_batch_dev(d, n, 16000, lab, None, None, None); torch.cuda.synchronize(); s2.close()
print("generic ok")
# round 2 paths: allocation-free entry point, non-default cepstrum parameters, donut groups of several clips in the
# warp-specialised float kernel (and its unaligned form), the call-flavoured generator, the staged long-signal classifier
lab = torch.empty(n, dtype=torch.int32, device="cuda"); ft = torch.empty((n, 40), device="cuda")
sj.infer_batch_noalloc_dev(d, n, 16000, lab, None, None, ft)
sj.infer_batch_noalloc_dev(d, 7, 5000, lab, None, None, ft)
torch.cuda.synchronize(); print("noalloc ok")
prm = dsp_b200.CepstrumParams(window=dsp_b200.WINDOW_HANN_SYMMETRIC, spectrum=dsp_b200.SPECTRUM_POWER, log_floor=1e-20)
s3 = dsp_b200.ScrubJay(0, 16000, params=prm); s3.set_stream(torch.cuda.current_stream().cuda_stream)
s3.infer_batch_dev(d, n, 16000, lab, None, None, ft); torch.cuda.synchronize(); s3.close(); print("params ok")
m = 2500
big = torch.empty((m, 16000), dtype=torch.float32, device="cuda")
sj.synth_clips_dev(big, 0, m, 16000, flavour=dsp_b200.SYNTH_CALLS)
lab = torch.empty(m, dtype=torch.int32, device="cuda"); det = torch.empty((m, 4), device="cuda")
sj.donut_classify_batch_dev(big, m, 16000, 1, lab, det)
odd = big.view(-1)[1:1 + 300 * 16001].view(300, 16001)  # unaligned rows
sj.donut_classify_batch_dev(odd, 300, 16001, 1, lab, det)
torch.cuda.synchronize(); print("donut ws ok", lab[:300].sum().item())
long_sig = big.view(-1)[:2 * 200000].view(2, 200000).contiguous()
sj.donut_classify_batch_dev(long_sig, 2, 200000, 1, lab, det); torch.cuda.synchronize(); print("donut long ok", det[:2, 0].tolist())
