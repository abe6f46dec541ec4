"""GPU test of scrubjay_infer_batch_noalloc_dev (include/scrubjay_cuda.h): the allocation-free entry point added
(no counterpart in the reference).  Its frame scratch is a static device array inside
the library; the outputs must equal those of the existing scrubjay_infer_batch_dev -- the same arithmetic in the same
order, so the bar is bit-identity (tighter than the 1e-4 relative tolerance of tests/test_cepstrum_gpu.py, which is
asserted as well) -- on two calls in a row (the second call finds the first call's contents in the shared array), and
sizes beyond the stated maximum must be refused with a clear error.  Named test_zz_* so that pytest collects it
last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FEAT_RTOL, FEAT_ATOL = 1e-4, 1e-5


def _ctx():
    import torch
    import dsp_b200
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    return sj


def _outputs(n):
    import torch
    return (torch.full((n,), -9, dtype=torch.int32, device="cuda"), torch.full((n,), -9.0, dtype=torch.float32, device="cuda"),
            torch.full((n,), -9.0, dtype=torch.float32, device="cuda"), torch.full((n, 40), -9.0, dtype=torch.float32, device="cuda"))


def _same(a, b, what):
    for x, y, name in zip(a, b, ("labels", "conf", "score", "feats")):
        x, y = x.cpu().numpy(), y.cpu().numpy()
        if name == "feats":
            assert (np.abs(x - y) <= FEAT_RTOL * np.abs(y) + FEAT_ATOL).all(), "%s: %s outside the tolerance" % (what, name)
        assert np.array_equal(x, y), "%s: %s differ from scrubjay_infer_batch_dev" % (what, name)


def test_noalloc_entry_point_equals_existing_entry_point_on_two_calls_in_a_row():
    import torch
    sj = _ctx()
    # two different batches (different clip counts, lengths and contents): the second call runs over what the first left
    # in the shared static array
    shapes = [(300, 16000), (37, 20481)]
    clips = []
    for i, (n, ns) in enumerate(shapes):
        k = (ns + 15999) // 16000
        d = torch.empty((n * k, 16000), dtype=torch.float32, device="cuda")
        sj.synth_clips_dev(d, 1000 * (i + 1), n * k, 16000)
        clips.append(d.view(n, k * 16000)[:, :ns].contiguous())
    want = []
    for (n, ns), d in zip(shapes, clips):
        o = _outputs(n)
        sj.set_latency_mode(False)  # the existing entry point, one-kernel path
        sj.infer_batch_dev(d, n, ns, *o)
        torch.cuda.synchronize()
        want.append(o)
    sj.set_latency_mode(True)
    free = []
    for rep in range(2):  # the two calls in a row, then the same pair again
        for (n, ns), d, w in zip(shapes, clips, want):
            o = _outputs(n)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            sj.infer_batch_noalloc_dev(d, n, ns, *o)
            torch.cuda.synchronize()
            free.append(torch.cuda.mem_get_info()[0] - free0)
            _same(o, w, "call %d, %d clips x %d samples" % (rep, n, ns))
    # no device memory is taken by a call once the kernels have been loaded (the first pair may trigger lazy module loading)
    assert all(f >= 0 for f in free[2:]), "free device memory fell across a no-alloc call: %s" % free
    # optional outputs may be NULL, like scrubjay_infer_batch_dev
    n, ns = shapes[0]
    lab = _outputs(n)[0]
    sj.infer_batch_noalloc_dev(clips[0], n, ns, lab)
    torch.cuda.synchronize()
    assert torch.equal(lab, want[0][0])
    sj.close()


def test_noalloc_limit_is_enforced_with_a_clear_error():
    import torch
    import dsp_b200
    sj = _ctx()
    lim = dsp_b200.NOALLOC_MAX_TOTAL_FRAMES
    assert lim == 16384
    # exactly at the limit: 1024 one-second clips (16 frames each)
    n, ns = lim // 16, 16000
    d = torch.empty((n, ns), dtype=torch.float32, device="cuda")
    sj.synth_clips_dev(d, 77, n, ns)
    a, b = _outputs(n), _outputs(n)
    sj.infer_batch_noalloc_dev(d, n, ns, *a)
    sj.set_latency_mode(False)
    sj.infer_batch_dev(d, n, ns, *b)
    torch.cuda.synchronize()
    _same(a, b, "at the limit")
    # one clip more: refused, nothing written
    d2 = torch.empty((n + 1, ns), dtype=torch.float32, device="cuda")
    o = _outputs(n + 1)
    with pytest.raises(dsp_b200.ScrubJayError, match="exceed the static scratch of 16384 frames"):
        sj.infer_batch_noalloc_dev(d2, n + 1, ns, *o)
    torch.cuda.synchronize()
    assert (o[0] == -9).all() and (o[3] == -9.0).all()
    # one long clip beyond the limit, and the degenerate arguments of the existing entry point
    with pytest.raises(dsp_b200.ScrubJayError, match="exceed the static scratch"):
        sj.infer_batch_noalloc_dev(d2, 1, (lim + 1) * 1024, *o)
    with pytest.raises(dsp_b200.ScrubJayError):
        sj.infer_batch_noalloc_dev(d2, 4, 0, *o)
    sj.infer_batch_noalloc_dev(d2, 0, ns, *o)  # empty batch: a no-op
    sj.close()
