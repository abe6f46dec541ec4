#!/usr/bin/env python3
"""bench.py -- clips/s of the batched cepstrum (MFCC mean/std) + RBF-SVM hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--clips C] [--impl ours|reference]
  (N > 1: launched by torchrun, one rank per GPU; RANK/LOCAL_RANK/WORLD_SIZE from the env)

Workload (BASELINE.json configs[1]): C = 1,000,000 synthetic 1 s / 16 kHz mono float32 clips per GPU
(64 GB, resident in HBM; inputs >> 126 MB L2, so no L2 flush is needed between steps).  One step =
one pass of libscrubjay_cuda's fused kernel over the whole resident batch -> label, confidence and
decision value per clip.  Clips are independent units: N GPUs = N shards, no data-path collective;
the only exchange is the final label/score all-gather (8 B/clip), done inside every timed step.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      HBM roofline of the fused kernel: algorithmic bytes (64,008 B/clip: 16000 fp32 in,
                int32 label + fp32 score out) / CUDA-event kernel time, vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the CPU restatement of the reference path (oracle/, float build; aubio/onnxruntime
                are not available, so kind = "port") timed on this box's host cores on a bounded sample
  e2e           same metric through the host-pointer C-ABI call scrubjay_infer_batch: pinned host
                clips -> H2D -> kernel -> D2H labels/conf/score, all inside the timed region
--impl reference times the CPU path alone on all host threads (rank 0 only).
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

N_SAMPLES = 16000
BYTES_PER_CLIP = N_SAMPLES * 4 + 4 + 4  # fp32 clip in, int32 label + fp32 score out
FP32_PIPE_CYCLES_PER_CLIP = 16 * (643 * 2 + 142)  # per scheduler: packed f32x2 instructions hold the pipe 2 cycles, scalar ones 1
FRAMES_PER_CLIP = 16
SEED = 0x5C2B
METRIC = "clips/s (1 s @16 kHz) cepstrum+SVM"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic_bytes_per_clip():
    """dram read+write bytes per clip of the fused kernel from the committed ncu --set full capture"""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return None


def bind_to_gpu_numa_node(index):
    """Bind this rank to the CPUs of its GPU's NUMA node BEFORE any pinned host allocation, so that the staging buffers
    of the host-pointer path are first-touched next to the GPU's PCIe root (all eight ranks sat on
    node 0).  Returns the node, or None where sysfs names none (single-socket boxes, VMs: nothing to bind)."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(index).pci_bus_id
        dom = torch.cuda.get_device_properties(index).pci_domain_id
        dev = torch.cuda.get_device_properties(index).pci_device_id
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


# ---------------------------------------------------------------- CPU path (oracle) timing
def cpu_run(oracle, clips, threads):
    """oracle.infer_batch over `clips`, split across `threads` OS threads (ctypes releases the GIL;
    the oracle handle is read-only).  Returns seconds."""
    import numpy as np
    n = clips.shape[0]
    bounds = [(n * i) // threads for i in range(threads + 1)]
    out = [None] * threads

    def work(i):
        if bounds[i + 1] > bounds[i]:
            out[i] = oracle.infer_batch(clips[bounds[i]:bounds[i + 1]])

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    dt = time.perf_counter() - t0
    labels = np.concatenate([o[0] for o in out if o is not None])
    return dt, labels


def cpu_baseline(clips_np, target_seconds=12.0):
    """time the float 'port' build of the oracle on a bounded sample sized for ~target_seconds"""
    import numpy as np
    from oracle_lib import CepstrumOracle
    oracle = CepstrumOracle(16000, "f32")
    threads = host_threads()
    dt1, _ = cpu_run(oracle, clips_np[:32], 1)  # single-thread rate, for the per-core figure
    probe = clips_np[:max(threads * 4, 64)]
    dt, _ = cpu_run(oracle, probe, threads)
    rate = probe.shape[0] / dt
    n = int(min(clips_np.shape[0], max(probe.shape[0], rate * target_seconds)))
    dt, _ = cpu_run(oracle, clips_np[:n], threads)
    return {"value": n / dt, "unit": "clips/s", "cores": threads, "kind": "port", "one_thread_value": 32 / dt1,
            "sample": "%d of the step's clips (same bytes the GPU consumed), %.1f s wall, oracle/cepstrum_oracle.c "
                      "float build -O3 (restated aubio+onnxruntime: the reference's own libraries are unavailable)"
                      % (n, dt)}


def donut_cpu_reference(clips_np, per_proc=384):
    """the UNMODIFIED reference classify() (sync/lib/classifier.cpp built into oracle/_ref by oracle/Makefile) on ALL host
    cores: it keeps static buffers (classifier.cpp:230-232), so it is not thread-safe -- one worker PROCESS per core
    (tools/donut_cpu_worker.py: no torch, no CUDA), each over its own slice of a bounded sample of the step's clips
    (BASELINE.md section 4).  Throughput = clips / the slowest worker's classify() loop (start-up excluded)."""
    import subprocess
    import tempfile
    import numpy as np
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdonut_ref_f32.so")):
        return None
    procs = host_threads()
    n = min(clips_np.shape[0], procs * per_proc)
    per = n // procs
    if per < 1:
        return None
    n = per * procs
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "clips.npy")
        np.save(path, clips_np[:n])
        worker = os.path.join(ROOT, "tools", "donut_cpu_worker.py")
        t0 = time.perf_counter()
        ps = [subprocess.Popen([sys.executable, worker, path, str(i * per), str(per)], stdout=subprocess.PIPE)
              for i in range(procs)]
        outs = [json.loads(p.communicate()[0].decode().strip().splitlines()[-1]) for p in ps]
        wall = time.perf_counter() - t0
    outs.sort(key=lambda o: o["first"])
    labels = [l for o in outs for l in o["labels"]]
    slow = max(o["seconds"] for o in outs)
    return {"value": n / slow, "unit": "clips/s", "cores": procs, "kind": "reference",
            "one_process_value": per / statistics.median(o["seconds"] for o in outs),
            "sample": "%d of the step's clips, %d worker processes x %d clips, slowest loop %.1f s (%.1f s wall with start-up), "
                      "unmodified sync/lib/classifier.cpp + PlainFFT.cpp -O3 (static buffers: one process per core)"
                      % (n, procs, per, slow, wall)}, labels


# ---------------------------------------------------------------- clocks sampler
class ClockSampler:
    """samples SM clock and throttle reasons (NVML) while the timed region runs"""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unsampled"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


# ---------------------------------------------------------------- reference arm (CPU only)
def run_reference(args, rank):
    if rank != 0:
        return 0
    from oracle_lib import CepstrumOracle
    from synth_np import synth_clips
    oracle = CepstrumOracle(16000, "f32")
    threads = host_threads()
    per_step = threads * args.ref_clips_per_thread
    clips = synth_clips(0, per_step, N_SAMPLES, SEED)
    for _ in range(args.warmup):
        cpu_run(oracle, clips, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_run(oracle, clips, threads)
    dt = time.perf_counter() - t0
    value = per_step * args.steps / dt
    sample = ("each step = %d synthetic clips (%d per thread) of the 1M-clip workload, restated reference path "
              "(oracle/cepstrum_oracle.c float build, -O3); aubio/onnxruntime not available so the reference "
              "binary itself cannot be built" % (per_step, args.ref_clips_per_thread))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "batched cepstrum features + SVM on 1M synthetic 1 s 16 kHz mono clips (bounded sample)",
                       "clips_per_step": per_step, "n_samples": N_SAMPLES},
            "frames_per_s": value * FRAMES_PER_CLIP,
            "cpu_baseline": {"value": value, "unit": "clips/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------- BASELINE configs 3, 4, 5 at their named sizes
def run_config3(sj, clips, labels, C, dev):
    """config 3: donut Butterworth IIR + FFT band-energy path on 16 384 synthetic streams x 60 s, one B200: every full
    second of every stream classified from zero state, like the firmware (sync/sync.cpp:190-212) -> decision tracks"""
    import torch
    import dsp_b200
    streams, secs = 16384, 60
    if streams * secs > C:
        secs = C // streams
    if secs < 1:
        return None
    n = streams * secs
    track = torch.empty((streams, secs), dtype=torch.int8, device=dev)
    out = {"streams": streams, "seconds_per_stream": secs, "clips": n, "bytes": n * N_SAMPLES * 4,
           "api": "scrubjay_donut_classify_streams_dev (per-second decision tracks)"}
    for name, variant in (("sync_f32", dsp_b200.DONUT_SYNC_F32), ("desktop_f64", dsp_b200.DONUT_DESKTOP_F64)):
        ms = []
        for _ in range(3):
            sj.donut_classify_streams_dev(clips, streams, secs * N_SAMPLES, secs * N_SAMPLES, variant, track)
            ms.append(sj.last_kernel_ms())
        ms = min(ms[1:])
        out[name] = {"ms": ms, "clips_per_s": n / ms * 1e3, "frames_per_s": 142 * n / ms * 1e3,
                     "audio_seconds_per_s": n / ms * 1e3, "scrub_jay_seconds": int((track == 2).sum().item())}
    return out


def run_config4(sj, clips, C, conf, dist, dev, rank, world, barrier, total=10_000_000):
    """config 4: clip-sharded cepstrum + SVM on 10 M clips over N GPUs with the label / score gather: every rank walks its
    10 M / N clips in resident passes of <= C clips (the resident batch is reused for every pass: inputs >> L2, so reuse
    buys nothing), results stored straight into every rank's complete 10 M-entry arrays by the kernel (peer stores)."""
    import torch
    from sharded_torch import PeerGather
    per = total // world
    g = PeerGather(sj, per, dist)
    passes = [(o, min(C, per - o)) for o in range(0, per, C)]

    def step():
        for o, n in passes:
            sj.gather_bind(g.peer_labels, g.peer_score, g.base + o)
            sj.infer_batch_gather_dev(clips, n, N_SAMPLES, conf)

    step()
    barrier()
    best = None
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = float(t.item()) if best is None else min(best, float(t.item()))
    filled = int((g.labels >= 0).sum().item())
    out = {"total_clips": per * world, "clips_per_gpu": per, "passes_per_gpu": len(passes), "ms": best,
           "clips_per_s": per * world / best * 1e3, "frames_per_s": per * world * FRAMES_PER_CLIP / best * 1e3,
           "gathered_entries_on_rank0": filled, "gather": "fused peer stores into every rank's 10 M-entry arrays",
           "timing": "CUDA events around all passes of one 10 M-clip step, max over ranks"}
    torch.cuda.synchronize()
    g.close()
    sj.gather_bind([], None, 0)
    return out


def run_config5(sj, dist, dev, rank, world, barrier, channels=4096, secs=3600, W=16, S=16):
    """config 5: 4096 channels x 1 h of continuous 16 kHz audio, sliding-frame cepstrum + SVM (tumbling 1 s windows),
    channels sharded over the GPUs (512 channels = 118 GB per GPU at N = 8), per-window labels / scores gathered by the
    kernel itself"""
    import torch
    from sharded_torch import PeerGather
    nch = channels // world
    ns = secs * N_SAMPLES
    d = torch.empty((nch * secs, N_SAMPLES), dtype=torch.float32, device=dev)
    sj.synth_clips_dev(d, rank * nch * secs, nch * secs, N_SAMPLES, SEED)
    nw = sj.stream_windows(ns, W, S)
    g = PeerGather(sj, nch * nw, dist)
    best = None
    for i in range(3):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sj.stream_infer_gather_dev(d, nch, ns, ns, win_frames=W, step_frames=S)
        e1.record()
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if i > 0:
            best = float(t.item()) if best is None else min(best, float(t.item()))
    hops = (ns + 1023) // 1024
    out = {"channels": nch * world, "seconds_per_channel": secs, "channels_per_gpu": nch, "bytes_per_gpu": nch * ns * 4, "ms": best,
           "windows_per_s": nch * world * nw / best * 1e3, "frames_per_s": nch * world * hops / best * 1e3,
           "audio_hours_per_s": nch * world * secs / 3600.0 / best * 1e3, "positive_windows_rank0_view": int(g.labels.sum().item()),
           "gather": "fused peer stores", "timing": "CUDA events incl. the gather, max over ranks"}
    torch.cuda.synchronize()
    g.close()
    del d
    return out


# ---------------------------------------------------------------- our arm
def run_ours(args, rank, local_rank, world):
    import numpy as np
    import torch
    import dsp_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no GPU (libscrubjay_cuda has no CPU fallback; use --impl reference for the CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_node = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    dist = None
    # libraries (NCCL's version banner) write to stdout; the contract is ONE JSON line there, so fd 1 points at
    # stderr until the line is printed
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    C = args.clips
    sj = dsp_b200.ScrubJay(device=local_rank, samplerate=16000)
    stream = torch.cuda.current_stream()
    sj.set_stream(stream.cuda_stream)
    clips = torch.empty((C, N_SAMPLES), dtype=torch.float32, device=dev)
    labels = torch.empty(C, dtype=torch.int32, device=dev)
    conf = torch.empty(C, dtype=torch.float32, device=dev)
    score = torch.empty(C, dtype=torch.float32, device=dev)
    # weak scaling: rank r owns global clips [r*C, (r+1)*C) of the counter-based synthetic set
    sj.synth_clips_dev(clips, rank * C, C, N_SAMPLES, SEED)
    # the path's only exchange: the final label/score gather, 8 B per clip.  Default: fused into the kernel (every
    # result is stored into every rank's arrays through NVLink peer mappings, tools/sharded_torch.py: PeerGather);
    # --gather nccl (or a box without CUDA IPC / peer access) runs two NCCL all-gathers after the kernel instead.
    peer, gather_mode = None, "none (1 GPU)"
    if dist is not None:
        gather_mode = "nccl all-gather after the kernel"
        if args.gather == "p2p":
            from sharded_torch import PeerGather
            ok = torch.ones(1, dtype=torch.int32, device=dev)
            try:
                peer = PeerGather(sj, C, dist)
            except Exception as e:  # every rank must take the same path
                sys.stderr.write("bench.py: rank %d: peer gather unavailable (%s)\n" % (rank, e))
                ok.zero_()
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 1:
                gather_mode = "fused into the kernel: peer stores over NVLink (CUDA IPC mappings)"
            else:
                if peer is not None:
                    peer.close()
                peer = None
                gather_mode += " (peer mapping failed on some rank)"
        if peer is None:
            all_labels = torch.empty(world * C, dtype=torch.int32, device=dev)
            all_score = torch.empty(world * C, dtype=torch.float32, device=dev)

    def step():
        if peer is not None:
            sj.infer_batch_gather_dev(clips, C, N_SAMPLES, conf)
        else:
            sj.infer_batch_dev(clips, C, N_SAMPLES, labels, conf, score, None)
            if dist is not None:
                dist.all_gather_into_tensor(all_labels, labels)
                dist.all_gather_into_tensor(all_score, score)

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = sj.launch_count
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    kms = []
    with ClockSampler(local_rank) as clk:
        ev[0].record(stream)
        for i in range(args.steps):
            step()
            ev[i + 1].record(stream)
        barrier()
    total_ms = ev[0].elapsed_time(ev[args.steps])
    launches = sj.launch_count - launches0
    gather_check = None
    if peer is not None:
        # The fused gather has no completion signal of its own: a consumer needs the cross-rank barrier that follows the
        # step.  Check it the hard way: poison every rank's arrays, barrier, ONE gather step, barrier, then
        # compare the complete arrays with an NCCL all-gather of the plain call's outputs.
        peer.labels.fill_(-7)
        peer.score.fill_(float("nan"))
        barrier()
        sj.infer_batch_gather_dev(clips, C, N_SAMPLES, conf)
        barrier()
        sj.infer_batch_dev(clips, C, N_SAMPLES, labels, conf, score, None)
        ref_l = torch.empty(world * C, dtype=torch.int32, device=dev)
        ref_s = torch.empty(world * C, dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(ref_l, labels)
        dist.all_gather_into_tensor(ref_s, score)
        torch.cuda.synchronize()
        assert torch.equal(peer.labels, ref_l), "fused gather: labels differ from the NCCL all-gather of the plain outputs"
        assert torch.equal(peer.score, ref_s), "fused gather: scores differ from the NCCL all-gather of the plain outputs"
        # per-step time WITH the barrier a consumer needs (the headline loop has one barrier after K steps)
        tb0 = time.perf_counter()
        for _ in range(5):
            sj.infer_batch_gather_dev(clips, C, N_SAMPLES, conf)
            barrier()
        tb = torch.tensor([(time.perf_counter() - tb0) / 5 * 1e3], dtype=torch.float64, device=dev)
        dist.all_reduce(tb, op=dist.ReduceOp.MAX)
        gather_check = {"verified": "poisoned arrays, one gather step, full labels + score arrays equal to an NCCL all-gather of the plain call on every rank",
                        "ms_per_step_with_barrier_each_step": float(tb.item()),
                        "headline_timing": "K steps back to back, one barrier + synchronize after the last (the per-step barrier is excluded)"}
    # kernel-only time, measured live: events bracketing the kernel alone (no gather), same stream
    for i in range(min(args.steps, 5)):
        sj.infer_batch_dev(clips, C, N_SAMPLES, labels, conf, score, None)
        kms.append(sj.last_kernel_ms())
    kernel_ms = statistics.mean(kms)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = world * C * args.steps / (total_ms * 1e-3)

    # ---- e2e: host-pointer C-ABI call, pinned host clips, H2D + kernel + D2H inside the timed region
    E = min(args.e2e_clips, C)

    def e2e_run(call, host_t, n, steps):
        """`steps` calls of `call` on the first n clips of host_t between a barrier and a synchronize; seconds, max over ranks"""
        barrier()
        t0 = time.perf_counter()
        out = None
        for _ in range(steps):
            out = call(host_t.data_ptr(), n_clips=n, n_samples=N_SAMPLES)
        torch.cuda.synchronize()
        own = time.perf_counter() - t0
        t = torch.tensor([own], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), own, out

    def balanced_shards(own_s, steps):
        """clips per step and rank in proportion to the ranks' own end-to-end rates under concurrent load (the PCIe links of
        a box are not all alike: DESIGN.md section 4), same total as equal shards; equal shards when the rates agree to 5 %"""
        if dist is None:
            return [E]
        r = torch.zeros(world, dtype=torch.float64, device=dev)
        r[rank] = E * steps / own_s
        dist.all_reduce(r)
        rates = r.cpu().numpy()
        fake = os.environ.get("SJ_BENCH_FAKE_RATE_FACTORS")  # test hook: exercise the unequal-shard path on a box with equal links
        if fake:
            rates = rates * np.array([float(x) for x in fake.split(",")][:world])
        if rates.max() / rates.min() < 1.05:
            return [E] * world
        sh = np.floor(world * E * rates / rates.sum() / 256.0).astype(np.int64) * 256
        sh = np.minimum(sh, C)
        return [int(x) for x in sh]

    host = torch.empty((E, N_SAMPLES), dtype=torch.float32, pin_memory=True)
    host.copy_(clips[:E])
    torch.cuda.synchronize()
    for _ in range(max(1, min(args.warmup, 2))):
        sj.infer_batch(host.data_ptr(), n_clips=E, n_samples=N_SAMPLES)
    e2e_s, own_s, (h_labels, h_conf, h_score) = e2e_run(sj.infer_batch, host, E, args.e2e_steps)
    e2e_equal = world * E * args.e2e_steps / e2e_s
    assert np.array_equal(h_labels, labels[:E].cpu().numpy()), "host-path labels differ from device-path labels"
    shards = balanced_shards(own_s, args.e2e_steps) if args.e2e_balance else [E] * world
    e2e_value, Er = e2e_equal, E
    if shards != [E] * world:
        Er = shards[rank]
        if Er > E:
            host = torch.empty((Er, N_SAMPLES), dtype=torch.float32, pin_memory=True)
            host.copy_(clips[:Er])
            torch.cuda.synchronize()
        sj.infer_batch(host.data_ptr(), n_clips=Er, n_samples=N_SAMPLES)
        bal_s, _, (h_labels, h_conf, h_score) = e2e_run(sj.infer_batch, host, Er, args.e2e_steps)
        assert np.array_equal(h_labels, labels[:Er].cpu().numpy()), "host-path labels differ from device-path labels"
        e2e_balanced = sum(shards) * args.e2e_steps / bal_s
        if e2e_balanced > e2e_equal:
            e2e_value = e2e_balanced
        else:  # the rebalanced run was not faster: keep the equal-shard figure as the headline and say so
            shards = [E] * world
            Er = E

    # ---- int16 PCM ingest (the reference's on-disk format; SURVEY 8(d): reported separately, 32 008 B/clip)
    pcm = None
    if args.pcm16:
        Cp = min(C, args.pcm16_clips)
        pcm_dev = torch.empty((Cp, N_SAMPLES), dtype=torch.int16, device=dev)
        for i in range(0, Cp, 65536):
            j = min(Cp, i + 65536)
            pcm_dev[i:j] = torch.clamp(torch.round(clips[i:j] * 32768.0), -32768, 32767).to(torch.int16)
        kp = []
        for i in range(4):
            sj.infer_batch_pcm16_dev(pcm_dev, Cp, N_SAMPLES, labels[:Cp], conf[:Cp], score[:Cp], None)
            kp.append(sj.last_kernel_ms())
        kp = min(kp[1:])
        Ep = min(Er, Cp)
        hpcm = torch.empty((Ep, N_SAMPLES), dtype=torch.int16, pin_memory=True)
        hpcm.copy_(pcm_dev[:Ep])
        torch.cuda.synchronize()
        sj.infer_batch_pcm16(hpcm.data_ptr(), n_clips=Ep, n_samples=N_SAMPLES)
        pcm_s, _, _ = e2e_run(sj.infer_batch_pcm16, hpcm, Ep, args.e2e_steps)
        tot = torch.tensor([float(Ep)], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(tot)
        pcm_e2e = float(tot.item()) * args.e2e_steps / pcm_s
        pgbs = (N_SAMPLES * 2 + 8) * Cp / (kp * 1e-3) / 1e9
        ppk, ppk_src = measured_peak_gbs()
        ptr = (ncu_traffic_bytes_per_clip() or {}).get("pcm16_dram_bytes_per_clip")
        pcm = {"clips_per_s_per_gpu_resident": Cp / (kp * 1e-3), "kernel_ms": kp, "clips": Cp,
               "bytes_per_clip": N_SAMPLES * 2 + 8, "achieved_gbs": pgbs,
               "roofline": {"bound": "hbm", "achieved": pgbs, "peak": ppk, "unit": "GB/s", "frac": pgbs / ppk,
                            "traffic": ptr * Cp if ptr else None, "peak_source": ppk_src, "kernel_ms": kp,
                            "kernel": "cepstrum_kernel<12 warps, TMA, fast-16k, PCM16>", "bytes_per_clip": N_SAMPLES * 2 + 8},
               "e2e": {"value": pcm_e2e, "unit": "clips/s",
                       "h2d_bytes_per_step": E * N_SAMPLES * 2, "d2h_bytes_per_step": E * 12,
                       "api": "scrubjay_infer_batch_pcm16 (host int16 PCM, pinned)"},
               "note": "outputs bit-identical to the float path on s/32768 (tests/test_pcm16_gpu.py)"}
        del pcm_dev

    # ---- donut path (BASELINE config 3: Butterworth IIR + FFT band energies), the same resident clips, per GPU
    donut = None
    if args.donut:
        import dsp_b200
        Cd = min(C, args.donut_clips)
        det = torch.empty((Cd, 4), dtype=torch.float32, device=dev)
        donut = {"clips": Cd, "bytes_per_clip": N_SAMPLES * 4 + 4, "frames_per_clip": 142}
        for name, variant in (("sync_f32", dsp_b200.DONUT_SYNC_F32), ("desktop_f64", dsp_b200.DONUT_DESKTOP_F64)):
            kd = []
            for i in range(3):
                sj.donut_classify_batch_dev(clips, Cd, N_SAMPLES, variant, labels[:Cd], det)
                kd.append(sj.last_kernel_ms())
            kd = min(kd[1:])
            gbs = (N_SAMPLES * 4 + 4) * Cd / (kd * 1e-3) / 1e9
            pk, pk_src = measured_peak_gbs()
            tr = (ncu_traffic_bytes_per_clip() or {}).get("donut_" + name + "_dram_bytes_per_clip")
            donut[name] = {"kernel_ms": kd, "clips_per_s_per_gpu": Cd / (kd * 1e-3), "frames_per_s_per_gpu": 142 * Cd / (kd * 1e-3),
                           "achieved_gbs": gbs,
                           "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk, "unit": "GB/s", "frac": gbs / pk,
                                        "traffic": tr * Cd if tr else None, "peak_source": pk_src, "kernel_ms": kd,
                                        "kernel": "donut_ws_kernel<%s> (%d filter + 8 transform + 4 epilogue warps)" % (("float", 4) if name == "sync_f32" else ("double", 3)),
                                        "bytes_per_clip": N_SAMPLES * 4 + 4,
                                        "note": "FP32 / FP64 issue of the exact (no-FMA) recurrence and FFT binds, not HBM"},
                           "positives": int(labels[:Cd].sum().item())}
        if args.cpu_baseline and world == 1:
            ref = donut_cpu_reference(host[:min(E, 32768)].numpy())
            if ref is not None:
                nr = len(ref[1])
                # a batch large enough for the warp-specialised kernel (small float batches take the round-1 kernel), then the
                # same clips alone (round-1 kernel): both must agree with the reference
                nb = min(Cd, max(nr, 65536))
                sj.donut_classify_batch_dev(clips, nb, N_SAMPLES, dsp_b200.DONUT_SYNC_F32, labels[:nb], det[:nb])
                same = int((labels[:nr].cpu().numpy() == np.array(ref[1], dtype=np.int32)).sum())
                sj.donut_classify_batch_dev(clips, nr, N_SAMPLES, dsp_b200.DONUT_SYNC_F32, labels[:nr], det[:nr])
                same_small = int((labels[:nr].cpu().numpy() == np.array(ref[1], dtype=np.int32)).sum())
                donut["cpu_reference"] = dict(ref[0], labels_equal_to_gpu="%d of %d" % (same, len(ref[1])),
                                              labels_equal_to_gpu_small_batch_kernel="%d of %d" % (same_small, len(ref[1])))
        donut["note"] = ("fused donut kernels (IIR x2 bit-exact, spectrogram x2, band sums, verdict); instruction-bound "
                         "(exact no-FMA replay of the reference recurrence and FFT), see DESIGN.md section 6")
        del det

    # ---- BASELINE configs 3 / 4 at their named sizes (config 5 runs last: it needs the clips' memory)
    config3 = run_config3(sj, clips, labels, C, dev) if (args.configs and world == 1) else None
    config4 = run_config4(sj, clips, C, conf, dist, dev, rank, world, barrier) if (args.configs and world >= 2 and peer is not None) else None

    # ---- BASELINE config 1 (single-clip inference): latency of one 1 s clip through the host-pointer call
    single = None
    if rank == 0:
        one = torch.empty((1, N_SAMPLES), dtype=torch.float32, pin_memory=True)
        one.copy_(clips[:1])
        torch.cuda.synchronize()
        ws, ks = [], []
        for i in range(24):
            t0 = time.perf_counter()
            sj.infer_batch(one.data_ptr(), n_clips=1, n_samples=N_SAMPLES)
            ws.append((time.perf_counter() - t0) * 1e6)
        for i in range(8):
            sj.infer_batch_dev(clips, 1, N_SAMPLES, labels, conf, score, None)
            ks.append(sj.last_kernel_ms() * 1e3)
        single = {"host_call_us": statistics.median(ws[4:]), "kernels_us": statistics.median(ks[2:]),
                  "api": "scrubjay_infer_batch, 1 clip in pinned host memory: zero-copy path (the kernels read the clip over PCIe and write "
                         "the results to mapped host memory: 2 launches + 1 synchronisation); frame-parallel small-batch kernels"}
        pageable = one.numpy().copy()
        wp = []
        for i in range(24):
            t0 = time.perf_counter()
            sj.infer_batch(pageable)
            wp.append((time.perf_counter() - t0) * 1e6)
        single["host_call_pageable_us"] = statistics.median(wp[4:])  # H2D copy + 2 kernels + D2H copies

    config5 = None
    if args.configs and world >= 8 and peer is not None:
        peer.close()
        peer = None
        del clips, host
        torch.cuda.empty_cache()
        config5 = run_config5(sj, dist, dev, rank, world, barrier)
        clips = None

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        achieved = BYTES_PER_CLIP * C / (kernel_ms * 1e-3) / 1e9
        traffic = ncu_traffic_bytes_per_clip()
        cpu = cpu_baseline(host[:min(E, 65536)].numpy()) if (args.cpu_baseline and world == 1) else None  # N = 1 only (host is kept at N = 1)
        if cpu and single and cpu.get("one_thread_value"):
            single["cpu_port_us"] = 1e6 / cpu["one_thread_value"]
        clk_sum = clk.summary()
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        fp32_frac = (value / world) * FP32_PIPE_CYCLES_PER_CLIP / (sms * 4 * clk_sum["sm_mhz"] * 1e6) if clk_sum.get("sm_mhz") else None
        line = {
            "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "batched cepstrum features + SVM on 1M synthetic 1 s 16 kHz mono clips, per GPU",
                       "clips_per_gpu": C, "n_samples": N_SAMPLES, "frames_per_clip": FRAMES_PER_CLIP,
                       "sharding": "clip-sharded x%d, label/score gather (8 B/clip) inside each step" % world,
                       "gather": gather_mode,
                       "numa": "rank bound to NUMA node %s of its GPU" % numa_node if numa_node is not None else "no NUMA node reported for the GPU (nothing bound)",
                       "l2": "inputs (%.0f GB/GPU) >> L2, no flush needed" % (C * N_SAMPLES * 4 / 1e9),
                       "timing": "CUDA events on the launching stream, max over ranks"},
            "frames_per_s": value * FRAMES_PER_CLIP,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (traffic or {}).get("dram_bytes_per_clip", 0) * C or None,
                         "traffic_source": "ncu dram bytes per clip (%.0f B, profiles/roofline_traffic.json: a 200 k-clip --set full capture) "
                                           "x the %d clips of this launch" % ((traffic or {}).get("dram_bytes_per_clip", 0), C),
                         "fp32_frac": fp32_frac, "fp32_pipe_cycles_per_clip": FP32_PIPE_CYCLES_PER_CLIP,
                         "fp32_frac_note": "clips/s x FP32-pipe cycles per clip (16 frames x (643 packed x 2 + 142 scalar) warp-instruction "
                                           "cycles per scheduler, profiles/r2f_ncu_opcodes_cepstrum.txt; issue rates measured by "
                                           "tools/micro/fp32_issue.cu) / (SMs x 4 schedulers x the SM clock sampled during the run)",
                         "fp32_note": "binding resources are the FP32 pipe (58 %), instruction issue (56 %) and the shared-memory LSU (72 %), "
                                      "all per profiles/r2f_ncu_full_cepstrum.txt; HBM is at 29 %",
                         "peak_source": peak_src, "kernel": "cepstrum_kernel<12 warps, TMA, fast-16k>",
                         "kernel_ms": kernel_ms, "bytes_per_clip": BYTES_PER_CLIP,
                         "note": "shared-memory bandwidth and instruction issue of the FFT, not HBM, bind this kernel (see DESIGN.md): frac is of the HBM roofline"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "clips/s", "h2d_bytes_per_step": E * N_SAMPLES * 4,
                    "d2h_bytes_per_step": E * 12, "clips_per_step": E, "steps": args.e2e_steps,
                    "per": "rank (mean over the ranks: the shards add up to n_gpus x clips_per_step)",
                    "shard_clips": shards, "equal_shards_value": e2e_equal,
                    "sharding": ("clips per rank in proportion to each rank's own end-to-end rate measured under concurrent load "
                                 "(unequal PCIe links), same total" if shards != [E] * world else "equal shards"),
                    "api": "scrubjay_infer_batch (host pointers, pinned)"},
            "pcm16": pcm,
            "donut": donut,
            "config3": config3, "config4": config4, "config5": config5,
            "gather_check": gather_check,
            "single_clip": single,
            "gpu_launches": launches,
            "clocks": clk_sum,
        }
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    barrier()
    if peer is not None:
        peer.close()
    sj.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--clips", type=int, default=1000000, help="resident clips per GPU per step")
    ap.add_argument("--e2e-clips", type=int, default=65536, help="clips per end-to-end step (pinned host buffer)")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e-balance", dest="e2e_balance", action="store_false",
                    help="keep equal e2e shards even when the ranks' host links differ")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--ref-clips-per-thread", type=int, default=1024)
    ap.add_argument("--gather", choices=["p2p", "nccl"], default="p2p",
                    help="N > 1: fused peer-store gather inside the kernel (default) or NCCL all-gathers after it")
    ap.add_argument("--no-pcm16", dest="pcm16", action="store_false", help="skip the int16 PCM ingest measurement")
    ap.add_argument("--pcm16-clips", type=int, default=500000)
    ap.add_argument("--no-donut", dest="donut", action="store_false", help="skip the donut-path measurement")
    ap.add_argument("--donut-clips", type=int, default=500000)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-configs", dest="configs", action="store_false",
                    help="skip BASELINE configs 3 (N = 1), 4 (N >= 2) and 5 (N = 8) at their named sizes")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 1)
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank)
    return run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    sys.exit(main())
