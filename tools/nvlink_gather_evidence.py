#!/usr/bin/env python3
"""NVLink evidence for the fused compute + gather kernel, measured WITHOUT a profiler: the NVML NVLink
data-throughput counters of both GPUs, read before and after K launches of scrubjay_infer_batch_gather_dev on GPU 0 with
gather targets on GPU 0 (local) and GPU 1 (peer, over NVLink / NVSwitch).  One process, two devices (same-process peer
access; the multi-process form maps the same memory through CUDA IPC).  Expected payload: 2 x 4 bytes per clip per peer.
usage (2 GPUs): python tools/nvlink_gather_evidence.py [clips] [launches] > profiles/..."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pynvml
import torch
import dsp_b200

FI_TX, FI_RX = 138, 139  # NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX / _RX (KiB; scopeId UINT_MAX = all links)


ERR = []


def smi_counters(i):
    """fallback: `nvidia-smi nvlink -gt d -i <gpu>` (Data Tx / Data Rx per link, KiB), summed over links"""
    import re, subprocess
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(i)], capture_output=True, text=True, timeout=20).stdout
        tx = sum(int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
        rx = sum(int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
        if "Data Tx" not in out:
            ERR.append("nvidia-smi nvlink: " + out.strip()[:200])
            return None
        return [tx, rx]
    except Exception as e:
        ERR.append("nvidia-smi nvlink: %r" % (e,))
        return None


def counters(h, i=0):
    c = nvml_counters(h)
    return c if c is not None else smi_counters(i)


def nvml_counters(h):
    try:
        v = pynvml.nvmlDeviceGetFieldValues(h, [(FI_TX, 0xFFFFFFFF), (FI_RX, 0xFFFFFFFF)])
        out = []
        for x in v:
            if x.nvmlReturn != 0:
                ERR.append("nvmlDeviceGetFieldValues: nvmlReturn %d" % x.nvmlReturn)
                return None
            out.append(int(x.value.ullVal))
        return out
    except Exception as e:
        ERR.append("nvmlDeviceGetFieldValues: %r" % (e,))
        return None


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    assert torch.cuda.device_count() >= 2, "needs two GPUs"
    pynvml.nvmlInit()
    hs = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(2)]
    sj = dsp_b200.ScrubJay(device=0, samplerate=16000)
    torch.cuda.set_device(0)
    sj.set_stream(torch.cuda.current_stream().cuda_stream)
    clips = torch.empty((n, 16000), dtype=torch.float32, device="cuda:0")
    sj.synth_clips_dev(clips, 0, n, 16000)
    lab = [torch.full((n,), -1, dtype=torch.int32, device="cuda:%d" % d) for d in range(2)]
    sco = [torch.full((n,), float("nan"), dtype=torch.float32, device="cuda:%d" % d) for d in range(2)]
    res = {"clips": n, "launches": K}
    for name, targets in (("local_only", [0]), ("local_plus_peer", [0, 1])):
        sj.gather_bind([lab[d] for d in targets], [sco[d] for d in targets], 0)
        sj.infer_batch_gather_dev(clips, n, 16000)  # warm-up
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        time.sleep(0.3)
        c0 = [counters(h, i) for i, h in enumerate(hs)]
        ms = []
        for _ in range(K):
            sj.infer_batch_gather_dev(clips, n, 16000)
            ms.append(sj.last_kernel_ms())
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        time.sleep(0.3)
        c1 = [counters(h, i) for i, h in enumerate(hs)]
        r = {"kernel_ms_median": float(np.median(ms))}
        if None not in c0 and None not in c1:
            r["gpu0_nvlink_tx_kib"] = c1[0][0] - c0[0][0]
            r["gpu0_nvlink_rx_kib"] = c1[0][1] - c0[0][1]
            r["gpu1_nvlink_tx_kib"] = c1[1][0] - c0[1][0]
            r["gpu1_nvlink_rx_kib"] = c1[1][1] - c0[1][1]
        else:
            r["nvml"] = "NVLink throughput counters unavailable on this box: " + "; ".join(sorted(set(ERR)))[:400]
        res[name] = r
    res["expected_peer_payload_kib"] = K * n * 8 / 1024.0
    ok = torch.equal(lab[0].cpu(), lab[1].cpu()) and torch.equal(sco[0].cpu(), sco[1].cpu()) and int((lab[1] < 0).sum()) == 0
    res["peer_arrays_equal_local_arrays"] = bool(ok)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
