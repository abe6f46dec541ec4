"""numpy mirror of dsp_b200/csrc/synth_kernel.cu: the same counter-based synthetic audio, generated
on the host (used by `bench.py --impl reference`, which must not touch the GPU library, and by CPU
tests).  Integer hashing is bit-identical to the device generator; the chirp sine goes through
numpy's sin instead of sinpif, so samples agree to ~1e-7, not bit for bit."""
import numpy as np

U32 = np.uint32


def mix32(x):
    x = np.asarray(x, dtype=U32).copy()
    with np.errstate(over="ignore"):
        x ^= x >> U32(16)
        x *= U32(0x7FEB352D)
        x ^= x >> U32(15)
        x *= U32(0x846CA68B)
        x ^= x >> U32(16)
    return x


def hash3(seed, a, b):
    with np.errstate(over="ignore"):
        ha = mix32(np.asarray(a, U32) * U32(0x9E3779B9) + U32(0x85EBCA6B))
        hb = mix32(np.asarray(b, U32) * U32(0xC2B2AE35) + U32(0x27D4EB2F))
    return mix32(U32(seed) ^ ha ^ hb)


def u01(h):
    return (np.asarray(h, U32) >> U32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def synth_call_clips(first_clip, n_clips, n_samples, seed=0x5C2B, fs=16000.0):
    """flavour 1 of the device generator (SJ_SYNTH_CALLS): "donut"-shaped calls -- a faint noise floor and three
    frequency-modulated bursts around a common centre time: 600-2400 Hz and 5100-6900 Hz for 0.4 s, 2600-4900 Hz for
    0.14 s, amplitudes log-uniform per clip -- placed so that every donut variant's verdict (classifier.c:184 and the
    table of SURVEY C.5) fires on a sizeable fraction of the clips and misses on the rest."""
    f32 = np.float32
    fs = f32(fs)
    out = np.zeros((n_clips, n_samples), np.float32)
    idx = np.arange(n_samples, dtype=np.int64)
    # (centre Hz, deviation Hz, half duration s, log10 amplitude lo, span)
    spec = [(3750.0, 1150.0, 0.07, -1.5, 1.1), (1500.0, 900.0, 0.20, -1.5, 1.0), (6000.0, 900.0, 0.20, -2.7, 1.2)]
    fm = f32(15.0)
    with np.errstate(over="ignore"):
        for c in range(n_clips):
            cid = U32((first_clip + c) & 0xFFFFFFFF)
            hc = hash3(seed, cid, U32(0xCA11))
            na = np.exp2(f32(3.3219281) * (f32(-3.0) + f32(0.8) * u01(mix32(hc + U32(1))))).astype(f32)
            centre = f32(0.5) + f32(0.16) * (u01(mix32(hc + U32(2))) - f32(0.5))
            x = na * (f32(2.0) * u01(hash3(seed, cid, (U32(0x10000) + idx.astype(U32)))) - f32(1.0))
            for k, (fc, dev, half, l0, span) in enumerate(spec):
                h = hash3(seed, cid, U32(0x2000 + k))
                amp = np.exp2(f32(3.3219281) * (f32(l0) + f32(span) * u01(mix32(h + U32(1))))).astype(f32)
                start = int((centre - f32(half)) * fs)
                stop = int((centre + f32(half)) * fs)
                sel = (idx >= start) & (idx < stop) & (idx >= 0)
                t = (idx[sel] - start).astype(np.float32)
                beta = f32(dev) / (f32(6.2831855) * fm)
                mod = t * (fm / fs)
                mod = mod - np.floor(mod)
                cyc = t * (f32(fc) / fs) + beta * np.sin(2.0 * np.pi * mod.astype(np.float64)).astype(np.float32)
                cyc = cyc - np.floor(cyc)
                x[sel] += amp * np.sin(2.0 * np.pi * cyc.astype(np.float64)).astype(np.float32)
            out[c] = x
    return out


def synth_clips(first_clip, n_clips, n_samples, seed=0x5C2B, fs=16000.0, flavour=0):
    if flavour == 1:
        return synth_call_clips(first_clip, n_clips, n_samples, seed, fs)
    lo_hz = np.array([500.0, 2500.0, 5000.0, 1000.0], np.float32)
    hi_hz = np.array([2500.0, 5000.0, 7000.0, 3000.0], np.float32)
    out = np.zeros((n_clips, n_samples), np.float32)
    idx = np.arange(n_samples, dtype=np.int64)
    f32 = np.float32
    fs = f32(fs)
    with np.errstate(over="ignore"):
        for c in range(n_clips):
            cid = U32((first_clip + c) & 0xFFFFFFFF)
            hc = hash3(seed, cid, U32(0xA5A5))
            zero_from = n_samples
            if int(hc) % 20 == 0:
                zero_from = 0 if (int(hc) & 0x100) else int(u01(mix32(hc)) * f32(n_samples))
            x = f32(0.25) * (f32(2.0) * u01(hash3(seed, cid, (U32(0x10000) + idx.astype(U32)))) - f32(1.0))
            for k in range(3):
                h = hash3(seed, cid, U32(0x1000 + k))
                band = int(h) & 3
                a, b = u01(mix32(h + U32(1))), u01(mix32(h + U32(2)))
                f0 = lo_hz[band] + a * (hi_hz[band] - lo_hz[band])
                f1 = lo_hz[band] + b * (hi_hz[band] - lo_hz[band])
                dur = f32(0.1) + f32(0.6) * u01(mix32(h + U32(3)))
                st = u01(mix32(h + U32(4))) * f32(1.0 - 0.3)
                start = int(st * fs)
                stop = start + int(dur * fs)
                cf0 = f0 / fs
                ck = f32(0.5) * (f1 - f0) / (fs * dur * fs)
                amp = f32(0.2) * u01(mix32(h + U32(5)))
                sel = (idx >= start) & (idx < stop)
                t = (idx[sel] - start).astype(np.float32)
                cyc = t * (cf0 + ck * t)
                cyc = cyc - np.floor(cyc)
                x[sel] += amp * np.sin(2.0 * np.pi * cyc.astype(np.float64)).astype(np.float32)
            x[zero_from:] = 0
            out[c] = x
    return out
