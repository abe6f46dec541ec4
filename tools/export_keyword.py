#!/usr/bin/env python3
"""Export the constants of the reference's keyword ("stop") detector to one flat blob.

Reads  2fa/audio/word/c/mfcc_params.h   (HANN_WINDOW, MEL_FILTER, DCT_MATRIX + the MFCC_* defines)
       2fa/audio/word/c/model_params.h  (SCALER_*, DENSE*_KERNEL / _BIAS + INPUT_SIZE, DENSE*_UNITS)
       2fa/audio/word/c/stop_detector.c (MAX_FRAMES)
under the reference tree and writes dsp_b200/data/keyword_model.bin.  The headers are generated data files
(weights of a trained model); only the numbers are taken.

Blob "SJKW" v1, little-endian:
  u32 magic 0x574B4A53, version 1, sample_rate, n_fft, frame_len, hop, n_mels, n_mfcc, n_bins, max_frames,
      input_size, units1, units2, units3, units4, reserved
  f32 hann[frame_len], mel[n_mels*n_bins], dct[n_mfcc*n_mels], scaler_mean[input], scaler_scale[input],
      w1[input*units1], b1[units1], w2[units1*units2], b2[units2], w3[units2*units3], b3[units3],
      w4[units3*units4], b4[units4]
"""
import os
import re
import struct
import sys

MAGIC = 0x574B4A53


def parse_header(path):
    txt = open(path).read()
    defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(\w+)\s+(\d+)\s", txt)}
    arrays = {}
    for m in re.finditer(r"static\s+const\s+float\s+(\w+)\[(\d+)\]\s*=\s*\{(.*?)\};", txt, re.S):
        vals = [float(v.rstrip("fF")) for v in re.findall(r"[-+]?(?:\d+\.?\d*|\.\d+)(?:[eE][-+]?\d+)?[fF]?", m.group(3))]
        assert len(vals) == int(m.group(2)), (m.group(1), len(vals), m.group(2))
        arrays[m.group(1)] = vals
    return defs, arrays


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    d = os.path.join(ref, "2fa", "audio", "word", "c")
    pd, pa = parse_header(os.path.join(d, "mfcc_params.h"))
    md, ma = parse_header(os.path.join(d, "model_params.h"))
    sd, _ = parse_header(os.path.join(d, "stop_detector.c"))
    max_frames = sd["MAX_FRAMES"]
    u = [md["DENSE%d_UNITS" % i] for i in (1, 2, 3, 4)]
    inp = md["INPUT_SIZE"]
    assert inp == pd["MFCC_N_MFCC"] * max_frames
    hdr = struct.pack("<16I", MAGIC, 1, pd["MFCC_SAMPLE_RATE"], pd["MFCC_N_FFT"], pd["MFCC_FRAME_LENGTH"],
                      pd["MFCC_HOP_LENGTH"], pd["MFCC_N_MELS"], pd["MFCC_N_MFCC"], pd["MFCC_N_FREQ_BINS"], max_frames,
                      inp, u[0], u[1], u[2], u[3], 0)
    seq = [pa["HANN_WINDOW"], pa["MEL_FILTER"], pa["DCT_MATRIX"], ma["SCALER_MEAN"], ma["SCALER_SCALE"],
           ma["DENSE1_KERNEL"], ma["DENSE1_BIAS"], ma["DENSE2_KERNEL"], ma["DENSE2_BIAS"], ma["DENSE3_KERNEL"],
           ma["DENSE3_BIAS"], ma["DENSE4_KERNEL"], ma["DENSE4_BIAS"]]
    assert len(pa["MEL_FILTER"]) == pd["MFCC_N_MELS"] * pd["MFCC_N_FREQ_BINS"]
    assert len(ma["DENSE1_KERNEL"]) == inp * u[0]
    body = b"".join(struct.pack("<%df" % len(a), *a) for a in seq)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = os.path.join(root, "dsp_b200", "data", "keyword_model.bin")
    with open(out, "wb") as f:
        f.write(hdr + body)
    print("wrote %s: %d bytes" % (out, len(hdr) + len(body)))


if __name__ == "__main__":
    main()
