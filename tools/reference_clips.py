#!/usr/bin/env python3
"""BASELINE config 1: the reference's own cepstrum test material through the path.

cepstrum/scrubjay_infer.c:15 iterates `testing/`; per file (:156-177) it runs mfcc_stats (aubio source at the
file's NATIVE rate, :21,28-30) and ort_predict.  The decodable material there is five 96 kHz stereo WAVs whose
`P_` / `N_` prefixes are the only ground truth the reference holds for this path; the 16-bit WAVs of
cepstrum/data/{neg,pos} (the training set of train.py:42-60, label = folder) are added as further rows.

  --make-golden   (build container only: needs /root/reference + sklearn)  writes tests/golden/reference_clips.json:
                  per file the 40 features of an INDEPENDENT float64 numpy/scipy restatement (tools/make_golden.py:
                  mfcc_stats_np), the oracle's features, and sklearn's decision_function / predict / predict_proba on
                  the reference's own scrubjay_svm.joblib for those features.
  --gpu           (GPU box) parses the WAVs copied to oracle/_ref/cepstrum_testing/ by oracle/Makefile with
                  scrubjay_wav_parse, decodes them on the device (scrubjay_decode_pcm16_dev, SJ_MIX_MEAN = aubio's
                  down-mix), runs scrubjay_infer_batch_dev at the file's rate, and writes the per-file table
                  (JSON + markdown) next to the oracle's and the golden values.
"""
import argparse
import glob
import json
import os
import sys
import wave

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
REF = "/root/reference"
SHIPPED = os.path.join(ROOT, "oracle", "_ref", "cepstrum_testing")
GOLDEN = os.path.join(ROOT, "tests", "golden", "reference_clips.json")


def clip_dir():
    """the WAVs: oracle/_ref/cepstrum_testing (copied there by oracle/Makefile; travels to the GPU box) or, in the
    build container, the reference tree itself"""
    if os.path.isdir(SHIPPED) and glob.glob(os.path.join(SHIPPED, "*")):
        return SHIPPED
    return None


def list_clips():
    d = clip_dir()
    if d is None:
        return []
    return sorted(f for f in glob.glob(os.path.join(d, "*")) if f.lower().endswith(".wav"))


def truth_of(name):
    """ground truth the reference implies: P_/N_ prefix (testing/), or the training folder (data/neg -> neg__ prefix)"""
    if name.startswith("P_") or name.startswith("pos__"):
        return 1
    if name.startswith("N_") or name.startswith("neg__"):
        return 0
    return -1


def read_wav_mean(path):
    """aubio_source_do's mono down-mix restated on the host in float32: every channel s / 32768, summed in channel
    order, divided by the channel count"""
    w = wave.open(path, "rb")
    assert w.getsampwidth() == 2, path
    a = np.frombuffer(w.readframes(w.getnframes()), np.int16).reshape(-1, w.getnchannels())
    acc = np.zeros(a.shape[0], np.float32)
    for c in range(a.shape[1]):
        acc = acc + a[:, c].astype(np.float32) * np.float32(1.0 / 32768.0)
    return (acc / np.float32(a.shape[1])).astype(np.float32), w.getframerate(), a.shape[1]


def labels_from(d, proba):
    """the three candidate label rules of SURVEY B.3"""
    return {"predict_d_gt_0": int(d > 0), "vote_d_ge_0": int(d >= 0), "proba_argmax": int(proba[1] > proba[0])}


def make_golden():
    import warnings
    warnings.filterwarnings("ignore")
    import joblib
    from make_golden import mfcc_stats_np
    from oracle_lib import CepstrumOracle
    model = joblib.load(os.path.join(REF, "cepstrum", "scrubjay_svm.joblib"))
    rows = []
    for path in list_clips():
        name = os.path.basename(path)
        x, sr, ch = read_wav_mean(path)
        f_np, _ = mfcc_stats_np(x, sr)
        f_or = CepstrumOracle(sr).mfcc_stats(x)
        X = f_or.astype(np.float64)[None]
        d = float(model.decision_function(X)[0])
        proba = model.predict_proba(X)[0].tolist()
        rows.append({"file": name, "truth": truth_of(name), "samplerate": sr, "channels": ch, "samples": int(x.size),
                     "frames": int((x.size + 1023) // 1024),
                     "features_numpy_f64": [float(v) for v in f_np], "features_oracle": [float(v) for v in f_or],
                     "sklearn_decision_function": d, "sklearn_predict": int(model.predict(X)[0]),
                     "sklearn_predict_proba": proba, "labels": labels_from(d, proba)})
        print("%-60s sr %6d  d %+.5f  predict %d  truth %d  max|oracle - numpy| %.2e" %
              (name[:60], sr, d, rows[-1]["sklearn_predict"], rows[-1]["truth"], np.abs(f_np - f_or).max()))
    json.dump({"generator": "tools/reference_clips.py --make-golden", "rows": rows}, open(GOLDEN, "w"), indent=1)
    print("wrote", GOLDEN)


def run_gpu_rows():
    """the GPU path on every shipped WAV: [(name, sr, features[40], score, label, conf, n_samples)]"""
    import torch
    import dsp_b200
    out = []
    ctxs = {}
    for path in list_clips():
        raw = open(path, "rb").read()
        info = dsp_b200.wav_parse(raw, strict=False)  # aubio_source = libsndfile: any chunk order
        assert info is not None and info.bits_per_sample == 16, path
        sr = info.sample_rate
        if sr not in ctxs:
            ctxs[sr] = dsp_b200.ScrubJay(device=0, samplerate=sr)
            ctxs[sr].set_stream(torch.cuda.current_stream().cuda_stream)
        sj = ctxs[sr]
        pcm = np.frombuffer(raw, "<i2", count=info.frames * info.channels, offset=info.data_offset)
        d_pcm = torch.from_numpy(pcm.copy()).cuda()
        d_x = torch.empty(info.frames, dtype=torch.float32, device="cuda")
        sj.decode_pcm16_dev(d_pcm, info.frames, info.channels, dsp_b200.MIX_MEAN, d_x)
        lab = torch.empty(1, dtype=torch.int32, device="cuda")
        conf = torch.empty(1, dtype=torch.float32, device="cuda")
        score = torch.empty(1, dtype=torch.float32, device="cuda")
        feats = torch.empty((1, 40), dtype=torch.float32, device="cuda")
        sj.infer_batch_dev(d_x, 1, int(info.frames), lab, conf, score, feats)
        torch.cuda.synchronize()
        out.append({"file": os.path.basename(path), "samplerate": sr, "samples": int(info.frames),
                    "features": feats.cpu().numpy()[0], "score": float(score.item()), "label": int(lab.item()),
                    "conf": float(conf.item()), "mono": d_x.cpu().numpy()})
    return out


def gpu_table(out_json, out_md):
    from oracle_lib import CepstrumOracle
    gold = {r["file"]: r for r in json.load(open(GOLDEN))["rows"]}
    rows = []
    for g in run_gpu_rows():
        ora = CepstrumOracle(g["samplerate"])
        of = ora.mfcc_stats(g["mono"])  # the oracle consumes the exact samples the GPU decoded
        ol, oc, os_ = ora.predict(of)
        gd = gold.get(g["file"])
        err = np.abs(g["features"] - of) / (1e-4 * np.abs(of) + 1e-5)
        rows.append({"file": g["file"], "truth": truth_of(g["file"]), "samplerate": g["samplerate"], "samples": g["samples"],
                     "gpu_features": [float(v) for v in g["features"]], "gpu_decision_value": g["score"],
                     "gpu_label": g["label"], "gpu_conf": g["conf"], "oracle_decision_value": os_, "oracle_label": ol,
                     "feature_err_over_tolerance": float(err.max()),
                     "golden_sklearn_decision_function": gd["sklearn_decision_function"] if gd else None,
                     "golden_labels": gd["labels"] if gd else None,
                     "max_abs_gpu_minus_numpy_f64": float(np.abs(g["features"] - np.array(gd["features_numpy_f64"])).max()) if gd else None})
    json.dump({"rows": rows}, open(out_json, "w"), indent=1)
    with open(out_md, "w") as f:
        f.write("| file | truth (prefix / folder) | rate | frames | GPU d | oracle d | sklearn d (golden) | label GPU / (i) / (ii) / (iii) | "
                "max feature err (x tol) | agrees with truth |\n|---|---|---|---|---|---|---|---|---|---|\n")
        for r in rows:
            gl = r["golden_labels"] or {}
            f.write("| %s | %d | %d | %d | %+.5f | %+.5f | %s | %d / %s / %s / %s | %.3f | %s |\n" % (
                r["file"][:48], r["truth"], r["samplerate"], (r["samples"] + 1023) // 1024, r["gpu_decision_value"],
                r["oracle_decision_value"],
                "%+.5f" % r["golden_sklearn_decision_function"] if r["golden_sklearn_decision_function"] is not None else "-",
                r["gpu_label"], gl.get("vote_d_ge_0", "-"), gl.get("predict_d_gt_0", "-"), gl.get("proba_argmax", "-"),
                r["feature_err_over_tolerance"], "yes" if r["gpu_label"] == r["truth"] else "NO"))
    print(open(out_md).read())


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--make-golden", action="store_true")
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "reference_clips"))
    a = ap.parse_args()
    if a.make_golden:
        make_golden()
    if a.gpu:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        gpu_table(a.out + ".json", a.out + ".md")
