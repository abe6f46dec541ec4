"""Host-side WAV ingest (scrubjay_wav_parse) -- no GPU needed.  Checked against Python's own RIFF walk on
synthetic files and, when the reference tree is mounted (this container only), against the reference's
fixtures: the frame count must equal what read_wav_file (donut-classifier/classifier.c:214-306) computes."""
import glob
import os
import struct

import numpy as np
import pytest

import dsp_b200


def _wav(pcm, rate=16000, channels=1, extra=b"", fmt_extra=b"", fmt=1):
    body = pcm.astype("<i2").tobytes()
    fmt_chunk = struct.pack("<HHIIHH", fmt, channels, rate, rate * channels * 2, channels * 2, 16) + fmt_extra
    chunks = b"fmt " + struct.pack("<I", len(fmt_chunk)) + fmt_chunk + extra + b"data" + struct.pack("<I", len(body)) + body
    return b"RIFF" + struct.pack("<I", 4 + len(chunks)) + b"WAVE" + chunks


def test_plain_and_chunked_wavs():
    pcm = (np.arange(3000) % 200 - 100).astype(np.int16)
    for ch in (1, 2, 3):
        for extra in (b"", b"LIST" + struct.pack("<I", 6) + b"abcdef", b"fact" + struct.pack("<I", 4) + b"\0\0\0\0" + b"LIST" + struct.pack("<I", 3) + b"xyz\0"):
            w = _wav(pcm[:3000 // ch * ch], 16000, ch, extra)
            info = dsp_b200.wav_parse(w)
            assert info is not None
            assert (info.channels, info.sample_rate, info.bits_per_sample, info.audio_format) == (ch, 16000, 16, 1)
            assert info.frames == 3000 // ch
            got = np.frombuffer(w, "<i2", count=info.frames * ch, offset=info.data_offset)
            assert np.array_equal(got, pcm[:3000 // ch * ch])


def test_extended_fmt_chunk_and_truncation():
    pcm = np.arange(100, dtype=np.int16)
    info = dsp_b200.wav_parse(_wav(pcm, fmt_extra=b"\0\0"))
    assert info is not None and info.frames == 100
    w = _wav(pcm)
    info = dsp_b200.wav_parse(w[:-20])  # file shorter than its data chunk says: what is there
    assert info is not None and info.frames == 90


def test_rejections_match_the_reference():
    pcm = np.zeros(10, np.int16)
    assert dsp_b200.wav_parse(b"RIFX" + _wav(pcm)[4:]) is None           # not RIFF (classifier.c:232-237)
    assert dsp_b200.wav_parse(_wav(pcm)[:8] + b"WAVX" + _wav(pcm)[12:]) is None  # not WAVE (:240-245)
    assert dsp_b200.wav_parse(_wav(pcm, fmt=3)) is None                  # AudioFormat != 1 (:264-269)
    assert dsp_b200.wav_parse(b"RIFF") is None
    w = _wav(pcm)
    assert dsp_b200.wav_parse(w[:12] + b"LIST" + w[16:]) is None          # "fmt " must come first (:248-253)


@pytest.mark.skipif(not os.path.isdir("/root/reference/donut-classifier"), reason="reference tree not mounted")
def test_reference_fixtures_parse_like_python_wave():
    import wave
    files = sorted(glob.glob("/root/reference/donut-classifier/**/*.wav", recursive=True))[:40]
    assert files
    seen = 0
    for f in files:
        data = open(f, "rb").read()
        info = dsp_b200.wav_parse(data)
        try:
            with wave.open(f, "rb") as w:
                ref = (w.getnchannels(), w.getframerate(), w.getsampwidth() * 8, w.getnframes())
        except Exception:
            continue
        assert info is not None, f
        assert (info.channels, info.sample_rate, info.bits_per_sample, info.frames) == ref, f
        seen += 1
    assert seen > 0


def test_lenient_parse_skips_chunks_in_front_of_fmt():
    """cepstrum/testing/P_1363v2-sj-short.WAV opens with a JUNK chunk: read_wav_file (classifier.c:248-253) would refuse
    it, libsndfile (behind aubio_source, scrubjay_infer.c:21) reads it.  strict=False follows the latter."""
    pcm = (np.arange(500) % 97 - 40).astype(np.int16)
    w = _wav(pcm, 96000, 2)
    junk = b"JUNK" + struct.pack("<I", 28) + b"\0" * 28
    w2 = w[:12] + junk + w[12:]
    assert dsp_b200.wav_parse(w2) is None
    info = dsp_b200.wav_parse(w2, strict=False)
    assert info is not None and (info.channels, info.sample_rate, info.frames) == (2, 96000, 250)
    got = np.frombuffer(w2, "<i2", count=500, offset=info.data_offset)
    assert np.array_equal(got, pcm)
    a, b = dsp_b200.wav_parse(w), dsp_b200.wav_parse(w, strict=False)
    assert (a.frames, a.data_offset) == (b.frames, b.data_offset)
