"""The keyword-detector oracle (oracle/keyword_oracle.c) against the committed golden fixture produced by the
UNMODIFIED reference sources (tools/make_keyword_golden.py) and, when oracle/_ref is present, against the
reference build itself.  Same compiler and libm => the restatement must agree exactly."""
import os

import numpy as np
import pytest

from keyword_lib import KeywordOracle, KeywordRef, test_signals as make_signals

GOLD = os.path.join(os.path.dirname(__file__), "golden", "keyword_golden.npz")
CASES = {"a": (6, 80000, 1), "b": (4, 16000, 2), "c": (2, 400, 3), "d": (2, 399, 4), "e": (2, 81000, 5), "f": (3, 23456, 6)}


@pytest.mark.parametrize("tag", sorted(CASES))
def test_oracle_matches_reference_golden(tag):
    g = np.load(GOLD)
    n, ns, seed = CASES[tag]
    x = make_signals(n, ns, seed)
    ora = KeywordOracle()
    for i in range(n):
        p, mf, nf = ora.classify(x[i])
        assert nf == g["nfr_" + tag][i]
        assert np.allclose(mf, g["mfcc_" + tag][i], rtol=1e-6, atol=1e-4), np.abs(mf - g["mfcc_" + tag][i]).max()
        assert abs(p - g["prob_" + tag][i]) <= 1e-6
    assert g["nfr_" + tag].max() <= 500 and (tag != "d" or g["nfr_d"].max() == 0)


@pytest.mark.skipif(not KeywordRef.available(), reason="oracle/_ref not built (no reference tree)")
def test_oracle_is_bit_identical_to_the_reference_build():
    ora, ref = KeywordOracle(), KeywordRef()
    x = make_signals(5, 40000, 77)
    for s in x:
        po, mo, no = ora.classify(s)
        pr, mr, nr = ref.classify(s)
        assert no == nr and np.array_equal(mo[:nr], mr[:nr]) and po == pr


def test_batch_helper():
    ora = KeywordOracle()
    x = make_signals(3, 16000, 9)
    assert np.array_equal(ora.classify_batch(x), np.array([ora.classify(s)[0] for s in x], np.float32))
