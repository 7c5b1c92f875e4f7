"""CPU: known answers for oracle/pipeline_ref.py (the checker of the device-side input pipeline and WER)."""
import numpy as np

from oracle import pipeline_ref as P


def test_levenshtein_known_answers():
    assert P.levenshtein("kitten", "sitting") == 3
    assert P.levenshtein("", "abc") == 3 and P.levenshtein("abc", "") == 3 and P.levenshtein("", "") == 0
    assert P.levenshtein("flaw", "lawn") == 2
    assert P.levenshtein(list("intention"), list("execution")) == 5


def test_word_error_terms_follow_jiwer_definition():
    # ids: 4 = '|' (word delimiter); "AB|C|DE" vs reference "AB|X|DE|F": one substitution + one deletion over 4 words
    hyp = [[5, 6, 4, 7, 4, 8, 9]]
    ref = [[5, 6, 4, 10, 4, 8, 9, 4, 11, -100, -100]]
    assert P.word_error_terms(hyp, ref, delim=4, unit="word") == [(2, 4)]
    assert P.split_words([4, 4, 5, 4, 4, 6, 4], 4) == [(5,), (6,)]  # repeated / leading / trailing delimiters give no empty words
    assert P.word_error_terms([[5, 6]], [[5, 7, -100]], unit="char") == [(1, 2)]


def test_normalize_and_pad_matches_feature_extractor_formula():
    rng = np.random.default_rng(0)
    waves = [rng.standard_normal(1000).astype(np.float32) * 3 + 1, rng.standard_normal(400).astype(np.float32)]
    x, m, f = P.normalize_and_pad(waves)
    assert x.shape == (2, 1000) and m.sum(1).tolist() == [1000, 400]
    for b, n in enumerate((1000, 400)):
        assert abs(x[b, :n].mean()) < 1e-5 and abs(x[b, :n].var() - 1) < 1e-4 and (x[b, n:] == 0).all()
    assert f.tolist() == [2, 1]  # floor((L-k)/s)+1 per conv layer: 1000 -> 199 -> 99 -> 49 -> 24 -> 11 -> 5 -> 2; 400 -> 79 -> ... -> 1
