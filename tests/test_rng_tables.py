"""CPU: the library's MT19937 table generator must equal CPython's random.seed(s) stream and
numpy's legacy np.random.seed(s) stream word-for-word (host code, no GPU needed)."""
import random

import numpy as np

from helpers import py_food_words, py_jitter_words


def test_food_stream_matches_cpython(salp_lib):
    from underwater_swimmer_rl_b200 import rng_tables
    tab = rng_tables.fill_words(5, 2048, 10_000, 3, rng_tables.FOOD_STREAM)
    for r in range(5):
        assert np.array_equal(tab[r], py_food_words(10_000 + 3 + r, 2048))
    # and the doubles the reference would see: random.random() == res53(w0, w1)
    rr = random.Random(10_003)
    d = np.array([rr.random() for _ in range(1024)])
    w = tab[0].astype(np.uint64)
    assert np.array_equal(d, ((w[0::2] >> 5) * 67108864.0 + (w[1::2] >> 6)) / 9007199254740992.0)


def test_jitter_stream_matches_numpy(salp_lib):
    from underwater_swimmer_rl_b200 import rng_tables
    tab = rng_tables.fill_words(4, 4096, 0, 0, rng_tables.JITTER_STREAM, threads=2)
    for r in range(4):
        assert np.array_equal(tab[r], py_jitter_words(r, 4096))
    d = np.random.RandomState(2).random_sample(2048)
    w = tab[2].astype(np.uint64)
    assert np.array_equal(d, ((w[0::2] >> 5) * 67108864.0 + (w[1::2] >> 6)) / 9007199254740992.0)


def test_large_seeds_and_threads_agree(salp_lib):
    from underwater_swimmer_rl_b200 import rng_tables
    a = rng_tables.fill_words(64, 64, 4_000_000_000, 17, rng_tables.FOOD_STREAM, threads=1)
    b = rng_tables.fill_words(64, 64, 4_000_000_000, 17, rng_tables.FOOD_STREAM, threads=8)
    assert np.array_equal(a, b)
    assert np.array_equal(a[5], py_food_words(4_000_000_000 + 17 + 5, 64))


def test_oracle_fill_words_matches_stdlib_numpy_and_product(salp_lib, oracle_lib):
    """oracle.salp_oracle.fill_words (used by bench.py's reference arm, which must not load the product library) ==
    CPython random / numpy legacy streams == the product's salp_fill_rng_words."""
    import numpy as np
    from helpers import py_food_words, py_jitter_words
    from underwater_swimmer_rl_b200 import rng_tables
    for kind, ref in ((0, py_food_words), (1, py_jitter_words)):
        t = oracle_lib.fill_words(5, 700, 10_000, 3, kind)   # 700 > 624: crosses a twist
        for r in range(5):
            assert (t[r] == ref(10_003 + r, 700)).all()
        assert (t == rng_tables.fill_words(5, 1024, 10_000, 3, kind)[:, :700]).all()
