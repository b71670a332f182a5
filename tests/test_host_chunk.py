"""CPU tests of the host-side mirror of BOSS::Chunk (constructor.Chunk): bit packing in the sdsl
int_vector layout, serialize/load round trip against the files the reference wrote (golden
sha256), and extend() == concatenation (boss_chunk.cpp:230-270, :330-386)."""
import hashlib
import os

import numpy as np
import pytest

import mtg_b200
import oracle_lib
from golden_cases import CASES, case_reads
from test_oracle_golden import W_width, load_golden


def chunk_from_arrays(g, k, bpc):
    sigma = int(g["sigma"])
    ww = W_width(sigma)
    wts = oracle_lib.pack_bits(g["weights"], bpc) if bpc else None
    return mtg_b200.Chunk(sigma, k, len(g["W"]), oracle_lib.pack_bits(g["W"], ww),
                          oracle_lib.pack_bits(g["last"], 1), g["F"], ww, wts, bpc)


@pytest.mark.parametrize("name", ["cfg1_k12_basic", "cfg3_k31_count8", "cfg5_protein_k12",
                                  "k21_saturating_w2", "k33_canonical_w16"])
def test_serialize_matches_reference_files_and_loads_back(name, tmp_path):
    case, g = CASES[name], load_golden(name)
    bpc = case.get("bpc", 0)
    c = chunk_from_arrays(g, case["K"] - 1, bpc)
    base = str(tmp_path / "x")
    c.serialize(base)
    for ext in ("", ".W", ".last", ".weights"):
        h = hashlib.sha256(open(base + ".dbg.chunk" + ext, "rb").read()).hexdigest()
        assert h == g["sha256"][ext], ext
    d = mtg_b200.Chunk.load(base)
    assert d.n_rows == c.n_rows and d.k == c.k and d.alph_size == c.alph_size
    assert np.array_equal(d.W, g["W"]) and np.array_equal(d.last, g["last"])
    assert np.array_equal(d.F, g["F"])
    if bpc:
        assert np.array_equal(d.weights, g["weights"])
    else:
        assert d.weights is None


def test_extend_is_concatenation_without_the_placeholder():
    a = load_golden("cfg3_k31_count8")
    c1 = chunk_from_arrays(a, 30, 8)
    c2 = chunk_from_arrays(a, 30, 8)
    c1.extend(c2)
    assert c1.n_rows == 2 * len(a["W"]) - 1
    assert np.array_equal(c1.W, np.concatenate([a["W"], a["W"][1:]]))
    assert np.array_equal(c1.last, np.concatenate([a["last"], a["last"][1:]]))
    assert np.array_equal(c1.weights, np.concatenate([a["weights"], a["weights"][1:]]))
    assert np.array_equal(c1.F, 2 * a["F"])
    with pytest.raises(ValueError):
        c1.extend(chunk_from_arrays(load_golden("cfg1_k12_basic"), 11, 0))


def test_load_rejects_a_corrupted_chunk(tmp_path):
    g = load_golden("k4_hand_w8")
    c = chunk_from_arrays(g, 3, 8)
    base = str(tmp_path / "y")
    c.serialize(base)
    with open(base + ".dbg.chunk.last", "wb") as f:
        f.write(np.uint64(3).tobytes() + np.uint64(5).tobytes())
    with pytest.raises(ValueError):
        mtg_b200.Chunk.load(base)


def test_word_wise_extend_equals_entry_wise():
    """Chunk.extend appends whole words through funnel shifts; compare with packing the concatenated
    entries from scratch for every width the path uses and ragged row counts."""
    import numpy as np
    from mtg_b200.constructor import Chunk
    rng = np.random.default_rng(0)
    for width in (1, 4, 6, 8, 12, 16, 31, 32):
        for n, m in ((1, 1), (1, 5), (17, 130), (64, 64), (65, 1), (1000, 4097), (129, 2)):
            a = rng.integers(0, 1 << width, n, dtype=np.uint64)
            b = rng.integers(0, 1 << width, m + 3, dtype=np.uint64)
            first = 1
            got = Chunk._append_bits(Chunk._pack(a, width), n, width, Chunk._pack(b, width), first, m)
            exp = Chunk._pack(np.concatenate([a, b[first:first + m]]), width)
            assert np.array_equal(got[:len(exp)], exp) and not got[len(exp):].any(), (width, n, m)
    # dirty padding bits in dst's last word are ignored
    a = rng.integers(0, 16, 5, dtype=np.uint64)
    words = Chunk._pack(a, 4).copy()
    words[0] |= np.uint64(0xFF) << np.uint64(40)
    b = rng.integers(0, 16, 9, dtype=np.uint64)
    got = Chunk._append_bits(words, 5, 4, Chunk._pack(b, 4), 1, 8)
    assert np.array_equal(got, Chunk._pack(np.concatenate([a, b[1:]]), 4))
