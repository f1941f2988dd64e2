"""Pins the oracle's range-ANS restatement (oracle/rans.c) against strings produced by the reference's own
compiled coder (tests/golden/rans.npz, made by tests/golden/make_golden_rans.py from
cpp_exts/rans/rans_interface.cpp) -- bit-exact -- and, where oracle/_ref/ans*.so is present, live."""
import os

import numpy as np
import pytest

from oracle import cdf as OC
from oracle import rans as OR

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "rans.npz"))
CASES = sorted({k.split(".")[0] for k in G.files})


def _case(name):
    return tuple(G[f"{name}.{k}"] for k in ("sym", "idx", "cdfs", "sizes", "offsets"))


@pytest.mark.parametrize("name", CASES)
def test_encode_matches_reference_string(name):
    sym, idx, cdfs, sizes, offsets = _case(name)
    assert OR.encode(sym, idx, cdfs, sizes, offsets) == G[f"{name}.string"].tobytes()


@pytest.mark.parametrize("name", CASES)
def test_decode_reference_string(name):
    sym, idx, cdfs, sizes, offsets = _case(name)
    assert np.array_equal(OR.decode(G[f"{name}.string"].tobytes(), idx, cdfs, sizes, offsets), sym)


@pytest.mark.parametrize("name", CASES)
def test_lanes_are_reference_strings_of_strided_subsets(name):
    sym, idx, cdfs, sizes, offsets = _case(name)
    for K in (3, 32):
        if f"{name}.lane0of{K}" not in G.files:
            continue
        lanes = OR.encode_lanes(sym, idx, cdfs, sizes, offsets, K)
        assert lanes[0] == G[f"{name}.lane0of{K}"].tobytes()
        assert lanes[K - 1] == G[f"{name}.lane{K - 1}of{K}"].tobytes()
        assert np.array_equal(OR.decode_lanes(lanes, idx, cdfs, sizes, offsets), sym)


def test_short_and_empty_strings_round_trip():
    # 0-2 symbols: the reference itself corrupts its heap here (see make_golden_rans.py), oracle only
    table = OC.get_scale_table().numpy()
    cdfs, offsets, sizes = OC.gc_update(table)
    for sym in ([], [0], [3, -2], [100000, -100000]):
        idx = np.full(len(sym), 5, dtype=np.int32)
        s = OR.encode(np.asarray(sym, np.int32), idx, cdfs, sizes, offsets)
        assert len(s) % 4 == 0 and len(s) >= 8
        assert np.array_equal(OR.decode(s, idx, cdfs, sizes, offsets), np.asarray(sym, np.int32))


def test_bypass_digits_cover_int_range():
    table = OC.get_scale_table().numpy()
    cdfs, offsets, sizes = OC.gc_update(table)
    sym = np.array([2**30, -2**30, 65535, -65536, 17, -17, 0] * 3, dtype=np.int32)
    idx = np.arange(sym.size, dtype=np.int32) % 64
    s = OR.encode(sym, idx, cdfs, sizes, offsets)
    assert np.array_equal(OR.decode(s, idx, cdfs, sizes, offsets), sym)


@pytest.mark.skipif(not OR.ref_available(), reason="oracle/_ref/ans*.so not built (reference absent)")
def test_live_against_compiled_reference():
    rng = np.random.default_rng(7)
    table = OC.get_scale_table().numpy()
    cdfs, offsets, sizes = OC.gc_update(table)
    for n in (16, 257, 4096):
        idx = rng.integers(0, 64, size=n).astype(np.int32)
        sym = np.round(rng.normal(size=n) * table[idx] * 1.5).astype(np.int32)
        s = OR.ref_encode(sym, idx, cdfs, sizes, offsets)
        assert OR.encode(sym, idx, cdfs, sizes, offsets) == s
        assert np.array_equal(OR.ref_decode(s, idx, cdfs, sizes, offsets), sym)
        assert np.array_equal(OR.decode(s, idx, cdfs, sizes, offsets), sym)
