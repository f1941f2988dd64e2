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


def test_oracle_model_compress_decompress_round_trip():
    """ScaleHyperprior.compress / decompress (models/google.py:313-333) restated on the oracle graph: the decoded
    reconstruction equals the eval-mode forward, and `bytes` are close to the likelihood estimate of the rate."""
    import importlib
    import torch
    from oracle import lmsfc as OL
    spec = importlib.import_module("l-msfc-v2-training_b200.spec").DEFAULT_SPEC
    torch.manual_seed(3)
    net = OL.OracleLMSFC(spec).eval()
    net.update()
    feats = OL.synth_features(spec, 2, 26, 38, seed=5)              # ragged pyramid: 26x38 -> 13x19 -> 7x10 -> 4x5
    out = net.compress(feats)
    assert len(out["strings"][0]) == 2 and len(out["strings"][1]) == 2 and out["y_hw"] == (2, 3) and out["shape"] == (1, 1)
    rec = net.decompress(out["strings"], out["shape"], out["y_hw"], out["mid_hw"], out["level_hw"])
    with torch.no_grad():
        fwd = net(feats, training=False)
    for k in spec.levels:
        assert rec["x_hat"][k].shape == feats[k].shape
        assert torch.equal(rec["x_hat"][k], fwd["x_hat"][k]), k
    # z strings decode to the forward's z_hat; rate: z is inside its tables at init, so bytes ~ -sum log2 L / 8 + flush
    bits_z = float(-torch.log2(fwd["likelihoods"]["z"]).sum())
    nbytes_z = sum(len(s) for s in out["strings"][1])
    assert bits_z / 8 <= nbytes_z <= bits_z / 8 * 1.02 + 2 * 8 + 8
