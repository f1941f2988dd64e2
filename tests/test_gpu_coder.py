"""Device range coder (csrc/rans.cu through the C ABI) against the oracle (oracle/rans.c) and against strings made
by the reference's own compiled coder (tests/golden/rans.npz): BIT-EXACT bytes and symbols."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import cdf as OC
from oracle import rans as OR

pytestmark = pytest.mark.gpu
PKG = "l-msfc-v2-training_b200"
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "rans.npz"))
CASES = sorted({k.split(".")[0] for k in G.files})
DEV = "cuda"


def coder():
    return importlib.import_module(PKG + ".coder")


def _case(name):
    return tuple(G[f"{name}.{k}"] for k in ("sym", "idx", "cdfs", "sizes", "offsets"))


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _gc_tables():
    table = OC.get_scale_table().numpy()
    cdfs, offsets, sizes = OC.gc_update(table)
    return table, cdfs, sizes, offsets


def _latent(rng, table, shape, outliers=0.001):
    idx = np.minimum(rng.integers(0, 64, size=shape), rng.integers(0, 64, size=shape)).astype(np.int32)
    sym = np.round(rng.normal(size=shape) * table[idx]).astype(np.int32)
    hit = rng.random(shape) < outliers
    sym[hit] = rng.integers(-70000, 70000, size=int(hit.sum())).astype(np.int32)
    return sym, idx


@pytest.mark.parametrize("name", CASES)
def test_bytes_equal_the_reference_string(name):
    sym, idx, cdfs, sizes, offsets = _case(name)
    s = coder().encode_batch(_t(sym[None]), _t(idx[None]), _t(cdfs), _t(sizes), _t(offsets))
    assert s == [G[f"{name}.string"].tobytes()]


@pytest.mark.parametrize("name", CASES)
def test_decodes_the_reference_string(name):
    sym, idx, cdfs, sizes, offsets = _case(name)
    out = coder().decode_batch([G[f"{name}.string"].tobytes()], _t(idx[None]), _t(cdfs), _t(sizes), _t(offsets), lanes=1)
    assert np.array_equal(out.cpu().numpy()[0], sym)


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("K", [3, 32])
def test_lanes_are_reference_strings_of_strided_subsets(name, K):
    sym, idx, cdfs, sizes, offsets = _case(name)
    if f"{name}.lane0of{K}" not in G.files:
        pytest.skip("case shorter than the lane count")
    c = coder()
    packed, off = c.encode_device(_t(sym[None]), _t(idx[None]), _t(cdfs), _t(sizes), _t(offsets), lanes=K)
    words = packed.cpu().numpy().view(np.uint32)
    assert words[off[0]:off[1]].tobytes() == G[f"{name}.lane0of{K}"].tobytes()
    assert words[off[K - 1]:off[K]].tobytes() == G[f"{name}.lane{K - 1}of{K}"].tobytes()
    lanes = OR.encode_lanes(sym, idx, cdfs, sizes, offsets, K)
    assert [words[off[k]:off[k + 1]].tobytes() for k in range(K)] == lanes
    s = c.encode_batch(_t(sym[None]), _t(idx[None]), _t(cdfs), _t(sizes), _t(offsets), lanes=K)
    out = c.decode_batch(s, _t(idx[None]), _t(cdfs), _t(sizes), _t(offsets))
    assert np.array_equal(out.cpu().numpy()[0], sym)


@pytest.mark.parametrize("shape,K", [((16, 256, 4, 4), 1), ((16, 256, 4, 4), 16), ((1, 256, 13, 19), 1),
                                     ((1, 256, 13, 19), 64), ((3, 7, 5, 3), 4), ((2, 1, 1, 1), 1), ((2, 1, 1, 3), 8)])
def test_batches_at_config_sizes_match_the_oracle(shape, K):
    # cfg 2/3 latents (16x256x4x4), cfg 4 (1x256x13x19), ragged and tiny items, more lanes than symbols
    rng = np.random.default_rng(sum(shape) + K)
    table, cdfs, sizes, offsets = _gc_tables()
    sym, idx = _latent(rng, table, shape)
    c = coder()
    s = c.encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=K)
    assert len(s) == shape[0]
    for i in range(shape[0]):
        if K == 1:
            assert s[i] == OR.encode(sym[i], idx[i], cdfs, sizes, offsets)
        else:
            assert s[i][8 + 4 * K:] == b"".join(OR.encode_lanes(sym[i], idx[i], cdfs, sizes, offsets, K))
    out = c.decode_batch(s, _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1 if K == 1 else None)
    assert out.shape == tuple(shape) and np.array_equal(out.cpu().numpy(), sym)
    if K == 1:   # and the oracle decodes the device's strings
        assert np.array_equal(OR.decompress(s, idx, cdfs, sizes, offsets), sym)


def test_large_frame_round_trip_and_rate():
    # video pipeline: many frames at once, 8 x 256 x 13 x 19 symbols; size-independent checks
    rng = np.random.default_rng(99)
    table, cdfs, sizes, offsets = _gc_tables()
    sym, idx = _latent(rng, table, (8, 256, 13, 19), outliers=0.0)
    c = coder()
    for K in (1, 32):
        s = c.encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=K)
        out = c.decode_batch(s, _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1 if K == 1 else None)
        assert np.array_equal(out.cpu().numpy(), sym)
        # the strings are within 1 % + the lane overhead of the model's entropy
        c32 = cdfs.astype(np.int64)
        v = sym - offsets[idx]
        p = (c32[idx, v + 1] - c32[idx, v]) / 65536.0
        ideal = -np.log2(p).sum() / 8
        total = sum(len(x) for x in s)
        assert total <= 1.01 * ideal + 8 * (12 * K + 8), (total, ideal)


def test_index_outside_the_tables_raises():
    table, cdfs, sizes, offsets = _gc_tables()
    sym = np.zeros((1, 16), np.int32)
    idx = np.full((1, 16), 64, np.int32)
    with pytest.raises(ValueError):
        coder().encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets))


def test_build_indexes_matches_the_reference_formula():
    # entropy_models.py:684-689
    table = OC.get_scale_table()
    g = torch.Generator().manual_seed(3)
    scales = torch.exp(torch.rand(4, 32, 9, 7, generator=g) * 9 - 3)
    scales.view(-1)[:64] = table            # exactly on the thresholds
    scales.view(-1)[64] = 0.0
    want = torch.full(scales.shape, len(table) - 1).int()
    sc = torch.clamp(scales, min=0.11)
    for s in table[:-1]:
        want -= (sc <= s).int()
    got = coder().gc_build_indexes(scales.to(DEV), table, 0.11)
    assert torch.equal(got.cpu(), want)


def test_modules_compress_like_the_reference():
    # EntropyBottleneck.compress / decompress and GaussianConditional.compress / decompress
    # (entropy_models.py:518-548, 235-325) on top of the device coder
    em = importlib.import_module(PKG + ".entropy_models")
    torch.manual_seed(0)
    eb = em.EntropyBottleneck(32).to(DEV)
    eb.update(force=True)
    x = (torch.randn(3, 32, 5, 4, device=DEV) * 3).round()
    strings = eb.compress(x)
    med = eb._get_medians().detach().reshape(1, -1, 1, 1)
    sym = torch.round(x - med).int().cpu().numpy()
    idx = np.broadcast_to(np.arange(32, dtype=np.int32).reshape(1, -1, 1, 1), sym.shape)
    tabs = [t.cpu().numpy() for t in (eb._quantized_cdf, eb._cdf_length, eb._offset)]
    assert strings == OR.compress(sym, idx, *tabs)
    back = eb.decompress(strings, x.shape[2:])
    assert torch.equal(back, torch.round(x - med) + med)

    gc = em.GaussianConditional(None).to(DEV)
    gc.update_scale_table(OC.get_scale_table().tolist())
    scales = torch.exp(torch.rand(2, 32, 6, 5, device=DEV) * 6 - 2)
    y = (torch.randn(2, 32, 6, 5, device=DEV) * scales)
    indexes = gc.build_indexes(scales)
    strings = gc.compress(y, indexes)
    tabs = [t.cpu().numpy() for t in (gc._quantized_cdf, gc._cdf_length, gc._offset)]
    assert strings == OR.compress(torch.round(y).int().cpu().numpy(), indexes.cpu().numpy(), *tabs)
    assert torch.equal(gc.decompress(strings, indexes), torch.round(y))


# ---------------------------------------------------------------------------- decoder index (coder.encode_batch(index=S))
@pytest.mark.parametrize("shape,S", [((1, 256, 13, 19), 1024), ((1, 256, 13, 19), 128), ((3, 256, 4, 4), 512),
                                     ((2, 5000), 4096), ((2, 4096), 1024), ((1, 129), 128), ((2, 100), 128)])
def test_indexed_strings_are_reference_strings_plus_a_trailer(shape, S):
    """The bytes in front of the index are the reference's string; the stock decoder (oracle restatement of
    rans_interface.cpp:215-288, which reads n symbols and never looks further) decodes the whole thing; the device decodes
    it segment-parallel to the same symbols.  Escapes (bypass digits) fall on and around checkpoints."""
    rng = np.random.default_rng(sum(shape) + S)
    table, cdfs, sizes, offsets = _gc_tables()
    sym, idx = _latent(rng, table, shape, outliers=0.01)
    flat = sym.reshape(shape[0], -1)
    flat[:, ::S] = 90000                        # an escape exactly on every checkpoint
    flat[:, S - 1::S] = -90000                  # and right in front of it
    c = coder()
    plain = c.encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets))
    s = c.encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets), index=S)
    n = flat.shape[1]
    ncp = (n + S - 1) // S - 1
    for a, b in zip(plain, s):
        assert b[:len(a)] == a
        assert len(b) == len(a) + (12 * ncp + 12 if ncp else 0)
    assert plain == [OR.encode(sym[i], idx[i], cdfs, sizes, offsets) for i in range(shape[0])]
    assert np.array_equal(OR.decompress(s, idx, cdfs, sizes, offsets), sym)          # stock decoder: trailer ignored
    out = c.decode_batch(s, _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1)
    assert np.array_equal(out.cpu().numpy(), sym)
    out = c.decode_batch(s, _t(idx), _t(cdfs), _t(sizes), _t(offsets))               # lane count read from the string: 1
    assert np.array_equal(out.cpu().numpy(), sym)


def test_corrupt_index_is_detected():
    rng = np.random.default_rng(5)
    table, cdfs, sizes, offsets = _gc_tables()
    sym, idx = _latent(rng, table, (1, 8192))
    c = coder()
    s = c.encode_batch(_t(sym), _t(idx), _t(cdfs), _t(sizes), _t(offsets), index=1024)[0]
    w = np.frombuffer(s, dtype=np.uint32).copy()
    ncp = int(w[-3])
    assert ncp == 7 and w[-1] == c.INDEX_MAGIC
    bad = w.copy()
    bad[w.size - 3 - 3 * ncp + 3 * 2] ^= 0x10          # state of checkpoint 3, low word
    with pytest.raises(ValueError, match="decoder index"):
        c.decode_batch([bad.tobytes()], _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1)
    bad = w.copy()
    bad[w.size - 3 - 3 * ncp + 3 * 4 + 2] += 1          # stream position of checkpoint 5
    with pytest.raises(ValueError, match="decoder index"):
        c.decode_batch([bad.tobytes()], _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1)
    bad = w.copy()
    bad[-3] = ncp + 1                                   # count disagrees with n: not an index, decoded as a plain string
    out = c.decode_batch([bad.tobytes()], _t(idx), _t(cdfs), _t(sizes), _t(offsets), lanes=1)
    assert np.array_equal(out.cpu().numpy(), sym)


def test_modules_carry_the_index():
    em = importlib.import_module(PKG + ".entropy_models")
    gc = em.GaussianConditional(None).to(DEV)
    gc.update_scale_table(OC.get_scale_table().tolist())
    torch.manual_seed(1)
    scales = torch.exp(torch.rand(2, 256, 13, 19, device=DEV) * 6 - 2)
    y = torch.randn(2, 256, 13, 19, device=DEV) * scales
    indexes = gc.build_indexes(scales)
    plain = gc.compress(y, indexes)
    gc.coder_index = 1024
    s = gc.compress(y, indexes)
    assert all(b[:len(a)] == a and len(b) > len(a) for a, b in zip(plain, s))
    assert torch.equal(gc.decompress(s, indexes), torch.round(y))
    gc.coder_index = 0
    assert torch.equal(gc.decompress(s, indexes), torch.round(y))       # decoding needs no setting: the string says it
