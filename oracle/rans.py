"""Oracle (test infrastructure, see oracle/__init__.py): range-ANS entropy coding.

ctypes front of oracle/rans.c, the C restatement of
cpp_exts/rans/rans_interface.cpp:108-288 + third_party/ryg_rans/rans64.h, and of the callers
entropy_models.py:235-268 (compress: one string per batch item over the flattened item) and
:270-325 (decompress).  `ref_encode` / `ref_decode` call the reference's own compiled `ans`
module (oracle/_ref/ans*.so, built by oracle/Makefile from /root/reference sources) when present.

The product's interleaved container (K lanes per item, lane k = the reference string of
symbols[k::K]) is restated by `encode_lanes` / `decode_lanes` on top of the same routine.
"""
import ctypes
import glob
import importlib.util
import os

import numpy as np

from . import cdf as _cdf

_HERE = os.path.dirname(os.path.abspath(__file__))
_L = None


def _lib():
    global _L
    if _L is None:
        lib = _cdf._lib()
        P, L, I = ctypes.c_void_p, ctypes.c_long, ctypes.c_int
        lib.lmsfc_oracle_rans_encode.argtypes = [P, P, L, L, L, P, I, P, P, P, L]
        lib.lmsfc_oracle_rans_encode.restype = L
        lib.lmsfc_oracle_rans_decode.argtypes = [P, P, L, L, L, P, I, P, P, P]
        lib.lmsfc_oracle_rans_decode.restype = None
        _L = lib
    return _L


def _i32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int32))


def encode_lane(symbols, indexes, cdfs, sizes, offsets, first=0, step=1, count=None) -> bytes:
    """RansEncoder.encode_with_indexes (rans_interface.cpp:209-220) on symbols[first::step][:count]."""
    s, ix, c, sz, of = _i32(symbols).ravel(), _i32(indexes).ravel(), _i32(cdfs), _i32(sizes), _i32(offsets)
    if count is None:
        count = (s.size - first + step - 1) // step if s.size > first else 0
    cap = 2 * count + 32
    words = np.empty(cap, dtype=np.uint32)
    n = _lib().lmsfc_oracle_rans_encode(s.ctypes.data, ix.ctypes.data, first, step, count, c.ctypes.data,
                                        c.shape[1], sz.ctypes.data, of.ctypes.data, words.ctypes.data, cap)
    if n < 0:
        raise RuntimeError("oracle rans scratch too small")
    return words[cap - n:].tobytes()


def decode_lane(string: bytes, indexes, cdfs, sizes, offsets, out, first=0, step=1, count=None):
    """RansDecoder.decode_with_indexes (rans_interface.cpp:222-288) into out[first::step][:count]."""
    ix, c, sz, of = _i32(indexes).ravel(), _i32(cdfs), _i32(sizes), _i32(offsets)
    if count is None:
        count = (ix.size - first + step - 1) // step if ix.size > first else 0
    words = np.frombuffer(string, dtype=np.uint32)
    words = np.concatenate([words, np.zeros(4, np.uint32)])   # the reference reads past the end at most one word
    _lib().lmsfc_oracle_rans_decode(words.ctypes.data, ix.ctypes.data, first, step, count, c.ctypes.data,
                                    c.shape[1], sz.ctypes.data, of.ctypes.data, out.ctypes.data)


def encode(symbols, indexes, cdfs, sizes, offsets) -> bytes:
    return encode_lane(symbols, indexes, cdfs, sizes, offsets)


def decode(string, indexes, cdfs, sizes, offsets):
    out = np.zeros(_i32(indexes).size, dtype=np.int32)
    decode_lane(string, indexes, cdfs, sizes, offsets, out)
    return out


def encode_lanes(symbols, indexes, cdfs, sizes, offsets, lanes: int):
    """lane k of K codes symbols[k::K] as one reference string."""
    return [encode_lane(symbols, indexes, cdfs, sizes, offsets, k, lanes) for k in range(lanes)]


def decode_lanes(strings, indexes, cdfs, sizes, offsets):
    out = np.zeros(_i32(indexes).size, dtype=np.int32)
    for k, s in enumerate(strings):
        decode_lane(s, indexes, cdfs, sizes, offsets, out, k, len(strings))
    return out


def compress(symbols, indexes, cdfs, sizes, offsets):
    """entropy_models.py:256-268: symbols / indexes [B, ...] -> list of B strings."""
    s, ix = _i32(symbols), _i32(indexes)
    return [encode(s[i], ix[i], cdfs, sizes, offsets) for i in range(s.shape[0])]


def decompress(strings, indexes, cdfs, sizes, offsets):
    """entropy_models.py:309-323 -> int32 array shaped like indexes."""
    ix = _i32(indexes)
    out = np.zeros(ix.shape, dtype=np.int32)
    for i, s in enumerate(strings):
        out[i] = decode(s, ix[i], cdfs, sizes, offsets).reshape(ix[i].shape)
    return out


# ------------------------------------------------------------------ the reference's own compiled coder
_REF = None


def ref_available():
    return bool(glob.glob(os.path.join(_HERE, "_ref", "ans*.so")))


def _ref():
    global _REF
    if _REF is None:
        path = glob.glob(os.path.join(_HERE, "_ref", "ans*.so"))[0]
        spec = importlib.util.spec_from_file_location("ans", path)
        _REF = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_REF)
    return _REF


def ref_encode(symbols, indexes, cdfs, sizes, offsets) -> bytes:
    return _ref().RansEncoder().encode_with_indexes(
        _i32(symbols).ravel().tolist(), _i32(indexes).ravel().tolist(), _i32(cdfs).tolist(),
        _i32(sizes).tolist(), _i32(offsets).tolist())


def ref_decode(string, indexes, cdfs, sizes, offsets):
    return np.asarray(_ref().RansDecoder().decode_with_indexes(
        string, _i32(indexes).ravel().tolist(), _i32(cdfs).tolist(), _i32(sizes).tolist(),
        _i32(offsets).tolist()), dtype=np.int32)
