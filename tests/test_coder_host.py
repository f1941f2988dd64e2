"""Host-side logic of the coder containers (no GPU): the per-item lane container of coder.py and the codec's .bin v2."""
import importlib
import struct

import numpy as np
import pytest

PKG = "l-msfc-v2-training_b200"


def test_item_container_parses_and_plain_strings_pass_through():
    c = importlib.import_module(PKG + ".coder")
    lanes = [np.arange(3, dtype=np.uint32), np.arange(2, dtype=np.uint32) + 7, np.arange(4, dtype=np.uint32) + 50]
    s = struct.pack("<2I", c.LANE_MAGIC, 3) + np.array([3, 2, 4], np.uint32).tobytes() + b"".join(x.tobytes() for x in lanes)
    k, counts, words = c._split_item(s)
    assert k == 3 and counts.tolist() == [3, 2, 4] and words.tolist() == np.concatenate(lanes).tolist()
    plain = np.array([1, 2, 3, 4, 5], np.uint32).tobytes()           # a reference string: no header
    k, counts, words = c._split_item(plain)
    assert k == 1 and counts.tolist() == [5] and words.size == 5
    # a header whose counts do not add up is not a container: taken as a plain string
    bad = struct.pack("<2I", c.LANE_MAGIC, 3) + np.array([3, 2, 9], np.uint32).tobytes() + b"\0" * 36
    assert c._split_item(bad)[0] == 1
    with pytest.raises(ValueError):
        c._split_item(b"abc")                                          # not a whole number of words
    with pytest.raises(ValueError):
        c._split_item(b"abcd")                                         # shorter than the coder's final state


def test_precision_other_than_16_is_refused():
    import torch
    c = importlib.import_module(PKG + ".coder")
    z = torch.zeros(1, 4, dtype=torch.int32)
    with pytest.raises(ValueError):
        c.encode_batch(z, z, z, z, z, precision=12)
    with pytest.raises(ValueError):
        c.decode_batch([b"\0" * 8], z, z, z, z, precision=12)


def test_host_tensors_are_refused_not_coded_on_the_cpu():
    import torch
    c = importlib.import_module(PKG + ".coder")
    z = torch.zeros(1, 4, dtype=torch.int32)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        c.encode_batch(z, z, z, z, z)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        c.decode_batch([b"\0" * 8], z, z, z, z)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        c.gc_build_indexes(torch.ones(1, 4), torch.ones(3), 0.11)


def test_bin_v2_round_trip_and_bad_magic(tmp_path):
    codec = importlib.import_module(PKG + ".codec").LMSFCCodec
    s = {"y_strings": [b"\x01" * 12, b"\x02" * 20], "z_strings": [b"\x03" * 8, b"\x04" * 8], "shape": (2, 3), "y_hw": (7, 10),
         "mid_hw": (4, 5), "keys": ["p2", "p3", "p4", "p5"], "level_hw": [(50, 75), (25, 38), (13, 19), (7, 10)]}
    path = tmp_path / "feat-img0.bin"
    codec._write_bin(str(path), s)
    r = codec._read_bin(str(path))
    assert r["y_strings"] == s["y_strings"] and r["z_strings"] == s["z_strings"] and r["keys"] == s["keys"]
    assert tuple(r["shape"]) == (2, 3) and tuple(r["y_hw"]) == (7, 10) and tuple(r["mid_hw"]) == (4, 5)
    assert [tuple(h) for h in r["level_hw"]] == s["level_hw"]
    assert path.stat().st_size == 5 + 4 + 24 + 1 + 4 * (1 + 2 + 8) + 2 * 8 + 12 + 20 + 8 + 8
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"LMSF\x01" + b"\0" * 64)                          # the v1 (raw symbol) container is not read as v2
    with pytest.raises(ValueError):
        codec._read_bin(str(bad))
