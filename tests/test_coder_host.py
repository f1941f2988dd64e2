"""Host-side logic of the coder containers (no GPU): the per-item lane container of coder.py and the codec's .bin v3 (v2 still read)."""
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


def test_bin_v3_round_trip_v2_still_read_and_bad_magic(tmp_path):
    import struct
    codec = importlib.import_module(PKG + ".codec").LMSFCCodec
    s = {"y_strings": [b"\x01" * 12, b"\x02" * 20], "z_strings": [b"\x03" * 8, b"\x04" * 8], "shape": (2, 3), "y_hw": (7, 10),
         "mid_hw": (4, 5), "keys": ["p2", "p3", "p4", "p5"], "level_hw": [(50, 75), (25, 38), (13, 19), (7, 10)],
         "input_size": [(200, 300)], "org_input_size": {"height": 199, "width": 298}}
    path = tmp_path / "feat-img0.bin"
    codec._write_bin(str(path), s)
    r = codec._read_bin(str(path))
    assert r["y_strings"] == s["y_strings"] and r["z_strings"] == s["z_strings"] and r["keys"] == s["keys"]
    assert tuple(r["shape"]) == (2, 3) and tuple(r["y_hw"]) == (7, 10) and tuple(r["mid_hw"]) == (4, 5)
    assert [tuple(h) for h in r["level_hw"]] == s["level_hw"]
    # v3 carries what a decode-only run needs (image_split_inference.py:140-145 otherwise back-fills with a warning)
    assert r["input_size"] == [(200, 300)] and r["org_input_size"] == {"height": 199, "width": 298}
    # a v2 file (round 1) is still readable
    v2 = tmp_path / "old.bin"
    with open(v2, "wb") as f:
        f.write(b"LMSF" + struct.pack("<B", 2) + struct.pack("<I", 2) + struct.pack("<6I", 2, 3, 7, 10, 4, 5) + struct.pack("<B", 4))
        for k, hw in zip(s["keys"], s["level_hw"]):
            f.write(struct.pack("<B", len(k)) + k.encode() + struct.pack("<2I", *hw))
        for ys, zs in zip(s["y_strings"], s["z_strings"]):
            f.write(struct.pack("<2I", len(ys), len(zs)) + ys + zs)
    r2 = codec._read_bin(str(v2))
    assert r2["y_strings"] == s["y_strings"] and [tuple(h) for h in r2["level_hw"]] == s["level_hw"]
    bad = tmp_path / "bad.bin"
    bad.write_bytes(b"LMSF\x01" + b"\0" * 64)                          # the v1 (raw symbol) container is not read
    with pytest.raises(ValueError):
        codec._read_bin(str(bad))


def test_decoder_index_trailer_is_found_only_when_it_fits_the_symbol_count():
    """coder._strip_index: the trailer `ncp x {state lo, state hi, words written}, ncp, S, "CKPT"` behind a reference string
    is accepted when every field agrees with n (ncp = ceil(n / S) - 1); anything else stays a plain string."""
    c = importlib.import_module(PKG + ".coder")
    L = importlib.import_module(PKG + "._lib")
    body = np.arange(100, dtype=np.uint32) + 1000
    n, S = 5000, 1024
    ncp = (n + S - 1) // S - 1
    assert ncp == 4 == L.lib.lmsfc_rans_checkpoints(n, 1, S)
    table = (np.arange(3 * ncp, dtype=np.uint32) + 7).reshape(ncp, 3)
    s = np.concatenate([body, table.reshape(-1), np.array([ncp, S, c.INDEX_MAGIC], np.uint32)])
    w, interval, tab = c._strip_index(s, n)
    assert interval == S and w.tolist() == body.tolist() and tab.tolist() == table.tolist()
    for other_n in (4096, 5121, 100):                       # a different symbol count: ncp does not match -> plain string
        w, interval, tab = c._strip_index(s, other_n)
        assert interval == 0 and tab is None and w.size == s.size
    bad = s.copy(); bad[-2] = 1000                           # interval not a power of two
    assert c._strip_index(bad, n)[2] is None
    bad = s.copy(); bad[-1] ^= 1                             # no magic
    assert c._strip_index(bad, n)[2] is None
    assert c._strip_index(s[-8:], n)[2] is None              # too short to hold the table it announces
    # checkpoint counts of the C ABI: the longest lane decides
    assert L.lib.lmsfc_rans_checkpoints(1024, 1, 1024) == 0 and L.lib.lmsfc_rans_checkpoints(1025, 1, 1024) == 1
    assert L.lib.lmsfc_rans_checkpoints(63232, 1, 1024) == 61 and L.lib.lmsfc_rans_checkpoints(63232, 64, 256) == 3
    assert L.lib.lmsfc_rans_checkpoints(0, 1, 1024) == 0
    import torch
    z = torch.zeros(1, 4, dtype=torch.int32)
    with pytest.raises(ValueError, match="lanes = 1"):
        c.encode_batch(z, z, z, z, z, lanes=4, index=1024)


def test_mean_scale_spec_and_codec_option():
    spec_m = importlib.import_module(PKG + ".spec")
    sp, ms = spec_m.DEFAULT_SPEC, spec_m.MEAN_SCALE_SPEC
    assert ms.mean_scale and not sp.mean_scale and ms.mid == 2 * ms.M
    by_name = {l[0]: l for l in ms.conv_layers(64, 64)}
    assert by_name["h_s.0"][3:5] == (ms.N, ms.M) and by_name["h_s.2"][3:5] == (ms.M, ms.mid) and by_name["h_s.4"][3:5] == (ms.mid, 2 * ms.M)
    assert ms.flops_fwd(64, 64) > sp.flops_fwd(64, 64)
    assert ms.gate_names() == sp.gate_names()
    with pytest.raises(ValueError):
        spec_m.LMSFCSpec(hyperprior="median")
    # the oracle builds the same graph from the same spec (reference modules: LeakyReLU, 2M-channel h_s, chunk)
    import torch
    from oracle import lmsfc as OL
    net = OL.OracleLMSFC(ms)
    assert isinstance(net.h_a[1], torch.nn.LeakyReLU) and net.h_s[4].out_channels == 2 * ms.M and net.h_s[2].out_channels == ms.mid
    feats = OL.synth_features(ms, 1, 16, 16)
    with torch.no_grad():
        out = net(feats, training=False)
    assert out["latents"]["means"].shape == out["latents"]["y"].shape == out["latents"]["scales"].shape
