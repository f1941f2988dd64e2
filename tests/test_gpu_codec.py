"""The `lmsfc` codec through the compressai_vision duck-type: a stand-in for
ImageSplitInference.__call__ (image_split_inference.py:93-175) drives encode -> decode, in memory and through
the `.bin` that decode-only runs glob for; the reconstruction must equal the model's own eval-mode forward."""
import importlib
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
PKG = "l-msfc-v2-training_b200"


class _FakeVisionModel:
    split_layer_list = ["p2", "p3", "p4", "p5"]
    device = "cuda"


def _features(B, H2, W2, device):
    from oracle import lmsfc as OL
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    f = OL.synth_features(spec, B, H2, W2, device=device)
    return {k: v.to(torch.bfloat16).float() for k, v in f.items()}


@pytest.mark.parametrize("B,H2,W2,where", [(1, 26, 38, "cuda"), (3, 16, 16, "cpu")])
def test_encode_decode_round_trip(tmp_path, B, H2, W2, where):
    codec_m = importlib.import_module(PKG + ".codec")
    torch.manual_seed(0)
    vm = _FakeVisionModel()
    reg = getattr(codec_m, "CODECS", None)
    cls = reg["lmsfc"] if reg is not None else codec_m.LMSFCCodec
    codec = cls(eval_encode="bpp", encoder_config={"qp": 3}, device="cuda", vision_model=vm, dataset=None)
    assert codec.qp_value == 3 and codec.eval_encode_type == "bpp"
    feats = _features(B, H2, W2, where)            # the video pipeline hands CPU tensors, the image pipeline device ones
    inp = {"data": dict(feats), "input_size": [(4 * H2, 4 * W2)], "org_input_size": {"height": 4 * H2, "width": 4 * W2}}
    res = codec.encode(inp, "", "", "")
    assert len(res["bytes"]) == B and all(b > 0 for b in res["bytes"])
    dec = codec.decode(res["bitstream"], "", "")
    assert list(dec["data"].keys()) == vm.split_layer_list
    # reference: the model's eval forward (dequantize mode)
    with torch.no_grad():
        out = codec.model({k: v.cuda() for k, v in feats.items()})
    for k in vm.split_layer_list:
        assert dec["data"][k].shape == feats[k].shape
        assert torch.equal(dec["data"][k], out["x_hat"][k].float()), k
    # `bytes` are real string lengths; the strings decode, with the ORACLE's coder, to the symbols of the eval forward
    import numpy as np
    from oracle import rans as OR
    m = codec.model
    st = res["bitstream"]
    assert res["bytes"] == [len(a) + len(b) for a, b in zip(st["y_strings"], st["z_strings"])]
    eb, gc = m.entropy_bottleneck, m.gaussian_conditional
    med = eb._get_medians().detach().reshape(1, -1, 1, 1)
    z_sym = torch.round(out["latents"]["z_hat"] - med).int().cpu().numpy()
    z_idx = np.broadcast_to(np.arange(z_sym.shape[1], dtype=np.int32).reshape(1, -1, 1, 1), z_sym.shape)
    tabs = [t.cpu().numpy() for t in (eb._quantized_cdf, eb._cdf_length, eb._offset)]
    assert np.array_equal(OR.decompress(st["z_strings"], z_idx, *tabs), z_sym)
    y_sym = torch.round(out["latents"]["y_hat"]).int().cpu().numpy()
    sc = torch.clamp(out["latents"]["scales"].float().cpu(), min=0.11)
    y_idx = torch.full(sc.shape, len(gc.scale_table) - 1).int()
    for t in gc.scale_table.cpu()[:-1]:
        y_idx -= (sc <= t).int()
    tabs = [t.cpu().numpy() for t in (gc._quantized_cdf, gc._cdf_length, gc._offset)]
    assert st["y_strings"] == OR.compress(y_sym, y_idx.numpy(), *tabs)
    assert np.array_equal(OR.decompress(st["y_strings"], y_idx.numpy(), *tabs), y_sym)
    # through the .bin a decode-only run would pick up ("{bitstream_name}-{file_prefix}*.bin")
    res2 = codec.encode(inp, str(tmp_path), "feat", "img0")
    assert os.path.basename(res2["bitstream"]) == "feat-img0.bin" and os.path.getsize(res2["bitstream"]) > 0
    dec2 = codec.decode(res2["bitstream"], str(tmp_path), "img0", org_img_size={"height": 1, "width": 2})
    assert dec2["org_input_size"] == {"height": 1, "width": 2}
    for k in vm.split_layer_list:
        assert torch.equal(dec2["data"][k], dec["data"][k])


@pytest.mark.parametrize("lanes", [16, "auto"])
def test_interleaved_lanes_round_trip(tmp_path, lanes):
    codec_m = importlib.import_module(PKG + ".codec")
    torch.manual_seed(0)
    vm = _FakeVisionModel()
    codec = codec_m.LMSFCCodec(eval_encode="bitrate", encoder_config={"qp": 1, "coder_lanes": lanes}, device="cuda",
                               vision_model=vm, dataset=None)
    H2, W2 = (20, 28) if lanes == 16 else (72, 104)      # auto: y has 256 x 5 x 7 symbols -> 8 lanes, z 256 x 2 x 2 -> 1
    feats = _features(2, H2, W2, "cuda")
    inp = {"data": dict(feats), "input_size": [(4 * H2, 4 * W2)], "org_input_size": {"height": 4 * H2, "width": 4 * W2}}
    res = codec.encode(inp, str(tmp_path), "feat", "v")
    dec = codec.decode(res["bitstream"], str(tmp_path), "v")
    with torch.no_grad():
        out = codec.model(feats)
    for k in vm.split_layer_list:
        assert torch.equal(dec["data"][k], out["x_hat"][k].float()), k


def test_frames_encoded_together_decode_one_at_a_time():
    """h_s(z_hat) picks the CDF rows through hard thresholds; its bits must not depend on the batch it was
    computed in (tile width / split-K choices follow the tile count).  Encode 4 frames in one call, decode each alone."""
    model_m = importlib.import_module(PKG + ".model")
    torch.manual_seed(0)
    net = model_m.LMSFC().cuda().eval()
    net.update(force=True)
    feats = _features(4, 40, 56, "cuda")
    with torch.no_grad():
        enc = net.compress(feats)
        whole = net.decompress(enc["strings"], enc["shape"], enc["y_hw"], enc["mid_hw"], enc["level_hw"])
        for i in range(4):
            one = net.decompress([[enc["strings"][0][i]], [enc["strings"][1][i]]], enc["shape"], enc["y_hw"], enc["mid_hw"],
                                 enc["level_hw"])
            for k in net.spec.levels:
                a, b = one["x_hat"][k].float(), whole["x_hat"][k][i:i + 1].float()
                assert float((a - b).abs().max()) <= 2e-2 * float(b.abs().max()), k      # g_s itself may tile differently


def test_mean_scale_hyperprior_round_trip(tmp_path):
    """encoder_config {"hyperprior": "mean-scale"} (models/google.py:337-405): symbols are round(y - means), the decoder
    adds the means back; decode(encode(x)) equals the eval forward bit for bit, the y strings are the ones the oracle's
    coder makes from those symbols, and the oracle graph run on the same weights reconstructs the same pyramid."""
    import numpy as np
    from oracle import lmsfc as OL
    from oracle import rans as OR
    codec_m = importlib.import_module(PKG + ".codec")
    torch.manual_seed(0)
    vm = _FakeVisionModel()
    codec = codec_m.LMSFCCodec(eval_encode="bpp", encoder_config={"qp": 3, "hyperprior": "mean-scale", "coder_index": 256},
                               device="cuda", vision_model=vm, dataset=None)
    m = codec.model
    assert m.spec.mean_scale and m.h_s["4"].out_channels == 2 * m.spec.M
    with torch.no_grad():                     # random-init means are ~0: move the last layer so that they matter
        m.h_s["4"].bias[m.spec.M:].uniform_(-3.0, 3.0)
    m.refresh_shadows()
    feats = _features(2, 26, 38, "cuda")
    inp = {"data": dict(feats), "input_size": [(104, 152)], "org_input_size": {"height": 104, "width": 152}}
    res = codec.encode(inp, str(tmp_path), "feat", "ms")
    dec = codec.decode(res["bitstream"], str(tmp_path), "ms")
    with torch.no_grad():
        out = m(feats)
    means = out["latents"]["means"]
    assert means is not None and float(means.abs().max()) > 1.0
    for k in vm.split_layer_list:
        assert torch.equal(dec["data"][k], out["x_hat"][k].float()), k
    gc = m.gaussian_conditional
    y_sym = torch.round(out["latents"]["y_hat"] - means).int().cpu().numpy()
    sc = torch.clamp(out["latents"]["scales"].float().cpu(), min=0.11)
    y_idx = torch.full(sc.shape, len(gc.scale_table) - 1).int()
    for t in gc.scale_table.cpu()[:-1]:
        y_idx -= (sc <= t).int()
    tabs = [t.cpu().numpy() for t in (gc._quantized_cdf, gc._cdf_length, gc._offset)]
    with torch.no_grad():
        enc = m.compress(feats)
    assert np.array_equal(OR.decompress(enc["strings"][0], y_idx.numpy(), *tabs), y_sym)
    # the oracle graph (reference modules, fp32) on the same weights, eval mode: same pyramid within bf16 tolerance
    ref = OL.OracleLMSFC(m.spec).cuda().eval()
    ref.load_state_dict({k: v for k, v in m.state_dict().items() if k in ref.state_dict()}, strict=False)
    with torch.no_grad():
        o = ref(feats, training=False)
    for k in vm.split_layer_list:
        a, b = out["x_hat"][k].float(), o["x_hat"][k]
        assert float((a - b).norm() / b.norm()) < 5e-2, k          # single latents may round to the neighbouring integer


class _FakeJDE:
    split_layer_list = [36, 61, 74]          # cfgs/vision_model/default.yaml:33
    device = "cuda"


def test_jde_three_scale_video_encode_in_frame_chunks(tmp_path):
    """BASELINE configs[4]: the JDE split hands three tensors of 256 / 512 / 1024 channels, and the video pipeline gives
    ONE encode call all frames of the sequence concatenated on N, on the CPU (video_split_inference.py:150-159,251-263).
    One .bin, per-frame byte counts, reconstruction == the model's eval forward frame by frame."""
    codec_m = importlib.import_module(PKG + ".codec")
    spec_m = importlib.import_module(PKG + ".spec")
    from oracle import lmsfc as OL
    torch.manual_seed(0)
    codec = codec_m.LMSFCCodec(eval_encode="bitrate", encoder_config={"qp": 2, "frame_chunk": 2, "coder_lanes": "auto"},
                               device="cuda", vision_model=_FakeJDE(), dataset=None)
    assert codec.spec == spec_m.JDE_SPEC
    T, H, W = 5, 20, 28
    f = OL.synth_features(spec_m.JDE_SPEC, T, H, W)
    feats = {k: v.to(torch.bfloat16).float() for k, v in f.items()}            # CPU tensors
    assert [tuple(v.shape[1:]) for v in feats.values()] == [(256, 20, 28), (512, 10, 14), (1024, 5, 7)]
    inp = {"data": dict(zip(_FakeJDE.split_layer_list, feats.values())), "input_size": [(8 * H, 8 * W)],
           "org_input_size": {"height": 8 * H, "width": 8 * W}}
    res = codec.encode(inp, str(tmp_path), "seq", "TVD-01")
    assert len(res["bytes"]) == T and all(b > 0 for b in res["bytes"])
    dec = codec.decode(res["bitstream"], str(tmp_path), "TVD-01")
    assert list(dec["data"].keys()) == ["36", "61", "74"] and dec["input_size"] == [(8 * H, 8 * W)]
    assert dec["org_input_size"] == {"height": 8 * H, "width": 8 * W}
    with torch.no_grad():
        out = codec.model({k: v.cuda() for k, v in feats.items()})
    for key, lv in zip(["36", "61", "74"], spec_m.JDE_SPEC.levels):
        a, b = dec["data"][key], out["x_hat"][lv].float()
        assert a.shape == feats[lv].shape
        assert float((a - b).abs().max()) <= 2e-2 * float(b.abs().max()), lv      # g_s may tile chunks of 2 differently from 5


def test_codec_refuses_cpu_device():
    codec_m = importlib.import_module(PKG + ".codec")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        codec_m.LMSFCCodec(eval_encode="bpp", encoder_config={}, device="cpu", vision_model=None, dataset=None)
