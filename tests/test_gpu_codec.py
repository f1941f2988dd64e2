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
    bits = -(torch.log2(out["likelihoods"]["y"]).flatten(1).sum(1) + torch.log2(out["likelihoods"]["z"]).flatten(1).sum(1))
    assert [int(torch.ceil(b / 8)) for b in bits.cpu()] == res["bytes"]
    # through the .bin a decode-only run would pick up ("{bitstream_name}-{file_prefix}*.bin")
    res2 = codec.encode(inp, str(tmp_path), "feat", "img0")
    assert os.path.basename(res2["bitstream"]) == "feat-img0.bin" and os.path.getsize(res2["bitstream"]) > 0
    dec2 = codec.decode(res2["bitstream"], str(tmp_path), "img0", org_img_size={"height": 1, "width": 2})
    assert dec2["org_input_size"] == {"height": 1, "width": 2}
    for k in vm.split_layer_list:
        assert torch.equal(dec2["data"][k], dec["data"][k])


def test_codec_refuses_cpu_device():
    codec_m = importlib.import_module(PKG + ".codec")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        codec_m.LMSFCCodec(eval_encode="bpp", encoder_config={}, device="cpu", vision_model=None, dataset=None)
