"""`lmsfc` codec for the compressai_vision split-inference pipelines.

Registry entry: `@register_codec("lmsfc")` (compressai_vision/registry/registers.py:106-113); constructed by
`CODECS[conf.type](**yaml_kwargs, device=..., vision_model=..., dataset=...)`
(compressai_vision/config/config.py:140-150).  Duck-type served (compressai_vision/codecs/base.py:53-135,
pipelines/base.py:219-245, split_inference/image_split_inference.py:100-175):

    encode(input, codec_output_dir, bitstream_name, file_prefix, img_input=False) -> {"bytes": [...], "bitstream": ...}
    decode(bitstream, codec_output_dir, file_prefix, org_img_size=None, img_input=False) -> {"data": {...}, ...}
    qp_value, eval_encode_type

`encode` runs g_a / h_a and the quantiser on the B200 kernels and returns the integer symbols of (y, z) plus the
side information needed to decode; `decode` runs h_s / g_s.  Entropy CODING (rANS, entropy_models.py:235-325) is the
next scope row (§8f-1): until it lands the `.bin` written for decode-only runs carries the raw int16 symbols and
`bytes` is the entropy-model estimate ceil(-sum log2 L / 8) per frame, which is what the rate term of training optimises.
"""
import math
import os
import struct
from typing import Dict

import torch
import torch.nn as nn

from . import functional as Fn
from .model import LMSFC
from .spec import DEFAULT_SPEC, LAMBDAS

try:  # the real registry when compressai_vision is importable (it needs hydra / detectron2)
    from compressai_vision.registry import register_codec
except Exception:  # same decorator semantics, local table
    CODECS: Dict[str, type] = {}

    def register_codec(name: str):
        def decorator(cls):
            CODECS[name] = cls
            return cls
        return decorator

MAGIC = b"LMSF"


@register_codec("lmsfc")
class LMSFCCodec(nn.Module):
    """kwargs (the codec yaml, cfgs/codec/*.yaml style): `eval_encode` ("bpp" | "bitrate"), `encoder_config`
    {"qp": index into spec.LAMBDAS, "checkpoint": path or None}, `device`, `vision_model`, `dataset`."""

    def __init__(self, **kwargs):
        super().__init__()
        enc = dict(kwargs.get("encoder_config") or {})
        self.qp = enc.get("qp", 3)
        self.eval_encode = kwargs.get("eval_encode", "bpp")
        self.device_name = str(kwargs.get("device", "cuda"))
        if not self.device_name.startswith("cuda"):
            raise RuntimeError("the lmsfc codec runs on a CUDA device only (no CPU fallback)")
        self.spec = kwargs.get("spec", DEFAULT_SPEC)
        self.lmbda = LAMBDAS[int(self.qp) % len(LAMBDAS)] if isinstance(self.qp, int) else float(self.qp)
        self.model = LMSFC(self.spec).to(self.device_name).eval()
        ckpt = enc.get("checkpoint")
        if ckpt:
            sd = torch.load(ckpt, map_location=self.device_name)
            self.model.load_state_dict(sd.get("state_dict", sd))
        vm = kwargs.get("vision_model")
        self.layer_keys = list(getattr(vm, "split_layer_list", self.spec.levels))

    @property
    def qp_value(self):
        return self.qp

    @property
    def eval_encode_type(self):
        return self.eval_encode

    # ------------------------------------------------------------------ encode
    @torch.no_grad()
    def encode(self, input: Dict, codec_output_dir: str = "", bitstream_name: str = "", file_prefix: str = "",
               img_input=False) -> Dict:
        if img_input:
            raise NotImplementedError("lmsfc codes feature pyramids (split inference), not images")
        data = input["data"]
        keys = list(data.keys())
        if len(keys) != len(self.spec.levels):
            raise ValueError(f"expected {len(self.spec.levels)} pyramid levels, got {keys}")
        dev = self.device_name
        ps = tuple(Fn.nhwc(data[k].to(dev, non_blocking=True).float(), Fn.BF16) for k in keys)   # video pipeline hands CPU tensors
        m = self.model
        y, z, mid_hw = m.analysis(ps)
        z_hat, z_lik = m.entropy_bottleneck(z, training=False)
        scales = m.hyper_synthesis(getattr(z_hat, Fn._BF16_ATTR), mid_hw, y.shape[2:])
        y_hat, y_lik = m.gaussian_conditional(y, scales, training=False)
        med = m.entropy_bottleneck.quantiles[:, 0, 1].reshape(1, -1, 1, 1)
        y_sym = torch.round(y_hat).to(torch.int16)
        z_sym = torch.round(z_hat - med).to(torch.int16)
        bits = -(torch.log2(y_lik).flatten(1).sum(1) + torch.log2(z_lik).flatten(1).sum(1))
        nbytes = [int(math.ceil(float(b) / 8.0)) for b in bits.cpu()]
        stream = {"y": y_sym, "z": z_sym, "keys": keys, "level_hw": [tuple(p.shape[2:]) for p in ps],
                  "mid_hw": tuple(mid_hw), "input_size": input.get("input_size"), "org_input_size": input.get("org_input_size")}
        if codec_output_dir and bitstream_name:
            os.makedirs(codec_output_dir, exist_ok=True)
            path = os.path.join(codec_output_dir, f"{bitstream_name}-{file_prefix}.bin" if file_prefix else f"{bitstream_name}.bin")
            self._write_bin(path, stream)
            return {"bytes": nbytes, "bitstream": path}
        return {"bytes": nbytes, "bitstream": stream}

    # ------------------------------------------------------------------ decode
    @torch.no_grad()
    def decode(self, bitstream, codec_output_dir: str = "", file_prefix: str = "", org_img_size: Dict = None,
               img_input=False) -> Dict:
        if img_input:
            raise NotImplementedError("lmsfc codes feature pyramids (split inference), not images")
        stream = self._read_bin(bitstream) if isinstance(bitstream, (str, os.PathLike)) else bitstream
        dev = self.device_name
        m = self.model
        med = m.entropy_bottleneck.quantiles[:, 0, 1].reshape(1, -1, 1, 1)
        z_hat = stream["z"].to(dev).float() + med
        y_hat = stream["y"].to(dev).float()
        # h_s(z_hat) only conditions the entropy decoder of y (models/google.py:326-333); with the symbol container
        # there is nothing to decode, so reconstruction needs y_hat alone.  z travels for format stability.
        del z_hat
        x_hat = m.synthesis(Fn.nhwc(y_hat, Fn.BF16), stream["level_hw"])
        out = {"data": {k: t.float().contiguous() for k, t in zip(stream["keys"], x_hat)}}
        if stream.get("input_size") is not None:
            out["input_size"] = stream["input_size"]
        if org_img_size is not None or stream.get("org_input_size") is not None:
            out["org_input_size"] = org_img_size if org_img_size is not None else stream["org_input_size"]
        return out

    # ------------------------------------------------------------------ container (raw symbols until rANS lands)
    @staticmethod
    def _write_bin(path, s):
        with open(path, "wb") as f:
            y, z = s["y"].cpu().contiguous(), s["z"].cpu().contiguous()
            f.write(MAGIC + struct.pack("<B", 1))
            f.write(struct.pack("<4I", *y.shape) + struct.pack("<4I", *z.shape))
            f.write(struct.pack("<B", len(s["keys"])))
            for k, hw in zip(s["keys"], s["level_hw"]):
                kb = str(k).encode()
                f.write(struct.pack("<B", len(kb)) + kb + struct.pack("<2I", *hw))
            f.write(struct.pack("<2I", *s["mid_hw"]))
            f.write(y.numpy().tobytes() + z.numpy().tobytes())

    @staticmethod
    def _read_bin(path):
        import numpy as np
        with open(path, "rb") as f:
            if f.read(4) != MAGIC or struct.unpack("<B", f.read(1))[0] != 1:
                raise ValueError(f"{path}: not an lmsfc v1 bitstream")
            ys, zs = struct.unpack("<4I", f.read(16)), struct.unpack("<4I", f.read(16))
            keys, level_hw = [], []
            for _ in range(struct.unpack("<B", f.read(1))[0]):
                n = struct.unpack("<B", f.read(1))[0]
                keys.append(f.read(n).decode())
                level_hw.append(struct.unpack("<2I", f.read(8)))
            mid_hw = struct.unpack("<2I", f.read(8))
            ny, nz = int(np.prod(ys)), int(np.prod(zs))
            y = torch.from_numpy(np.frombuffer(f.read(2 * ny), dtype=np.int16).copy()).reshape(ys)
            z = torch.from_numpy(np.frombuffer(f.read(2 * nz), dtype=np.int16).copy()).reshape(zs)
        return {"y": y, "z": z, "keys": keys, "level_hw": level_hw, "mid_hw": mid_hw}
