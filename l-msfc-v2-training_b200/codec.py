"""`lmsfc` codec for the compressai_vision split-inference pipelines.

Registry entry: `@register_codec("lmsfc")` (compressai_vision/registry/registers.py:106-113); constructed by
`CODECS[conf.type](**yaml_kwargs, device=..., vision_model=..., dataset=...)`
(compressai_vision/config/config.py:140-150).  Duck-type served (compressai_vision/codecs/base.py:53-135,
pipelines/base.py:219-245, split_inference/image_split_inference.py:100-175):

    encode(input, codec_output_dir, bitstream_name, file_prefix, img_input=False) -> {"bytes": [...], "bitstream": ...}
    decode(bitstream, codec_output_dir, file_prefix, org_img_size=None, img_input=False) -> {"data": {...}, ...}
    qp_value, eval_encode_type

`encode` = ScaleHyperprior.compress (models/google.py:313-324) on the B200 kernels: g_a, h_a, the device range coder
for z, h_s on the decoded z_hat, build_indexes, the range coder for y; `decode` = decompress (:326-333) + g_s.  The
per-frame strings are the reference's rANS strings (coder.py; `coder_lanes` > 1 selects the interleaved container),
`bytes` is their real length, and the `.bin` written for decode-only runs holds a small header and the strings.
"""
import json
import os
import struct
from typing import Dict

import torch
import torch.nn as nn

from .model import LMSFC
from .spec import DEFAULT_SPEC, JDE_SPEC, LAMBDAS, LMSFCSpec

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
        vm = kwargs.get("vision_model")
        split = list(getattr(vm, "split_layer_list", []) or [])
        spec = kwargs.get("spec")
        if spec is None and enc.get("level_channels"):       # explicit pyramid: keys from the vision model, channels from the yaml
            ch = tuple(int(c) for c in enc["level_channels"])
            keys = tuple(str(k) for k in split) if len(split) == len(ch) else tuple(f"l{i}" for i in range(len(ch)))
            spec = LMSFCSpec(levels=keys, level_weights=(1.0,) * len(ch), level_channels=ch,
                             base_stride=int(enc.get("base_stride", 4)))
        if spec is None:                                     # by the number of split tensors: 3 = JDE (cfgs/vision_model/default.yaml:33)
            spec = JDE_SPEC if len(split) == 3 else DEFAULT_SPEC
        if enc.get("hyperprior"):                            # "scale" (default) | "mean-scale" (models/google.py:337-405)
            import dataclasses
            spec = dataclasses.replace(spec, hyperprior=str(enc["hyperprior"]), hyper_mid=int(enc.get("hyper_mid", 0)))
        self.spec = spec
        self.frame_chunk = int(enc.get("frame_chunk", 8))    # frames per device batch of a video-pipeline call
        self.lmbda = LAMBDAS[int(self.qp) % len(LAMBDAS)] if isinstance(self.qp, int) else float(self.qp)
        self.model = LMSFC(self.spec).to(self.device_name).eval()
        ckpt = enc.get("checkpoint")
        if ckpt:
            sd = torch.load(ckpt, map_location=self.device_name)
            self.model.load_state_dict(sd.get("state_dict", sd))
        lanes = enc.get("coder_lanes", 1)          # 1: the reference's bytes; K: K strings per item; "auto": by latent size
        self.model.coder_lanes_auto = (lanes == "auto")
        if lanes != "auto":
            self.model.entropy_bottleneck.coder_lanes = int(lanes)
            self.model.gaussian_conditional.coder_lanes = int(lanes)
        index = int(enc.get("coder_index", 0))     # S > 0: decoder index every S symbols behind each reference string
        self.model.entropy_bottleneck.coder_index = index
        self.model.gaussian_conditional.coder_index = index
        self._tables_ready = False      # CDF tables are built on first use (after any checkpoint load)
        self.layer_keys = split or list(self.spec.levels)

    @property
    def qp_value(self):
        return self.qp

    @property
    def eval_encode_type(self):
        return self.eval_encode

    # ------------------------------------------------------------------ encode
    def _ensure_tables(self):
        if not self._tables_ready:
            self.model.update(force=True)          # CompressionModel.update (models/base.py:90-113)
            self._tables_ready = True

    @torch.no_grad()
    def encode(self, input: Dict, codec_output_dir: str = "", bitstream_name: str = "", file_prefix: str = "",
               img_input=False) -> Dict:
        if img_input:
            raise NotImplementedError("lmsfc codes feature pyramids (split inference), not images")
        data = input["data"]
        keys = list(data.keys())
        if len(keys) != len(self.spec.levels):
            raise ValueError(f"expected {len(self.spec.levels)} pyramid levels, got {keys}")
        for k, c in zip(keys, self.spec.channels):
            if data[k].dim() != 4 or data[k].shape[1] != c:
                raise ValueError(f"level {k}: expected [N,{c},H,W], got {tuple(data[k].shape)}")
        self._ensure_tables()
        dev = self.device_name
        # The image pipeline hands one frame on the device; the video pipeline hands ALL frames of a sequence concatenated
        # on N, on the CPU (video_split_inference.py:150-159, 251-263): they go through the device in chunks, every frame
        # keeps its own strings, and the whole call is one bitstream.
        frames = int(data[keys[0]].shape[0])
        y_strings, z_strings, out = [], [], None
        for f0 in range(0, frames, max(1, self.frame_chunk)):
            sl = slice(f0, min(frames, f0 + max(1, self.frame_chunk)))
            feats = {lv: data[k][sl].to(dev, non_blocking=True).float() for lv, k in zip(self.spec.levels, keys)}
            out = self.model.compress(feats)
            y_strings += out["strings"][0]
            z_strings += out["strings"][1]
        nbytes = [len(a) + len(b) for a, b in zip(y_strings, z_strings)]
        stream = {"y_strings": y_strings, "z_strings": z_strings, "shape": out["shape"], "y_hw": out["y_hw"],
                  "mid_hw": out["mid_hw"], "level_hw": out["level_hw"], "keys": keys,
                  "input_size": input.get("input_size"), "org_input_size": input.get("org_input_size")}
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
        self._ensure_tables()
        frames = len(stream["y_strings"])
        parts = []
        for f0 in range(0, frames, max(1, self.frame_chunk)):
            sl = slice(f0, min(frames, f0 + max(1, self.frame_chunk)))
            rec = self.model.decompress([stream["y_strings"][sl], stream["z_strings"][sl]], stream["shape"], stream["y_hw"],
                                        stream["mid_hw"], stream["level_hw"])
            parts.append({lv: rec["x_hat"][lv].float() for lv in self.spec.levels})
        out = {"data": {k: torch.cat([p[lv] for p in parts], 0).contiguous() for k, lv in zip(stream["keys"], self.spec.levels)}}
        if stream.get("input_size") is not None:
            out["input_size"] = stream["input_size"]
        if org_img_size is not None or stream.get("org_input_size") is not None:
            out["org_input_size"] = org_img_size if org_img_size is not None else stream["org_input_size"]
        return out

    # ------------------------------------------------------------------ container: header + per-frame rANS strings
    @staticmethod
    def _write_bin(path, s):
        """v3: "LMSF", u8 3, u32 header length, JSON header (latent / level sizes, level keys, input_size and
        org_input_size of the pipeline -- decode-only runs need them, image_split_inference.py:140-145), then per frame
        u32 len(y), u32 len(z), the two rANS strings."""
        plain = lambda v: None if v is None else json.loads(json.dumps(v, default=lambda o: list(o) if hasattr(o, "__iter__") else int(o)))
        head = json.dumps({"frames": len(s["y_strings"]), "shape": list(s["shape"]), "y_hw": list(s["y_hw"]),
                           "mid_hw": list(s["mid_hw"]), "keys": [str(k) for k in s["keys"]],
                           "level_hw": [list(hw) for hw in s["level_hw"]], "input_size": plain(s.get("input_size")),
                           "org_input_size": plain(s.get("org_input_size"))}).encode()
        with open(path, "wb") as f:
            f.write(MAGIC + struct.pack("<B", 3) + struct.pack("<I", len(head)) + head)
            for ys, zs in zip(s["y_strings"], s["z_strings"]):
                f.write(struct.pack("<2I", len(ys), len(zs)) + ys + zs)

    @staticmethod
    def _read_bin(path):
        with open(path, "rb") as f:
            if f.read(4) != MAGIC:
                raise ValueError(f"{path}: not an lmsfc bitstream")
            version = struct.unpack("<B", f.read(1))[0]
            if version == 3:
                h = json.loads(f.read(struct.unpack("<I", f.read(4))[0]).decode())
                ys, zs = [], []
                for _ in range(h["frames"]):
                    ly, lz = struct.unpack("<2I", f.read(8))
                    ys.append(f.read(ly))
                    zs.append(f.read(lz))
                isz = h.get("input_size")
                return {"y_strings": ys, "z_strings": zs, "shape": tuple(h["shape"]), "y_hw": tuple(h["y_hw"]),
                        "mid_hw": tuple(h["mid_hw"]), "keys": h["keys"], "level_hw": [tuple(hw) for hw in h["level_hw"]],
                        "input_size": [tuple(x) for x in isz] if isz is not None else None,
                        "org_input_size": h.get("org_input_size")}
            if version != 2:
                raise ValueError(f"{path}: not an lmsfc v2 / v3 bitstream")
            frames = struct.unpack("<I", f.read(4))[0]
            dims = struct.unpack("<6I", f.read(24))
            keys, level_hw = [], []
            for _ in range(struct.unpack("<B", f.read(1))[0]):
                n = struct.unpack("<B", f.read(1))[0]
                keys.append(f.read(n).decode())
                level_hw.append(struct.unpack("<2I", f.read(8)))
            ys, zs = [], []
            for _ in range(frames):
                ly, lz = struct.unpack("<2I", f.read(8))
                ys.append(f.read(ly))
                zs.append(f.read(lz))
        return {"y_strings": ys, "z_strings": zs, "shape": dims[0:2], "y_hw": dims[2:4], "mid_hw": dims[4:6],
                "keys": keys, "level_hw": level_hw}
