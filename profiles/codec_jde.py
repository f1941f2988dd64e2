"""BASELINE configs[4] (JDE-1088x608: three split tensors, 256 x 76 x 136 / 512 x 38 x 68 / 1024 x 19 x 34) through the
`lmsfc` codec: per-frame encode / decode latency incl. the CDF table build, for one frame and for a 16-frame video call in
chunks of 8 (the video pipeline hands all frames at once, on the CPU: video_split_inference.py:150-159,251-263).
Random-init weights (no checkpoint offline): bytes are not meaningful, times are.   python profiles/codec_jde.py [out.txt]"""
import importlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"


class JDE:
    split_layer_list = [36, 61, 74]
    device = "cuda"


def wall(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) * 1e3 / iters


def main():
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
    codec_m = importlib.import_module(PKG + ".codec")
    spec = importlib.import_module(PKG + ".spec").JDE_SPEC
    data = importlib.import_module(PKG + ".data")
    torch.manual_seed(0)
    print("# lmsfc codec at BASELINE configs[4] (JDE 1088x608, 3 scales 256/512/1024 ch); ms per frame, wall clock incl. host work", file=out)
    for variant, cfg in (("K=1 (reference strings)", {"coder_lanes": 1}), ("K=1 + decoder index 1024", {"coder_lanes": 1, "coder_index": 1024}),
                         ("lanes auto", {"coder_lanes": "auto"})):
        codec = codec_m.LMSFCCodec(eval_encode="bitrate", encoder_config={"qp": 3, "frame_chunk": 8, **cfg}, device="cuda",
                                   vision_model=JDE(), dataset=None)
        t0 = time.perf_counter()
        codec.model.update(force=True)
        torch.cuda.synchronize()
        t_tab = (time.perf_counter() - t0) * 1e3
        codec._tables_ready = True
        t_tab2 = wall(lambda: codec.model.update(force=True), iters=3, warm=1)
        for frames, where in ((1, "cuda"), (16, "cpu")):
            f = data.synth_features(spec, frames, 76, 136, device=where)
            feats = {k: v.to(torch.bfloat16).float() for k, v in f.items()}
            inp = {"data": dict(zip(JDE.split_layer_list, feats.values())), "input_size": [(608, 1088)],
                   "org_input_size": {"height": 608, "width": 1088}}
            res = codec.encode(inp, "", "", "")
            enc = wall(lambda: codec.encode(inp, "", "", ""), iters=3, warm=1) / frames
            dec = wall(lambda: codec.decode(res["bitstream"], "", ""), iters=3, warm=1) / frames
            print(f"{variant:28s} {frames:2d} frame(s) from {where:4s}: encode {enc:7.3f}  decode {dec:7.3f}  bytes/frame "
                  f"{sum(res['bytes']) // frames:7d}   table build (update): first {t_tab:.1f} ms, warm {t_tab2:.1f} ms", file=out, flush=True)
        del codec


if __name__ == "__main__":
    main()
