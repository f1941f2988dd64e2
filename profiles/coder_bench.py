"""Entropy-coding latency of the device range coder (csrc/rans.cu) at the latent sizes of BASELINE.json's configs,
beside the reference's own CPU coder (oracle/_ref/ans*.so = rans_interface.cpp compiled as is, driven the way
EntropyModel.compress drives it, `.tolist()` included: entropy_models.py:256-268) on the box's host cores.

    python profiles/coder_bench.py [out.json]

kernel_us: CUDA events around the launches only (symbols / indexes / tables resident in HBM);
e2e_us: coder.encode_batch / decode_batch, host strings out / in (D2H / H2D inside the timed region).
Algorithmic bytes per symbol: encode reads 4 (symbol) + 4 (index) + 8 (two CDF entries), writes the string;
decode reads 4 (index) + the string + the searched CDF entries, writes 4.
"""
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"


def product_tables(dev="cuda"):
    """GaussianConditional CDF tables built by the product (entropy_models.update on the device CDF kernel)."""
    em = importlib.import_module(PKG + ".entropy_models")
    model_m = importlib.import_module(PKG + ".model")
    gc = em.GaussianConditional(None).to(dev)
    gc.update_scale_table(model_m.get_scale_table().tolist())
    return gc.scale_table.cpu().numpy(), gc._quantized_cdf, gc._cdf_length, gc._offset


def latent(rng, table, shape):
    idx = np.minimum(rng.integers(0, 64, size=shape), rng.integers(0, 64, size=shape))
    idx = np.minimum(idx, rng.integers(0, 64, size=shape)).astype(np.int32)
    sym = np.round(rng.normal(size=shape) * table[idx]).astype(np.int32)
    return sym, idx


def ev_time(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def wall(fn, iters=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) * 1e6 / iters


def main():
    c = importlib.import_module(PKG + ".coder")
    L = importlib.import_module(PKG + "._lib")
    ptr, stream = L.ptr, L.stream
    dev = "cuda"
    table, tc, ts, to = product_tables(dev)
    cdfs, sizes, offsets = (t.cpu().numpy() for t in (tc, ts, to))
    try:        # the CPU arm only: the reference's own compiled coder (oracle/_ref), like bench.py's cpu_baseline leg
        from oracle import rans as OR
    except Exception:
        OR = None
    rng = np.random.default_rng(5)
    rows = []
    for name, shape in (("cfg2 y 16x256x4x4", (16, 256, 4, 4)), ("cfg4 y 1x256x13x19", (1, 256, 13, 19)),
                        ("cfg4 video 32 frames", (32, 256, 13, 19))):
        sym, idx = latent(rng, table, shape)
        ds, di = torch.from_numpy(sym).to(dev), torch.from_numpy(idx).to(dev)
        items, n = shape[0], int(np.prod(shape[1:]))
        for K in (1, 16, 64, 256):
            T = items * K
            cap = int(L.lib.lmsfc_rans_lane_cap(n, K))
            scratch = torch.empty(T * cap, dtype=torch.int32, device=dev)
            packed = torch.empty(T * cap, dtype=torch.int32, device=dev)
            meta = torch.zeros(2 * T + 2, dtype=torch.int32, device=dev)

            def enc():
                L.rans_encode(ptr(ds), ptr(di), items, n, K, ptr(tc), tc.shape[0], tc.shape[1], ptr(ts), ptr(to),
                              ptr(scratch), cap, ptr(meta[:T]), ptr(meta[T:]), ptr(packed), T * cap, ptr(meta[2 * T + 1:]),
                              stream())
            enc_us = ev_time(enc)
            total_words = int(meta[2 * T].item())
            lane_off = meta[T:2 * T].clone()
            out = torch.empty_like(ds)
            status = torch.zeros(1, dtype=torch.int32, device=dev)

            def dec():
                L.rans_decode(ptr(packed), total_words, ptr(lane_off), ptr(di), items, n, K, ptr(tc), tc.shape[0],
                              tc.shape[1], ptr(ts), ptr(to), ptr(out), ptr(status), stream())
            dec_us = ev_time(dec)
            assert torch.equal(out, ds) and int(status.item()) == 0
            strings = c.encode_batch(ds, di, tc, ts, to, lanes=K)
            e2e_enc = wall(lambda: c.encode_batch(ds, di, tc, ts, to, lanes=K))
            e2e_dec = wall(lambda: c.decode_batch(strings, di, tc, ts, to, lanes=1 if K == 1 else None))
            nsym = items * n
            rows.append({"case": name, "symbols": nsym, "lanes_per_item": K, "string_bytes": sum(map(len, strings)),
                         "encode_kernel_us": round(enc_us, 1), "decode_kernel_us": round(dec_us, 1),
                         "encode_e2e_us": round(e2e_enc, 1), "decode_e2e_us": round(e2e_dec, 1),
                         "encode_Msym_s": round(nsym / enc_us, 1), "decode_Msym_s": round(nsym / dec_us, 1),
                         "encode_GBps_algorithmic": round((16 * nsym + 4 * total_words) / enc_us / 1e3, 2)})
            print(rows[-1], flush=True)
        # reference-format strings (K = 1) with the decoder index every S symbols behind them (coder.encode_batch(index=S))
        for S in (4096, 1024, 256):
            ncp = int(L.lib.lmsfc_rans_checkpoints(n, 1, S))
            if ncp == 0:
                continue
            plain = c.encode_batch(ds, di, tc, ts, to)
            strings = c.encode_batch(ds, di, tc, ts, to, index=S)
            e2e_enc = wall(lambda: c.encode_batch(ds, di, tc, ts, to, index=S))
            e2e_dec = wall(lambda: c.decode_batch(strings, di, tc, ts, to, lanes=1))
            assert torch.equal(c.decode_batch(strings, di, tc, ts, to, lanes=1), ds)
            # kernel alone: the tables as decode_batch uploads them
            body = [np.frombuffer(a, dtype=np.uint32) for a in plain]
            off = np.concatenate([[0], np.cumsum([b.size for b in body])])
            tab = np.stack([np.frombuffer(b, dtype=np.uint32)[a.size: a.size + 3 * ncp].reshape(ncp, 3)
                            for a, b in zip(body, strings)]).astype(np.int64)
            tab[:, :, 2] = off[1:, None] - tab[:, :, 2]
            d_packed = torch.from_numpy(np.concatenate(body).view(np.int32)).to(dev)
            d_off = torch.from_numpy(off[:-1].astype(np.uint32).view(np.int32)).to(dev)
            d_ck = torch.from_numpy(tab.astype(np.uint32).reshape(-1).view(np.int32)).to(dev)
            out = torch.empty_like(ds)
            status = torch.zeros(1, dtype=torch.int32, device=dev)

            def dec_idx():
                L.rans_decode_indexed(ptr(d_packed), int(off[-1]), ptr(d_off), ptr(di), items, n, 1, ptr(tc), tc.shape[0],
                                      tc.shape[1], ptr(ts), ptr(to), ptr(out), ptr(status), S, ncp, ptr(d_ck), stream())
            dec_us = ev_time(dec_idx)
            assert torch.equal(out, ds) and int(status.item()) == 0
            rows.append({"case": name, "symbols": items * n, "lanes_per_item": 1, "index_interval": S,
                         "string_bytes": sum(map(len, strings)), "plain_bytes": sum(map(len, plain)),
                         "decode_kernel_us": round(dec_us, 1), "encode_e2e_us": round(e2e_enc, 1),
                         "decode_e2e_us": round(e2e_dec, 1), "decode_Msym_s": round(items * n / dec_us, 1)})
            print(rows[-1], flush=True)
        if OR is not None and OR.ref_available():
            # the reference arm: compressai's own compiled coder driven as EntropyModel.compress does (1 thread)
            def ref_enc():
                return [OR.ref_encode(sym[i], idx[i], cdfs, sizes, offsets) for i in range(items)]
            t = time.perf_counter(); s_ref = ref_enc(); enc_ref = (time.perf_counter() - t) * 1e6
            t = time.perf_counter()
            for i in range(items):
                OR.ref_decode(s_ref[i], idx[i], cdfs, sizes, offsets)
            dec_ref = (time.perf_counter() - t) * 1e6
            assert s_ref == c.encode_batch(ds, di, tc, ts, to, lanes=1)
            rows.append({"case": name, "impl": "reference (oracle/_ref/ans, 1 host thread, incl. .tolist())",
                         "symbols": items * n, "encode_e2e_us": round(enc_ref, 1), "decode_e2e_us": round(dec_ref, 1)})
            print(rows[-1], flush=True)
    # whole codec (SURVEY.md section 8 cfg 4: one OpenImages-seg frame, P2 = 200 x 304): encode = g_a, h_a, z coder, h_s,
    # build_indexes, y coder; decode = z coder, h_s, build_indexes, y coder, g_s.  Random-init weights, CDFs from update().
    codec_m = importlib.import_module(PKG + ".codec")
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    torch.manual_seed(0)
    for frames, H2, W2 in ((1, 200, 304), (8, 200, 304)):
        feats = {k: v.to(torch.bfloat16).float()
                 for k, v in importlib.import_module(PKG + ".data").synth_features(spec, frames, H2, W2, device=dev).items()}
        inp = {"data": feats, "input_size": [(4 * H2, 4 * W2)], "org_input_size": {"height": 4 * H2, "width": 4 * W2}}
        for K, S in ((1, 0), (1, 1024), (64, 0)):
            codec = codec_m.LMSFCCodec(eval_encode="bitrate", encoder_config={"qp": 3, "coder_lanes": K, "coder_index": S},
                                       device=dev, vision_model=None, dataset=None)
            res = codec.encode(inp, "", "", "")
            enc_us = wall(lambda: codec.encode(inp, "", "", ""), iters=5, warm=2)
            dec_us = wall(lambda: codec.decode(res["bitstream"], "", ""), iters=5, warm=2)
            rows.append({"case": f"lmsfc codec, {frames} frame(s) {4 * H2}x{4 * W2}", "lanes_per_item": K, "index_interval": S,
                         "bytes_per_frame": int(np.mean(res["bytes"])), "encode_ms_per_frame": round(enc_us / 1e3 / frames, 3),
                         "decode_ms_per_frame": round(dec_us / 1e3 / frames, 3)})
            print(rows[-1], flush=True)
            del codec
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
