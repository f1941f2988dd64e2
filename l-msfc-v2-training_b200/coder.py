"""Host side of the device range coder (csrc/rans.cu), behind EntropyModel.compress / decompress.

Replaces `compressai.ans.RansEncoder().encode_with_indexes` / `RansDecoder().decode_with_indexes` and the
per-item `.tolist()` loops around them (compressai/entropy_models/entropy_models.py:256-268, :309-323): the whole
batch is coded by one launch (one warp per string), and only the finished bytes cross PCIe.

Strings
  lanes = 1   the item's string is byte-identical to the reference's (one rANS string over the flattened item).
  lanes = K   container: u32 K, K x u32 word counts, then the K lane strings back to back; lane k is the
              reference string of symbols[k::K].  Costs 12 bytes per lane (length + the 8-byte final state) and
              buys K-way parallelism inside one item; `decode_batch` reads K from the container.
  index = S   (lanes = 1) the reference string FOLLOWED by a decoder index: for every S-th symbol the coder state and
              stream position in front of it (12 bytes each), then `u32 count, u32 S, u32 "CKPT"`.  The stock decoder
              reads n symbols from the front and never looks at what follows (rans_interface.cpp:215-288), so the string
              stays decodable by the reference; `decode_batch` finds the index and decodes the segments in parallel
              (one serial string otherwise costs ~0.2 us per symbol on the device).
"""
import struct
from typing import List

import numpy as np
import torch

from . import _lib as L
from ._lib import ptr, stream

LANE_MAGIC = 0x4C414E45  # "LANE": first word of a multi-lane item string (a reference string never needs it: K = 1)
INDEX_MAGIC = 0x434B5054  # "CKPT": last word of a string that carries a decoder index behind the reference string


def _tables(cdf, cdf_length, offset, dev):
    cdf = cdf.to(dev, torch.int32).contiguous()
    return cdf, cdf_length.to(dev, torch.int32).reshape(-1).contiguous(), offset.to(dev, torch.int32).reshape(-1).contiguous()


def _check_status(status, who):
    s = int(status.item())
    if s == 1:
        raise ValueError(f"{who}: index outside the CDF tables (or an empty / over-long CDF row)")
    if s == 3:
        raise ValueError(f"{who}: the decoder index does not match the string (corrupt index or string)")
    if s:
        raise RuntimeError(f"{who}: out of string space (status {s})")


def encode_device(symbols, indexes, cdf, cdf_length, offset, lanes: int = 1, index: int = 0):
    """-> (packed uint32 device tensor [total words], lane_offsets int64 numpy [items*lanes + 1]) and, with index = S > 0,
    the checkpoint table uint32 numpy [items*lanes, ncp, 3] ({state low, state high, words written}) as third element.
    symbols / indexes: int32 [items, ...] on a CUDA device."""
    if symbols.device.type != "cuda":
        raise RuntimeError("coder: symbols must live on a CUDA device (l-msfc-v2-training_b200 has no CPU fallback)")
    dev = symbols.device
    items = symbols.shape[0]
    sym = symbols.reshape(items, -1).to(torch.int32).contiguous()
    idx = indexes.reshape(items, -1).to(dev, torch.int32).contiguous()
    n = sym.shape[1] if items else 0
    cdf, sizes, offs = _tables(cdf, cdf_length, offset, dev)
    T = items * lanes
    cap = int(L.lib.lmsfc_rans_lane_cap(n, lanes))
    scratch = torch.empty(max(T * cap, 1), dtype=torch.int32, device=dev)
    packed = torch.empty(max(T * cap, 1), dtype=torch.int32, device=dev)
    meta = torch.zeros(2 * T + 2, dtype=torch.int32, device=dev)     # lane_words [T] | lane_offsets [T+1] | status
    lane_words, lane_offsets, status = meta[:T], meta[T:2 * T + 1], meta[2 * T + 1:]
    ncp = int(L.lib.lmsfc_rans_checkpoints(n, lanes, index)) if index else 0
    ckpt = torch.zeros((T, ncp, 3), dtype=torch.int32, device=dev) if ncp else None
    if index:
        L.rans_encode_indexed(ptr(sym), ptr(idx), items, n, lanes, ptr(cdf), cdf.shape[0], cdf.shape[1], ptr(sizes), ptr(offs),
                              ptr(scratch), cap, ptr(lane_words), ptr(lane_offsets), ptr(packed), T * cap, ptr(status),
                              int(index), ncp, ptr(ckpt), stream())
    else:
        L.rans_encode(ptr(sym), ptr(idx), items, n, lanes, ptr(cdf), cdf.shape[0], cdf.shape[1], ptr(sizes), ptr(offs),
                      ptr(scratch), cap, ptr(lane_words), ptr(lane_offsets), ptr(packed), T * cap, ptr(status), stream())
    host = meta.cpu().numpy()
    if host[-1] == 1:
        raise ValueError("rans_encode: index outside the CDF tables (or an empty / over-long CDF row)")
    if host[-1]:
        raise RuntimeError(f"rans_encode: out of string space (status {host[-1]})")
    off = host[T:2 * T + 1].astype(np.int64) & 0xFFFFFFFF
    body = packed[: int(off[-1])] if T else packed[:0]
    if index:
        table = ckpt.cpu().numpy().view(np.uint32) if ncp else np.zeros((T, 0, 3), np.uint32)
        return body, off, table
    return body, off


def encode_batch(symbols, indexes, cdf, cdf_length, offset, precision: int = 16, lanes: int = 1, index: int = 0) -> List[bytes]:
    """One string per batch item (entropy_models.py:256-268).  index = S: the decoder index behind every string
    (lanes = 1 only; strings shorter than S symbols get none)."""
    if precision != 16:
        raise ValueError("the range coder works at 16-bit precision (rans_interface.cpp:50)")
    if index and lanes != 1:
        raise ValueError("the decoder index belongs to reference-format strings (lanes = 1); lanes > 1 is parallel already")
    items = symbols.shape[0]
    table = None
    if index:
        packed, off, table = encode_device(symbols, indexes, cdf, cdf_length, offset, lanes, index)
    else:
        packed, off = encode_device(symbols, indexes, cdf, cdf_length, offset, lanes)
    words = packed.cpu().numpy().view(np.uint32)
    out = []
    for i in range(items):
        lo, hi = off[i * lanes], off[(i + 1) * lanes]
        body = words[lo:hi].tobytes()
        if lanes == 1:
            if table is not None and table.shape[1]:
                body += table[i].tobytes() + struct.pack("<3I", table.shape[1], int(index), INDEX_MAGIC)
            out.append(body)
        else:
            counts = np.diff(off[i * lanes:(i + 1) * lanes + 1]).astype(np.uint32)
            out.append(struct.pack("<2I", LANE_MAGIC, lanes) + counts.tobytes() + body)
    return out


def _split_item(s: bytes):
    """-> (lanes, per-lane word counts, payload words)."""
    w = np.frombuffer(s, dtype=np.uint32)
    if len(s) % 4 or w.size < 2:
        raise ValueError("Invalid strings or indexes parameters")
    if w.size >= 3 and w[0] == LANE_MAGIC and 1 < int(w[1]) <= (w.size - 2) // 3 and \
            int(w[2:2 + int(w[1])].astype(np.int64).sum()) == w.size - 2 - int(w[1]):
        k = int(w[1])
        return k, w[2:2 + k].astype(np.int64), w[2 + k:]
    return 1, np.array([w.size], dtype=np.int64), w


def _strip_index(w, n):
    """w: uint32 words of a single-lane string for n symbols -> (reference string words, interval, table [ncp,3] | None).
    The trailer is accepted only when all of its fields agree with n (count = lmsfc_rans_checkpoints(n, 1, S))."""
    if w.size >= 8 and w[-1] == INDEX_MAGIC:
        S, ncp = int(w[-2]), int(w[-3])
        if S >= 128 and S & (S - 1) == 0 and ncp >= 1 and ncp == (n + S - 1) // S - 1 and w.size >= 3 * ncp + 3 + 2:
            body = w[: w.size - 3 * ncp - 3]
            return body, S, w[body.size: body.size + 3 * ncp].reshape(ncp, 3)
    return w, 0, None


def decode_batch(strings, indexes, cdf, cdf_length, offset, precision: int = 16, lanes=None):
    """strings: one per batch item -> int32 tensor shaped like `indexes` on its device (entropy_models.py:309-323).
    lanes = 1 takes every string as a plain reference string; None reads the lane count from the container."""
    if precision != 16:
        raise ValueError("the range coder works at 16-bit precision (rans_interface.cpp:50)")
    dev = cdf.device if indexes.device.type != "cuda" else indexes.device
    if dev.type != "cuda":
        raise RuntimeError("coder: CDF tables / indexes must live on a CUDA device (no CPU fallback)")
    items = len(strings)
    idx = indexes.reshape(items, -1).to(dev, torch.int32).contiguous()
    n = idx.shape[1] if items else 0
    if lanes == 1:
        parts = [(1, np.array([len(s) // 4], dtype=np.int64), np.frombuffer(s, dtype=np.uint32)) for s in strings]
        if any(len(s) % 4 or len(s) < 8 for s in strings):
            raise ValueError("Invalid strings or indexes parameters")
    else:
        parts = [_split_item(s) for s in strings]
    lanes = parts[0][0] if parts else 1
    if any(p[0] != lanes for p in parts):
        raise ValueError("decode_batch: items were coded with different lane counts")
    interval, tables = 0, None
    if lanes == 1 and parts:                         # decoder index behind the reference string (encode_batch(index=S))
        stripped = [_strip_index(p[2], n) for p in parts]
        if all(t[2] is not None for t in stripped) and len({t[1] for t in stripped}) == 1:
            interval, tables = stripped[0][1], [t[2] for t in stripped]
        parts = [(1, np.array([t[0].size], dtype=np.int64), t[0]) for t in stripped]
    counts = np.concatenate([p[1] for p in parts]) if parts else np.zeros(0, np.int64)
    off = np.concatenate([[0], np.cumsum(counts)])
    words = np.concatenate([p[2] for p in parts]) if parts else np.zeros(0, np.uint32)
    out = torch.zeros(idx.shape, dtype=torch.int32, device=dev)
    if items and n:
        pieces = [words.view(np.int32), off[:-1].astype(np.uint32).view(np.int32)]
        ncp = 0
        if tables is not None:
            # third word: words written when the checkpoint was taken -> absolute position of the segment's first word
            tab = np.stack(tables).astype(np.int64)                  # [items, ncp, 3]
            ncp = tab.shape[1]
            tab[:, :, 2] = off[1:, None] - tab[:, :, 2]
            if (tab[:, :, 2] < off[:-1, None] + 2).any():
                raise ValueError("Invalid strings or indexes parameters")
            pieces.append(tab.astype(np.uint32).reshape(-1).view(np.int32))
        d = torch.from_numpy(np.concatenate(pieces)).to(dev)
        packed, lane_off = d[: words.size], d[words.size: words.size + items * lanes]
        cdf, sizes, offs = _tables(cdf, cdf_length, offset, dev)
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        if ncp:
            ck = d[words.size + items * lanes:]
            L.rans_decode_indexed(ptr(packed), words.size, ptr(lane_off), ptr(idx), items, n, lanes, ptr(cdf), cdf.shape[0],
                                  cdf.shape[1], ptr(sizes), ptr(offs), ptr(out), ptr(status), interval, ncp, ptr(ck), stream())
        else:
            L.rans_decode(ptr(packed), words.size, ptr(lane_off), ptr(idx), items, n, lanes, ptr(cdf), cdf.shape[0],
                          cdf.shape[1], ptr(sizes), ptr(offs), ptr(out), ptr(status), stream())
        _check_status(status, "rans_decode")
    return out.reshape(indexes.shape)


def gc_build_indexes(scales, scale_table, scale_bound: float):
    """GaussianConditional.build_indexes (entropy_models.py:684-689) -> int32, same shape."""
    if scales.device.type != "cuda":
        raise RuntimeError("coder: scales must live on a CUDA device (no CPU fallback)")
    s = scales.detach().float().contiguous()
    table = scale_table.to(s.device, torch.float32).contiguous()
    out = torch.empty(s.shape, dtype=torch.int32, device=s.device)
    L.gc_build_indexes(ptr(s), s.numel(), ptr(table), table.numel(), float(scale_bound), ptr(out), stream())
    return out
