"""Writes tests/golden/rans.npz: strings produced by the REFERENCE's own compiled range coder
(oracle/_ref/ans*.so, built by oracle/Makefile from
/root/reference/compressai/compressai/cpp_exts/rans/rans_interface.cpp) for seeded inputs.
Run in the authoring container only (the reference does not travel):  python tests/golden/make_golden_rans.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cdf as OC   # noqa: E402
from oracle import rans as OR  # noqa: E402


def gc_case(rng, n, outlier_rate):
    table = OC.get_scale_table().numpy()
    cdfs, offsets, sizes = OC.gc_update(table)
    idx = rng.integers(0, len(table), size=n).astype(np.int32)
    # small scales dominate a trained latent: skew the draw
    idx = np.minimum(idx, rng.integers(0, len(table), size=n)).astype(np.int32)
    sym = np.round(rng.normal(size=n) * table[idx]).astype(np.int32)
    out = rng.random(n) < outlier_rate
    sym[out] = (rng.integers(-70000, 70000, size=int(out.sum()))).astype(np.int32)
    return sym, idx, cdfs, sizes, offsets


def eb_case(rng, channels, n_per):
    rows, width = channels, 21
    cdfs = np.zeros((rows, width + 2), dtype=np.int32)
    sizes = np.zeros(rows, dtype=np.int32)
    offsets = np.zeros(rows, dtype=np.int32)
    for r in range(rows):
        ln = int(rng.integers(3, width + 1))
        pmf = rng.random(ln).astype(np.float32) ** 3 + 1e-4
        pmf /= pmf.sum()
        tail = np.float32(1e-4)
        c = OC.pmf_to_quantized_cdf(np.concatenate([pmf, [tail]]))
        cdfs[r, : c.size] = c.astype(np.int32)
        sizes[r] = ln + 2
        offsets[r] = -(ln // 2)
    idx = np.repeat(np.arange(rows, dtype=np.int32), n_per)
    sym = np.zeros(idx.size, dtype=np.int32)
    for r in range(rows):
        ln = sizes[r] - 2
        pm = np.diff(cdfs[r, : ln + 1]).astype(np.float64)
        pm /= pm.sum()
        sym[idx == r] = rng.choice(ln, size=n_per, p=pm) + offsets[r]
    hit = rng.random(sym.size) < 0.01
    sym[hit] += rng.integers(-40, 40, size=int(hit.sum())).astype(np.int32)
    return sym, idx, cdfs, sizes, offsets


def main():
    assert OR.ref_available(), "build oracle/_ref first (make -C oracle)"
    rng = np.random.default_rng(20240611)
    cases = {}
    cases["gc_5000"] = gc_case(rng, 5000, 0.002)
    cases["gc_outliers"] = gc_case(rng, 700, 0.3)
    # the reference's flush() writes 2 words below a buffer of `#records` words (rans_interface.cpp:184-201):
    # strings of 1-2 symbols corrupt its heap, so the shortest reference-generated case has 8 symbols
    cases["gc_8"] = gc_case(rng, 8, 0.0)
    cases["eb_64x37"] = eb_case(rng, 64, 37)
    out = {}
    for name, (sym, idx, cdfs, sizes, offsets) in cases.items():
        s = OR.ref_encode(sym, idx, cdfs, sizes, offsets)
        back = OR.ref_decode(s, idx, cdfs, sizes, offsets)
        assert np.array_equal(back, sym), name
        assert OR.encode(sym, idx, cdfs, sizes, offsets) == s, name
        for k, v in zip(("sym", "idx", "cdfs", "sizes", "offsets"), (sym, idx, cdfs, sizes, offsets)):
            out[f"{name}.{k}"] = v
        out[f"{name}.string"] = np.frombuffer(s, dtype=np.uint8)
        # lanes: reference strings of the strided subsets (the product's interleaved container)
        for K in (3, 32):
            if sym.size >= K:
                for k in (0, K - 1):
                    sk = OR.ref_encode(sym[k::K], idx[k::K], cdfs, sizes, offsets)
                    out[f"{name}.lane{k}of{K}"] = np.frombuffer(sk, dtype=np.uint8)
        print(name, sym.size, "symbols ->", len(s), "bytes")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "rans.npz"), **out)


if __name__ == "__main__":
    main()
