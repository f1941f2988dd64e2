"""One encode + decode of a cfg-4 latent (1 x 256 x 13 x 19) per lane count, for `ncu -k regex:rans_` captures."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from profiles.coder_bench import latent, product_tables  # noqa: E402

c = importlib.import_module("l-msfc-v2-training_b200.coder")
table, tc, ts, to = product_tables()
sym, idx = latent(np.random.default_rng(5), table, (1, 256, 13, 19))
ds, di = torch.from_numpy(sym).cuda(), torch.from_numpy(idx).cuda()
for K in [int(a) for a in sys.argv[1:]] or [1, 64]:
    for _ in range(2):
        s = c.encode_batch(ds, di, tc, ts, to, lanes=K)
        out = c.decode_batch(s, di, tc, ts, to, lanes=1 if K == 1 else None)
    assert torch.equal(out, ds)
print("ok")
