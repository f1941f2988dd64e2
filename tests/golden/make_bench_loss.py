"""Writes tests/golden/bench_loss.json: the loss of the FIRST train step of bench.py's configuration (seed-1234
weights, seed-1234 synthetic batch-16 features, lambda 0.013) computed by the oracle graph (oracle/lmsfc.py, fp32,
CPU).  bench.py and tests/test_gpu_step.py compare the product's first-step loss with it (rel <= 1e-2).  The
quantisation noise differs between the two (torch's generator here, Philox on the device): its effect on the
loss is measured below by redrawing it.

    python tests/golden/make_bench_loss.py
"""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
PKG = "l-msfc-v2-training_b200"


def main():
    from oracle import lmsfc as OL
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    # the product model draws its initial weights from torch.manual_seed(1234) in bench.py; the oracle graph declares
    # its parameters in the same order with the same initialisers -> identical state_dict for the same seed?  Not
    # guaranteed (GEMM-layout storage), so the weights are copied from a product-model instance built on the CPU.
    model_m = importlib.import_module(PKG + ".model")
    torch.manual_seed(1234)
    net = model_m.LMSFC(spec)                      # construction is host-only; no kernel runs
    ref = OL.OracleLMSFC(spec)
    missing = ref.load_state_dict({k: v for k, v in net.state_dict().items() if k in ref.state_dict()}, strict=False)
    assert not missing.missing_keys, missing
    feats = OL.synth_features(spec, 16, 64, 64, seed=1234)
    feats = {k: v.to(torch.bfloat16).float() for k, v in feats.items()}
    vals = []
    with torch.no_grad():
        for draw in range(3):
            torch.manual_seed(100 + draw)
            out = ref(feats)
            crit = ref.criterion(out, feats, 0.013)
            vals.append({k: float(v) for k, v in crit.items()})
            print(draw, vals[-1])
    spread = max(v["loss"] for v in vals) / min(v["loss"] for v in vals) - 1
    res = dict(vals[0], noise_spread_rel=spread, seed=1234, batch=16, H2=64, W2=64, lmbda=0.013,
               how="oracle/lmsfc.py fp32 CPU, weights of LMSFC() under torch.manual_seed(1234), bf16-rounded synthetic features")
    with open(os.path.join(ROOT, "tests", "golden", "bench_loss.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(res)


if __name__ == "__main__":
    main()
