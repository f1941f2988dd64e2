"""GPU parity of the whole L-MSFC train step against the oracle graph (oracle/lmsfc.py = the
reference's modules composed per spec.py) on identical weights, inputs and injected noise.

north_star tolerances: loss rel <= 1e-2, bpp within 0.5 %; gradients within bf16 tolerance
(normalised max error per parameter tensor <= 6e-2, cosine >= 0.999 over all parameters).
"""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
PKG = "l-msfc-v2-training_b200"


@pytest.fixture(autouse=True)
def _oracle_in_true_fp32():
    """The oracle runs on the same GPU: cuDNN / cuBLAS must not silently use TF32 for it."""
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


# per parameter tensor: cosine with / relative L2 distance from the fp32 gradient.  Measured on a B200 (profiles/
# r2_parity_bench_configs.txt): the worst tensors are always the hyper-analysis ones (|g| ~1e-4 of the total norm; the
# ReLU / |y| kinks flip between bf16 and fp32 activations) at relL2 0.06-0.10; everything outside h_a / h_s stays
# below 0.02
COS_MIN, RELL2_MAX, RELL2_MAX_MAIN = 0.99, 0.15, 0.06


def _mods():
    return [importlib.import_module(f"{PKG}.{m}") for m in ("spec", "model", "train")]


def _pair(B, H2, W2, seed=0, spec=None):
    from oracle import lmsfc as OL
    spec_m, model_m, _ = _mods()
    torch.manual_seed(seed)
    spec = spec or spec_m.DEFAULT_SPEC
    net = model_m.LMSFC(spec).to(DEV)
    ref = OL.OracleLMSFC(spec).to(DEV)
    sd = net.state_dict()
    missing = ref.load_state_dict({k: v for k, v in sd.items() if k in ref.state_dict()}, strict=False)
    assert not missing.missing_keys, missing
    feats = OL.synth_features(spec, B, H2, W2, device=DEV)
    feats = {k: v.to(torch.bfloat16).float() for k, v in feats.items()}   # bf16-representable inputs
    with torch.no_grad():
        o = ref(feats)
    zs, ys = o["latents"]["z"].shape, o["latents"]["y"].shape
    g = torch.Generator().manual_seed(seed + 1)
    nz = (torch.rand(zs, generator=g) - 0.5).to(DEV)
    ny = (torch.rand(ys, generator=g) - 0.5).to(DEV)
    return spec, net, ref, feats, nz, ny


# (16, 64, 64) = BASELINE configs[1], the configuration bench.py is quoted on (every tile / split-K choice of the cost
# model, the full-wave weight-gradient kernels); (1, 128, 192) = configs[0]; (1, 200, 304) = configs[3] (y 13 x 19,
# ragged tiles at every level)
# ("jde", 1, 76, 136) = configs[4]: the 3-scale 256 / 512 / 1024-channel JDE split (spec.JDE_SPEC), one 1088x608 frame
# ("ms", ...) / ("ms16", ...): the same pyramid under the mean-scale hyperprior (spec.MEAN_SCALE_SPEC, models/google.py:337-405)
@pytest.mark.parametrize("B,H2,W2", [(2, 32, 32), (1, 26, 38), (16, 64, 64), (1, 128, 192), (1, 200, 304), ("jde", 76, 136),
                                     ("jde2", 20, 28), ("ms", 32, 32), ("ms1", 26, 38), ("ms16", 64, 64)])
def test_forward_loss_and_gradients(B, H2, W2):
    from oracle import lmsfc as OL
    jde = isinstance(B, str) and B.startswith("jde")
    ms = isinstance(B, str) and B.startswith("ms")
    if jde:
        B = 2 if B == "jde2" else 1
    if ms:
        B = {"ms": 2, "ms1": 1, "ms16": 16}[B]
    sm = _mods()[0]
    spec, net, ref, feats, nz, ny = _pair(B, H2, W2, spec=sm.JDE_SPEC if jde else (sm.MEAN_SCALE_SPEC if ms else None))
    _, _, train_m = _mods()
    lmbda = 0.013
    net.build_arena()
    crit_mod = importlib.import_module(f"{PKG}.losses").RateDistortionLoss(lmbda=lmbda, spec=spec)
    out = net(feats, nz, ny)
    crit = crit_mod(out, feats)
    crit["loss"].backward()
    # oracle: EB noise in its [C,1,B*h*w] layout (entropy_models.py:490-498)
    nz_o = nz.permute(1, 0, 2, 3).reshape(nz.shape[1], 1, -1)
    o = ref(feats, nz_o, ny)
    rc = ref.criterion(o, feats, lmbda)
    rc["loss"].backward()
    torch.cuda.synchronize()
    rel = lambda a, b: abs(float(a) - float(b)) / abs(float(b))
    print("loss", float(crit["loss"]), float(rc["loss"]), "bpp", float(crit["bpp_loss"]), float(rc["bpp_loss"]),
          "mse", float(crit["mse_loss"]), float(rc["mse_loss"]))
    assert rel(crit["loss"], rc["loss"]) < 1e-2
    assert rel(crit["bpp_loss"], rc["bpp_loss"]) < 5e-3
    assert rel(crit["mse_loss"], rc["mse_loss"]) < 1e-2
    for k in spec.levels:
        a, b = out["x_hat"][k].float(), o["x_hat"][k]
        assert a.shape == b.shape
        assert float((a - b).abs().max() / b.abs().max()) < 3e-2, k
    rp = dict(ref.named_parameters())
    dot = na = nb = 0.0
    rows = []
    for n, p in net.named_parameters():
        if n.endswith("quantiles"):
            assert rp[n].grad is None
            continue
        ga, gb = p.grad.float(), rp[n].grad.float()
        d, a2, b2 = float((ga * gb).sum()), float((ga * ga).sum()), float((gb * gb).sum())
        rows.append((d / max((a2 * b2) ** 0.5, 1e-30), n, float((ga - gb).norm() / gb.norm().clamp_min(1e-30)), b2 ** 0.5))
        dot += d; na += a2; nb += b2
    cos = dot / (na ** 0.5 * nb ** 0.5)
    rows.sort()
    print("grad cosine", cos, "norms", na ** 0.5, nb ** 0.5)
    for r in rows[:5]:
        print("  lowest per-tensor cosine %.5f  %-34s relL2 %.3e  |g| %.3e" % r)
    worst = sorted(rows, key=lambda r: -r[2])[:5]
    for r in worst:
        print("  largest per-tensor relL2  %.3e  %-34s cosine %.5f  |g| %.3e" % (r[2], r[1], r[0], r[3]))
    # whole-gradient agreement, and every tensor individually (ReLU / |y| kinks that flip between
    # bf16 and fp32 activations move single small tensors of the hyper path by a few percent)
    assert cos > 0.9999 and abs(na ** 0.5 / nb ** 0.5 - 1) < 1e-2
    assert rows[0][0] > COS_MIN, rows[0]
    assert worst[0][2] < RELL2_MAX, worst[0]
    main = [r for r in sorted(rows, key=lambda r: -r[2]) if not r[1].startswith(("h_a.", "h_s."))]
    for r in main[:3]:
        print("  largest relL2 outside the hyper nets  %.3e  %-34s cosine %.5f  |g| %.3e" % (r[2], r[1], r[0], r[3]))
    assert main[0][2] < RELL2_MAX_MAIN, main[0]


@pytest.mark.parametrize("variant", ["scale", "mean-scale"])
def test_train_steps_track_the_oracle(variant):
    """Three optimiser steps (clip 1.0 + Adam 1e-4, aux Adam 1e-3) on both sides, same noise."""
    from oracle import lmsfc as OL
    spec, net, ref, feats, nz, ny = _pair(2, 32, 32, seed=3, spec=_mods()[0].MEAN_SCALE_SPEC if variant == "mean-scale" else None)
    _, _, train_m = _mods()
    ts = train_m.TrainStep(net, lmbda=0.013)
    opt, aux_opt = OL.configure_optimizers(ref)
    nz_o = nz.permute(1, 0, 2, 3).reshape(nz.shape[1], 1, -1)
    for it in range(3):
        a = ts.step(feats, nz, ny)
        b = OL.train_step(ref, opt, aux_opt, feats, 0.013, 1.0, nz_o, ny)
        print(it, {k: (float(a[k]), float(b[k])) for k in a})
        for k in ("loss", "bpp_loss", "mse_loss", "aux_loss"):
            assert abs(float(a[k]) - float(b[k])) <= 1e-2 * abs(float(b[k])), (it, k)
    q = dict(ref.named_parameters())["entropy_bottleneck.quantiles"]
    torch.testing.assert_close(net.entropy_bottleneck.quantiles.detach(), q.detach(), rtol=1e-4, atol=1e-5)


def test_cuda_graph_step_matches_eager_and_trains():
    spec, net, ref, feats, nz, ny = _pair(2, 32, 32, seed=5)
    _, _, train_m = _mods()
    ts = train_m.TrainStep(net, lmbda=0.013)
    first = float(ts.step(feats)["loss"])
    ts.capture(feats)
    losses = [float(ts.step(feats)["loss"]) for _ in range(20)]
    print("eager", first, "graph", losses[:3], "...", losses[-1])
    assert all(l == l and l < 1e6 for l in losses)
    assert losses[-1] < first                     # it trains
    assert len(set(losses)) > 10                  # fresh noise / weights every replay


def test_pipelined_graph_step_equals_the_sequential_order():
    """capture(pipelined=True) rotates the iteration ([Adam of the previous backward || forward] -> backward); after
    flush() the weights and the loss sequence must be those of the eager loop with the same number of iterations
    (differences: float atomics order in the gradient reductions only)."""
    spec_m, model_m, train_m = _mods()
    from oracle import lmsfc as OL
    spec = spec_m.DEFAULT_SPEC
    feats = OL.synth_features(spec, 2, 32, 32, device=DEV)
    runs = []
    for pipelined in (False, True):
        torch.manual_seed(11)
        net = model_m.LMSFC(spec).to(DEV)
        ts = train_m.TrainStep(net, lmbda=0.013)
        losses = []
        if pipelined:
            ts.capture(feats, warmup=1, pipelined=True)      # 1 eager iteration + 2 forward passes (priming), 1 update pending
            for _ in range(3):
                losses.append(float(ts.replay()["loss"]))
            assert ts.pending
            ts.flush()
            assert not ts.pending
        else:
            for i in range(6):
                l = float(ts.step(feats)["loss"])
                if i >= 3:
                    losses.append(l)
        torch.cuda.synchronize()
        runs.append((losses, net.arena.net_params.clone(), net.arena.aux_params.clone(), int(ts.opt["net"].step_count)))
    (la, pa, qa, sa), (lb, pb, qb, sb) = runs
    print("sequential", la, "pipelined", lb)
    assert sa == sb == 6
    for x, y in zip(la, lb):
        assert abs(x - y) <= 2e-3 * abs(x), (la, lb)
    rel = float((pa - pb).norm() / pa.norm())
    assert rel < 1e-4, rel
    torch.testing.assert_close(qa, qb, rtol=1e-3, atol=1e-4)


def test_bench_configuration_first_loss_is_the_pinned_oracle_value():
    """bench.py's exact set-up (seed-1234 weights, seed-1234 batch-16 features, lambda 0.013, in-kernel Philox noise):
    the first step's loss against tests/golden/bench_loss.json (written by tests/golden/make_bench_loss.py from the
    oracle graph in fp32).  bench.py asserts the same value before it times anything."""
    import json
    import os
    spec_m, model_m, train_m = _mods()
    data_m = importlib.import_module(f"{PKG}.data")
    pin = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "bench_loss.json")))
    torch.manual_seed(1234)
    net = model_m.LMSFC(spec_m.DEFAULT_SPEC).to(DEV)
    ts = train_m.TrainStep(net, lmbda=pin["lmbda"])
    f = data_m.synth_features(spec_m.DEFAULT_SPEC, pin["batch"], pin["H2"], pin["W2"], seed=pin["seed"])
    feats = {k: v.to(torch.bfloat16).to(DEV) for k, v in f.items()}
    out = ts.step(feats)
    torch.cuda.synchronize()
    print({k: float(v) for k, v in out.items()}, pin)
    assert abs(float(out["loss"]) - pin["loss"]) <= 1e-2 * pin["loss"]
    assert abs(float(out["bpp_loss"]) - pin["bpp_loss"]) <= 5e-3 * pin["bpp_loss"]
    assert abs(float(out["mse_loss"]) - pin["mse_loss"]) <= 1e-2 * pin["mse_loss"]
