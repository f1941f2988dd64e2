"""The reference training script's loop on the B200 modules, unchanged: compressai/examples/train.py:74-118
(configure_optimizers -> net_aux_optimizer, train_one_epoch), :279 (ReduceLROnPlateau), :283-290 / :310-321
(checkpoint save / resume).  Plus the captured step's own checkpoint, its LR schedule and flush() semantics.
"""
import importlib
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
PKG = "l-msfc-v2-training_b200"


def _m(name):
    return importlib.import_module(f"{PKG}.{name}")


def _feats(B=2, H2=32, W2=32, seed=1234):
    return {k: v.to(DEV) for k, v in _m("data").synth_features(_m("spec").DEFAULT_SPEC, B, H2, W2, seed=seed).items()}


def _reference_loop(model, criterion, optimizer, aux_optimizer, loader, clip_max_norm):
    """train_one_epoch, examples/train.py:85-118, line for line (`d` is the pyramid dict here)."""
    model.train()
    losses = []
    for i, d in enumerate(loader):
        optimizer.zero_grad()
        aux_optimizer.zero_grad()
        out_net = model(d)
        out_criterion = criterion(out_net, d)
        out_criterion["loss"].backward()
        if clip_max_norm > 0:
            torch.nn.utils.clip_grad_norm_(model.parameters(), clip_max_norm)
        optimizer.step()
        aux_loss = model.aux_loss()
        aux_loss.backward()
        aux_optimizer.step()
        losses.append((out_criterion["loss"].item(), out_criterion["mse_loss"].item(), out_criterion["bpp_loss"].item(),
                       aux_loss.item()))
    return losses


def _build(seed):
    torch.manual_seed(seed)
    net = _m("model").LMSFC(_m("spec").DEFAULT_SPEC).to(DEV)
    # configure_optimizers, examples/train.py:74-82
    conf = {"net": {"type": "Adam", "lr": 1e-4}, "aux": {"type": "Adam", "lr": 1e-3}}
    opt = _m("optim").net_aux_optimizer(net, conf)
    return net, opt["net"], opt["aux"]


def test_reference_train_loop_runs_unchanged_and_resumes(tmp_path):
    from oracle import lmsfc as OL
    spec = _m("spec").DEFAULT_SPEC
    net, optimizer, aux_optimizer = _build(7)
    assert isinstance(optimizer, torch.optim.Optimizer) and isinstance(aux_optimizer, torch.optim.Optimizer)
    ref = OL.OracleLMSFC(spec).to(DEV)
    ref.load_state_dict({k: v for k, v in net.state_dict().items() if k in ref.state_dict()}, strict=False)
    ropt, raux = OL.configure_optimizers(ref)
    criterion = _m("losses").RateDistortionLoss(lmbda=0.013)
    lr_scheduler = torch.optim.lr_scheduler.ReduceLROnPlateau(optimizer, "min", patience=0)   # train.py:279
    loader = [_feats(seed=100 + i) for i in range(3)]
    got = _reference_loop(net, criterion, optimizer, aux_optimizer, loader, clip_max_norm=1.0)
    want = [OL.train_step(ref, ropt, raux, d, 0.013, 1.0) for d in loader]
    for g, w in zip(got, want):                      # different noise draws: the rate term differs by its sampling error
        print("loop", g, {k: float(v) for k, v in w.items()})
        assert abs(g[0] - float(w["loss"])) <= 2e-2 * abs(float(w["loss"]))
        assert abs(g[1] - float(w["mse_loss"])) <= 2e-2 * abs(float(w["mse_loss"]))
        assert abs(g[3] - float(w["aux_loss"])) <= 1e-3 * abs(float(w["aux_loss"]))
    # torch's clip ran first: the optimiser's own clip is disabled by default, as in the reference
    assert optimizer.clip_max_norm == 0.0
    # an LR schedule drives the fused kernel: two non-improving epochs -> lr * 0.1
    lr_scheduler.step(1.0)
    lr_scheduler.step(2.0)
    assert abs(optimizer.param_groups[0]["lr"] - 1e-5) < 1e-12
    # ---- save_checkpoint, train.py:310-321
    path = os.path.join(tmp_path, "checkpoint.pth.tar")
    torch.save({"epoch": 0, "state_dict": net.state_dict(), "loss": got[-1][0], "optimizer": optimizer.state_dict(),
                "aux_optimizer": aux_optimizer.state_dict(), "lr_scheduler": lr_scheduler.state_dict()}, path)
    # ---- resume, train.py:283-290
    net2, optimizer2, aux_optimizer2 = _build(99)            # different init: everything must come from the checkpoint
    lr_scheduler2 = torch.optim.lr_scheduler.ReduceLROnPlateau(optimizer2, "min", patience=0)
    checkpoint = torch.load(path, map_location=DEV)
    net2.load_state_dict(checkpoint["state_dict"])
    optimizer2.load_state_dict(checkpoint["optimizer"])
    aux_optimizer2.load_state_dict(checkpoint["aux_optimizer"])
    lr_scheduler2.load_state_dict(checkpoint["lr_scheduler"])
    assert abs(optimizer2.param_groups[0]["lr"] - 1e-5) < 1e-12
    assert int(optimizer2.step_count) == 3 and int(aux_optimizer2.step_count) == 3
    net2.entropy_bottleneck._rng(torch.device(DEV, 0)).state.copy_(net.entropy_bottleneck._rng(torch.device(DEV, 0)).state)
    nxt = [_feats(seed=200)]
    a = _reference_loop(net, criterion, optimizer, aux_optimizer, nxt, 1.0)[0]
    b = _reference_loop(net2, criterion, optimizer2, aux_optimizer2, nxt, 1.0)[0]
    print("resumed", a, b)
    for x, y in zip(a, b):
        assert abs(x - y) <= 1e-5 * abs(x)
    rel = float((net.arena.net_params - net2.arena.net_params).norm() / net.arena.net_params.norm())
    assert rel < 1e-6, rel
    torch.testing.assert_close(net.arena.aux_params, net2.arena.aux_params, rtol=1e-6, atol=1e-7)


def test_lr_change_after_capture_reaches_the_kernel():
    """lr used to be frozen into the captured kernel arguments."""
    feats = _feats()
    deltas = []
    for scale in (1.0, 0.5):
        torch.manual_seed(21)
        net = _m("model").LMSFC(_m("spec").DEFAULT_SPEC).to(DEV)
        ts = _m("train").TrainStep(net, lmbda=0.013)
        ts.capture(feats, warmup=1, pipelined=True)
        ts.flush()
        before = net.arena.net_params.clone()
        for g in ts.opt["net"].param_groups:          # what ReduceLROnPlateau does
            g["lr"] = g["lr"] * scale
        ts.replay()
        ts.flush()
        torch.cuda.synchronize()
        deltas.append(net.arena.net_params - before)
    ratio = float(deltas[1].norm() / deltas[0].norm())
    err = float((deltas[1] - 0.5 * deltas[0]).norm() / deltas[0].norm())
    print("update norm ratio", ratio, "err", err)
    assert abs(ratio - 0.5) < 5e-3 and err < 1e-2


def test_replay_flush_replay_flush_equals_the_eager_loop():
    """after flush() the next replay must not apply a second (momentum-only) update, and the second flush
    must apply the last backward pass."""
    feats = _feats()
    runs = []
    for mode in ("eager", "pipelined"):
        torch.manual_seed(31)
        net = _m("model").LMSFC(_m("spec").DEFAULT_SPEC).to(DEV)
        ts = _m("train").TrainStep(net, lmbda=0.013)
        if mode == "eager":
            for _ in range(3 + 4):                   # capture() below runs 3 iterations before the first replay
                ts.step(feats)
        else:
            ts.capture(feats, warmup=1, pipelined=True)
            ts.replay(); ts.flush()
            ts.replay(); ts.flush()
            ts.replay(); ts.replay(); ts.flush()
            ts.flush()                               # nothing pending: a no-op
            assert not ts.pending
        torch.cuda.synchronize()
        runs.append((net.arena.net_params.clone(), net.arena.aux_params.clone(), int(ts.opt["net"].step_count),
                     int(ts.opt["aux"].step_count)))
    (pa, qa, sa, ta), (pb, qb, sb, tb) = runs
    print("steps", sa, sb, ta, tb)
    assert sa == sb == 7 and ta == tb == 7
    rel = float((pa - pb).norm() / pa.norm())
    assert rel < 1e-4, rel
    torch.testing.assert_close(qa, qb, rtol=1e-3, atol=1e-4)


def test_train_step_checkpoint_resumes_the_captured_run(tmp_path):
    feats = _feats()
    torch.manual_seed(41)
    net = _m("model").LMSFC(_m("spec").DEFAULT_SPEC).to(DEV)
    ts = _m("train").TrainStep(net, lmbda=0.013)
    ts.capture(feats, warmup=1, pipelined=True)
    for _ in range(3):
        ts.replay()
    path = os.path.join(tmp_path, "step.pt")
    torch.save({"state_dict": net.state_dict(), "train_step": ts.state_dict()}, path)      # flushes the pending update
    torch.manual_seed(43)
    net2 = _m("model").LMSFC(_m("spec").DEFAULT_SPEC).to(DEV)
    ts2 = _m("train").TrainStep(net2, lmbda=0.013)
    ck = torch.load(path, map_location=DEV)
    net2.load_state_dict(ck["state_dict"])
    ts2.load_state_dict(ck["train_step"])
    ts2.capture(feats, warmup=0, pipelined=True, resume=True)
    la = [float(ts.replay()["loss"]) for _ in range(3)]
    lb = [float(ts2.replay()["loss"]) for _ in range(3)]
    ts.flush(); ts2.flush()
    torch.cuda.synchronize()
    print("original", la, "resumed", lb)
    assert int(ts.opt["net"].step_count) == int(ts2.opt["net"].step_count)
    for x, y in zip(la, lb):
        assert abs(x - y) <= 2e-3 * abs(x)
    rel = float((net.arena.net_params - net2.arena.net_params).norm() / net.arena.net_params.norm())
    assert rel < 1e-4, rel
