"""Feature-shard dataset / loader (data.py): host logic on CPU, the feed into a captured train step on GPU."""
import importlib

import pytest
import torch

PKG = "l-msfc-v2-training_b200"


def _dump(tmp_path, name, frames, H2, W2, as_list=False):
    """A feature dump like pipelines/base.py:118-152 writes ({"data": {...}, "input_size": ...})."""
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    g = torch.Generator().manual_seed(hash(name) % 1000)
    data = {}
    for k, (h, w) in zip(spec.levels, spec.level_shapes(H2, W2)):
        # value encodes (level, frame, y, x) so crops can be checked for alignment
        yy = torch.arange(h).view(1, 1, h, 1).float()
        xx = torch.arange(w).view(1, 1, 1, w).float()
        fr = torch.arange(frames).view(frames, 1, 1, 1).float()
        t = (fr * 4 + yy / 64 + xx / 8192).expand(frames, spec.F, h, w).clone()
        data[k] = list(t) if as_list else t
    torch.save({"data": data, "input_size": [(4 * H2, 4 * W2)], "org_input_size": {"height": 4 * H2, "width": 4 * W2}},
               tmp_path / f"{name}.h5")


def test_crops_are_aligned_across_levels(tmp_path):
    data_m = importlib.import_module(PKG + ".data")
    _dump(tmp_path, "seqA", 3, 100, 152)
    _dump(tmp_path, "seqB", 2, 50, 75, as_list=True)       # ragged pyramid: 50x75 -> 25x38 -> 13x19 -> 7x10
    ds = data_m.FeatureFolder(str(tmp_path), crop=(128, 128))
    assert len(ds) == 5
    shapes = ds.level_crop_shapes()
    assert shapes == [(32, 32), (16, 16), (8, 8), (4, 4)]
    out = {k: torch.zeros(1, 256, h, w) for k, (h, w) in zip(ds.spec.levels, shapes)}
    for i in range(10):
        y0, x0 = ds.crop_into(i, out, 0)
        for d, k in enumerate(reversed(ds.spec.levels)):
            t = out[k][0, 0]
            frame = ds.samples[i % len(ds)][1]
            assert abs(float(t[0, 0]) - (frame * 4 + (y0 << d) / 64 + (x0 << d) / 8192)) < 1e-5, (k, i)
            assert abs(float(t[-1, -1] - t[0, 0]) - ((t.shape[0] - 1) / 64 + (t.shape[1] - 1) / 8192)) < 1e-5


def test_ranks_read_disjoint_samples_and_bad_crop_raises(tmp_path):
    data_m = importlib.import_module(PKG + ".data")
    _dump(tmp_path, "seqA", 5, 64, 64)
    a = data_m.FeatureFolder(str(tmp_path), rank=0, world=2)
    b = data_m.FeatureFolder(str(tmp_path), rank=1, world=2)
    assert not set(a.samples) & set(b.samples) and len(a) + len(b) == 5
    with pytest.raises(ValueError):
        data_m.FeatureFolder(str(tmp_path), crop=(250, 256))
    with pytest.raises(FileNotFoundError):
        data_m.FeatureFolder(str(tmp_path / "nothing"))


def test_loader_batches_on_a_background_thread(tmp_path):
    data_m = importlib.import_module(PKG + ".data")
    _dump(tmp_path, "seqA", 3, 64, 96)
    ds = data_m.FeatureFolder(str(tmp_path), crop=(128, 128))
    ld = data_m.ShardLoader(ds, batch=4, ring=3, pin=False)
    try:
        seen = 0
        for _ in range(5):
            slot, host = ld.next_host()
            assert host["p2"].shape == (4, 256, 32, 32) and host["p2"].dtype == torch.bfloat16
            assert host["p2"].is_contiguous(memory_format=torch.channels_last)
            assert host["p5"].shape == (4, 256, 4, 4)
            seen += 1
            ld.release(slot)
        assert seen == 5 and ld.bytes_per_batch() == 4 * 256 * (32 * 32 + 16 * 16 + 8 * 8 + 4 * 4) * 2
    finally:
        ld.close()


@pytest.mark.gpu
def test_loader_feeds_the_captured_step(tmp_path):
    """run_steps (copy stream -> graph input buffers) gives the same losses as stepping on the same batches directly."""
    data_m = importlib.import_module(PKG + ".data")
    model_m = importlib.import_module(PKG + ".model")
    train_m = importlib.import_module(PKG + ".train")
    _dump(tmp_path, "seqA", 4, 32, 32)
    torch.manual_seed(0)

    def run(direct):
        torch.manual_seed(0)
        net = model_m.LMSFC().cuda()
        ts = train_m.TrainStep(net, lmbda=0.013)
        ds = data_m.FeatureFolder(str(tmp_path), crop=(128, 128), seed=7)
        ld = data_m.ShardLoader(ds, batch=2, shuffle=False)
        try:
            slot, first = ld.next_host()
            feats = {k: v.cuda().float() for k, v in first.items()}
            ts.step(feats)
            ts.capture(feats, warmup=1, pipelined=False, n_buffers=2)
            ld.release(slot)
            if direct:
                rows = []
                for i in range(4):
                    slot, host = ld.next_host()
                    out = ts.replay({k: v.cuda() for k, v in host.items()}, buffer=0)
                    rows.append(out["scalars"].cpu().clone())
                    ld.release(slot)
                return torch.stack(rows)
            return data_m.run_steps(ts, ld, 4).clone()
        finally:
            ld.close()
    a, b = run(True), run(False)
    assert torch.isfinite(a).all() and torch.allclose(a, b, rtol=1e-3, atol=1e-5), (a, b)


def test_synthetic_features_equal_the_oracles():
    """bench.py's product arm and the CPU arm must see the same synthetic inputs."""
    from oracle import lmsfc as OL
    data_m = importlib.import_module(PKG + ".data")
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    a = data_m.synth_features(spec, 2, 13, 19, seed=77)
    b = OL.synth_features(spec, 2, 13, 19, seed=77)
    assert list(a) == list(b) and all(torch.equal(a[k], b[k]) for k in a)
    assert float(max(v.max() for v in a.values())) <= 28.40 and float(min(v.min() for v in a.values())) >= -26.43
