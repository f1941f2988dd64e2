"""CPU: the oracle restatement (oracle/) against fixtures generated from the REAL
reference by tests/golden/make_golden.py, plus the reference's own known-answer
tests (compressai/tests/test_ops.py:104-118, test_layers.py:135-172)."""
import hashlib
import importlib

import numpy as np
import pytest
import torch

from oracle import cdf as OC
from oracle import lmsfc as OL
from oracle import ops as O

T = lambda a: torch.from_numpy(np.asarray(a))


def close(a, b, rtol=1e-5, atol=1e-6):
    np.testing.assert_allclose(np.asarray(a), np.asarray(b), rtol=rtol, atol=atol)


# ------------------------------------------------------------------ integer CDF
def test_cdf_kat_reference_tests():
    # compressai/tests/test_ops.py:104-106
    assert OC.pmf_to_quantized_cdf([0.1, 0.2, 0.0, 0.0], 16).tolist() == [0, 21845, 65534, 65535, 65536]
    assert OC.pmf_to_quantized_cdf([0.5, 0.25, 0.25, 1e-9], 16).tolist() == [0, 32768, 49151, 65535, 65536]
    # test_ops.py:108-118 error behaviour
    for bad in ([-0.1, 0.5], [float("inf"), 0.5], [float("nan"), 0.5], [0.0, 0.0]):
        with pytest.raises(ValueError):
            OC.pmf_to_quantized_cdf(bad, 16)


def test_cdf_golden_bit_exact(golden):
    g = golden("cdf")
    offs = g["pmf_offsets"]
    pos = 0
    for i in range(len(offs) - 1):
        p = g["pmf_flat"][offs[i]:offs[i + 1]]
        want = g["cdf_flat"][pos:pos + len(p) + 1]
        pos += len(p) + 1
        got = OC.pmf_to_quantized_cdf(p, 16)
        assert np.array_equal(got, want), i
    pos = 0
    for i in range(10):
        p = g["pmf_flat"][offs[i]:offs[i + 1]]
        want = g["cdf12_flat"][pos:pos + len(p) + 1]
        pos += len(p) + 1
        assert np.array_equal(OC.pmf_to_quantized_cdf(p, 12), want)


def test_cdf_against_compiled_reference(golden):
    if not OC.ref_available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(3)
    for n in (1, 2, 9, 300):
        p = rng.random(n).astype(np.float32)
        p[rng.random(n) < 0.3] = 0
        p[0] += 1e-3
        assert np.array_equal(OC.pmf_to_quantized_cdf(p), OC.ref_pmf_to_quantized_cdf(p))


# ------------------------------------------------------------------ small ops
def test_lower_bound_and_ste(golden):
    g = golden("modules")
    x = T(g["lb_x"]).requires_grad_()
    y = O.lower_bound(x, 0.25)
    y.backward(T(g["lb_g"]))
    close(y.detach(), g["lb_y"]); close(x.grad, g["lb_gx"])
    x = T(g["ste_x"]).requires_grad_()
    y = O.quantize_ste(x)
    y.backward(T(g["lb_g"]))
    close(y.detach(), g["ste_y"]); close(x.grad, g["ste_gx"])


@pytest.mark.parametrize("name,inv", [("gdn", False), ("igdn", True)])
def test_gdn(golden, name, inv):
    g = golden("modules")
    x = T(g[f"{name}_x"]).requires_grad_()
    b = T(g[f"{name}_beta"]).requires_grad_()
    gm = T(g[f"{name}_gamma"]).requires_grad_()
    y = O.gdn(x, b, gm, inverse=inv)
    y.backward(T(g[f"{name}_gy"]))
    close(y.detach(), g[f"{name}_y"], 1e-5, 1e-6)
    close(x.grad, g[f"{name}_gx"], 1e-4, 1e-5)
    close(b.grad, g[f"{name}_gbeta"], 1e-4, 1e-5)
    close(gm.grad, g[f"{name}_ggamma"], 1e-4, 1e-5)


def test_gdn_closed_form_at_init(golden):
    # compressai/tests/test_layers.py:135-159: y == x / sqrt(1 + .1 * x^2)
    g = golden("modules")
    m = OL.GDNParams(4)
    x = T(g["gdn_init_x"])
    y = m(x)
    close(y.detach(), g["gdn_init_y"])
    close(y.detach(), x / torch.sqrt(1 + 0.1 * x ** 2), 1e-5, 1e-6)
    # init constants SURVEY §8(b1)
    assert abs(m.gamma[0, 0].item() - 0.316228) < 1e-6 and abs(m.gamma[0, 1].item() - 3.8147e-06) < 1e-10


# ------------------------------------------------------------------ entropy models
def _eb_params(g):
    mats = [T(g[f"eb_p__matrix{i}"]).requires_grad_() for i in range(5)]
    bias = [T(g[f"eb_p__bias{i}"]).requires_grad_() for i in range(5)]
    fac = [T(g[f"eb_p__factor{i}"]).requires_grad_() for i in range(4)]
    q = T(g["eb_p_quantiles"]).requires_grad_()
    return mats, bias, fac, q


def test_entropy_bottleneck_forward_backward(golden):
    g = golden("modules")
    mats, bias, fac, q = _eb_params(g)
    x = T(g["eb_x"]).requires_grad_()
    out, lik = O.eb_forward(x, mats, bias, fac, q, True, T(g["eb_noise"]))
    close(out.detach(), g["eb_out"]); close(lik.detach(), g["eb_lik"], 1e-5, 1e-9)
    (torch.sum(torch.log(lik) * T(g["eb_glik"])) + torch.sum(out * T(g["eb_gout"]))).backward()
    close(x.grad, g["eb_gx"], 1e-4, 1e-6)
    for i in range(5):
        close(mats[i].grad, g[f"eb_g__matrix{i}"], 1e-4, 1e-6)
        close(bias[i].grad, g[f"eb_g__bias{i}"], 1e-4, 1e-6)
    for i in range(4):
        close(fac[i].grad, g[f"eb_g__factor{i}"], 1e-4, 1e-6)
    assert q.grad is None and g["eb_g_quantiles"].size == 0
    out, lik = O.eb_forward(T(g["eb_x"]), mats, bias, fac, q, False)
    close(out.detach(), g["eb_out_eval"]); close(lik.detach(), g["eb_lik_eval"], 1e-5, 1e-9)
    # bound + gradient rule at the tails
    xw = T(g["ebw_x"]).requires_grad_()
    out, lik = O.eb_forward(xw, mats, bias, fac, q, True, T(g["ebw_noise"]))
    torch.log(lik).sum().backward()
    close(lik.detach(), g["ebw_lik"], 1e-5, 1e-12); close(xw.grad, g["ebw_gx"], 1e-4, 1e-6)


def test_entropy_bottleneck_aux_and_update(golden):
    g = golden("modules")
    mats, bias, fac, q = _eb_params(g)
    aux = O.eb_aux_loss(mats, bias, fac, q, T(g["eb_target"]))
    aux.backward()
    close(aux.detach(), g["eb_aux"], 1e-6); close(q.grad, g["eb_aux_gq"], 1e-5, 1e-6)
    assert all(m.grad is None for m in mats)
    cdf, off, ln = OC.eb_update([m.detach() for m in mats], [b.detach() for b in bias],
                                [f.detach() for f in fac], q.detach())
    assert np.array_equal(cdf, g["eb_cdf"]) and np.array_equal(off, g["eb_offset"])
    assert np.array_equal(ln, g["eb_cdf_length"])


def test_gaussian_conditional(golden):
    g = golden("modules")
    y, sc, mu = (T(g[k]).requires_grad_() for k in ("gc_y", "gc_sc", "gc_mu"))
    out, lik = O.gc_forward(y, sc, mu, True, T(g["gc_noise"]))
    close(out.detach(), g["gc_out"]); close(lik.detach(), g["gc_lik"], 1e-5, 1e-10)
    (torch.sum(torch.log(lik) * T(g["gc_glik"])) + torch.sum(out * T(g["gc_gout"]))).backward()
    close(y.grad, g["gc_gy"], 1e-4, 1e-6); close(sc.grad, g["gc_gsc"], 1e-4, 1e-6)
    close(mu.grad, g["gc_gmu"], 1e-4, 1e-6)
    y.grad = sc.grad = None
    out, lik = O.gc_forward(y, sc, None, True, T(g["gc0_noise"]))
    torch.log(lik).sum().backward()
    close(lik.detach(), g["gc0_lik"], 1e-5, 1e-10)
    close(y.grad, g["gc0_gy"], 1e-4, 1e-6); close(sc.grad, g["gc0_gsc"], 1e-4, 1e-6)
    out, lik = O.gc_forward(y.detach(), sc.detach(), mu.detach(), False)
    close(out, g["gc_out_eval"]); close(lik, g["gc_lik_eval"], 1e-5, 1e-10)


def test_gaussian_conditional_tables(golden):
    g = golden("modules")
    cdf, off, ln = OC.gc_update(g["gc8_scale_table"])
    assert np.array_equal(cdf, g["gc8_cdf"]) and np.array_equal(off, g["gc8_offset"])
    assert np.array_equal(ln, g["gc8_cdf_length"])
    st = OC.get_scale_table()
    close(st, g["gc64_scale_table"], 1e-6)
    cdf, off, ln = OC.gc_update(T(g["gc64_scale_table"]))
    assert list(cdf.shape) == g["gc64_shape"].tolist()          # [64, 3133]
    assert hashlib.sha256(cdf.astype(np.int32).tobytes()).digest() == g["gc64_sha256"].tobytes()
    assert np.array_equal(cdf[[0, 13, 40, 63]], g["gc64_rows"])
    assert np.array_equal(off, g["gc64_offset"]) and np.array_equal(ln, g["gc64_cdf_length"])


def test_rd_loss(golden):
    g = golden("modules")
    tgt = T(g["rd_tgt"])
    r = O.rd_loss([T(g["rd_l1"]), T(g["rd_l2"])], [T(g["rd_xh"])], [tgt], 0.013,
                  tgt.shape[0] * tgt.shape[2] * tgt.shape[3])
    close([r["loss"].item(), r["mse_loss"].item(), r["bpp_loss"].item()], g["rd_out"], 1e-6)


# ------------------------------------------------------------------ whole step
def _load_oracle_net(pkg, g):
    F, N, M, B, H2, W2 = (int(v) for v in g["cfg"])
    spec = pkg.spec.LMSFCSpec(F=F, N=N, M=M)
    net = OL.OracleLMSFC(spec)
    sd = net.state_dict()
    for k in sd:
        if k.endswith("target"):
            continue
        sd[k] = T(g["w_" + k])
    net.load_state_dict(sd)
    feats = {k: T(g["in_" + k]) for k in spec.levels}
    return spec, net, feats


@pytest.mark.parametrize("tag", ["tiny", "odd"])
def test_train_step_against_reference_modules(pkg, golden, tag):
    """oracle/lmsfc.py (restated maths) == the same topology assembled from the
    reference's own modules: forward, loss, every gradient, clip norm, and the
    parameters after one Adam/aux-Adam step (examples/train.py:94-107)."""
    g = golden("step_" + tag)
    spec, net, feats = _load_oracle_net(pkg, g)
    F, N, M, B, H2, W2 = (int(v) for v in g["cfg"])
    ref_feats = OL.synth_features(spec, B, H2, W2, seed=1234)
    for k in spec.levels:
        close(ref_feats[k], g["in_" + k], 0, 0)
    opt, aux_opt = OL.configure_optimizers(net)
    opt.zero_grad(); aux_opt.zero_grad()
    out = net(feats, T(g["noise_z"]), T(g["noise_y"]), training=True)
    for k in ("y", "z", "y_hat", "z_hat", "scales"):
        close(out["latents"][k].detach(), g["lat_" + k], 1e-4, 1e-5)
    close(out["likelihoods"]["y"].detach(), g["lik_y"], 1e-4, 1e-8)
    close(out["likelihoods"]["z"].detach(), g["lik_z"], 1e-4, 1e-8)
    for k in spec.levels:
        close(out["x_hat"][k].detach(), g["xhat_" + k], 1e-4, 1e-4)
    crit = net.criterion(out, feats, float(g["lmbda"]))
    close([crit["loss"].item(), crit["mse_loss"].item(), crit["bpp_loss"].item()], g["loss"], 1e-5)
    crit["loss"].backward()
    for n, p in net.named_parameters():
        if n.endswith("quantiles"):
            assert p.grad is None
            continue
        ref = g["g_" + n]
        scale = max(1e-6, float(np.abs(ref).max()))
        np.testing.assert_allclose(p.grad.numpy(), ref, rtol=1e-3, atol=2e-4 * scale, err_msg=n)
    gn = torch.nn.utils.clip_grad_norm_(net.parameters(), 1.0)
    close(gn, g["grad_norm"], 1e-4)
    opt.step()
    aux = net.aux_loss(); aux.backward(); aux_opt.step()
    close(aux.detach(), g["aux_loss"], 1e-5)
    for n, p in net.named_parameters():
        np.testing.assert_allclose(p.detach().numpy(), g["w1_" + n], rtol=1e-4, atol=2e-6, err_msg=n)
