"""CPU-side checks (no GPU needed): the C-ABI library loads and exports every symbol that
include/b200lmsfc.h declares, argument validation answers without touching a device, the product
package never imports the oracle, and the frozen spec's FLOP counter reproduces SURVEY.md §8(d).
"""
import ctypes
import importlib
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "l-msfc-v2-training_b200"
HEADER = os.path.join(ROOT, "include", "b200lmsfc.h")
LIB = os.path.join(ROOT, PKG, "libb200lmsfc.so")


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lmsfc_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = _declared_symbols()
    for must in ("lmsfc_conv_forward", "lmsfc_conv_dgrad", "lmsfc_conv_wgrad", "lmsfc_gdn_forward", "lmsfc_gdn_backward",
                 "lmsfc_eb_forward", "lmsfc_eb_backward", "lmsfc_gc_forward", "lmsfc_gc_backward", "lmsfc_eb_aux_loss",
                 "lmsfc_mse_forward_backward", "lmsfc_adam_step", "lmsfc_pmf_to_quantized_cdf", "lmsfc_pmf_rows_to_cdf",
                 "lmsfc_last_error"):
        assert must in names, must


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(LIB)
    missing = [n for n in _declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing
    lib.lmsfc_version.restype = ctypes.c_int
    assert lib.lmsfc_version() >= 100


def test_binding_covers_the_header():
    L = importlib.import_module(PKG + "._lib")
    assert sorted(L.EXPORTS) == _declared_symbols()


def test_argument_errors_are_reported_without_a_device():
    """Bad arguments are rejected on the host before any CUDA call: return code 1 -> ValueError
    carrying lmsfc_last_error() (the reference raises ValueError for its bad-mode / bad-size cases,
    entropy_models.py:158-159)."""
    L = importlib.import_module(PKG + "._lib")
    with pytest.raises(ValueError, match="Invalid quantization mode"):
        L.gc_forward(None, None, None, None, 0, 7, 0.11, 1e-9, 0, 0, None, None, None, None, None, None)
    d = L.ConvDesc(0, 4, 2, 1, 8, 8, 4, 4, 256, 0, 256)          # k = 4 is not an L-MSFC kernel size
    with pytest.raises(ValueError, match="conv: k=4"):
        L.conv_forward(d, None, None, None, None, 0, None, None, None, None)
    d = L.ConvDesc(0, 5, 2, 1, 8, 8, 4, 4, 100, 0, 256)          # channels must be multiples of 64
    with pytest.raises(ValueError, match="channels"):
        L.conv_forward(d, None, None, None, None, 0, None, None, None, None)
    d = L.ConvDesc(0, 5, 2, 1, 8, 8, 5, 4, 256, 0, 256)          # output size must follow the stride
    with pytest.raises(ValueError, match="does not match"):
        L.conv_forward(d, None, None, None, None, 0, None, None, None, None)


def test_product_package_never_imports_the_oracle():
    """oracle/ is test infrastructure: no product module may import, call or link it."""
    pkg_dir = os.path.join(ROOT, PKG)
    bad = []
    for dp, _, files in os.walk(pkg_dir):
        if "build" in dp.split(os.sep):
            continue
        for f in files:
            if not f.endswith((".py", ".cu", ".cuh", ".h")) and f != "Makefile":
                continue
            txt = open(os.path.join(dp, f)).read()
            if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M) or "oracle/" in txt and f != "spec.py" and \
                    re.search(r"(include|open|CDLL).*oracle/", txt):
                bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_modules_refuse_cpu_tensors():
    """No CPU fallback: the drop-in modules raise on host tensors instead of computing there."""
    import torch
    layers = importlib.import_module(PKG + ".layers")
    em = importlib.import_module(PKG + ".entropy_models")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        layers.GDN(256)(torch.zeros(1, 256, 4, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        layers.conv(256, 256)(torch.zeros(1, 256, 8, 8))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        em.EntropyBottleneck(256)(torch.zeros(1, 256, 4, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        em.GaussianConditional(None)(torch.zeros(1, 256, 4, 4), torch.ones(1, 256, 4, 4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        em.EntropyBottleneck(256).update()


def test_state_dict_layout_matches_reference():
    """key names / shapes of GDN, EntropyBottleneck, GaussianConditional."""
    import torch
    layers = importlib.import_module(PKG + ".layers")
    em = importlib.import_module(PKG + ".entropy_models")
    sd = layers.GDN(8).state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {
        "beta": (8,), "gamma": (8, 8), "beta_reparam.pedestal": (1,), "beta_reparam.lower_bound.bound": (1,),
        "gamma_reparam.pedestal": (1,), "gamma_reparam.lower_bound.bound": (1,)}
    assert abs(float(sd["gamma"][0, 0]) - 0.316228) < 1e-5 and abs(float(sd["gamma"][0, 1]) - 3.8147e-06) < 1e-9
    sd = em.EntropyBottleneck(4).state_dict()
    want = {"_matrix0": (4, 3, 1), "_matrix1": (4, 3, 3), "_matrix2": (4, 3, 3), "_matrix3": (4, 3, 3), "_matrix4": (4, 1, 3),
            "_bias0": (4, 3, 1), "_bias1": (4, 3, 1), "_bias2": (4, 3, 1), "_bias3": (4, 3, 1), "_bias4": (4, 1, 1),
            "_factor0": (4, 3, 1), "_factor1": (4, 3, 1), "_factor2": (4, 3, 1), "_factor3": (4, 3, 1),
            "quantiles": (4, 1, 3), "_offset": (0,), "_quantized_cdf": (0,), "_cdf_length": (0,), "target": (3,),
            "likelihood_lower_bound.bound": (1,)}
    assert {k: tuple(v.shape) for k, v in sd.items()} == want
    assert len(list(em.EntropyBottleneck(4).parameters())) == 15          # tests/test_models.py:61-63
    assert torch.allclose(sd["quantiles"][0, 0], torch.tensor([-10.0, 0.0, 10.0]))
    sd = em.GaussianConditional(None).state_dict()
    for k in ("_offset", "_quantized_cdf", "_cdf_length", "scale_table", "scale_bound", "likelihood_lower_bound.bound",
              "lower_bound_scale.bound"):
        assert k in sd, k


def test_flop_counter_reproduces_survey():
    """SURVEY.md §8(d) formula applied to spec.py: 183,444 * P flop forward per feature-set (the
    survey's worked example, 183,732 * P, counted the hyper nets 0.16 % higher)."""
    spec = importlib.import_module(PKG + ".spec").DEFAULT_SPEC
    P = 256 * 256
    assert spec.flops_fwd(64, 64) == 183444 * P
    assert abs(spec.flops_fwd(64, 64) / (183732 * P) - 1) < 2e-3
    assert spec.flops_train(64, 64) == 3 * 183444 * P
    assert spec.level_shapes(26, 38) == [(26, 38), (13, 19), (7, 10), (4, 5)]
    assert spec.num_pixels(16, 64, 64) == 16 * 256 * 256
