"""Oracle (test infrastructure, see oracle/__init__.py): CDF-table build.

Restates entropy_models.py:204-212 (_pmf_to_cdf), :389-429 (EntropyBottleneck.
update) and :625-648 (GaussianConditional.update) on top of the C restatement
of cpp_exts/ops/ops.cpp:40-109 in oracle/cdf.c.  `ref_pmf_to_quantized_cdf`
calls the reference's own compiled `_CXX` (oracle/_ref/, built by
oracle/Makefile from /root/reference sources) when it is present.
"""
import ctypes
import importlib.util
import glob
import os
import subprocess

import numpy as np
import scipy.stats
import torch

from . import ops as O

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    """Compile oracle/cdf.c (and oracle/_ref when /root/reference exists)."""
    subprocess.run(["make", "-C", _HERE, "-s"], check=True, stdout=subprocess.DEVNULL)


def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "liblmsfc_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.lmsfc_oracle_pmf_to_quantized_cdf.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        _LIB.lmsfc_oracle_pmf_to_quantized_cdf.restype = ctypes.c_int
    return _LIB


def pmf_to_quantized_cdf(pmf, precision=16):
    """ops.cpp:40-109.  pmf: 1-D float sequence -> np.uint32 [n+1].
    Raises ValueError like the reference's std::domain_error -> pybind11."""
    p = np.ascontiguousarray(np.asarray(pmf, dtype=np.float32))
    out = np.zeros(p.size + 1, dtype=np.uint32)
    rc = _lib().lmsfc_oracle_pmf_to_quantized_cdf(
        p.ctypes.data, int(p.size), int(precision), out.ctypes.data)
    if rc == 1:
        raise ValueError("Invalid `pmf`, non-finite or negative element found")
    if rc == 2:
        raise ValueError("Invalid `pmf`: at least one element must have a non-zero probability.")
    if rc != 0:
        raise ValueError("pmf_to_quantized_cdf: no symbol to steal from")
    return out


def ref_available():
    return bool(glob.glob(os.path.join(_HERE, "_ref", "_CXX*.so")))


_REF = None


def ref_pmf_to_quantized_cdf(pmf, precision=16):
    """The reference's own compiled ops.cpp (oracle/_ref/_CXX*.so)."""
    global _REF
    if _REF is None:
        path = glob.glob(os.path.join(_HERE, "_ref", "_CXX*.so"))[0]
        spec = importlib.util.spec_from_file_location("_CXX", path)
        _REF = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_REF)
    return np.asarray(_REF.pmf_to_quantized_cdf([float(np.float32(v)) for v in pmf], precision),
                      dtype=np.uint32)


def pmf_rows_to_cdf(pmf, tail_mass, pmf_length, max_length, precision=16):
    """entropy_models.py:204-212."""
    cdf = np.zeros((len(pmf_length), max_length + 2), dtype=np.int32)
    for i in range(len(pmf_length)):
        prob = np.concatenate([pmf[i][: int(pmf_length[i])], tail_mass[i]])
        c = pmf_to_quantized_cdf(prob, precision)
        cdf[i, : c.size] = c.astype(np.int32)
    return cdf


def eb_update(matrices, biases, factors, quantiles):
    """entropy_models.py:389-429 -> (quantized_cdf, offset, cdf_length) int32."""
    with torch.no_grad():
        medians = quantiles[:, 0, 1]
        minima = torch.clamp(torch.ceil(medians - quantiles[:, 0, 0]).int(), min=0)
        maxima = torch.clamp(torch.ceil(quantiles[:, 0, 2] - medians).int(), min=0)
        offset = -minima
        pmf_start = medians - minima
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max().item())
        samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]
        lower = O.eb_logits_cumulative(samples - 0.5, matrices, biases, factors, True)
        upper = O.eb_logits_cumulative(samples + 0.5, matrices, biases, factors, True)
        sign = -torch.sign(lower + upper)
        pmf = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))[:, 0, :]
        tail = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
    cdf = pmf_rows_to_cdf(pmf.numpy(), tail.numpy(), pmf_length.numpy(), max_length)
    return cdf, offset.numpy().astype(np.int32), (pmf_length + 2).numpy().astype(np.int32)


def gc_pmf(scale_table, tail_mass=1e-9):
    """entropy_models.py:626-642 -> (pmf, tail, pmf_center, pmf_length, max_length)."""
    scale_table = torch.as_tensor(scale_table, dtype=torch.float32)
    multiplier = -scipy.stats.norm.ppf(tail_mass / 2)
    pmf_center = torch.ceil(scale_table * multiplier).int()
    pmf_length = 2 * pmf_center + 1
    max_length = int(torch.max(pmf_length).item())
    samples = torch.abs(torch.arange(max_length).int() - pmf_center[:, None]).float()
    s = scale_table.unsqueeze(1).float()
    upper = O.std_cumulative((0.5 - samples) / s)
    lower = O.std_cumulative((-0.5 - samples) / s)
    return upper - lower, 2 * lower[:, :1], pmf_center, pmf_length, max_length


def gc_update(scale_table, tail_mass=1e-9):
    """entropy_models.py:625-648 -> (quantized_cdf, offset, cdf_length) int32."""
    pmf, tail, center, length, max_length = gc_pmf(scale_table, tail_mass)
    cdf = pmf_rows_to_cdf(pmf.numpy(), tail.numpy(), length.numpy(), max_length)
    return cdf, (-center).numpy().astype(np.int32), (length + 2).numpy().astype(np.int32)


def get_scale_table(lo=0.11, hi=256, levels=64):
    """models/base.py:57-59."""
    import math
    return torch.exp(torch.linspace(math.log(lo), math.log(hi), levels))
