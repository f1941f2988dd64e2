"""Drop-in EntropyBottleneck / GaussianConditional (interfaces of
compressai/entropy_models/entropy_models.py:99-725) on fused sm_100a kernels.

Constructor signatures, parameter / buffer names and shapes, `forward` return values, `update`,
`loss`, error behaviour (ValueError cases) follow the reference; forward + backward run in
lmsfc_eb_* / lmsfc_gc_* (quantise + likelihood + bound + sum-ln fused), the integer CDF tables in
lmsfc_pmf_rows_to_cdf.  Entropy *coding* (compress / decompress / build_indexes,
SURVEY.md §8f-1) runs in csrc/rans.cu through coder.py: strings byte-identical to the reference's rANS.
"""
from typing import Any, List, Optional, Tuple, Union

import numpy as np
import scipy.stats
import os

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L
from . import functional as Fn
from ._lib import ptr, stream
from .ops import LowerBound


def pmf_to_quantized_cdf(pmf, precision: int = 16):
    """compressai.entropy_models.entropy_models.pmf_to_quantized_cdf (:89-92) -> IntTensor."""
    p = np.ascontiguousarray(torch.as_tensor(pmf, dtype=torch.float32).detach().cpu().numpy())
    out = np.zeros(p.size + 1, dtype=np.uint32)
    L.pmf_to_quantized_cdf(p.ctypes.data, int(p.size), int(precision), out.ctypes.data)
    return torch.from_numpy(out.astype(np.int64)).int()


class _EntropyCoder:
    """entropy_models.py:47-80: the `entropy_model.entropy_coder` object, list-based like the pybind11 `ans` coder it
    wraps there (`encode_with_indexes(symbols, indexes, cdfs, cdf_lengths, offsets) -> bytes` and back), here one
    string on the device coder.  `device` is where the lists are staged."""

    def __init__(self, method: str = "ans", device="cuda"):
        if method != "ans":
            raise ValueError(f'Unknown entropy coder "{method}" (available: ans)')
        self.name = method
        self.device = device

    def _tabs(self, cdfs, cdf_lengths, offsets):
        i32 = dict(dtype=torch.int32, device=self.device)
        width = max(len(r) for r in cdfs)
        cdf = torch.zeros((len(cdfs), width), **i32)
        for r, row in enumerate(cdfs):
            cdf[r, : len(row)] = torch.as_tensor(row, **i32)
        return cdf, torch.as_tensor(cdf_lengths, **i32), torch.as_tensor(offsets, **i32)

    def encode_with_indexes(self, symbols, indexes, cdfs, cdf_lengths, offsets) -> bytes:
        from . import coder
        i32 = dict(dtype=torch.int32, device=self.device)
        return coder.encode_batch(torch.as_tensor(symbols, **i32).reshape(1, -1), torch.as_tensor(indexes, **i32).reshape(1, -1),
                                  *self._tabs(cdfs, cdf_lengths, offsets))[0]

    def decode_with_indexes(self, string, indexes, cdfs, cdf_lengths, offsets):
        from . import coder
        idx = torch.as_tensor(indexes, dtype=torch.int32, device=self.device).reshape(1, -1)
        return coder.decode_batch([string], idx, *self._tabs(cdfs, cdf_lengths, offsets), lanes=1).reshape(-1).tolist()


class EntropyModel(nn.Module):
    """entropy_models.py:99-325."""

    def __init__(self, likelihood_bound: float = 1e-9, entropy_coder: Optional[str] = None,
                 entropy_coder_precision: int = 16):
        super().__init__()
        if entropy_coder is not None and entropy_coder != "ans":      # entropy_models.py:44-56: available_entropy_coders()
            raise ValueError(f'Invalid entropy coder "{entropy_coder}", choose from (ans)')
        self.entropy_coder_name = entropy_coder or "ans"
        self.entropy_coder = _EntropyCoder(self.entropy_coder_name)
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.likelihood_bound = float(likelihood_bound)
        self.use_likelihood_bound = likelihood_bound > 0
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())
        self.rng = None          # Fn.RNG, created lazily on the module's device (a model may share one)
        self.rng_offset = 0      # distinguishes the draws of different modules within one step
        self.rng_auto_advance = True   # False when the owning model advances a shared RNG once per step
        # float pmf of update() on the device (kernels eb_pmf / gc_pmf) instead of the host: for per-frame table refresh.
        # Off by default: the host evaluation reproduces the reference's CPU float ops bit for bit, the device one only
        # up to single counts of the 16-bit tables (inside the reference's +-2 tolerance for rebuilt tables).
        self.device_pmf = os.environ.get("LMSFC_DEVICE_PMF", "0") == "1"
        self.coder_lanes = 1     # strings per item (coder.py): 1 = the reference's bytes, K > 1 = interleaved container
        self.coder_index = 0     # S > 0 (lanes = 1): decoder index every S symbols behind the reference string (coder.py)

    @property
    def offset(self):
        return self._offset

    @property
    def quantized_cdf(self):
        return self._quantized_cdf

    @property
    def cdf_length(self):
        return self._cdf_length

    def _rng(self, device):
        if self.rng is None or self.rng.state.device != device:
            self.rng = Fn.RNG(device, seed=torch.initial_seed() & 0x7FFFFFFFFFFFFFFF)
        return self.rng

    def quantize(self, inputs, mode: str, means=None):
        """:155-180.  Stand-alone helper (the training path fuses it into the likelihood kernel)."""
        if mode not in ("noise", "dequantize", "symbols"):
            raise ValueError(f'Invalid quantization mode: "{mode}"')
        if mode == "noise":
            half = float(0.5)
            noise = torch.empty_like(inputs).uniform_(-half, half)
            return inputs + noise
        outputs = inputs.clone()
        if means is not None:
            outputs -= means
        outputs = torch.round(outputs)
        if mode == "dequantize":
            if means is not None:
                outputs += means
            return outputs
        return outputs.int()

    @staticmethod
    def dequantize(inputs, means=None, dtype: torch.dtype = torch.float):
        if means is not None:
            outputs = inputs.type_as(means)
            outputs += means
        else:
            outputs = inputs.type(dtype)
        return outputs

    def _pmf_to_cdf(self, pmf, tail_mass, pmf_length, max_length):
        """:204-212 -- one row per channel / scale, one thread block per row on the device."""
        rows = len(pmf_length)
        pmf = pmf.detach().float().contiguous()
        tail = tail_mass.detach().float().reshape(rows).contiguous()
        plen = pmf_length.detach().int().contiguous()
        max_length = int(max_length)
        dev = self._offset.device
        if dev.type != "cuda":
            raise RuntimeError("EntropyModel._pmf_to_cdf: move the module to a CUDA device first "
                               "(l-msfc-v2-training_b200 has no CPU fallback)")
        pmf_d, tail_d, plen_d = pmf.to(dev), tail.to(dev), plen.to(dev)
        cdf = torch.empty((rows, max_length + 2), dtype=torch.int32, device=dev)
        status = torch.zeros(rows, dtype=torch.int32, device=dev)
        L.pmf_rows_to_cdf(ptr(pmf_d), pmf_d.stride(0), ptr(tail_d), ptr(plen_d), rows,
                          self.entropy_coder_precision, ptr(cdf), max_length + 2, ptr(status), stream())
        bad = int(status.max().item())
        if bad == 1:
            raise ValueError("Invalid `pmf`, non-finite or negative element found")
        if bad == 2:
            raise ValueError("Invalid `pmf`: at least one element must have a non-zero probability.")
        if bad:
            raise ValueError("pmf_to_quantized_cdf: no symbol to steal from")
        return cdf

    def _check_cdf_size(self):
        if self._quantized_cdf.numel() == 0:
            raise ValueError("Uninitialized CDFs. Run update() first")
        if len(self._quantized_cdf.size()) != 2:
            raise ValueError(f"Invalid CDF size {self._quantized_cdf.size()}")

    def _check_offsets_size(self):
        if self._offset.numel() == 0:
            raise ValueError("Uninitialized offsets. Run update() first")
        if len(self._offset.size()) != 1:
            raise ValueError(f"Invalid offsets size {self._offset.size()}")

    def _check_cdf_length(self):
        if self._cdf_length.numel() == 0:
            raise ValueError("Uninitialized CDF lengths. Run update() first")
        if len(self._cdf_length.size()) != 1:
            raise ValueError(f"Invalid offsets size {self._cdf_length.size()}")

    def compress(self, inputs, indexes, means=None):
        """:235-268 -> list[bytes], one string per batch item (rANS, coder.py)."""
        from . import coder
        symbols = self.quantize(inputs, "symbols", means)
        if len(inputs.size()) < 2:
            raise ValueError("Invalid `inputs` size. Expected a tensor with at least 2 dimensions.")
        if inputs.size() != indexes.size():
            raise ValueError("`inputs` and `indexes` should have the same size.")
        self._check_cdf_size()
        self._check_cdf_length()
        self._check_offsets_size()
        return coder.encode_batch(symbols, indexes, self._quantized_cdf, self._cdf_length, self._offset,
                                  self.entropy_coder_precision, lanes=int(self.coder_lanes),
                                  index=int(self.coder_index) if int(self.coder_lanes) == 1 else 0)

    def decompress(self, strings, indexes, dtype: torch.dtype = torch.float, means=None):
        """:270-325."""
        from . import coder
        if not isinstance(strings, (tuple, list)):
            raise ValueError("Invalid `strings` parameter type.")
        if not len(strings) == indexes.size(0):
            raise ValueError("Invalid strings or indexes parameters")
        if len(indexes.size()) < 2:
            raise ValueError("Invalid `indexes` size. Expected a tensor with at least 2 dimensions.")
        self._check_cdf_size()
        self._check_cdf_length()
        self._check_offsets_size()
        if means is not None:
            if means.size()[:2] != indexes.size()[:2]:
                raise ValueError("Invalid means or indexes parameters")
            if means.size() != indexes.size():
                for i in range(2, len(indexes.size())):
                    if means.size(i) != 1:
                        raise ValueError("Invalid means parameters")
        outputs = coder.decode_batch(strings, indexes, self._quantized_cdf, self._cdf_length, self._offset,
                                     self.entropy_coder_precision, lanes=1 if int(self.coder_lanes) == 1 else None)
        return self.dequantize(outputs, means, dtype)


class EntropyBottleneck(EntropyModel):
    """entropy_models.py:328-548: factorised prior, per-channel 1-3-3-3-3-1 monotone network."""

    def __init__(self, channels: int, *args: Any, tail_mass: float = 1e-9, init_scale: float = 10,
                 filters: Tuple[int, ...] = (3, 3, 3, 3), **kwargs: Any):
        super().__init__(*args, **kwargs)
        self.channels = int(channels)
        self.filters = tuple(int(f) for f in filters)
        self.init_scale = float(init_scale)
        self.tail_mass = float(tail_mass)
        if self.filters != (3, 3, 3, 3):
            raise ValueError("EntropyBottleneck on the B200 kernels implements the reference's default filters=(3, 3, 3, 3) "
                             "(the per-channel 1-3-3-3-3-1 network is unrolled in registers, csrc/entropy.cu); every model "
                             "of the reference's zoo and of this path uses that default -- see INTEGRATION.md")
        filters = (1,) + self.filters + (1,)
        scale = self.init_scale ** (1 / (len(self.filters) + 1))
        channels = self.channels
        for i in range(len(self.filters) + 1):
            init = np.log(np.expm1(1 / scale / filters[i + 1]))
            matrix = torch.Tensor(channels, filters[i + 1], filters[i])
            matrix.data.fill_(init)
            self.register_parameter(f"_matrix{i:d}", nn.Parameter(matrix))
            bias = torch.Tensor(channels, filters[i + 1], 1)
            nn.init.uniform_(bias, -0.5, 0.5)
            self.register_parameter(f"_bias{i:d}", nn.Parameter(bias))
            if i < len(self.filters):
                factor = torch.Tensor(channels, filters[i + 1], 1)
                nn.init.zeros_(factor)
                self.register_parameter(f"_factor{i:d}", nn.Parameter(factor))
        self.quantiles = nn.Parameter(torch.Tensor(channels, 1, 3))
        init = torch.Tensor([-self.init_scale, 0, self.init_scale])
        self.quantiles.data = init.repeat(self.quantiles.size(0), 1, 1)
        target = np.log(2 / self.tail_mass - 1)
        self.register_buffer("target", torch.Tensor([-target, 0, target]))

    def _get_medians(self):
        return self.quantiles[:, :, 1:2]

    # column ranges of the 14 reference-named parameters inside the per-channel block of 58 floats
    PACK_COLUMNS = (("_matrix0", 0, 3), ("_bias0", 3, 6), ("_factor0", 6, 9),
                    ("_matrix1", 9, 18), ("_bias1", 18, 21), ("_factor1", 21, 24),
                    ("_matrix2", 24, 33), ("_bias2", 33, 36), ("_factor2", 36, 39),
                    ("_matrix3", 39, 48), ("_bias3", 48, 51), ("_factor3", 51, 54),
                    ("_matrix4", 54, 57), ("_bias4", 57, 58))
    PACK_WIDTH = 58

    def packed_parameters(self):
        """[C,58] per-channel parameter block read by the kernels.  Once the model's ParamArena is
        built the 14 reference-named parameters ARE strided views of that block (and their .grad of
        `_packed_grad`), so this is free; stand-alone modules concatenate (autograd routes the
        gradient back)."""
        if getattr(self, "_packed", None) is not None:
            return self._packed
        C = self.channels
        parts = []
        for i in range(len(self.filters) + 1):
            parts.append(getattr(self, f"_matrix{i:d}").reshape(C, -1))
            parts.append(getattr(self, f"_bias{i:d}").reshape(C, -1))
            if i < len(self.filters):
                parts.append(getattr(self, f"_factor{i:d}").reshape(C, -1))
        return torch.cat(parts, dim=1)

    def _logits_cumulative(self, inputs, stop_gradient: bool):
        """:436-455, host-side (parameter-sized grids in update()/tests only)."""
        logits = inputs
        for i in range(len(self.filters) + 1):
            matrix = getattr(self, f"_matrix{i:d}")
            if stop_gradient:
                matrix = matrix.detach()
            logits = torch.matmul(F.softplus(matrix), logits)
            bias = getattr(self, f"_bias{i:d}")
            if stop_gradient:
                bias = bias.detach()
            logits = logits + bias
            if i < len(self.filters):
                factor = getattr(self, f"_factor{i:d}")
                if stop_gradient:
                    factor = factor.detach()
                logits = logits + torch.tanh(factor) * torch.tanh(logits)
        return logits

    def update(self, force: bool = False, update_quantiles: bool = False) -> bool:
        """:389-429.  The float pmf is evaluated on the host exactly as the reference does on CPU
        (so the integer tables are bit-identical to `compressai`'s); the integer stage runs on
        the device."""
        if self._offset.numel() > 0 and not force:
            return False
        dev = self._offset.device
        cpu = {n: p.detach().cpu() for n, p in self.named_parameters(recurse=False)}
        q = cpu["quantiles"]
        medians = q[:, 0, 1]
        minima = torch.clamp(torch.ceil(medians - q[:, 0, 0]).int(), min=0)
        maxima = torch.clamp(torch.ceil(q[:, 0, 2] - medians).int(), min=0)
        pmf_start = medians - minima
        pmf_length = maxima + minima + 1
        max_length = int(pmf_length.max().item())
        if self.device_pmf:                      # the pmf grid by one kernel on the packed parameters
            C = self.channels
            packed = self.packed_parameters().detach().float().contiguous()
            pmf = torch.empty((C, max_length), dtype=torch.float32, device=dev)
            tail_mass = torch.empty(C, dtype=torch.float32, device=dev)
            start_d = pmf_start.float().to(dev)
            L.eb_pmf(ptr(packed), ptr(start_d), C, max_length, ptr(pmf), max_length, ptr(tail_mass), stream())
            self._offset = (-minima).to(dev)
            self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
            self._cdf_length = (pmf_length + 2).to(dev)
            return True
        samples = torch.arange(max_length)[None, :] + pmf_start[:, None, None]

        def logits(v):
            out = v
            for i in range(len(self.filters) + 1):
                out = torch.matmul(F.softplus(cpu[f"_matrix{i:d}"]), out) + cpu[f"_bias{i:d}"]
                if i < len(self.filters):
                    out = out + torch.tanh(cpu[f"_factor{i:d}"]) * torch.tanh(out)
            return out

        lower = logits(samples - 0.5)
        upper = logits(samples + 0.5)
        sign = -torch.sign(lower + upper)
        pmf = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))[:, 0, :]
        tail_mass = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
        self._offset = (-minima).to(dev)
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
        self._cdf_length = (pmf_length + 2).to(dev)
        return True

    def loss(self):
        """:431-434 (aux loss; gradient reaches `quantiles` only)."""
        return Fn.eb_aux_apply(self.quantiles, self.packed_parameters(), self.target)

    def forward(self, x, training: Optional[bool] = None, noise=None):
        """:471-516.  x [B,C,...] -> (outputs, likelihood), same shape.  `noise` (optional, shaped
        like x) injects the U(-.5,.5) draw for parity tests; otherwise Philox in the kernel."""
        if training is None:
            training = self.training
        if x.dim() < 2 or x.shape[1] != self.channels:
            raise ValueError(f"EntropyBottleneck({self.channels}) got input of shape {tuple(x.shape)}")
        shape = x.shape
        x4 = x.reshape(shape[0], shape[1], -1, 1)
        n4 = None if noise is None else noise.reshape(shape[0], shape[1], -1, 1)
        mode = "noise" if training else "dequantize"
        medians = self.quantiles[:, 0, 1].detach().contiguous()
        out, lik = Fn.eb_apply(x4, self.packed_parameters(), medians, n4, mode,
                               self.likelihood_bound if self.use_likelihood_bound else 0.0,
                               self._rng(x.device) if noise is None and training else None, self.rng_offset,
                               packed_grad=getattr(self, "_packed_grad", None))
        if noise is None and training and self.rng_auto_advance:
            self.rng.advance()
        out_r, lik_r = out.reshape(shape), lik.reshape(shape)
        b16 = getattr(out, Fn._BF16_ATTR)
        setattr(out_r, Fn._BF16_ATTR, b16.reshape(shape))
        return out_r, Fn.carry_attrs(lik, lik_r)

    @staticmethod
    def _build_indexes(size):
        dims = len(size)
        N, C = size[0], size[1]
        view_dims = np.ones((dims,), dtype=np.int64)
        view_dims[1] = -1
        indexes = torch.arange(C).view(*view_dims).int()
        return indexes.repeat(N, 1, *size[2:])

    @staticmethod
    def _extend_ndims(tensor, n):
        return tensor.reshape(-1, *([1] * n)) if n > 0 else tensor.reshape(-1)

    def compress(self, x):
        indexes = self._build_indexes(x.size()).to(x.device)
        medians = self._get_medians().detach()
        spatial_dims = len(x.size()) - 2
        medians = self._extend_ndims(medians, spatial_dims)
        medians = medians.expand(x.size(0), *([-1] * (spatial_dims + 1)))
        return super().compress(x, indexes, medians)

    def decompress(self, strings, size):
        output_size = (len(strings), self._quantized_cdf.size(0), *size)
        indexes = self._build_indexes(output_size).to(self._quantized_cdf.device)
        medians = self._extend_ndims(self._get_medians().detach(), len(size))
        medians = medians.expand(len(strings), *([-1] * (len(size) + 1)))
        return super().decompress(strings, indexes, medians.dtype, medians)


class GaussianConditional(EntropyModel):
    """entropy_models.py:551-689."""

    def __init__(self, scale_table: Optional[Union[List, Tuple]], *args: Any, scale_bound: float = 0.11,
                 tail_mass: float = 1e-9, **kwargs: Any):
        super().__init__(*args, **kwargs)
        if not isinstance(scale_table, (type(None), list, tuple)):
            raise ValueError(f'Invalid type for scale_table "{type(scale_table)}"')
        if isinstance(scale_table, (list, tuple)) and len(scale_table) < 1:
            raise ValueError(f'Invalid scale_table length "{len(scale_table)}"')
        if scale_table and (scale_table != sorted(scale_table) or any(s <= 0 for s in scale_table)):
            raise ValueError(f'Invalid scale_table "({scale_table})"')
        self.tail_mass = float(tail_mass)
        if scale_bound is None and scale_table:
            scale_bound = scale_table[0]
        if scale_bound <= 0:
            raise ValueError("Invalid parameters")
        self.scale_bound_value = float(scale_bound)
        self.lower_bound_scale = LowerBound(scale_bound)
        self.register_buffer("scale_table",
                             self._prepare_scale_table(scale_table) if scale_table else torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]) if scale_bound is not None else None)
        self.rng_offset = 1

    @staticmethod
    def _prepare_scale_table(scale_table):
        return torch.Tensor(tuple(float(s) for s in scale_table))

    @staticmethod
    def _standardized_quantile(quantile):
        return scipy.stats.norm.ppf(quantile)

    def update_scale_table(self, scale_table, force=False):
        if self._offset.numel() > 0 and not force:
            return False
        device = self.scale_table.device
        self.scale_table = self._prepare_scale_table(scale_table).to(device)
        self.update()
        return True

    def update(self):
        """:625-648; float pmf on the host (bit-identical to the reference's CPU path), integer
        stage on the device."""
        dev = self._offset.device
        table = self.scale_table.detach().cpu()
        multiplier = -self._standardized_quantile(self.tail_mass / 2)
        pmf_center = torch.ceil(table * multiplier).int()
        pmf_length = 2 * pmf_center + 1
        max_length = int(torch.max(pmf_length).item())
        if self.device_pmf:                      # the erfc grid by one kernel
            S = table.numel()
            pmf = torch.empty((S, max_length), dtype=torch.float32, device=dev)
            tail_mass = torch.empty(S, dtype=torch.float32, device=dev)
            table_d, center_d = self.scale_table.detach().float().contiguous().to(dev), pmf_center.to(dev)
            L.gc_pmf(ptr(table_d), ptr(center_d), S, max_length, ptr(pmf), max_length, ptr(tail_mass), stream())
            self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
            self._offset = (-pmf_center).to(dev)
            self._cdf_length = (pmf_length + 2).to(dev)
            return
        samples = torch.abs(torch.arange(max_length).int() - pmf_center[:, None]).float()
        samples_scale = table.unsqueeze(1).float()
        const = float(-(2 ** -0.5))
        upper = 0.5 * torch.erfc(const * ((0.5 - samples) / samples_scale))
        lower = 0.5 * torch.erfc(const * ((-0.5 - samples) / samples_scale))
        pmf = upper - lower
        tail_mass = 2 * lower[:, :1]
        self._quantized_cdf = self._pmf_to_cdf(pmf, tail_mass, pmf_length, max_length).to(dev)
        self._offset = (-pmf_center).to(dev)
        self._cdf_length = (pmf_length + 2).to(dev)

    def forward(self, inputs, scales, means=None, training: Optional[bool] = None, noise=None):
        """:669-682 -> (outputs, likelihood)."""
        if training is None:
            training = self.training
        if inputs.shape != scales.shape or (means is not None and means.shape != inputs.shape):
            raise ValueError("inputs, scales and means must have the same shape")
        shape = inputs.shape
        as4 = (lambda t: None if t is None else (t if t.dim() == 4 else t.reshape(shape[0], shape[1], -1, 1)))
        mode = "noise" if training else "dequantize"
        out, lik = Fn.gc_apply(as4(inputs), as4(scales), as4(means), as4(noise), mode, self.scale_bound_value,
                               self.likelihood_bound if self.use_likelihood_bound else 0.0,
                               self._rng(inputs.device) if noise is None and training else None, self.rng_offset)
        if noise is None and training and self.rng_auto_advance:
            self.rng.advance()
        if out.shape != shape:
            out_r, lik_r = out.reshape(shape), lik.reshape(shape)
            setattr(out_r, Fn._BF16_ATTR, getattr(out, Fn._BF16_ATTR).reshape(shape))
            out, lik = out_r, Fn.carry_attrs(lik, lik_r)
        return out, lik

    def build_indexes(self, scales):
        """:684-689."""
        from . import coder
        return coder.gc_build_indexes(scales, self.scale_table, self.scale_bound_value)
