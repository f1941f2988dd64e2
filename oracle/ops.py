"""Oracle (test infrastructure, see oracle/__init__.py): functional PyTorch-fp32
restatement of the reference's per-module maths.  Autograd supplies backward.

Reference = CompressAI 1.2.4 under /root/reference/compressai/compressai/.
"""
import math

import torch
import torch.nn.functional as F

PEDESTAL = 2.0 ** -36  # reparam_offset**2, ops/parametrizers.py:47-53


class _LowerBoundFn(torch.autograd.Function):
    """ops/bound_ops.py:36-57: fwd max(x,b); bwd g*[(x>=b) | (g<0)]."""

    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, g):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (g < 0)) * g, None


def lower_bound(x, bound):
    if not torch.is_tensor(bound):
        bound = torch.tensor([float(bound)], dtype=x.dtype, device=x.device)
    return _LowerBoundFn.apply(x, bound)


def nonneg_init(x, minimum=0.0):
    """ops/parametrizers.py:57-58."""
    ped = torch.tensor([PEDESTAL], dtype=x.dtype)
    return torch.sqrt(torch.max(x + ped, ped))


def nonneg_reparam(x, minimum=0.0):
    """ops/parametrizers.py:60-63: LowerBound(x, sqrt(min+ped))**2 - ped."""
    bound = (float(minimum) + PEDESTAL) ** 0.5
    out = lower_bound(x, torch.tensor([bound], dtype=torch.float32, device=x.device))
    ped = torch.tensor([PEDESTAL], dtype=torch.float32, device=x.device)
    return out ** 2 - ped


def gdn(x, beta, gamma, inverse=False, beta_min=1e-6):
    """layers/gdn.py:77-92.  x [B,C,H,W]; beta [C]; gamma [C,C] ([out,in])."""
    C = x.shape[1]
    b = nonneg_reparam(beta, beta_min)
    g = nonneg_reparam(gamma, 0.0).reshape(C, C, 1, 1)
    norm = F.conv2d(x ** 2, g, b)
    norm = torch.sqrt(norm) if inverse else torch.rsqrt(norm)
    return x * norm


def quantize_noise(x, noise):
    """entropy_models.py:161-165 with the uniform(-.5,.5) draw injected."""
    return x + noise


def quantize_dequantize(x, means=None):
    """entropy_models.py:167-176."""
    out = x.clone()
    if means is not None:
        out = out - means
    out = torch.round(out)
    if means is not None:
        out = out + means
    return out


def quantize_ste(x):
    """ops/ops.py:66-80."""
    return (torch.round(x) - x).detach() + x


def eb_logits_cumulative(v, matrices, biases, factors, stop_gradient=False):
    """entropy_models.py:436-455.  v [C,1,L]."""
    logits = v
    n = len(matrices)
    for i in range(n):
        m = matrices[i].detach() if stop_gradient else matrices[i]
        logits = torch.matmul(F.softplus(m), logits)
        b = biases[i].detach() if stop_gradient else biases[i]
        logits = logits + b
        if i < n - 1:
            f = factors[i].detach() if stop_gradient else factors[i]
            logits = logits + torch.tanh(f) * torch.tanh(logits)
    return logits


def eb_likelihood(v, matrices, biases, factors, bound=1e-9):
    """entropy_models.py:458-469 + :502-503.  v [C,1,L] (already quantised)."""
    lower = eb_logits_cumulative(v - 0.5, matrices, biases, factors)
    upper = eb_logits_cumulative(v + 0.5, matrices, biases, factors)
    sign = (-torch.sign(lower + upper)).detach()
    lik = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
    if bound > 0:
        lik = lower_bound(lik, bound)
    return lik


def eb_forward(x, matrices, biases, factors, quantiles, training=True, noise=None,
               bound=1e-9):
    """entropy_models.py:471-516.  x [B,C,...]; noise is drawn in the permuted
    [C,1,B*H*W] layout (entropy_models.py:490-498) and must be passed that way."""
    perm = list(range(x.dim()))
    perm[0], perm[1] = 1, 0
    xp = x.permute(*perm).contiguous()
    shape = xp.shape
    values = xp.reshape(xp.shape[0], 1, -1)
    if training:
        outputs = quantize_noise(values, noise.reshape(values.shape))
    else:
        outputs = quantize_dequantize(values, quantiles[:, :, 1:2])
    lik = eb_likelihood(outputs, matrices, biases, factors, bound)
    outputs = outputs.reshape(shape).permute(*perm).contiguous()
    lik = lik.reshape(shape).permute(*perm).contiguous()
    return outputs, lik


def eb_aux_loss(matrices, biases, factors, quantiles, target):
    """entropy_models.py:431-434."""
    logits = eb_logits_cumulative(quantiles, matrices, biases, factors, stop_gradient=True)
    return torch.abs(logits - target).sum()


def std_cumulative(x):
    """entropy_models.py:604-608."""
    return 0.5 * torch.erfc(-(2 ** -0.5) * x)


def gc_likelihood(y_hat, scales, means=None, scale_bound=0.11, bound=1e-9):
    """entropy_models.py:650-667 + :680-681."""
    values = y_hat - means if means is not None else y_hat
    s = lower_bound(scales, scale_bound)
    values = torch.abs(values)
    upper = std_cumulative((0.5 - values) / s)
    lower = std_cumulative((-0.5 - values) / s)
    lik = upper - lower
    if bound > 0:
        lik = lower_bound(lik, bound)
    return lik


def gc_forward(y, scales, means=None, training=True, noise=None, scale_bound=0.11,
               bound=1e-9):
    """entropy_models.py:669-682."""
    if training:
        y_hat = quantize_noise(y, noise)
    else:
        y_hat = quantize_dequantize(y, means)
    return y_hat, gc_likelihood(y_hat, scales, means, scale_bound, bound)


def rd_loss(likelihoods, x_hats, targets, lmbda, num_pixels, level_weights=None):
    """losses/rate_distortion.py:47-59, extended to a list of pyramid levels as
    frozen in spec.py: D = sum_l w_l * mean((x_hat_l - x_l)^2)."""
    bpp = sum(torch.log(l).sum() / (-math.log(2) * num_pixels) for l in likelihoods)
    if level_weights is None:
        level_weights = [1.0] * len(x_hats)
    mse = sum(w * F.mse_loss(a, b) for w, a, b in zip(level_weights, x_hats, targets))
    return {"bpp_loss": bpp, "mse_loss": mse, "loss": lmbda * 255 ** 2 * mse + bpp}
